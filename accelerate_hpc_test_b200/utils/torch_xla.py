"""TPU / torch_xla support is not part of this build (B200 only).  The module exists so that
``from accelerate.utils.torch_xla import install_xla`` style imports fail with a clear message instead of ImportError."""


def install_xla(upgrade: bool = False):
    raise NotImplementedError("torch_xla / TPU back-ends are not supported by the B200-native build.")
