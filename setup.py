"""Packaging for accelerate_hpc_test_b200.

``pip install -e .`` (or ``python setup.py build_py``) compiles the sm_100a kernel library in-tree first
(``accelerate_hpc_test_b200/_build.py``: plain nvcc, no torch headers, seconds per file, no GPU needed), then installs the Python package
with the same console scripts the reference installs (reference setup.py:70-78), so ``accelerate launch ...`` keeps working after the
import is switched.  Set ``B200_SKIP_KERNEL_BUILD=1`` to package the Python side only (CPU-only machines without nvcc).
"""

import os
import sys
from pathlib import Path

from setuptools import find_packages, setup
from setuptools.command.build_py import build_py

ROOT = Path(__file__).resolve().parent
PACKAGE = "accelerate_hpc_test_b200"


class BuildKernelsThenPython(build_py):
    def run(self):
        if not int(os.environ.get("B200_SKIP_KERNEL_BUILD", "0")):
            sys.path.insert(0, str(ROOT))
            from accelerate_hpc_test_b200._build import build

            print(f"kernel library: {build()}")
        super().run()


extras = {
    "quality": ["ruff"],
    "testing": ["pytest", "pytest-xdist", "pytest-timeout", "transformers", "safetensors", "scikit-learn", "scipy", "tqdm"],
    "trackers": ["tensorboard", "wandb", "comet-ml", "mlflow", "aim", "clearml", "dvclive", "swanlab", "trackio"],
    "rich": ["rich"],
}
extras["dev"] = sorted({dep for group in ("quality", "testing", "rich") for dep in extras[group]})

setup(
    name="accelerate-hpc-test-b200",
    version="0.2.0",
    description="Accelerate's training / big-model / launch API on hand-written Blackwell (sm_100a) kernels and NVLink collectives",
    long_description=(ROOT / "README.md").read_text(encoding="utf-8"),
    long_description_content_type="text/markdown",
    license="Apache-2.0",
    packages=find_packages(include=[PACKAGE, f"{PACKAGE}.*"]),
    package_data={PACKAGE: ["csrc/*.cu", "csrc/*.cuh", "lib/*.so"]},
    data_files=[("include", ["include/b200_ops.h"])],
    include_package_data=True,
    cmdclass={"build_py": BuildKernelsThenPython},
    entry_points={
        "console_scripts": [
            f"accelerate={PACKAGE}.commands.accelerate_cli:main",
            f"accelerate-config={PACKAGE}.commands.config:main",
            f"accelerate-estimate-memory={PACKAGE}.commands.estimate:main",
            f"accelerate-launch={PACKAGE}.commands.launch:main",
            f"accelerate-merge-weights={PACKAGE}.commands.merge:main",
        ]
    },
    python_requires=">=3.10",
    install_requires=["numpy>=1.17", "packaging>=20.0", "psutil", "pyyaml", "torch>=2.7", "safetensors>=0.4.3"],
    extras_require=extras,
    classifiers=[
        "Environment :: GPU :: NVIDIA CUDA :: 12",
        "Programming Language :: Python :: 3",
        "Topic :: Scientific/Engineering :: Artificial Intelligence",
    ],
)
