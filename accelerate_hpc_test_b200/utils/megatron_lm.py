"""Megatron-LM compatibility objects (reference src/accelerate/utils/megatron_lm.py).

Megatron-LM is not a dependency: its tp / cp / sequence-parallel degrees are expressed through
``ParallelismConfig`` and executed by the native engines.  The dummy dataloader / scheduler classes
are kept so scripts written against the reference import cleanly.
"""


class MegatronLMDummyDataLoader:
    """Placeholder carrying the dataset arguments of a Megatron-style data pipeline."""

    def __init__(self, **dataset_kwargs):
        self.dataset_args = dict(dataset_kwargs)

    def set_megatron_data_args(self):
        raise NotImplementedError("Megatron-LM data pipelines are not part of this build; use a torch DataLoader with `Accelerator.prepare`.")


class MegatronLMDummyScheduler:
    """Placeholder scheduler (``total_num_steps`` / ``warmup_num_steps``)."""

    def __init__(self, optimizer, total_num_steps=None, warmup_num_steps=0, **kwargs):
        self.optimizer = optimizer
        self.total_num_steps = total_num_steps
        self.warmup_num_steps = warmup_num_steps
        self.kwargs = kwargs


def megatron_plugin_to_parallelism_config(plugin):
    """The ``ParallelismConfig`` equivalent of a ``MegatronLMPlugin`` (tp / cp sizes; pp must be 1)."""
    from ..parallelism_config import ParallelismConfig

    kw = plugin.to_parallelism_kwargs()
    return ParallelismConfig(tp_size=kw["tp_size"] or 1, cp_size=kw["cp_size"] or 1)
