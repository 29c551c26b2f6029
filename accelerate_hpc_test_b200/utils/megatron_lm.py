"""Megatron-LM compatibility objects (reference src/accelerate/utils/megatron_lm.py).

Megatron-LM is not a dependency: its tp / cp / sequence-parallel degrees are expressed through
``ParallelismConfig`` and executed by the native engines.  The dummy dataloader / scheduler classes
are kept so scripts written against the reference import cleanly.
"""


class MegatronLMDummyDataLoader:
    """Placeholder carrying the dataset arguments of a Megatron-style data pipeline."""

    def __init__(self, **dataset_kwargs):
        self.dataset_args = dict(dataset_kwargs)

    _UNAVAILABLE = "Megatron-LM data pipelines are not part of this build; use a torch DataLoader with `Accelerator.prepare`."

    def set_megatron_data_args(self):
        raise NotImplementedError(self._UNAVAILABLE)

    def get_train_valid_test_datasets_provider(self, accelerator):
        raise NotImplementedError(self._UNAVAILABLE)

    def build_train_valid_test_data_iterators(self, accelerator):
        raise NotImplementedError(self._UNAVAILABLE)


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


# --------------------------------------------------------------------------------------- engine-style API
# The reference's Megatron path hands the whole step to a ``MegatronEngine`` (``model(**batch)`` runs forward,
# backward and the optimizer, reference utils/megatron_lm.py:924-1158; the optimizer / scheduler wrappers are
# therefore no-ops, :355-413; per-family ``*TrainStep`` classes supply batch slicing and the loss, :415-922).
# The same surface is kept here on top of the native engines: the model is a native (optionally tensor-parallel)
# module, the micro-batch loop accumulates into the engine's gradient buffers and the fused AdamW applies the update.
import torch

from ..optimizer import AcceleratedOptimizer
from ..scheduler import AcceleratedScheduler


class MegatronLMOptimizerWrapper(AcceleratedOptimizer):
    """``zero_grad`` / ``step`` are no-ops: ``MegatronEngine.forward`` already ran them (reference :355-372)."""

    def __init__(self, optimizer):
        super().__init__(optimizer, device_placement=False, scaler=None)
        if not hasattr(self.optimizer, "skipped_iter"):
            self.optimizer.skipped_iter = False

    def zero_grad(self, set_to_none=None):
        pass

    def step(self):
        pass

    @property
    def step_was_skipped(self):
        return bool(getattr(self.optimizer, "skipped_iter", False))


class MegatronLMSchedulerWrapper(AcceleratedScheduler):
    """The engine steps the schedule itself (reference :401-407)."""

    def __init__(self, scheduler, optimizers):
        super().__init__(scheduler, optimizers)

    def step(self, *args, **kwargs):
        return


class AbstractTrainStep:
    """Batch slicing, forward step and loss of one model family (reference :415-430)."""

    def __init__(self, name):
        self.name = name

    def get_batch_func(self, accelerator, megatron_dataset_flag=False):
        def get_batch(data_iterator):
            batch = next(data_iterator)
            return {k: (v.to(accelerator.device) if torch.is_tensor(v) else v) for k, v in batch.items()}

        return get_batch

    def get_loss_func(self, accelerator):
        def loss_func(output):
            loss = output.loss if hasattr(output, "loss") else output["loss"]
            return loss, {"lm loss": avg_losses_across_data_parallel_group([loss.detach()])[0]}

        return loss_func

    def get_forward_step_func(self):
        get_batch, loss_func = self._get_batch, self._loss_func

        def forward_step(data_iterator, model):
            batch = get_batch(data_iterator)
            return loss_func(model(**batch))

        return forward_step

    def bind(self, accelerator):
        self._get_batch = self.get_batch_func(accelerator)
        self._loss_func = self.get_loss_func(accelerator)
        return self


class GPTTrainStep(AbstractTrainStep):
    """Causal LM: ``input_ids`` (+ ``labels``, default = inputs) -> next-token loss (reference :574-716)."""

    def __init__(self, accelerator=None, args=None):
        super().__init__("GPTTrainStep")
        if accelerator is not None:
            self.bind(accelerator)

    def get_batch_func(self, accelerator, megatron_dataset_flag=False):
        base = super().get_batch_func(accelerator)

        def get_batch(data_iterator):
            b = base(data_iterator)
            if "labels" not in b and "shift_labels" not in b:
                b["labels"] = b["input_ids"]
            return {k: b[k] for k in ("input_ids", "labels", "shift_labels", "position_ids", "attention_mask") if k in b}

        return get_batch


class BertTrainStep(AbstractTrainStep):
    """Encoder: sequence-classification / MLM batches pass through unchanged (reference :432-572)."""

    def __init__(self, accelerator=None, args=None):
        super().__init__("BertTrainStep")
        if accelerator is not None:
            self.bind(accelerator)


class T5TrainStep(AbstractTrainStep):
    """Encoder-decoder batches (``input_ids``, ``decoder_input_ids``, ``labels``) pass through (reference :718-922)."""

    def __init__(self, accelerator=None, args=None):
        super().__init__("T5TrainStep")
        if accelerator is not None:
            self.bind(accelerator)

    # Megatron's masks are boolean "do NOT attend here" tensors of shape [batch, queries, keys]
    @staticmethod
    def attn_mask_postprocess(attention_mask):
        """Self-attention mask from a ``[batch, seq]`` keep-mask: position (i, j) is blocked unless both tokens are real."""
        both_real = attention_mask.unsqueeze(1) * attention_mask.unsqueeze(2)
        return both_real < 0.5

    @staticmethod
    def get_decoder_mask(seq_length, device):
        """Causal mask ``[1, seq, seq]``: everything above the diagonal is blocked."""
        return torch.ones((1, seq_length, seq_length), device=device).tril() < 0.5

    @staticmethod
    def get_enc_dec_mask(attention_mask, dec_seq_length, device):
        """Cross-attention mask ``[batch, decoder positions, encoder positions]``: padded encoder tokens are blocked for every query."""
        batch = attention_mask.shape[0]
        keys_real = torch.ones((batch, dec_seq_length, 1), device=device) * attention_mask.unsqueeze(1)
        return keys_real < 0.5


class MegatronEngine(torch.nn.Module):
    """``engine(**batch)`` = one full optimisation step over ``num_micro_batches`` slices of the batch.

    ``model`` is a prepared native module, ``optimizer`` / ``scheduler`` the prepared (accelerated) objects.  In
    ``train()`` mode ``forward`` returns the averaged loss after having stepped the optimizer and the schedule; in
    ``eval()`` mode it returns the loss without touching gradients (reference :1033-1158)."""

    def __init__(self, accelerator, model, optimizer, scheduler=None):
        super().__init__()
        self.accelerator = accelerator
        self.module = model
        self.base_model = model
        self.optimizer = optimizer
        self.scheduler = scheduler
        plugin = accelerator.state.megatron_lm_plugin
        custom = getattr(plugin, "custom_train_step_class", None) if plugin is not None else None
        if custom is not None:
            self.train_step_handler = custom(None, **(getattr(plugin, "custom_train_step_kwargs", None) or {}))
            if hasattr(self.train_step_handler, "bind"):
                self.train_step_handler.bind(accelerator)
        else:
            kind = (getattr(plugin, "model_type_name", None) or "gpt") if plugin is not None else "gpt"
            cls = {"gpt": GPTTrainStep, "bert": BertTrainStep, "t5": T5TrainStep}.get(kind)
            if cls is None:
                raise ValueError(f"Unsupported model type: {kind}")
            self.train_step_handler = cls(accelerator)
        self.num_micro_batches = int(getattr(plugin, "num_micro_batches", None) or 1) if plugin is not None else 1
        inner = getattr(optimizer, "optimizer", optimizer)
        inner.skipped_iter = False
        self.total_loss_dict = {}
        self.eval_total_loss_dict = {}
        self.iteration = 0

    def get_module_config(self):
        """The wrapped model's configuration with the training hooks Megatron's schedules look for: a loss-scaling function (identity
        here: bf16 needs none, fp16 scaling is the Accelerator's), the gradient-synchronisation switch of the native engine, and the
        micro-batch count."""
        import types

        config = getattr(self.module, "config", None) or getattr(getattr(self.module, "module", None), "config", None)
        described = types.SimpleNamespace(**(vars(config) if config is not None and hasattr(config, "__dict__") else {}))
        described.grad_scale_func = lambda loss: loss
        described.no_sync_func = getattr(self.module, "no_sync", None)
        described.num_micro_batches = self.num_micro_batches
        return described

    def get_batch_data_iterator(self, batch_data):
        """Split ``batch_data`` into ``num_micro_batches`` equal slices along dim 0 (reference :1008-1031)."""
        n = self.num_micro_batches
        if n <= 1:
            return iter([batch_data])
        chunks = []
        for i in range(n):
            part = {}
            for k, v in batch_data.items():
                if torch.is_tensor(v) and v.dim() > 0:
                    if v.shape[0] % n:
                        raise ValueError(f"batch dimension {v.shape[0]} of `{k}` is not divisible by num_micro_batches={n}")
                    m = v.shape[0] // n
                    part[k] = v[i * m : (i + 1) * m]
                else:
                    part[k] = v
            chunks.append(part)
        return iter(chunks)

    def train_step(self, **batch_data):
        it = self.get_batch_data_iterator(batch_data)
        forward_step = self.train_step_handler.get_forward_step_func()
        self._raw_optimizer().zero_grad()
        n = self.num_micro_batches
        losses = {}
        for _ in range(n):
            loss, reduced = forward_step(it, self.module)
            (loss / n).backward()
            for k, v in reduced.items():
                losses[k] = losses.get(k, 0.0) + v / n
        self._raw_optimizer().step()
        inner = getattr(self.optimizer, "optimizer", self.optimizer)
        skipped = bool(getattr(inner, "skipped_iter", False))
        if self.scheduler is not None and not skipped:
            sched = getattr(self.scheduler, "scheduler", self.scheduler)
            sched.step()
        return losses, skipped

    def _raw_optimizer(self):
        # the accelerated wrapper's step / zero_grad are no-ops for this engine: drive the inner optimizer
        opt = self.optimizer
        return opt.optimizer if isinstance(opt, MegatronLMOptimizerWrapper) else opt

    @torch.no_grad()
    def eval_step(self, **batch_data):
        it = self.get_batch_data_iterator(batch_data)
        forward_step = self.train_step_handler.get_forward_step_func()
        losses = {}
        for _ in range(self.num_micro_batches):
            _, reduced = forward_step(it, self.module)
            for k, v in reduced.items():
                losses[k] = losses.get(k, 0.0) + v / self.num_micro_batches
        return losses

    def forward(self, **batch_data):
        if self.module.training:
            losses, skipped = self.train_step(**batch_data)
            self.iteration += 1
            for k, v in losses.items():
                self.total_loss_dict[k] = self.total_loss_dict.get(k, 0.0) + v
        else:
            losses = self.eval_step(**batch_data)
            for k, v in losses.items():
                self.eval_total_loss_dict[k] = self.eval_total_loss_dict.get(k, 0.0) + v
        vals = list(losses.values())
        loss = vals[0] if len(vals) == 1 else sum(vals)
        return loss if torch.is_tensor(loss) else torch.tensor(loss)

    def train(self, mode: bool = True):
        self.module.train(mode)
        return self

    def eval(self):
        self.module.eval()
        return self

    def log_eval_results(self):
        out = {k: (v.item() if torch.is_tensor(v) else float(v)) for k, v in self.eval_total_loss_dict.items()}
        self.eval_total_loss_dict = {}
        return out

    def save_checkpoint(self, output_dir):
        self.accelerator.save_state(output_dir)

    def load_checkpoint(self, input_dir):
        self.accelerator.load_state(input_dir)


def avg_losses_across_data_parallel_group(losses):
    """Mean of every loss over the data-parallel ranks (reference :1215-1223)."""
    from ..state import PartialState

    t = torch.stack([l.detach().reshape(()).float() for l in losses])
    st = PartialState()
    if st.num_processes > 1 and torch.distributed.is_initialized():
        torch.distributed.all_reduce(t)
        t = t / st.num_processes
    return t


def gather_across_data_parallel_groups(tensor):
    """Concatenate ``tensor`` (or a nested structure of tensors) from all data-parallel ranks (reference :1226-1247)."""
    from .operations import gather

    return gather(tensor)


# --------------------------------------------------------------------------------------- function-style API
# The reference's ``Accelerator._prepare_megatron_lm`` is a sequence of five helpers (reference utils/megatron_lm.py: ``initialize``
# :97-170, ``prepare_data_loader`` :235-352, ``prepare_model_optimizer_scheduler`` :173-232, ``prepare_optimizer`` :375-381,
# ``prepare_scheduler`` :409-412) that scripts also call directly.  Here they do the same jobs on the native engines.
def initialize(accelerator, extra_args_provider=None, args_defaults=None):
    """Resolve the plugin's settings into one argument namespace (what Megatron's ``initialize_megatron`` produces), seed every
    generator from it and keep it on the plugin as ``megatron_args``.  ``args_defaults`` override the plugin's defaults;
    ``extra_args_provider`` may add ``argparse`` arguments, which are read from the command line."""
    import argparse

    from .random import set_seed

    plugin = accelerator.state.megatron_lm_plugin
    if plugin is None:
        raise ValueError("`initialize` needs an Accelerator created with a `MegatronLMPlugin` (or launched with `--use_megatron_lm`).")
    settings = dict(plugin.megatron_lm_default_args)
    settings.update(args_defaults or {})
    if extra_args_provider is not None:
        parser = extra_args_provider(argparse.ArgumentParser(allow_abbrev=False))
        extra, _ = parser.parse_known_args()
        settings.update({k: v for k, v in vars(extra).items() if v is not None})
    settings.setdefault("seed", 1234)
    settings.update(rank=accelerator.process_index, world_size=accelerator.num_processes, local_rank=accelerator.local_process_index,
                    tensor_model_parallel_size=plugin.tp_degree, pipeline_model_parallel_size=plugin.pp_degree,
                    context_parallel_size=plugin.context_parallel_size)
    args = argparse.Namespace(**settings)
    set_seed(args.seed)
    plugin.megatron_args = args
    return args


def prepare_data_loader(accelerator, dataloader):
    """An ordinary ``DataLoader`` is sharded over the data-parallel ranks only (ranks of one tensor / context parallel group read the
    same batches); with ``num_micro_batches`` it yields *whole-step* batches — ``MegatronEngine`` cuts them into micro-batches.
    ``MegatronLMDummyDataLoader`` (Megatron's own indexed-dataset pipeline) is not part of this build."""
    if isinstance(dataloader, MegatronLMDummyDataLoader):
        dataloader.set_megatron_data_args()
    plugin = accelerator.state.megatron_lm_plugin
    micro_batches = int(getattr(plugin, "num_micro_batches", 1) or 1)
    if micro_batches > 1 and getattr(dataloader, "batch_size", None) is not None and not getattr(dataloader, "_is_accelerate_prepared", False):
        from torch.utils.data import DataLoader

        rebuilt = {k: getattr(dataloader, k) for k in ("num_workers", "collate_fn", "pin_memory", "drop_last", "timeout", "worker_init_fn")}
        sampler = dataloader.sampler
        shuffle = type(sampler).__name__ == "RandomSampler"
        dataloader = DataLoader(dataloader.dataset, batch_size=dataloader.batch_size * micro_batches, shuffle=shuffle, **rebuilt)
    return accelerator.prepare_data_loader(dataloader)


def prepare_optimizer(accelerator, model):
    """The optimizer Megatron would build from its arguments: the fused AdamW (or SGD) over ``model``'s parameters, hyper-parameters
    from the plugin's resolved arguments (``set_optimizer_type`` records them when the script created an optimizer itself)."""
    from ..optim import FusedAdamW

    plugin = accelerator.state.megatron_lm_plugin
    args = plugin.megatron_lm_default_args
    params = [p for p in model.parameters() if p.requires_grad]
    if args.get("optimizer", "adam") == "sgd":
        return torch.optim.SGD(params, lr=args.get("lr", 1e-3), momentum=args.get("sgd_momentum", 0.0), weight_decay=args.get("weight_decay", 0.0))
    return FusedAdamW(params, lr=args.get("lr", 1e-3), betas=(args.get("adam_beta1", 0.9), args.get("adam_beta2", 0.999)),
                      eps=args.get("adam_eps", 1e-8), weight_decay=args.get("weight_decay", 0.01))


def prepare_scheduler(accelerator, optimizer, scheduler):
    """A real scheduler passes through; a ``MegatronLMDummyScheduler`` becomes the warm-up + decay schedule its step counts (and the
    plugin's ``lr_decay_style`` / ``min_lr``) describe."""
    import math

    if not isinstance(scheduler, MegatronLMDummyScheduler):
        return scheduler
    plugin = accelerator.state.megatron_lm_plugin
    inner = getattr(optimizer, "optimizer", optimizer)
    total, warm = scheduler.total_num_steps, scheduler.warmup_num_steps or 0
    style = getattr(plugin, "lr_decay_style", "linear") or "linear"
    peak = inner.param_groups[0]["lr"]
    floor = (getattr(plugin, "min_lr", 0) or 0) / peak if peak else 0.0

    def factor(step):
        if warm and step < warm:
            return step / max(1, warm)
        if not total or style == "constant":
            return 1.0
        progress = min(1.0, (step - warm) / max(1, total - warm))
        shape = 0.5 * (1.0 + math.cos(math.pi * progress)) if style == "cosine" else 1.0 - progress
        return floor + (1.0 - floor) * shape

    return torch.optim.lr_scheduler.LambdaLR(inner, factor)


def prepare_model_optimizer_scheduler(accelerator):
    """Model, optimizer and scheduler from the plugin alone, for scripts that let Megatron build everything: the model comes from
    ``custom_prepare_model_function(accelerator)`` or ``custom_model_provider_function()`` (this build has no Megatron model zoo to
    fall back on), the optimizer and schedule from the resolved arguments."""
    plugin = accelerator.state.megatron_lm_plugin
    if plugin.custom_prepare_model_function is not None:
        if plugin.custom_model_provider_function is None:
            raise ValueError("You must provide a `custom_model_provider_function` when using a `custom_prepare_model_function`.")
        model = plugin.custom_prepare_model_function(accelerator)
    elif plugin.custom_model_provider_function is not None:
        model = plugin.custom_model_provider_function()
    else:
        raise NotImplementedError("Building Megatron's own bert / gpt / t5 models needs Megatron-LM, which is not part of this build: pass a "
                                  "model to `prepare`, or set `custom_model_provider_function` on the `MegatronLMPlugin`.")
    model = accelerator.prepare_model(model)
    optimizer = accelerator.prepare_optimizer(prepare_optimizer(accelerator, model))
    if hasattr(model, "bind_optimizer"):
        model.bind_optimizer(optimizer)
    args = plugin.megatron_lm_default_args
    schedule = None
    if args.get("train_iters"):
        schedule = MegatronLMDummyScheduler(optimizer, total_num_steps=args["train_iters"], warmup_num_steps=args.get("lr_warmup_iters") or 0)
        schedule = accelerator.prepare_scheduler(prepare_scheduler(accelerator, optimizer, schedule))
    return model, optimizer, schedule


def model_provider_func(pre_process=True, post_process=True, add_encoder=True, add_decoder=True):
    """Megatron's model provider builds its own bert / gpt / t5 implementations; they are not part of this build (the native models are
    in ``accelerate_hpc_test_b200.models``).  Set ``custom_model_provider_function`` on the ``MegatronLMPlugin`` instead."""
    raise NotImplementedError("Megatron-LM's model zoo is not part of this build: set `custom_model_provider_function` on the "
                              "`MegatronLMPlugin`, or pass a model to `Accelerator.prepare`.")


def finish_mpu_init():
    """Megatron finishes its model-parallel set-up here (process groups, seeds).  The process groups of this build come from
    ``ParallelismConfig``'s device mesh at ``Accelerator`` construction, so what remains is the seeding — from the arguments that
    ``initialize`` resolved, if it ran."""
    from ..state import AcceleratorState
    from .random import set_seed

    plugin = AcceleratorState().megatron_lm_plugin if AcceleratorState._shared_state else None
    args = getattr(plugin, "megatron_args", None)
    set_seed(getattr(args, "seed", 1234))
