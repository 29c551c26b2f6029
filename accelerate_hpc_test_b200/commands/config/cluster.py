"""The ``accelerate config`` questionnaire for a local machine / cluster, as a declarative question flow.

Every question is an ``Ask`` record (where the answer goes, the prompt, how to parse it, its default and the
condition under which it is asked).  A ``Flow`` walks the records of each section in order, keeps the answers
in per-section mappings (``fsdp_config``, ``deepspeed_config``, …) and finally builds the ``ClusterConfig``.
Sections can be unit-tested by feeding scripted answers (``tests/test_config_questionnaire.py``).

Covered branches (parity target: reference ``commands/config/cluster.py:58-924``, CUDA + CPU paths only — this build
has no XPU/NPU/MLU/HPU/TPU back-ends, so those sub-questions do not exist):

  machine / multi-node rendezvous / debug · CPU-only and mpirun · torch dynamo · DeepSpeed (ZeRO stage, offload,
  config file, MoE, multi-node launcher, filters) · FSDP v1 and v2 · ParallelismConfig (dp_replicate / dp_shard /
  tp / cp / sp) · Megatron-LM · process count, GPU ids, NUMA affinity · mixed precision and the fp8 back-ends
  (native MXFP8, TransformerEngine, torchao, MS-AMP).
"""

from __future__ import annotations

import os
from dataclasses import dataclass, field
from typing import Any, Callable, Optional, Sequence

from ...utils.constants import (
    DEEPSPEED_MULTINODE_LAUNCHERS,
    FSDP2_STATE_DICT_TYPE,
    FSDP_AUTO_WRAP_POLICY,
    FSDP_BACKWARD_PREFETCH,
    FSDP_SHARDING_STRATEGY,
    FSDP_STATE_DICT_TYPE,
    TORCH_DYNAMO_MODES,
)
from ...utils.dataclasses import ComputeEnvironment, DistributedType
from . import config_utils as io
from .config_args import ClusterConfig

YES, NO = True, False
Cond = Callable[["Flow"], bool]


@dataclass
class Ask:
    """One question.  ``slot`` is ``"section.key"`` (``"top.key"`` for top-level fields, ``"tmp.key"`` for answers
    that only steer the flow)."""

    slot: str
    text: str
    kind: str = "str"  # bool | int | float | str | menu
    default: Any = None
    choices: Sequence = ()
    when: Optional[Cond] = None
    parse: Optional[Callable] = None  # menu: index -> value; others: str -> value (overrides `kind`)
    menu_default: int = 0


@dataclass
class Flow:
    top: dict = field(default_factory=dict)
    tmp: dict = field(default_factory=dict)
    dynamo_config: dict = field(default_factory=dict)
    deepspeed_config: dict = field(default_factory=dict)
    fsdp_config: dict = field(default_factory=dict)
    parallelism_config: dict = field(default_factory=dict)
    megatron_lm_config: dict = field(default_factory=dict)
    mpirun_config: dict = field(default_factory=dict)
    fp8_config: dict = field(default_factory=dict)

    # -- storage ---------------------------------------------------------------------------------
    def _bucket(self, slot: str):
        section, _, key = slot.partition(".")
        return getattr(self, section), key

    def get(self, slot: str, default=None):
        bucket, key = self._bucket(slot)
        return bucket.get(key, default)

    def put(self, slot: str, value) -> None:
        bucket, key = self._bucket(slot)
        bucket[key] = value

    # -- asking ----------------------------------------------------------------------------------
    def ask(self, q: Ask):
        if q.when is not None and not q.when(self):
            return None
        if q.kind == "menu":
            parse = q.parse or io.pick_from(list(q.choices))
            value = io._ask_options(q.text, list(q.choices), parse, default=q.menu_default)
        elif q.kind == "bool":
            hint = " [YES/no]: " if q.default else " [yes/NO]: "
            value = io._ask_field(q.text + hint, io._convert_yes_no_to_bool, default=bool(q.default), error_message=io.YES_NO_HINT)
        else:
            parse = q.parse or {"int": int, "float": float, "str": str}[q.kind]
            shown = "" if q.default is None else f" [{q.default}]"
            msg = {"int": "Please enter an integer.", "float": "Please enter a number."}.get(q.kind)
            value = io._ask_field(f"{q.text}{shown}: ", parse, default=q.default, error_message=msg)
        self.put(q.slot, value)
        return value

    def run(self, questions: Sequence[Ask]) -> "Flow":
        for q in questions:
            self.ask(q)
        return self

    # -- shorthands used by the conditions -----------------------------------------------------------
    @property
    def kind(self) -> DistributedType:
        return self.top.get("distributed_type", DistributedType.NO)

    def is_kind(self, *kinds) -> bool:
        return self.kind in kinds

    @property
    def on_cpu(self) -> bool:
        return bool(self.top.get("use_cpu", False))

    @property
    def multi_node(self) -> bool:
        return int(self.top.get("num_machines", 1)) > 1


def _set(slot: str, value) -> Callable[[Flow], bool]:
    """A pseudo-condition that stores a constant and asks nothing (keeps derived values inside the tables)."""

    def apply(flow: Flow) -> bool:
        flow.put(slot, value(flow) if callable(value) else value)
        return False

    return apply


def _derive(slot: str, value) -> Ask:
    return Ask(slot, "", when=_set(slot, value))


# =========================================================================== sections
_DISTRIBUTED = (DistributedType.MULTI_GPU, DistributedType.MULTI_CPU)

MACHINE = [
    Ask("top.distributed_type", "Which type of machine are you using?", "menu",
        choices=["No distributed training", "multi-CPU", "multi-GPU"], parse=io._convert_distributed_mode),
    Ask("top.num_machines", "How many different machines will you use (use more than 1 for multi-node training)?", "int", 1,
        when=lambda f: f.is_kind(*_DISTRIBUTED)),
    Ask("top.machine_rank", "What is the rank of this machine?", "menu", when=lambda f: f.multi_node, parse=int),
    Ask("top.main_process_ip", "What is the IP address of the machine that will host the main process?", when=lambda f: f.multi_node),
    Ask("top.main_process_port", "What is the port you will use to communicate with the main process?", "int",
        when=lambda f: f.multi_node),
    Ask("top.same_network", "Are all the machines on the same local network? Answer `no` if nodes are on the cloud and/or on "
        "different network hosts", "bool", YES, when=lambda f: f.multi_node),
    Ask("top.rdzv_backend", "What rendezvous backend will you use? ('static', 'c10d', ...)", "str", "static",
        when=lambda f: f.multi_node and not f.top["same_network"]),
    Ask("top.debug", "Should distributed operations be checked while running for errors? This can avoid timeout issues but "
        "will be slower.", "bool", NO, when=lambda f: f.is_kind(*_DISTRIBUTED)),
    Ask("top.use_cpu", "Do you want to run your training on CPU only (even if a GPU is available)?", "bool", NO,
        when=lambda f: f.is_kind(DistributedType.NO)),
    _derive("top.use_cpu", lambda f: True if f.is_kind(DistributedType.MULTI_CPU) else f.top.get("use_cpu", False)),
    Ask("tmp.use_mpirun", "Do you want accelerate to launch mpirun?", "bool", NO,
        when=lambda f: f.is_kind(DistributedType.MULTI_CPU) and f.multi_node),
    Ask("mpirun_config.mpirun_hostfile", "Please enter the path to the hostfile to use with mpirun", "str", "~/hostfile",
        when=lambda f: f.tmp.get("use_mpirun"), parse=lambda s: os.path.expanduser(s.strip())),
    Ask("mpirun_config.mpirun_ccl", "Enter the number of oneCCL worker threads", "str", "1", when=lambda f: f.tmp.get("use_mpirun")),
]


def _machine_rank_choices(flow: Flow) -> None:
    """The rank menu lists 0..num_machines-1; built once the machine count is known."""
    for q in MACHINE:
        if q.slot == "top.machine_rank":
            q.choices = list(range(int(flow.top.get("num_machines", 1))))


DYNAMO = [
    Ask("tmp.use_dynamo", "Do you wish to optimize your script with torch dynamo?", "bool", NO),
    Ask("dynamo_config.dynamo_backend", "Which dynamo backend would you like to use?", "menu",
        choices=[b.lower() for b in io.DYNAMO_BACKENDS], parse=io._convert_dynamo_backend, menu_default=2,
        when=lambda f: f.tmp["use_dynamo"]),
    Ask("tmp.dynamo_custom", "Do you want to customize the defaults sent to torch.compile?", "bool", NO,
        when=lambda f: f.tmp["use_dynamo"]),
    Ask("dynamo_config.dynamo_mode", "Which mode do you want to use?", "menu", choices=TORCH_DYNAMO_MODES,
        when=lambda f: f.tmp.get("dynamo_custom")),
    Ask("dynamo_config.dynamo_use_fullgraph", "Do you want the fullgraph mode or it is ok to break model into several subgraphs?",
        "bool", NO, when=lambda f: f.tmp.get("dynamo_custom")),
    Ask("dynamo_config.dynamo_use_dynamic", "Do you want to enable dynamic shape tracing?", "bool", NO,
        when=lambda f: f.tmp.get("dynamo_custom")),
    Ask("dynamo_config.dynamo_use_regional_compilation", "Do you want to enable regional compilation?", "bool", NO,
        when=lambda f: f.tmp.get("dynamo_custom")),
]

_OFFLOAD_DEVICES = ["none", "cpu", "nvme"]


def _gpu_capable(f: Flow) -> bool:
    return f.is_kind(DistributedType.MULTI_GPU, DistributedType.NO) and not f.on_cpu


def _ds_manual(f: Flow) -> bool:
    return f.is_kind(DistributedType.DEEPSPEED) and not f.tmp.get("ds_from_file")


DEEPSPEED = [
    Ask("tmp.use_deepspeed", "Do you want to use DeepSpeed (ZeRO semantics; runs on the native sharded engine)?", "bool", NO,
        when=_gpu_capable),
    _derive("top.distributed_type", lambda f: DistributedType.DEEPSPEED if f.tmp.get("use_deepspeed") else f.kind),
    Ask("tmp.ds_from_file", "Do you want to specify a json file to a DeepSpeed config?", "bool", NO,
        when=lambda f: f.is_kind(DistributedType.DEEPSPEED)),
    Ask("deepspeed_config.deepspeed_config_file", "Please enter the path to the json DeepSpeed config file", "str", "none",
        when=lambda f: f.tmp.get("ds_from_file")),
    Ask("deepspeed_config.zero_stage", "What should be your DeepSpeed's ZeRO optimization stage?", "menu", choices=[0, 1, 2, 3],
        parse=int, menu_default=2, when=_ds_manual),
    Ask("deepspeed_config.offload_optimizer_device", "Where to offload optimizer states?", "menu", choices=_OFFLOAD_DEVICES,
        when=lambda f: _ds_manual(f) and f.deepspeed_config["zero_stage"] >= 2),
    Ask("deepspeed_config.offload_param_device", "Where to offload parameters?", "menu", choices=_OFFLOAD_DEVICES,
        when=lambda f: _ds_manual(f) and f.deepspeed_config["zero_stage"] >= 2),
    Ask("deepspeed_config.offload_param_nvme_path", "Nvme Path to offload parameters?", "str", "/nvme",
        when=lambda f: f.deepspeed_config.get("offload_param_device") == "nvme"),
    Ask("deepspeed_config.offload_optimizer_nvme_path", "Nvme Path to offload optimizer states?", "str", "/nvme",
        when=lambda f: f.deepspeed_config.get("offload_optimizer_device") == "nvme"),
    Ask("deepspeed_config.gradient_accumulation_steps", "How many gradient accumulation steps you're passing in your script?",
        "int", 1, when=_ds_manual),
    Ask("tmp.ds_clip", "Do you want to use gradient clipping?", "bool", NO, when=_ds_manual),
    Ask("deepspeed_config.gradient_clipping", "What is the gradient clipping value?", "float", 1.0, when=lambda f: f.tmp.get("ds_clip")),
    Ask("deepspeed_config.zero3_save_16bit_model", "Do you want to save 16-bit model weights when using ZeRO Stage-3?", "bool", NO,
        when=lambda f: _ds_manual(f) and f.deepspeed_config["zero_stage"] == 3),
    Ask("deepspeed_config.zero3_init_flag", "Do you want to enable `deepspeed.zero.Init` when using ZeRO Stage-3 for constructing "
        "massive models?", "bool", NO, when=lambda f: f.is_kind(DistributedType.DEEPSPEED)),
    Ask("tmp.ds_moe", "Do you want to enable Mixture-of-Experts training (MoE)?", "bool", NO,
        when=lambda f: f.is_kind(DistributedType.DEEPSPEED)),
    Ask("deepspeed_config.deepspeed_moe_layer_cls_names", "Specify the comma-separated list of transformers MoE layer class names "
        "(case-sensitive), e.g : `MixtralSparseMoeBlock`, `Qwen2MoeSparseMoeBlock` ...", when=lambda f: f.tmp.get("ds_moe")),
    Ask("deepspeed_config.deepspeed_multinode_launcher", "Which Type of launcher do you want to use?", "menu",
        choices=DEEPSPEED_MULTINODE_LAUNCHERS, when=lambda f: f.is_kind(DistributedType.DEEPSPEED) and f.multi_node),
    Ask("deepspeed_config.deepspeed_hostfile", "DeepSpeed configures multi-node compute resources with hostfile. Each row is of the "
        "format `hostname slots=[num_gpus]`. Please specify the location of hostfile",
        when=lambda f: f.deepspeed_config.get("deepspeed_multinode_launcher") not in (None, DEEPSPEED_MULTINODE_LAUNCHERS[1])),
    Ask("tmp.ds_exclude", "Do you want to specify exclusion filter string?", "bool", NO,
        when=lambda f: "deepspeed_hostfile" in f.deepspeed_config),
    Ask("deepspeed_config.deepspeed_exclusion_filter", "DeepSpeed exclusion filter string", when=lambda f: f.tmp.get("ds_exclude")),
    Ask("tmp.ds_include", "Do you want to specify inclusion filter string?", "bool", NO,
        when=lambda f: "deepspeed_hostfile" in f.deepspeed_config),
    Ask("deepspeed_config.deepspeed_inclusion_filter", "DeepSpeed inclusion filter string", when=lambda f: f.tmp.get("ds_include")),
]


def _fsdp(f: Flow) -> bool:
    return f.is_kind(DistributedType.FSDP)


def _fsdp_v(version: int) -> Cond:
    return lambda f: _fsdp(f) and f.fsdp_config.get("fsdp_version") == version


FSDP = [
    Ask("tmp.use_fsdp", "Do you want to use FullyShardedDataParallel?", "bool", NO,
        when=lambda f: f.is_kind(DistributedType.MULTI_GPU)),
    _derive("top.distributed_type", lambda f: DistributedType.FSDP if f.tmp.get("use_fsdp") else f.kind),
    Ask("fsdp_config.fsdp_version", "What should be your FSDP version?", "menu", choices=[1, 2], parse=lambda i: int(i) + 1,
        menu_default=1, when=_fsdp),
    Ask("fsdp_config.fsdp_reshard_after_forward", "What should be your sharding strategy?", "menu", choices=FSDP_SHARDING_STRATEGY,
        when=_fsdp_v(1)),
    Ask("fsdp_config.fsdp_reshard_after_forward", "Do you want to enable resharding after forward?", "bool", YES, when=_fsdp_v(2)),
    Ask("fsdp_config.fsdp_offload_params", "Do you want to offload parameters and gradients to CPU?", "bool", NO, when=_fsdp),
    Ask("fsdp_config.fsdp_auto_wrap_policy", "What should be your auto wrap policy?", "menu", choices=FSDP_AUTO_WRAP_POLICY, when=_fsdp),
    Ask("tmp.fsdp_no_split", "Do you want to use the model's `_no_split_modules` to wrap. Only applicable for 🤗 Transformers", "bool",
        NO, when=lambda f: f.fsdp_config.get("fsdp_auto_wrap_policy") == FSDP_AUTO_WRAP_POLICY[0]),
    Ask("fsdp_config.fsdp_transformer_layer_cls_to_wrap", "Specify the comma-separated list of transformer layer class names "
        "(case-sensitive) to wrap ,e.g, :`LlamaDecoderLayer`, `BertLayer` ...",
        when=lambda f: f.fsdp_config.get("fsdp_auto_wrap_policy") == FSDP_AUTO_WRAP_POLICY[0] and not f.tmp.get("fsdp_no_split")),
    Ask("fsdp_config.fsdp_min_num_params", "What should be your FSDP's minimum number of parameters for Default Auto Wrapping "
        "Policy?", "int", 100000000, when=lambda f: f.fsdp_config.get("fsdp_auto_wrap_policy") == FSDP_AUTO_WRAP_POLICY[1]),
    Ask("fsdp_config.fsdp_backward_prefetch", "What should be your FSDP's backward prefetch policy?", "menu",
        choices=FSDP_BACKWARD_PREFETCH, when=_fsdp_v(1)),
    Ask("fsdp_config.fsdp_state_dict_type", "What should be your FSDP's state dict type?", "menu", choices=FSDP_STATE_DICT_TYPE,
        menu_default=2, when=_fsdp_v(1)),
    Ask("fsdp_config.fsdp_state_dict_type", "What should be your FSDP's state dict type?", "menu", choices=FSDP2_STATE_DICT_TYPE,
        menu_default=0, when=_fsdp_v(2)),
    Ask("fsdp_config.fsdp_forward_prefetch", "Do you want to enable FSDP's forward prefetch policy?", "bool", NO, when=_fsdp_v(1)),
    Ask("fsdp_config.fsdp_use_orig_params", "Do you want to enable FSDP's `use_orig_params` feature?", "bool", YES, when=_fsdp_v(1)),
    Ask("fsdp_config.fsdp_cpu_ram_efficient_loading", "Do you want to enable CPU RAM efficient model loading? Only applicable for 🤗 "
        "Transformers models.", "bool", YES, when=_fsdp),
    Ask("fsdp_config.fsdp_sync_module_states", "Do you want each individually wrapped FSDP unit to broadcast module parameters from "
        "rank 0 at the start?", "bool", YES, when=lambda f: _fsdp_v(1)(f) and not f.fsdp_config["fsdp_cpu_ram_efficient_loading"]),
    _derive("fsdp_config.fsdp_sync_module_states",
            lambda f: True if f.fsdp_config.get("fsdp_cpu_ram_efficient_loading") else f.fsdp_config.get("fsdp_sync_module_states")),
    Ask("fsdp_config.fsdp_activation_checkpointing", "Do you want to enable FSDP activation checkpointing?", "bool", NO, when=_fsdp),
]

_PC = "parallelism_config.parallelism_config_"


def _pc(f: Flow) -> bool:
    return bool(f.tmp.get("use_parallelism_config"))


PARALLELISM = [
    Ask("tmp.use_parallelism_config", "Do you want to use the parallelism config?", "bool", NO, when=_fsdp_v(2)),
    Ask(_PC + "dp_replicate_size", "What is the data parallelism replicate size?", "int", 1, when=_pc),
    Ask(_PC + "dp_shard_size", "What is the FSDP shard size?", "int", 1, when=_pc),
    Ask(_PC + "tp_size", "What is the tensor parallelism size?", "int", 1, when=_pc),
    Ask(_PC + "cp_size", "What is the context parallelism size?", "int", 1, when=_pc),
    Ask(_PC + "cp_comm_strategy", "What is the context parallelism communication strategy?", "menu",
        choices=["allgather", "alltoall"], when=lambda f: _pc(f) and f.parallelism_config["parallelism_config_cp_size"] > 1),
    Ask(_PC + "sp_size", "What is the sequence parallelism (Ulysses) size?", "int", 1,
        when=lambda f: _pc(f) and f.parallelism_config["parallelism_config_cp_size"] == 1),
    Ask(_PC + "sp_seq_length_is_variable", "Does the sequence length change between batches?", "bool", YES,
        when=lambda f: f.parallelism_config.get("parallelism_config_sp_size", 1) > 1),
    Ask(_PC + "sp_seq_length", "What is the (fixed) sequence length?", "int",
        when=lambda f: f.parallelism_config.get("parallelism_config_sp_seq_length_is_variable") is False),
]

_MG = "megatron_lm_config.megatron_lm_"


def _megatron(f: Flow) -> bool:
    return f.is_kind(DistributedType.MEGATRON_LM)


MEGATRON = [
    Ask("tmp.use_megatron", "Do you want to use Megatron-LM ?", "bool", NO, when=lambda f: f.is_kind(DistributedType.MULTI_GPU)),
    _derive("top.distributed_type", lambda f: DistributedType.MEGATRON_LM if f.tmp.get("use_megatron") else f.kind),
    Ask(_MG + "tp_degree", "What is the Tensor Parallelism degree/size?", "int", 1, when=_megatron),
    Ask(_MG + "sequence_parallelism", "Do you want to enable Sequence Parallelism?", "bool", YES,
        when=lambda f: _megatron(f) and f.megatron_lm_config["megatron_lm_tp_degree"] > 1),
    Ask(_MG + "pp_degree", "What is the Pipeline Parallelism degree/size?", "int", 1, when=_megatron),
    Ask(_MG + "num_micro_batches", "What is the number of micro-batches?", "int", 1,
        when=lambda f: _megatron(f) and f.megatron_lm_config["megatron_lm_pp_degree"] > 1),
    Ask(_MG + "recompute_activations", "Do you want to enable selective activation recomputation?", "bool", YES, when=_megatron),
    Ask(_MG + "use_distributed_optimizer", "Do you want to use distributed optimizer which shards optimizer state and gradients "
        "across data parallel ranks?", "bool", YES, when=_megatron),
    Ask(_MG + "gradient_clipping", "What is the gradient clipping value based on global L2 Norm (0 to disable)?", "float", 1.0,
        when=_megatron),
]

_MANY = (DistributedType.MULTI_GPU, DistributedType.MULTI_CPU, DistributedType.FSDP, DistributedType.DEEPSPEED,
         DistributedType.MEGATRON_LM)


RESOURCES = [
    Ask("top.num_processes", "How many processes should be used for distributed training (one per device)?", "int", 1,
        when=lambda f: f.is_kind(*_MANY)),
    Ask("top.gpu_ids", "What GPU(s) (by id) should be used for training on this machine as a comma-separated list?", "str", "all",
        when=lambda f: not f.on_cpu and not f.is_kind(DistributedType.MULTI_CPU)),
    Ask("top.enable_cpu_affinity", "Would you like to enable numa efficiency? (Currently only supported on NVIDIA hardware).", "bool",
        NO, when=lambda f: f.is_kind(DistributedType.NO, DistributedType.MULTI_GPU) and not f.on_cpu),
]

_TE = lambda f: f.fp8_config.get("backend") == "TE"  # noqa: E731
_FP8 = lambda f: str(f.top.get("mixed_precision")) == "fp8"  # noqa: E731

PRECISION = [
    Ask("top.mixed_precision", "Do you wish to use mixed precision?", "menu", choices=["no", "fp16", "bf16", "fp8"],
        parse=lambda i: str(io._convert_mixed_precision(i)),
        when=lambda f: not f.on_cpu and not f.tmp.get("ds_from_file")),
    Ask("fp8_config.backend", "Which FP8 backend do you want to use?", "menu",
        choices=["native (MXFP8 block-scaled GEMMs of this build)", "te", "torchao", "msamp"],
        parse=lambda i: ["NATIVE", "TE", "AO", "MSAMP"][int(i)], when=_FP8),
    Ask("fp8_config.recipe", "Which native fp8 recipe do you want to use?", "menu", choices=["mxfp8", "tensorwise"],
        when=lambda f: f.fp8_config.get("backend") == "NATIVE"),
    Ask("fp8_config.use_autocast_during_eval", "Do you want to use FP8 autocast during eval mode? Generally better metrics are found "
        "when this is disabled", "bool", NO, when=_TE),
    Ask("fp8_config.margin", "What margin should be used for gradient scaling?", "int", 0, when=_TE),
    Ask("fp8_config.interval", "What interval should be used for for how often the scaling factor is recomputed?", "int", 1, when=_TE),
    Ask("fp8_config.fp8_format", "Which weight format should be used?", "menu", choices=["HYBRID", "E4M3", "E5M2"], when=_TE),
    Ask("fp8_config.amax_history_length", "What length of history should be used for the amax scaling factor computation?", "int",
        1024, when=_TE),
    Ask("fp8_config.amax_compute_algorithm", "Which algorithm should be used for the amax scaling factor computation?", "menu",
        choices=["max", "most_recent"], when=_TE),
    Ask("tmp.te_override", "Do you want to to execute `fprop`, `dgrad`, and `wgrad` GEMMS in higher precision?", "bool", NO, when=_TE),
    Ask("tmp.te_fprop", "Should `fprop` be executed in higher precision?", "bool", NO, when=lambda f: f.tmp.get("te_override")),
    Ask("tmp.te_dgrad", "Should `dgrad` be executed in higher precision?", "bool", NO, when=lambda f: f.tmp.get("te_override")),
    Ask("tmp.te_wgrad", "Should `wgrad` be executed in higher precision?", "bool", NO, when=lambda f: f.tmp.get("te_override")),
    Ask("fp8_config.use_mxfp8_block_scaling", "Do you want to use MXFP8 1x32 block scaling (Blackwell)?", "bool", YES, when=_TE),
    Ask("fp8_config.optimization_level", "Which optimization level should be used?", "menu", choices=["O1", "O2"], menu_default=1,
        when=lambda f: f.fp8_config.get("backend") == "MSAMP"),
    Ask("fp8_config.enable_fsdp_float8_all_gather", "Do you want to enable FSDP2 float8 all gather? This is recommended for better "
        "performance if using FSDP2.", "bool", YES, when=lambda f: f.fp8_config.get("backend") == "AO"),
    Ask("fp8_config.pad_inner_dim", "Do you want to pad the inner dimension of weight matrices before float8 matmuls? This is required "
        "for _scaled_mm which has strong constraints on the matrix dimensions.", "bool", YES,
        when=lambda f: f.fp8_config.get("backend") == "AO"),
]


# =========================================================================== driver
def run_questionnaire(flow: Flow | None = None) -> Flow:
    """Ask every section in order; returns the filled ``Flow`` (exposed for tests)."""
    flow = flow or Flow()
    flow.ask(MACHINE[0])
    flow.ask(MACHINE[1])
    _machine_rank_choices(flow)
    flow.run(MACHINE[2:])
    for section in (DYNAMO, DEEPSPEED, FSDP, PARALLELISM, MEGATRON, RESOURCES, PRECISION):
        flow.run(section)
    if flow.tmp.get("te_override") is not None and flow.fp8_config.get("backend") == "TE":
        flow.fp8_config["override_linear_precision"] = tuple(bool(flow.tmp.get(k, False)) for k in ("te_fprop", "te_dgrad", "te_wgrad"))
    return flow


def build_config(flow: Flow) -> ClusterConfig:
    top = flow.top
    if flow.on_cpu:
        precision = "no"
    elif flow.tmp.get("ds_from_file"):
        precision = None  # the DeepSpeed json decides
    else:
        precision = str(top.get("mixed_precision", "no"))
    if top.get("dynamo_config") is None and precision == "no" and flow.tmp.get("use_dynamo") and not flow.on_cpu:
        print("Torch dynamo used without mixed precision requires TF32 to be efficient. Accelerate will enable it by default when launching your scripts.")
    return ClusterConfig(
        compute_environment=ComputeEnvironment.LOCAL_MACHINE,
        distributed_type=flow.kind,
        mixed_precision=precision,
        use_cpu=flow.on_cpu,
        debug=bool(top.get("debug", False)),
        num_processes=int(top.get("num_processes", 1)),
        machine_rank=int(top.get("machine_rank", 0)),
        num_machines=int(top.get("num_machines", 1)),
        gpu_ids=top.get("gpu_ids"),
        main_process_ip=top.get("main_process_ip"),
        main_process_port=top.get("main_process_port"),
        rdzv_backend=top.get("rdzv_backend", "static"),
        same_network=bool(top.get("same_network", True)),
        main_training_function="main",
        enable_cpu_affinity=bool(top.get("enable_cpu_affinity", False)),
        downcast_bf16=False,
        tpu_use_cluster=False,
        tpu_use_sudo=False,
        dynamo_config=flow.dynamo_config,
        deepspeed_config=flow.deepspeed_config,
        fsdp_config={k: v for k, v in flow.fsdp_config.items() if v is not None},
        parallelism_config=flow.parallelism_config,
        megatron_lm_config=flow.megatron_lm_config,
        mpirun_config=flow.mpirun_config,
        fp8_config=flow.fp8_config or None,
    )


def get_cluster_input() -> ClusterConfig:
    return build_config(run_questionnaire())
