"""Engine-style (Megatron-LM) API on the native stack: ``engine(**batch)`` is a full optimisation step over
micro-batches (reference utils/megatron_lm.py:924-1158, tests rely on examples/by_feature/megatron_lm_gpt_pretraining.py)."""

import torch

from accelerate_hpc_test_b200 import Accelerator
from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
from accelerate_hpc_test_b200.optim import FusedAdamW
from accelerate_hpc_test_b200.utils import (GPTTrainStep, MegatronEngine, MegatronLMOptimizerWrapper, MegatronLMSchedulerWrapper,
                                            avg_losses_across_data_parallel_group)


def test_engine_step_equals_manual_accumulation():
    cfg = LlamaConfig.tiny(head_dim=32, hidden_size=64, intermediate_size=128, vocab_size=128, num_hidden_layers=2)
    torch.manual_seed(0)
    a = LlamaForCausalLM(cfg, dtype=torch.float32)
    torch.manual_seed(0)
    b = LlamaForCausalLM(cfg, dtype=torch.float32)
    acc = Accelerator(cpu=True)
    oa = FusedAdamW(a.parameters(), lr=1e-2)
    ob = FusedAdamW(b.parameters(), lr=1e-2)
    sched = torch.optim.lr_scheduler.LambdaLR(oa, lambda s: 1.0 / (1 + s))
    engine = MegatronEngine(acc, a, MegatronLMOptimizerWrapper(oa), MegatronLMSchedulerWrapper(sched, oa))
    engine.num_micro_batches = 2
    assert isinstance(engine.train_step_handler, GPTTrainStep)
    ids = torch.randint(0, cfg.vocab_size, (4, 16))
    engine.train()
    loss = engine(input_ids=ids)
    # manual: two micro-batches, averaged
    ob.zero_grad()
    ref = 0.0
    for i in range(2):
        out = b(ids[2 * i : 2 * i + 2], labels=ids[2 * i : 2 * i + 2])
        (out.loss / 2).backward()
        ref += out.loss.item() / 2
    ob.step()
    assert abs(float(loss) - ref) < 1e-5
    for p, q in zip(a.parameters(), b.parameters()):
        assert torch.allclose(p, q, atol=1e-6)
    assert abs(oa.param_groups[0]["lr"] - 0.5e-2) < 1e-9  # the engine stepped the schedule
    engine.eval()
    ev = engine(input_ids=ids)
    assert torch.isfinite(ev) and engine.log_eval_results()
    assert avg_losses_across_data_parallel_group([torch.tensor(2.0)]).item() == 2.0


def test_model_configs_are_translated_into_megatron_arguments():
    import types

    import pytest

    from accelerate_hpc_test_b200.models import LlamaConfig, LlamaForCausalLM
    from accelerate_hpc_test_b200.utils import MegatronLMPlugin
    from accelerate_hpc_test_b200.utils.dataclasses import MODEL_CONFIGS_TO_MEGATRON_PARSERS

    assert {"megatron-bert", "bert", "gpt2", "t5", "llama"} <= set(MODEL_CONFIGS_TO_MEGATRON_PARSERS)

    class GPT2LMHeadModel:
        config = types.SimpleNamespace(n_layer=12, n_embd=768, n_head=12, n_positions=1024, vocab_size=50257, return_dict=True)

    plugin = MegatronLMPlugin()
    MODEL_CONFIGS_TO_MEGATRON_PARSERS["gpt2"](plugin, GPT2LMHeadModel(), {"input_ids": torch.zeros(4, 512, dtype=torch.long)})
    args = plugin.megatron_lm_default_args
    assert args["model_type_name"] == "gpt" and args["num_layers"] == 12 and args["hidden_size"] == 768 and args["seq_length"] == 512
    assert args["tokenizer_type"] == "GPT2BPETokenizer" and args["pretraining_flag"] is True and args["orig_vocab_size"] == 50257
    # an explicit decoder length wins over the batch, and over a differing generic one (with a warning)
    plugin = MegatronLMPlugin(seq_length=256, decoder_seq_length=128)
    with pytest.warns(UserWarning, match="Using `decoder_seq_length`"):
        MODEL_CONFIGS_TO_MEGATRON_PARSERS["gpt2"](plugin, GPT2LMHeadModel(), None)
    assert plugin.seq_length == 128

    class BertForMaskedLM:
        config = types.SimpleNamespace(num_hidden_layers=2, hidden_size=64, num_attention_heads=4, max_position_embeddings=128,
                                       num_labels=2, vocab_size=1000, return_dict=True)

    class BertForSequenceClassification(BertForMaskedLM):
        pass

    plugin = MegatronLMPlugin()
    MODEL_CONFIGS_TO_MEGATRON_PARSERS["megatron-bert"](plugin, BertForMaskedLM(), None)
    assert plugin.megatron_lm_default_args["pretraining_flag"] is True and plugin.seq_length == 128  # falls back to the position table
    plugin = MegatronLMPlugin()
    MODEL_CONFIGS_TO_MEGATRON_PARSERS["bert"](plugin, BertForSequenceClassification(), None)
    assert plugin.megatron_lm_default_args["pretraining_flag"] is False and plugin.megatron_lm_default_args["num_labels"] == 2

    class T5ForConditionalGeneration:
        config = types.SimpleNamespace(num_layers=6, d_model=512, num_heads=8, vocab_size=32128, return_dict=True)

    plugin = MegatronLMPlugin()
    batch = {"input_ids": torch.zeros(2, 96, dtype=torch.long), "labels": torch.zeros(2, 24, dtype=torch.long)}
    MODEL_CONFIGS_TO_MEGATRON_PARSERS["t5"](plugin, T5ForConditionalGeneration(), batch)
    args = plugin.megatron_lm_default_args
    assert (args["encoder_seq_length"], args["decoder_seq_length"], args["max_position_embeddings"]) == (96, 24, 1024)

    # the native Llama: grouped-query attention, SwiGLU width and RoPE base come along
    cfg = LlamaConfig.tiny(head_dim=32, hidden_size=64, intermediate_size=128, vocab_size=128, num_attention_heads=4, num_key_value_heads=2)
    model = LlamaForCausalLM(cfg, dtype=torch.float32)
    plugin = MegatronLMPlugin()
    MODEL_CONFIGS_TO_MEGATRON_PARSERS["llama"](plugin, model, {"input_ids": torch.zeros(2, 48, dtype=torch.long)})
    args = plugin.megatron_lm_default_args
    assert args["model_type_name"] == "gpt" and args["num_query_groups"] == 2 and args["ffn_hidden_size"] == 128 and args["swiglu"] is True
    assert args["seq_length"] == 48 and args["normalization"] == "RMSNorm"


def test_glm4_moe_config_is_translated_with_its_routing_and_layer_layout():
    import types

    from accelerate_hpc_test_b200.utils import MegatronLMPlugin
    from accelerate_hpc_test_b200.utils.dataclasses import MODEL_CONFIGS_TO_MEGATRON_PARSERS, parse_glm4_moe_config

    assert MODEL_CONFIGS_TO_MEGATRON_PARSERS["glm4_moe"] is parse_glm4_moe_config

    class Glm4MoeForCausalLM:
        config = types.SimpleNamespace(
            model_type="glm4_moe", num_hidden_layers=6, hidden_size=256, num_attention_heads=8, num_key_value_heads=2, head_dim=32,
            vocab_size=4096, max_position_embeddings=8192, intermediate_size=512, moe_intermediate_size=128, n_routed_experts=16,
            num_experts_per_tok=4, routed_scaling_factor=2.5, first_k_dense_replace=1, use_qk_norm=False, rope_theta=1e6,
            partial_rotary_factor=0.5, eos_token_id=2, bf16=True, tie_word_embeddings=False, return_dict=True)

    plugin = MegatronLMPlugin()
    parse_glm4_moe_config(plugin, Glm4MoeForCausalLM(), {"input_ids": torch.zeros(2, 640, dtype=torch.long)})
    a = plugin.megatron_lm_default_args
    assert a["model_type_name"] == "gpt" and a["pretraining_flag"] is False and a["original_model_type"] == "glm4_moe"
    assert a["seq_length"] == 640 == plugin.seq_length                       # no config length, no plugin length: the batch decides
    assert a["moe_layer_freq"] == [0, 1, 1, 1, 1, 1] and a["num_experts"] == 16 and a["moe_router_topk"] == 4
    assert a["moe_ffn_hidden_size"] == a["moe_shared_expert_intermediate_size"] == 128 and a["ffn_hidden_size"] == 512
    assert a["moe_router_score_function"] == "sigmoid" and a["moe_router_topk_scaling_factor"] == 2.5 and a["moe_grouped_gemm"] is True
    assert a["num_query_groups"] == 2 and a["kv_channels"] == 32 and a["qk_layernorm"] is False and a["add_qkv_bias"] is True
    assert a["rotary_percent"] == 0.5 and a["rotary_base"] == 1e6 and a["norm_epsilon"] == 1e-3
    assert a["untie_embeddings_and_output_weights"] is True and a["tokenizer_type"] == "HuggingFaceTokenizer" and "fp8" not in a

    # the config's own maximum wins over the family field; an explicit plugin length wins over both; fp8 parameters ride along
    class WithLength(Glm4MoeForCausalLM):
        config = types.SimpleNamespace(**{**vars(Glm4MoeForCausalLM.config), "max_sequence_length": 2048, "fp8_param": True,
                                          "fp8": "hybrid", "fp8_param_gather": True, "fp8_recipe": "mxfp8"})

    plugin = MegatronLMPlugin(decoder_seq_length=1024)
    parse_glm4_moe_config(plugin, WithLength(), None)
    assert plugin.seq_length == 2048 and plugin.megatron_lm_default_args["fp8_recipe"] == "mxfp8"
    plugin = MegatronLMPlugin(seq_length=512)
    parse_glm4_moe_config(plugin, WithLength(), None)
    assert plugin.seq_length == 512


def test_plugin_records_batch_geometry_optimizer_and_schedule():
    import types

    import pytest

    from accelerate_hpc_test_b200.utils import MegatronLMDummyScheduler, MegatronLMPlugin

    plugin = MegatronLMPlugin(num_micro_batches=4, train_samples=1000, lr_warmup_samples=10)
    plugin.set_training_args(micro_batch_size=2, dp_degree=8)
    assert plugin.global_batch_size == 64 and plugin.megatron_lm_default_args["micro_batch_size"] == 2
    layer = torch.nn.Linear(2, 2)
    plugin.set_optimizer_type(torch.optim.AdamW(layer.parameters(), lr=3e-4, betas=(0.9, 0.95), eps=1e-6, weight_decay=0.1))
    args = plugin.megatron_lm_default_args
    assert (args["optimizer"], args["adam_beta2"], args["adam_eps"], args["lr"], args["weight_decay"]) == ("adam", 0.95, 1e-6, 3e-4, 0.1)
    plugin.set_optimizer_type(torch.optim.SGD(layer.parameters(), lr=0.1, momentum=0.9))
    assert args["optimizer"] == "sgd" and args["sgd_momentum"] == 0.9
    with pytest.raises(ValueError, match="rmsprop is not supported by Megatron-LM"):
        plugin.set_optimizer_type(torch.optim.RMSprop(layer.parameters()))
    with pytest.warns(UserWarning) as caught:
        plugin.set_scheduler_args(MegatronLMDummyScheduler(None, total_num_steps=800, warmup_num_steps=80))
    assert {"train_samples", "lr_warmup_samples"} <= {word for w in caught for word in ("train_samples", "lr_warmup_samples") if word in str(w.message)}
    assert (plugin.train_iters, plugin.lr_warmup_iters, plugin.train_samples, plugin.lr_warmup_samples) == (100, 10, None, 0)
    assert args["train_iters"] == 100 and args["lr_decay_style"] == "linear" and args["min_lr"] == 0
    plugin.set_mixed_precision("bf16")
    assert args["bf16"] is True and args["DDP_impl"] == "local"
    plugin.set_tensorboard_logging_options()
    assert args["log_timers_to_tensorboard"] is True and args["log_validation_ppl_to_tensorboard"] is True
    model = types.SimpleNamespace(config=types.SimpleNamespace(model_type="mamba"))
    with pytest.raises(ValueError, match="not supports mamba model"):
        plugin.set_network_size_args(model)
    cfg = LlamaConfig.tiny(head_dim=32, hidden_size=64, intermediate_size=128, vocab_size=128)
    plugin.set_network_size_args(LlamaForCausalLM(cfg, dtype=torch.float32), {"input_ids": torch.zeros(2, 24, dtype=torch.long)})
    assert args["hidden_size"] == 64 and args["seq_length"] == 24 and args["model_type_name"] == "gpt"


def _function_style_api():
    from torch.utils.data import DataLoader

    from accelerate_hpc_test_b200.utils import (MegatronLMDummyDataLoader, MegatronLMDummyScheduler, MegatronLMPlugin, megatron_lm_initialize,
                                                megatron_lm_prepare_data_loader, megatron_lm_prepare_model_optimizer_scheduler,
                                                megatron_lm_prepare_optimizer, megatron_lm_prepare_scheduler)

    cfg = LlamaConfig.tiny(head_dim=32, hidden_size=64, intermediate_size=128, vocab_size=128, num_hidden_layers=2)

    def provider():
        torch.manual_seed(0)
        return LlamaForCausalLM(cfg, dtype=torch.float32)

    plugin = MegatronLMPlugin(num_micro_batches=2, lr_decay_style="cosine", min_lr=1e-4, custom_model_provider_function=provider,
                              other_megatron_args={"seed": 77})
    acc = Accelerator(megatron_lm_plugin=plugin)
    assert acc.distributed_type.value == "MEGATRON_LM"
    args = megatron_lm_initialize(acc, args_defaults={"lr": 1e-2, "weight_decay": 0.0})
    assert args.seed == 77 and args.world_size == 2 and args.rank == acc.process_index and plugin.megatron_args is args
    assert args.tensor_model_parallel_size == 1 and args.lr == 1e-2
    first = torch.rand(1)
    megatron_lm_initialize(acc)
    assert torch.equal(torch.rand(1), first)                      # seeded from the arguments, the same on every call and rank

    data = [{"input_ids": torch.randint(0, 128, (16,), generator=torch.Generator().manual_seed(i))} for i in range(32)]
    collate = lambda rows: {"input_ids": torch.stack([r["input_ids"] for r in rows])}  # noqa: E731
    loader = megatron_lm_prepare_data_loader(acc, DataLoader(data, batch_size=2, collate_fn=collate))
    batch = next(iter(loader))
    assert batch["input_ids"].shape == (4, 16)                    # whole-step batches: micro batch x num_micro_batches
    assert len(loader) == 32 // 4 // 2                            # ... sharded over the two data-parallel ranks
    try:
        megatron_lm_prepare_data_loader(acc, MegatronLMDummyDataLoader(data_path="corpus"))
    except NotImplementedError as error:
        assert "Megatron-LM data pipelines are not part of this build" in str(error)
    else:
        raise AssertionError("the indexed-dataset pipeline must be refused")

    plugin.megatron_lm_default_args.update(lr=1e-2, weight_decay=0.0, train_iters=10, lr_warmup_iters=2)
    model, optimizer, scheduler = megatron_lm_prepare_model_optimizer_scheduler(acc)
    assert type(optimizer.optimizer).__name__ == "FusedAdamW" and optimizer.optimizer.param_groups[0]["lr"] == 0.0     # warm-up starts at zero
    rates = []
    for _ in range(5):
        out = model(input_ids=batch["input_ids"], labels=batch["input_ids"])
        acc.backward(out.loss)
        optimizer.step()
        scheduler.step()
        optimizer.zero_grad()
        rates.append(optimizer.optimizer.param_groups[0]["lr"])
    assert max(rates) == 1e-2 and rates[-1] < 1e-2 and min(rates[1:]) >= 1e-4        # warm-up, cosine decay, never below min_lr
    # the pieces on their own
    bare = megatron_lm_prepare_optimizer(acc, provider())
    assert bare.param_groups[0]["lr"] == 1e-2 and bare.param_groups[0]["weight_decay"] == 0.0
    real = torch.optim.lr_scheduler.ConstantLR(bare, factor=1.0)
    assert megatron_lm_prepare_scheduler(acc, bare, real) is real
    built = megatron_lm_prepare_scheduler(acc, bare, MegatronLMDummyScheduler(bare, total_num_steps=4, warmup_num_steps=0))
    assert isinstance(built, torch.optim.lr_scheduler.LambdaLR)
    plugin.custom_model_provider_function = None
    try:
        megatron_lm_prepare_model_optimizer_scheduler(acc)
    except NotImplementedError as error:
        assert "custom_model_provider_function" in str(error)
    else:
        raise AssertionError("without a provider there is no model to build")


def test_function_style_api_two_processes():
    from accelerate_hpc_test_b200 import debug_launcher

    debug_launcher(_function_style_api, num_processes=2)


def _prepare_describes_the_job():
    from torch.utils.data import DataLoader

    from accelerate_hpc_test_b200.utils import MegatronLMPlugin

    plugin = MegatronLMPlugin(num_micro_batches=2)
    acc = Accelerator(megatron_lm_plugin=plugin)
    cfg = LlamaConfig.tiny(head_dim=32, hidden_size=64, intermediate_size=128, vocab_size=128, num_hidden_layers=2)
    model = LlamaForCausalLM(cfg, dtype=torch.float32)
    optimizer = torch.optim.AdamW(model.parameters(), lr=2e-3, betas=(0.9, 0.98), weight_decay=0.05)
    data = [torch.randint(0, 128, (16,)) for _ in range(16)]
    model, optimizer, loader = acc.prepare(model, optimizer, DataLoader(data, batch_size=4))
    described = plugin.megatron_lm_default_args
    assert (described["micro_batch_size"], described["data_parallel_size"], described["global_batch_size"]) == (4, 2, 16)
    assert (described["num_layers"], described["hidden_size"], described["model_type_name"]) == (2, 64, "gpt")
    assert (described["optimizer"], described["lr"], described["adam_beta2"], described["weight_decay"]) == ("adam", 2e-3, 0.98, 0.05)
    ids = next(iter(loader))
    acc.backward(model(input_ids=ids, labels=ids).loss)
    optimizer.step()


def test_prepare_records_the_megatron_description_two_processes():
    from accelerate_hpc_test_b200 import debug_launcher

    debug_launcher(_prepare_describes_the_job, num_processes=2)


def test_t5_masks_engine_config_and_unavailable_pieces():
    import pytest

    from accelerate_hpc_test_b200.utils.megatron_lm import MegatronLMDummyDataLoader, T5TrainStep, finish_mpu_init, model_provider_func

    keep = torch.tensor([[1, 1, 0], [1, 1, 1]])
    blocked = T5TrainStep.attn_mask_postprocess(keep)
    assert blocked.dtype == torch.bool and blocked[0].int().tolist() == [[0, 0, 1], [0, 0, 1], [1, 1, 1]] and not blocked[1].any()
    assert T5TrainStep.get_decoder_mask(3, "cpu")[0].int().tolist() == [[0, 1, 1], [0, 0, 1], [0, 0, 0]]
    cross = T5TrainStep.get_enc_dec_mask(keep, dec_seq_length=2, device="cpu")
    assert cross.shape == (2, 2, 3) and cross[0].int().tolist() == [[0, 0, 1], [0, 0, 1]] and not cross[1].any()

    cfg = LlamaConfig.tiny(head_dim=32, hidden_size=64, intermediate_size=128, vocab_size=128, num_hidden_layers=1)
    model = LlamaForCausalLM(cfg, dtype=torch.float32)
    engine = MegatronEngine(Accelerator(cpu=True), model, MegatronLMOptimizerWrapper(FusedAdamW(model.parameters(), lr=1e-3)))
    engine.num_micro_batches = 3
    described = engine.get_module_config()
    assert described.hidden_size == 64 and described.num_micro_batches == 3 and described.grad_scale_func(2.0) == 2.0

    loader = MegatronLMDummyDataLoader(data_path="corpus", seq_length=128)
    assert loader.dataset_args == {"data_path": "corpus", "seq_length": 128}
    for call in (loader.set_megatron_data_args, lambda: loader.get_train_valid_test_datasets_provider(None),
                 lambda: loader.build_train_valid_test_data_iterators(None)):
        with pytest.raises(NotImplementedError, match="not part of this build"):
            call()
    with pytest.raises(NotImplementedError, match="custom_model_provider_function"):
        model_provider_func()
    finish_mpu_init()
    first = torch.rand(1)
    finish_mpu_init()
    assert torch.equal(torch.rand(1), first)
