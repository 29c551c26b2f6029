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
