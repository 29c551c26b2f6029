from .llama import CausalLMOutput, LlamaConfig, LlamaForCausalLM, LlamaModel, LlamaDecoderLayer, B200Linear, B200RMSNorm
