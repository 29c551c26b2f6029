"""Big-model inference benchmark: for each placement of the same checkpoint — everything on the first device, a balanced device map
over the visible GPUs, part of the layers offloaded to CPU, part to disk — measure load time, time to the first token, time per
generated token and peak memory, and check that every placement produces exactly the tokens of the unsplit model.  The counterpart of
the reference's ``benchmarks/big_model_inference/big_model_inference.py`` (which times hub checkpoints; there is no hub here, so the
checkpoint is written first from a seeded native Llama).

    python benchmarks/big_model_inference.py --model tiny --cpu                    # what tests/test_benchmarks.py runs
    python benchmarks/big_model_inference.py --model llama3-8b --new-tokens 32     # on a B200 node
"""

import argparse
import json
import os
import resource
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch


def config_for(name, layers):
    from accelerate_hpc_test_b200.models import LlamaConfig

    cfg = {"tiny": lambda: LlamaConfig.tiny(head_dim=32, hidden_size=128, intermediate_size=256, vocab_size=512, num_attention_heads=4,
                                            num_key_value_heads=2, num_hidden_layers=4),
           "llama3-1b": LlamaConfig.llama3_1b, "llama3-8b": LlamaConfig.llama3_8b}[name]()
    if layers:
        cfg.num_hidden_layers = layers
    return cfg


def write_checkpoint(cfg, dtype, folder, shard_bytes):
    """A seeded model saved the way ``save_model`` does it: sharded safetensors + index."""
    from accelerate_hpc_test_b200 import Accelerator
    from accelerate_hpc_test_b200.models import LlamaForCausalLM

    torch.manual_seed(0)
    model = LlamaForCausalLM(cfg, dtype=dtype, device="cpu").eval()
    Accelerator(cpu=True).save_model(model, folder, max_shard_size=shard_bytes, safe_serialization=True)
    return model


@torch.no_grad()
def generate(model, ids, new_tokens, sync):
    """Greedy decoding without a KV cache (the native model has none): every step re-runs the prefix, which is what offloaded
    placements are judged on — how fast weights stream in per forward."""
    first = None
    t0 = time.perf_counter()
    for _ in range(new_tokens):
        logits = model(ids, return_logits=True).logits[:, -1]
        ids = torch.cat([ids, logits.argmax(-1, keepdim=True).to(ids.device)], dim=1)
        if first is None:
            sync()
            first = time.perf_counter() - t0
    sync()
    total = time.perf_counter() - t0
    return ids, first, (total - first) / max(new_tokens - 1, 1)


def placements(cfg, device, n_gpus):
    """name -> device_map.  Layer-granular maps so that cpu / disk really hold whole decoder layers."""
    n = cfg.num_hidden_layers
    first = 0 if device.type == "cuda" else "cpu"

    def layered(where_of_layer):
        dm = {"model.embed_tokens": first, "model.norm": first, "lm_head": first}
        dm.update({f"model.layers.{i}": where_of_layer(i) for i in range(n)})
        return dm

    out = {"one_device": {"": first}}
    if n_gpus > 1:
        out["balanced"] = "balanced"
    out["half_on_cpu"] = layered(lambda i: first if i < n // 2 else "cpu") if device.type == "cuda" else None
    out["half_on_disk"] = layered(lambda i: first if i < n // 2 else "disk")
    return {k: v for k, v in out.items() if v is not None}


def main(argv=None):
    p = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    p.add_argument("--model", default="tiny", choices=["tiny", "llama3-1b", "llama3-8b"])
    p.add_argument("--layers", type=int, default=0)
    p.add_argument("--cpu", action="store_true")
    p.add_argument("--prompt-tokens", type=int, default=32)
    p.add_argument("--new-tokens", type=int, default=8)
    p.add_argument("--shard-size", default="2GB")
    p.add_argument("--out", default=None)
    args = p.parse_args(argv)

    from accelerate_hpc_test_b200 import init_empty_weights, load_checkpoint_and_dispatch
    from accelerate_hpc_test_b200.models import LlamaForCausalLM
    from accelerate_hpc_test_b200.utils import release_memory

    use_cuda = torch.cuda.is_available() and not args.cpu
    device = torch.device("cuda", 0) if use_cuda else torch.device("cpu")
    dtype = torch.bfloat16 if use_cuda else torch.float32
    sync = torch.cuda.synchronize if use_cuda else (lambda: None)
    cfg = config_for(args.model, args.layers)
    rows = []
    with tempfile.TemporaryDirectory() as folder:
        reference = write_checkpoint(cfg, dtype, folder, args.shard_size)
        torch.manual_seed(1)
        prompt = torch.randint(0, cfg.vocab_size, (1, args.prompt_tokens))
        expected = None
        if not use_cuda:                       # the unsplit model as the oracle (on GPU the one_device row plays that part)
            expected, _, _ = generate(reference, prompt, args.new_tokens, sync)
        reference = release_memory(reference)
        for name, device_map in placements(cfg, device, torch.cuda.device_count() if use_cuda else 0).items():
            if use_cuda:
                torch.cuda.empty_cache()
                torch.cuda.reset_peak_memory_stats()
            t0 = time.perf_counter()
            with init_empty_weights():
                model = LlamaForCausalLM(cfg, dtype=dtype)
            model = load_checkpoint_and_dispatch(model, folder, device_map=device_map, dtype=dtype, no_split_module_classes=["LlamaDecoderLayer"],
                                                 offload_folder=os.path.join(folder, f"offload_{name}")).eval()
            sync()
            load_s = time.perf_counter() - t0
            tokens, first_s, per_token_s = generate(model, prompt.to(device), args.new_tokens, sync)
            tokens = tokens.cpu()
            if expected is None:
                expected = tokens
            row = {"placement": name, "load_s": round(load_s, 3), "first_token_s": round(first_s, 4), "per_token_s": round(per_token_s, 4),
                   "same_tokens": bool(torch.equal(tokens, expected)),
                   "host_peak_gb": round(resource.getrusage(resource.RUSAGE_SELF).ru_maxrss / 2**20, 2)}
            if use_cuda:
                row["gpu_peak_gb"] = [round(torch.cuda.max_memory_allocated(i) / 2**30, 2) for i in range(torch.cuda.device_count())]
            rows.append(row)
            print(json.dumps(row), flush=True)
            model = release_memory(model)
    summary = {"model": args.model, "layers": cfg.num_hidden_layers, "dtype": str(dtype).split(".")[-1], "device": device.type,
               "prompt_tokens": args.prompt_tokens, "new_tokens": args.new_tokens, "rows": rows}
    if args.out:
        with open(args.out, "w") as f:
            json.dump(summary, f, indent=1)
    assert all(r["same_tokens"] for r in rows), "a placement changed the generated tokens"
    return summary


if __name__ == "__main__":
    main()
