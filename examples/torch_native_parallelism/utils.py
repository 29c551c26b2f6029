"""Shared helpers of the N-D parallelism examples: synthetic token data, an analytic FLOPs-per-token model for Llama-shaped
transformers, and ``PerformanceTracker`` — tokens/s, steps/s, TFLOP/s per device and peak memory measured after a warm-up, on the
device (CUDA events) when there is one.  (Counterpart of the reference's ``examples/torch_native_parallelism/utils.py``.)"""

import time

import torch


def synthetic_tokens(count: int, sequence_length: int, vocab_size: int, seed: int = 0) -> list:
    g = torch.Generator().manual_seed(seed)
    return [{"input_ids": torch.randint(0, vocab_size, (sequence_length,), generator=g)} for _ in range(count)]


def collate(rows: list) -> dict:
    ids = torch.stack([row["input_ids"] for row in rows])
    return {"input_ids": ids, "labels": ids.clone()}


def get_model_flops_per_token(config, sequence_length: int) -> float:
    """Training FLOPs per token (forward + backward = 3 x forward) of a decoder-only transformer with grouped-query attention and a
    gated MLP: 2 FLOPs per weight per token for every matrix, plus the attention scores / values, which grow with the sequence
    (halved for the causal mask)."""
    h, layers, vocab = config.hidden_size, config.num_hidden_layers, config.vocab_size
    head_dim = getattr(config, "head_dim", None) or h // config.num_attention_heads
    q_width = config.num_attention_heads * head_dim
    kv_width = config.num_key_value_heads * head_dim
    projections = h * (q_width + 2 * kv_width) + q_width * h            # q, k, v and output projections
    mlp = 3 * h * config.intermediate_size                              # gate, up, down
    attention = 2 * q_width * sequence_length / 2                       # QK^T and PV over a causal window: S/2 keys on average
    forward = 2 * layers * (projections + mlp + attention) + 2 * h * vocab
    return 3.0 * forward


class PerformanceTracker:
    """``tracker.step(tokens_in_this_step)`` after every optimizer step; the first ``warmup_steps`` are not measured.

    ``metrics()`` returns ``tokens_per_second`` (this process; multiply by the data-parallel size for the job), ``steps_per_second``,
    ``tflops_per_device`` when ``flops_per_token`` is known, and the peak allocated / reserved memory in GB on a GPU."""

    def __init__(self, warmup_steps: int = 3, flops_per_token: float = None, device=None):
        self.warmup_steps, self.flops_per_token = warmup_steps, flops_per_token
        self.device = torch.device(device) if device is not None else torch.device("cuda" if torch.cuda.is_available() else "cpu")
        self.cuda = self.device.type == "cuda"
        self.steps = self.tokens = 0
        self._seen = 0
        self._start = self._start_event = None

    def step(self, tokens: int) -> None:
        self._seen += 1
        if self._seen == self.warmup_steps:
            if self.cuda:
                torch.cuda.reset_peak_memory_stats(self.device)
                self._start_event = torch.cuda.Event(enable_timing=True)
                self._start_event.record()
            self._start = time.perf_counter()
        elif self._seen > self.warmup_steps:
            self.steps += 1
            self.tokens += tokens

    @property
    def measuring(self) -> bool:
        return self._start is not None and self.steps > 0

    def metrics(self) -> dict:
        if not self.measuring:
            return {}
        if self.cuda:
            end = torch.cuda.Event(enable_timing=True)
            end.record()
            end.synchronize()
            seconds = self._start_event.elapsed_time(end) / 1e3
        else:
            seconds = time.perf_counter() - self._start
        out = {"steps": self.steps, "total_tokens": self.tokens, "seconds": seconds,
               "tokens_per_second": self.tokens / seconds, "steps_per_second": self.steps / seconds}
        if self.flops_per_token:
            out["tflops_per_device"] = self.tokens * self.flops_per_token / seconds / 1e12
        if self.cuda:
            out["peak_allocated_gb"] = torch.cuda.max_memory_allocated(self.device) / 2**30
            out["peak_reserved_gb"] = torch.cuda.max_memory_reserved(self.device) / 2**30
        return out

    def summary(self) -> str:
        m = self.metrics()
        if not m:
            return "no measured steps yet"
        text = f"{m['tokens_per_second']:,.0f} tokens/s, {m['steps_per_second']:.2f} steps/s over {m['steps']} steps"
        if "tflops_per_device" in m:
            text += f", {m['tflops_per_device']:.1f} TFLOP/s per device"
        if "peak_allocated_gb" in m:
            text += f", peak {m['peak_allocated_gb']:.1f} GB allocated / {m['peak_reserved_gb']:.1f} GB reserved"
        return text
