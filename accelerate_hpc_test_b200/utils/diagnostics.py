"""Device-side timing, exposed-communication accounting, collective shadow checks and fault injection.

The reference has none of this (its profiling is ``ProfileKwargs`` -> ``torch.profiler``, its only
distributed debug aid is ``ACCELERATE_DEBUG_MODE`` shape verification, recovery is torchelastic's restart).  A framework that
owns its kernels and its cross-GPU memory needs more:

* :class:`DeviceTimer` — named regions timed with CUDA events on the launching stream (never wall clock), optional NVTX
  ranges, summary reduced as the **max over ranks**;
* :class:`CommTrace` — every place where the compute stream waits for a communication stream is bracketed by two events; the
  time between them is communication that was *not* hidden behind compute ("exposed comm ms/step", the second half of the
  BASELINE.json metric);
* shadow mode (``ACCELERATE_B200_COMM_SHADOW=1``) — the engines re-run a collective through torch.distributed next to the
  peer-memory kernel and compare;  poison mode (``ACCELERATE_B200_COMM_POISON=1``) — recycled communication buffers are filled
  with NaN so that a read of stale data cannot go unnoticed;
* fault injection (``ACCELERATE_B200_FAULT_INJECT="kill:rank=1,step=3"``) — makes a rank die at a given optimizer step of the
  first launch attempt, to exercise ``--max_restarts`` + ``save_state`` / ``load_state`` recovery.
"""

from __future__ import annotations

import os
import time
from collections import OrderedDict
from contextlib import contextmanager
from typing import Optional

import torch


def _cuda() -> bool:
    return torch.cuda.is_available()


class DeviceTimer:
    """``with timer.region("fwd"): ...`` — CUDA-event timing of named regions on the current stream.

    On CPU the regions fall back to ``perf_counter`` (tests, gloo runs).  ``summary()`` synchronises once, and returns
    ``{name: {"ms": total, "calls": n, "ms_per_call": ...}}``; with ``reduce=True`` every number is the max over ranks.
    """

    def __init__(self, nvtx: Optional[bool] = None, enabled: bool = True):
        self.enabled = enabled
        self.nvtx = bool(int(os.environ.get("ACCELERATE_B200_NVTX", "0"))) if nvtx is None else nvtx
        self._pending = []  # (name, start_event, end_event)
        self._acc = OrderedDict()

    @contextmanager
    def region(self, name: str):
        if not self.enabled:
            yield
            return
        if _cuda():
            if self.nvtx:
                torch.cuda.nvtx.range_push(name)
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            try:
                yield
            finally:
                e.record()
                if self.nvtx:
                    torch.cuda.nvtx.range_pop()
                self._pending.append((name, s, e))
        else:
            t0 = time.perf_counter()
            try:
                yield
            finally:
                self._add(name, (time.perf_counter() - t0) * 1e3)

    def _add(self, name, ms):
        tot, n = self._acc.get(name, (0.0, 0))
        self._acc[name] = (tot + ms, n + 1)

    def _drain(self):
        if self._pending:
            torch.cuda.synchronize()
            for name, s, e in self._pending:
                self._add(name, s.elapsed_time(e))
            self._pending = []

    def summary(self, reduce: bool = True, reset: bool = True) -> dict:
        self._drain()
        names = list(self._acc)
        vals = torch.tensor([self._acc[n][0] for n in names], dtype=torch.float64)
        if reduce and names and torch.distributed.is_available() and torch.distributed.is_initialized():
            dev = torch.device("cuda", torch.cuda.current_device()) if torch.distributed.get_backend() == "nccl" else torch.device("cpu")
            t = vals.to(dev)
            torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
            vals = t.cpu()
        out = {n: {"ms": float(v), "calls": self._acc[n][1], "ms_per_call": float(v) / max(1, self._acc[n][1])} for n, v in zip(names, vals)}
        if reset:
            self._acc = OrderedDict()
        return out


class CommTrace:
    """Accounting of the time the compute stream spends waiting for communication streams.

    The engines call :meth:`wait_event` / :meth:`wait_stream` instead of the raw stream methods.  When tracing is on, the
    wait is bracketed by two timing events recorded on the waiting stream; nothing else runs on that stream between them, so
    their distance is exactly the stall.  :meth:`exposed_ms` synchronises and returns the accumulated stall.
    """

    def __init__(self, enabled: Optional[bool] = None):
        self.enabled = bool(int(os.environ.get("ACCELERATE_B200_TRACE_COMM", "0"))) if enabled is None else enabled
        self._pairs = []

    def wait_event(self, stream: "torch.cuda.Stream", event: "torch.cuda.Event", tag: str = ""):
        if not self.enabled:
            stream.wait_event(event)
            return
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        stream.wait_event(event)
        b.record(stream)
        self._pairs.append((tag, a, b))

    def wait_stream(self, stream: "torch.cuda.Stream", other: "torch.cuda.Stream", tag: str = ""):
        if not self.enabled:
            stream.wait_stream(other)
            return
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        stream.wait_stream(other)
        b.record(stream)
        self._pairs.append((tag, a, b))

    def exposed_ms(self, by_tag: bool = False, reset: bool = True):
        if not self._pairs:
            return {} if by_tag else 0.0
        torch.cuda.synchronize()
        per = {}
        for tag, a, b in self._pairs:
            per[tag] = per.get(tag, 0.0) + a.elapsed_time(b)
        if reset:
            self._pairs = []
        return per if by_tag else float(sum(per.values()))


# ---------------------------------------------------------------------------------------------- debug switches
def comm_shadow_enabled() -> bool:
    return os.environ.get("ACCELERATE_B200_COMM_SHADOW", "0") == "1"


def comm_poison_enabled() -> bool:
    return os.environ.get("ACCELERATE_B200_COMM_POISON", "0") == "1"


def shadow_check(name: str, ours: torch.Tensor, reference: torch.Tensor, rtol: float = 0.0, atol: float = 0.0):
    """Raise with a precise diagnostic when a peer-memory collective disagrees with its torch.distributed twin."""
    a, b = ours.float(), reference.float()
    bad = ~torch.isclose(a, b, rtol=rtol, atol=atol, equal_nan=True)
    if bool(bad.any()):
        idx = int(bad.flatten().nonzero()[0])
        rank = torch.distributed.get_rank() if torch.distributed.is_initialized() else 0
        raise RuntimeError(
            f"[comm shadow] {name}: rank {rank}: {int(bad.sum())} of {bad.numel()} elements differ from the torch.distributed "
            f"result (first at flat index {idx}: ours {a.flatten()[idx].item()!r} vs reference {b.flatten()[idx].item()!r})"
        )


# ---------------------------------------------------------------------------------------------- fault injection
class FaultInjector:
    """``ACCELERATE_B200_FAULT_INJECT="kill:rank=R,step=S[,code=C][,every_attempt=1]"``.

    ``maybe_fail(step)`` is called by ``AcceleratedOptimizer.step`` with the number of optimizer steps taken so far.  By
    default the fault fires only on the first launch attempt (``TORCHELASTIC_RESTART_COUNT == 0``), so a job started with
    ``--max_restarts 1`` dies once, restarts and must finish from its last checkpoint."""

    def __init__(self, spec: Optional[str] = None):
        spec = os.environ.get("ACCELERATE_B200_FAULT_INJECT", "") if spec is None else spec
        self.kind = None
        self.args = {}
        if spec:
            kind, _, rest = spec.partition(":")
            self.kind = kind.strip()
            for kv in filter(None, rest.split(",")):
                k, _, v = kv.partition("=")
                self.args[k.strip()] = v.strip()
            if self.kind not in ("kill", "raise", "hang"):
                raise ValueError(f"unknown fault kind {self.kind!r} in ACCELERATE_B200_FAULT_INJECT (kill | raise | hang)")

    @property
    def armed(self) -> bool:
        return self.kind is not None

    def maybe_fail(self, step: int, rank: Optional[int] = None):
        if not self.armed:
            return
        if rank is None:
            rank = int(os.environ.get("RANK", "0"))
        if int(self.args.get("rank", 0)) != rank or int(self.args.get("step", 0)) != step:
            return
        if not int(self.args.get("every_attempt", 0)) and int(os.environ.get("TORCHELASTIC_RESTART_COUNT", "0")) > 0:
            return
        msg = f"[fault injection] rank {rank}: injected `{self.kind}` at optimizer step {step}"
        print(msg, flush=True)
        if self.kind == "kill":
            os._exit(int(self.args.get("code", 17)))
        if self.kind == "raise":
            raise RuntimeError(msg)
        if self.kind == "hang":
            time.sleep(float(self.args.get("seconds", 3600)))


_injector: Optional[FaultInjector] = None


def fault_injector() -> FaultInjector:
    global _injector
    if _injector is None:
        _injector = FaultInjector()
    return _injector


def _reset_fault_injector():
    global _injector
    _injector = None


# --------------------------------------------------------------------------- clock record for every timed number
class ClockRecord:
    """``nvidia-smi`` clocks / power / throttle reasons of one GPU sampled every 200 ms while a timed region runs.

    Every kernel- or step-level number this repository reports carries ``summary()`` next to it, so a reader can
    tell a power-capped (normal on a 1 kW part) run from a thermally throttled or clock-locked one::

        with sample_clocks() as clk:
            ms = timed(fn)
        row = {"ms": ms, "clocks": clk.summary()}
    """

    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    REASONS = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")

    def __init__(self, gpu_index: Optional[int] = None, period_ms: int = 200):
        self.gpu_index = gpu_index
        self.period_ms = period_ms
        self._proc = None
        self._lines: list = []
        self._thread = None

    def __enter__(self):
        import subprocess
        import threading

        idx = self.gpu_index
        if idx is None:
            idx = torch.cuda.current_device() if _cuda() else 0
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            if vis:
                try:
                    idx = int(vis.split(",")[idx])
                except (ValueError, IndexError):
                    pass
        try:
            self._proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits", "-lms",
                                           str(self.period_ms), "-i", str(idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except (OSError, ValueError):
            self._proc = None
            return self

        def pump():
            for line in self._proc.stdout:
                self._lines.append(line.strip())

        self._thread = threading.Thread(target=pump, daemon=True)
        self._thread.start()
        return self

    def __exit__(self, *exc):
        if self._proc is not None:
            self._proc.terminate()
            try:
                self._proc.wait(timeout=5)
            except Exception:  # noqa: BLE001
                self._proc.kill()
        return False

    def summary(self) -> dict:
        import statistics

        if self._proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"], "samples": 0}
        sm, smax, power, seen = [], [], [], set()
        for line in self._lines:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                smax.append(float(f[2]))
                power.append(float(f[3]))
            except ValueError:
                continue
            for name, val in zip(self.REASONS, f[4:8]):
                if val.lower().startswith("active"):
                    seen.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(seen)}


def sample_clocks(gpu_index: Optional[int] = None) -> ClockRecord:
    return ClockRecord(gpu_index)


class ClockProbe:
    """Instantaneous SM clock / power / throttle-reason samples through NVML, for timed regions much shorter than the 200 ms
    period ``ClockRecord`` polls at (micro-benchmarks of single kernels).  ``sample()`` costs a few microseconds of host time and
    nothing on the device: call it while the measured work is in flight (after the closing event has been recorded, before the
    synchronize), then put ``summary()`` next to the number::

        probe = ClockProbe()
        for _ in range(iters):
            start.record(); fn(); stop.record(); probe.sample(); torch.cuda.synchronize()
        row["clocks"] = probe.summary()
    """

    _REASON_BITS = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}

    def __init__(self, gpu_index: Optional[int] = None):
        self._nvml = None
        self._handle = None
        self._samples: list = []
        try:
            import pynvml

            pynvml.nvmlInit()
            idx = gpu_index
            if idx is None:
                idx = torch.cuda.current_device() if _cuda() else 0
                vis = os.environ.get("CUDA_VISIBLE_DEVICES")
                if vis:
                    try:
                        idx = int(vis.split(",")[idx])
                    except (ValueError, IndexError):
                        pass
            self._handle = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self._max = pynvml.nvmlDeviceGetMaxClockInfo(self._handle, pynvml.NVML_CLOCK_SM)
            self._nvml = pynvml
        except Exception:  # noqa: BLE001 - no NVML on this host
            self._nvml = None

    def sample(self) -> None:
        if self._nvml is None:
            return
        n = self._nvml
        try:
            mhz = n.nvmlDeviceGetClockInfo(self._handle, n.NVML_CLOCK_SM)
            watts = n.nvmlDeviceGetPowerUsage(self._handle) / 1000.0
            try:
                mask = n.nvmlDeviceGetCurrentClocksEventReasons(self._handle)
            except AttributeError:
                mask = n.nvmlDeviceGetCurrentClocksThrottleReasons(self._handle)
            self._samples.append((mhz, watts, mask))
        except Exception:  # noqa: BLE001
            pass

    def reset(self) -> None:
        self._samples = []

    def summary(self) -> dict:
        import statistics

        if self._nvml is None or not self._samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable"], "samples": 0}
        mask = 0
        for _, _, m in self._samples:
            mask |= m
        return {"sm_mhz": statistics.median(s[0] for s in self._samples), "sm_max_mhz": self._max,
                "power_w_max": round(max(s[1] for s in self._samples), 1), "samples": len(self._samples),
                "reasons": sorted(name for name, bit in self._REASON_BITS.items() if mask & bit)}
