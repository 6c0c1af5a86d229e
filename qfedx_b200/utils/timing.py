"""Device timing helpers: CUDA events on the current stream, max over ranks, L2 flush, and an
nvidia-smi clock sampler (the recipe in /opt/skills/guides/B200_PROFILING.md)."""
from __future__ import annotations

import contextlib
import statistics
import subprocess
import threading
import time

import torch

_L2_FLUSH = None


def flush_l2(nbytes: int = 256 << 20) -> None:
    """Write a buffer larger than the 126 MB L2 so the next timed launch starts cold."""
    global _L2_FLUSH
    if not torch.cuda.is_available():
        return
    if _L2_FLUSH is None or _L2_FLUSH.numel() < nbytes:
        _L2_FLUSH = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    _L2_FLUSH.zero_()


class CudaTimer:
    """Accumulates milliseconds between start()/stop() pairs using CUDA events."""

    def __init__(self):
        self.pairs = []

    def start(self):
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        self._s = ev

    def stop(self):
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        self.pairs.append((self._s, ev))

    def total_ms(self) -> float:
        torch.cuda.synchronize()
        return sum(s.elapsed_time(e) for s, e in self.pairs)


def max_over_ranks(value: float) -> float:
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return value
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.tensor([value], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


class ClockSampler:
    """Samples SM clock and throttle reasons with nvidia-smi while a timed region runs."""
    QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int = 0, period_s: float = 0.2):
        self.gpu_index, self.period_s = gpu_index, period_s
        self.samples, self._stop, self._thr = [], threading.Event(), None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                                      "-i", str(self.gpu_index)], capture_output=True, text=True, timeout=5).stdout
                parts = [p.strip() for p in out.strip().split(",")]
                if len(parts) >= 7:
                    self.samples.append(parts)
            except Exception:
                pass
            self._stop.wait(self.period_s)

    def __enter__(self):
        self._thr = threading.Thread(target=self._run, daemon=True)
        self._thr.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        self._thr.join(timeout=6)

    def summary(self) -> dict:
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for p in self.samples:
            try:
                sm.append(float(p[0])); mx.append(float(p[1]))
            except ValueError:
                continue
            for name, val in zip(names, p[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}
