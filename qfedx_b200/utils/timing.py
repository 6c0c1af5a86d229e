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
    """Samples SM clock and throttle reasons while a timed region runs (the clocks line of the profiling recipe).

    Reads NVML in-process (pynvml from nvidia-ml-py: no fork per sample, so a 50 ms period does not disturb a
    latency-bound timed region); falls back to one ``nvidia-smi`` subprocess per sample when NVML cannot be loaded."""
    QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int = 0, period_s: float = 0.2):
        self.gpu_index, self.period_s = gpu_index, period_s
        self.samples, self._stop, self._thr = [], threading.Event(), None
        self.source = "nvidia-smi"
        self._nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._handle = pynvml.nvmlDeviceGetHandleByIndex(self._physical_index(pynvml, gpu_index))
            self._nvml = pynvml
            self.source = "nvml"
        except Exception:
            self._nvml = None

    @staticmethod
    def _physical_index(pynvml, idx: int) -> int:
        """torch's device index counts CUDA_VISIBLE_DEVICES entries; NVML counts physical devices."""
        import os
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            ids = [v.strip() for v in vis.split(",") if v.strip()]
            if idx < len(ids) and ids[idx].isdigit():
                return int(ids[idx])
        return idx

    def _sample_nvml(self):
        n = self._nvml
        sm = n.nvmlDeviceGetClockInfo(self._handle, n.NVML_CLOCK_SM)
        mx = n.nvmlDeviceGetMaxClockInfo(self._handle, n.NVML_CLOCK_SM)
        try:
            pw = n.nvmlDeviceGetPowerUsage(self._handle) / 1000.0
        except Exception:
            pw = 0.0
        r = n.nvmlDeviceGetCurrentClocksEventReasons(self._handle) if hasattr(n, "nvmlDeviceGetCurrentClocksEventReasons") \
            else n.nvmlDeviceGetCurrentClocksThrottleReasons(self._handle)
        bit = lambda name_new, name_old: getattr(n, name_new, getattr(n, name_old, 0))
        flags = [bit("nvmlClocksEventReasonHwSlowdown", "nvmlClocksThrottleReasonHwSlowdown"),
                 bit("nvmlClocksEventReasonHwThermalSlowdown", "nvmlClocksThrottleReasonHwThermalSlowdown"),
                 bit("nvmlClocksEventReasonSwThermalSlowdown", "nvmlClocksThrottleReasonSwThermalSlowdown"),
                 bit("nvmlClocksEventReasonSwPowerCap", "nvmlClocksThrottleReasonSwPowerCap")]
        return [str(sm), str(mx), str(pw)] + ["Active" if (f and (r & f)) else "Not Active" for f in flags]

    def _run(self):
        while not self._stop.is_set():
            try:
                if self._nvml is not None:
                    self.samples.append(self._sample_nvml())
                else:
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
                "reasons": sorted(reasons), "samples": len(sm), "source": self.source}
