"""Tracing helpers: NVTX ranges around the phases of a round (visible in Nsight timelines) and a
phase timer that records CUDA-event durations per phase for the JSONL log.  The reference has no
tracing; its roadmap asks for wall-clock and GPU-usage metrics
(/root/reference/ROADMAP.md:111-115)."""
from __future__ import annotations

import contextlib
from collections import defaultdict

import torch


@contextlib.contextmanager
def nvtx_range(name: str):
    on = torch.cuda.is_available()
    if on:
        torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        if on:
            torch.cuda.nvtx.range_pop()


class PhaseTimer:
    """``with timer.phase('train'): ...`` records device time per phase (CUDA events; host time on CPU)."""

    def __init__(self):
        self._events = defaultdict(list)
        self._host = defaultdict(float)

    @contextlib.contextmanager
    def phase(self, name: str):
        if torch.cuda.is_available():
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with nvtx_range(name):
                s.record()
                yield
                e.record()
            self._events[name].append((s, e))
        else:
            import time
            t0 = time.perf_counter()
            yield
            self._host[name] += (time.perf_counter() - t0) * 1e3

    def summary_ms(self) -> dict:
        out = dict(self._host)
        if self._events:
            torch.cuda.synchronize()
            for name, pairs in self._events.items():
                out[name] = out.get(name, 0.0) + sum(s.elapsed_time(e) for s, e in pairs)
        self._events.clear()
        self._host.clear()
        return out
