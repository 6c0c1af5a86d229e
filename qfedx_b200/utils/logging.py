"""Per-round metrics: JSONL file + optional TensorBoard + console.

The reference only prints (/root/reference/src/CFed/Classical_FL.py:119-125,151-155) and plans
MLflow (/root/reference/ROADMAP.md:92-93); the metrics named in ROADMAP.md:111-115 (accuracy,
epsilon, wall-clock, communication MB per round) are the columns written here.
"""
from __future__ import annotations

import json
import time
from pathlib import Path


class RoundLogger:
    def __init__(self, path=None, tensorboard_dir=None, print_every: int = 5, quiet: bool = False):
        self.path = Path(path) if path else None
        if self.path:
            self.path.parent.mkdir(parents=True, exist_ok=True)
            self._fh = open(self.path, "a")
        else:
            self._fh = None
        self.print_every, self.quiet = print_every, quiet
        self.rows = []
        self._tb = None
        if tensorboard_dir:
            try:
                from torch.utils.tensorboard import SummaryWriter
                self._tb = SummaryWriter(tensorboard_dir)
            except Exception:  # tensorboard is optional
                self._tb = None

    def log(self, rnd: int, **metrics) -> None:
        row = {"round": rnd, "time": time.time(), **metrics}
        self.rows.append(row)
        if self._fh:
            self._fh.write(json.dumps(row) + "\n")
            self._fh.flush()
        if self._tb:
            for k, v in metrics.items():
                if isinstance(v, (int, float)):
                    self._tb.add_scalar(k, v, rnd)
        if not self.quiet and (rnd % self.print_every == 0):
            shown = ", ".join(f"{k}={v:.4f}" if isinstance(v, float) else f"{k}={v}" for k, v in metrics.items()
                              if isinstance(v, (int, float)))
            print(f"round {rnd}: {shown}")

    def close(self):
        if self._fh:
            self._fh.close()
        if self._tb:
            self._tb.close()
