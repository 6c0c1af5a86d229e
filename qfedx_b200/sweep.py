"""Experiment grid with multi-seed aggregation (/root/reference/ROADMAP.md:98-121: qubits, depth, client fraction,
noise multiplier sigma, clip C, Dirichlet alpha ... "5 seeds; report mean +- std", accuracy-vs-epsilon tables, and the
three arms of ROADMAP.md:4: centralized, federated without privacy, federated with DP + secure aggregation).

``run_sweep(preset, overrides, grid, seeds)``: ``grid`` is a list of "dotted.key=v1,v2,..." strings; every point of
the Cartesian product is run for seeds base_seed .. base_seed + seeds - 1 and summarised as one row."""
from __future__ import annotations

import itertools
import math
from typing import Dict, List, Sequence

import yaml

from .config import load_config

METRICS = ("final_accuracy", "final_loss", "final_auc", "epsilon", "comm_mb_total", "wall_clock_s", "gpu_hours")


def parse_grid(grid: Sequence[str]) -> Dict[str, list]:
    out = {}
    for item in grid or ():
        key, _, raw = item.partition("=")
        out[key.strip()] = [yaml.safe_load(v) for v in raw.split(",")]
    return out


def _mean_std(vals):
    vals = [float(v) for v in vals if v is not None and v == v]
    if not vals:
        return None, None
    m = sum(vals) / len(vals)
    sd = math.sqrt(sum((v - m) ** 2 for v in vals) / (len(vals) - 1)) if len(vals) > 1 else 0.0
    return m, sd


def run_point(preset, overrides, point: Dict[str, object], seeds: int, arm: str = "configured") -> dict:
    from .cli import run_experiment
    results = []
    base = load_config(preset, list(overrides))
    for s in range(seeds):
        ov = list(overrides) + [f"{k}={yaml.safe_dump(v, default_flow_style=True).strip().splitlines()[0]}" for k, v in point.items()]
        ov += [f"seed={base.seed + s}", f"fl.seed={base.fl.seed + s}", "fl.quiet=true"]
        if arm == "centralized":
            ov += ["train.centralized=true", "fl.clip=0.0", "fl.sigma=0.0", "fl.secagg=false", "fl.client_fraction=1.0"]
        elif arm == "no-privacy":
            ov += ["fl.clip=0.0", "fl.sigma=0.0", "fl.secagg=false"]
        results.append(run_experiment(load_config(preset, ov)))
    row = {"arm": arm, **{k: v for k, v in point.items()}, "seeds": seeds}
    for m in METRICS:
        mean, sd = _mean_std([r.get(m) for r in results])
        row[m] = mean
        row[m + "_std"] = sd
    return row


def run_sweep(preset=None, overrides: Sequence[str] = (), grid: Sequence[str] = (), seeds: int = 3,
              baselines: bool = False) -> List[dict]:
    g = parse_grid(grid)
    keys = list(g)
    rows = []
    for combo in itertools.product(*[g[k] for k in keys]) if keys else [()]:
        point = dict(zip(keys, combo))
        if baselines:
            rows.append(run_point(preset, overrides, point, seeds, "centralized"))
            rows.append(run_point(preset, overrides, point, seeds, "no-privacy"))
        rows.append(run_point(preset, overrides, point, seeds, "configured"))
    return rows


def format_table(rows: List[dict]) -> str:
    """Markdown table: one row per grid point and arm, metrics as mean +- std over the seeds."""
    if not rows:
        return "(no rows)"
    keys = [k for k in rows[0] if k not in ("arm", "seeds") and not k.endswith("_std") and k not in METRICS]
    shown = [m for m in METRICS if any(r.get(m) is not None for r in rows)]
    head = ["arm"] + keys + shown
    lines = ["| " + " | ".join(head) + " |", "|" + "---|" * len(head)]
    for r in rows:
        cells = [str(r["arm"])] + [str(r.get(k)) for k in keys]
        for m in shown:
            cells.append("-" if r.get(m) is None else f"{r[m]:.4g} +- {r[m + '_std']:.2g}")
        lines.append("| " + " | ".join(cells) + " |")
    return "\n".join(lines)
