"""Checkpoint / resume.  The reference saves only datasets and discards the trained model
(/root/reference/src/CFed/Classical_FL.py:157,214); its plan is "checkpoint theta every K rounds"
(/root/reference/ROADMAP.md:90-91).  A checkpoint is one ``torch.save`` dict: round, model
``state_dict`` (the reference's exchange format), flat parameters, RNG state, accountant state,
history and config; writes are atomic (temp file + rename)."""
from __future__ import annotations

import os
from pathlib import Path

import torch


def save_checkpoint(path, *, rnd, model_state, flat_params=None, accountant=None, history=None,
                    config=None, rng=None) -> str:
    path = Path(path)
    path.parent.mkdir(parents=True, exist_ok=True)
    payload = {"round": rnd, "model": model_state, "flat": flat_params, "accountant": accountant,
               "history": history, "config": config, "rng": rng, "format": "qfedx_b200/1"}
    tmp = path.with_suffix(path.suffix + ".tmp")
    torch.save(payload, tmp)
    os.replace(tmp, path)
    return str(path)


def load_checkpoint(path, map_location="cpu") -> dict:
    return torch.load(path, map_location=map_location, weights_only=False)


def latest_checkpoint(directory) -> str | None:
    d = Path(directory)
    if not d.exists():
        return None
    cands = sorted(d.glob("round_*.pt"))
    return str(cands[-1]) if cands else None
