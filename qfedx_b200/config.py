"""Experiment configuration: dataclasses <-> YAML, dotted overrides, one preset per BASELINE.json config.

The reference keeps its settings in dict literals edited by hand
(/root/reference/src/CFed/Classical_FL.py:161-173) and plans Hydra (/root/reference/ROADMAP.md:16,70);
Hydra is not installed, so this is a small equivalent: ``load_config(path_or_preset, overrides)``.
"""
from __future__ import annotations

import copy
from dataclasses import asdict, dataclass, field, fields, is_dataclass
from pathlib import Path
from typing import Any, Dict, Sequence, Tuple

import yaml

from .fl.server import FLConfig
from .noise import NoiseConfig


@dataclass
class DataConfig:
    dataset: str = "mnist"           # mnist (idx files or the synthetic fallback) | iris (4 features -> 4 qubits)
    raw_folder: str = "./dataset/raw"
    processed_folder: str = "./dataset/processed"
    digits: Tuple[int, ...] = (0, 1, 2)
    val_split: float = 0.1
    num_clients: int = 4
    partition_type: str = "iid"      # iid | non_iid
    alpha: float = 0.5
    synthetic_fallback: bool = True
    synthetic_scale: float = 1.0
    remap_labels: bool = True
    reducer: str = "block"           # block (7x7 means, as the reference's demo) | pca


@dataclass
class ModelConfig:
    family: str = "vqc"              # vqc | tinycnn
    n_qubits: int = 4
    n_layers: int = 2
    entangler: str = "cnot"
    grad_method: str = "adjoint"     # adjoint | param_shift
    engine: str = "auto"             # auto | registers | passes | sharded | torch | graph (tinycnn: step-per-node engine)


@dataclass
class TrainConfig:
    local_epochs: int = 5
    batch_size: int = 32
    lr: float = 0.05
    optimizer: str = "adam"          # adam | sgd | spsa (torch engine)
    momentum: float = 0.9
    centralized: bool = False        # pool every client's data into one: the centralized baseline (ROADMAP.md:4)


@dataclass
class ExperimentConfig:
    name: str = "vqc4"
    seed: int = 42
    device: str = "auto"
    data: DataConfig = field(default_factory=DataConfig)
    model: ModelConfig = field(default_factory=ModelConfig)
    train: TrainConfig = field(default_factory=TrainConfig)
    fl: FLConfig = field(default_factory=FLConfig)
    noise: NoiseConfig = field(default_factory=NoiseConfig)


PRESET_DIR = Path(__file__).resolve().parent.parent / "configs"


def _from_dict(cls, d: Dict[str, Any]):
    kwargs = {}
    for f in fields(cls):
        if f.name not in d:
            continue
        v = d[f.name]
        default = getattr(cls(), f.name) if not is_dataclass(f.type) else None
        if is_dataclass(default) or (isinstance(v, dict) and is_dataclass(getattr(cls(), f.name))):
            kwargs[f.name] = _from_dict(type(getattr(cls(), f.name)), v)
        elif isinstance(getattr(cls(), f.name), tuple) and isinstance(v, list):
            kwargs[f.name] = tuple(v)
        else:
            kwargs[f.name] = v
    unknown = set(d) - {f.name for f in fields(cls)}
    if unknown:
        raise KeyError(f"unknown config keys for {cls.__name__}: {sorted(unknown)}")
    return cls(**kwargs)


def apply_overrides(d: Dict[str, Any], overrides: Sequence[str]) -> Dict[str, Any]:
    d = copy.deepcopy(d)
    for item in overrides or ():
        key, _, raw = item.partition("=")
        node = d
        parts = key.split(".")
        for p in parts[:-1]:
            node = node.setdefault(p, {})
        node[parts[-1]] = yaml.safe_load(raw)
    return d


def load_config(source: str | None = None, overrides: Sequence[str] = ()) -> ExperimentConfig:
    """``source``: a YAML path, a preset name (configs/<name>.yaml) or None for defaults."""
    d: Dict[str, Any] = {}
    if source:
        path = Path(source)
        if not path.exists():
            path = PRESET_DIR / f"{source}.yaml"
        if not path.exists():
            raise FileNotFoundError(f"no config file or preset named {source!r}")
        d = yaml.safe_load(path.read_text()) or {}
    return _from_dict(ExperimentConfig, apply_overrides(d, overrides))


def dump_config(cfg: ExperimentConfig) -> str:
    def clean(x):
        if isinstance(x, dict):
            return {k: clean(v) for k, v in x.items()}
        if isinstance(x, tuple):
            return list(x)
        return x
    return yaml.safe_dump(clean(asdict(cfg)), sort_keys=False)
