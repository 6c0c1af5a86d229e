"""TinyCNN: the reference's classical model (/root/reference/src/CFed/Classical_FL.py:21-38).

conv5x5(1->16, pad 2) - ReLU - maxpool2 - conv5x5(16->32, pad 2) - ReLU - maxpool2 - flatten(1568)
- fc 64 - ReLU - dropout 0.5 - fc num_classes; 113,859 parameters for 3 classes.  The state_dict
keys and shapes match the reference so checkpoints interchange.  The eager ``forward`` is the
numerics reference; federated training uses the client-batched CUDA path in
``qfedx_b200.ops.tinycnn`` through ``qfedx_b200.fl.trainers.TinyCNNBatchedTrainer``.
"""
from __future__ import annotations

from collections import OrderedDict

import torch
import torch.nn as nn
import torch.nn.functional as F

PARAM_SHAPES = OrderedDict([
    ("conv1.weight", (16, 1, 5, 5)), ("conv1.bias", (16,)),
    ("conv2.weight", (32, 16, 5, 5)), ("conv2.bias", (32,)),
    ("fc1.weight", (64, 1568)), ("fc1.bias", (64,)),
    ("fc2.weight", None), ("fc2.bias", None),
])


def param_layout(num_classes: int = 3):
    """[(name, shape, offset)] of the flat parameter vector, in state_dict order."""
    out, off = [], 0
    for name, shape in PARAM_SHAPES.items():
        if shape is None:
            shape = (num_classes, 64) if name.endswith("weight") else (num_classes,)
        n = 1
        for s in shape:
            n *= s
        out.append((name, shape, off))
        off += n
    return out, off


class TinyCNN(nn.Module):
    def __init__(self, num_classes: int = 3, p_drop: float = 0.5):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 16, 5, padding=2)
        self.conv2 = nn.Conv2d(16, 32, 5, padding=2)
        self.fc1 = nn.Linear(32 * 7 * 7, 64)
        self.fc2 = nn.Linear(64, num_classes)
        self.dropout = nn.Dropout(p_drop)
        self.num_classes = num_classes

    def forward(self, x):
        h = F.max_pool2d(F.relu(self.conv1(x)), 2)
        h = F.max_pool2d(F.relu(self.conv2(h)), 2)
        h = F.relu(self.fc1(h.flatten(1)))
        return self.fc2(self.dropout(h))


def flatten_state(sd) -> torch.Tensor:
    return torch.cat([sd[k].reshape(-1).float() for k in sd.keys()])


def unflatten_state(flat: torch.Tensor, like_sd) -> "OrderedDict[str, torch.Tensor]":
    out, off = OrderedDict(), 0
    for k, v in like_sd.items():
        n = v.numel()
        out[k] = flat[off:off + n].reshape(v.shape).to(v.dtype)
        off += n
    return out
