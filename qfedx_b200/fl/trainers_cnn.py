"""Client-batched TinyCNN trainer (K8).  Two engines:

* ``engine="persistent"`` (default): ONE launch per round - a group of CTAs per client stays resident and walks
  epochs x minibatches itself (csrc/tinycnn_persist.cu); conv2 on tcgen05 with split bf16 operands (fp32-accurate).
* ``engine="graph"``: every SGD step of every selected client is a sequence of nine kernels, the whole round
  (epochs x batches) replayed from a single CUDA graph; ``conv2`` picks fp32 CUDA-core or bf16-operand tcgen05 kernels.

Semantics follow the reference's client_update (/root/reference/src/CFed/Classical_FL.py:40-64):
fresh SGD(momentum) state per round, full passes over shuffled minibatches of 32, dropout 0.5
before the last layer, cross-entropy loss; and its evaluate_model (:83-102).
"""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from ..models.tinycnn import TinyCNN, flatten_state, unflatten_state
from ..ops import _lib
from .dataset import FederatedDataset

KERNELS_PER_STEP = 8


class TinyCNNBatchedTrainer:
    def __init__(self, model: TinyCNN, data: FederatedDataset, epochs: int = 5, batch_size: int = 32,
                 lr: float = 0.01, momentum: float = 0.9, p_drop: float = 0.5, shuffle: bool = True, seed: int = 0,
                 use_graph: bool = True, generator=None, conv2: str = "tcgen05", shuffle_mode: str = "array",
                 overlap_steps: bool = True, engine: str = "persistent", ctas_per_client: int = 0):
        if data.device.type != "cuda":
            raise _lib.QfxError("TinyCNNBatchedTrainer runs on CUDA only; use TorchLocalTrainer on CPU")
        if batch_size > 32:
            raise _lib.QfxError("the client-batched TinyCNN kernels support batch_size <= 32")
        if engine not in ("persistent", "graph"):
            raise _lib.QfxError(f"unknown TinyCNN engine {engine!r}")
        self.engine, self.ctas_per_client = engine, int(ctas_per_client)
        self.model, self.data = model, data
        self.epochs, self.batch_size, self.lr, self.momentum, self.p_drop = epochs, batch_size, lr, momentum, p_drop
        self.shuffle, self.seed, self.use_graph, self.generator = shuffle, seed, use_graph, generator
        # 'array': one device sort per round gives explicit permutations (fastest here: a single dependent load in the
        # kernels); 'feistel': in-kernel keyed permutation, no array and nothing outside the CUDA graph
        self.shuffle_mode = shuffle_mode
        self.n_classes = model.num_classes
        check_labels(data.y, self.n_classes, "client data")
        self.conv2_mode = {"fp32": 0, "tcgen05": 1}[conv2]   # tcgen05: bf16 operands, fp32 accumulate
        # pipelined steps: the conv2 update of step k runs beside conv1 forward of step k+1 (one more launch per step)
        self.overlap_steps = bool(overlap_steps)
        self.kernels_per_step = KERNELS_PER_STEP + (1 if self.overlap_steps else 0)
        self._sizes = np.asarray(self.data.sizes)
        self.n_wrap = 0
        self._template = {k: v.detach().clone() for k, v in model.state_dict().items()}
        self.dim = sum(v.numel() for v in self._template.values())
        self.nd = (self.dim + 63) // 64 * 64
        dev = data.device
        self.X = data.X.contiguous().float()
        self.labels = data.y.to(torch.int32).contiguous()
        self.launches = 0
        self._graphs = {}
        self._perm = None
        self._eval_scratch = None
        self._stream = torch.cuda.Stream(device=dev)
        self._round_dev = torch.zeros(1, dtype=torch.int32, device=dev)   # dropout key of the replayed graph

    @property
    def weights(self) -> torch.Tensor:
        return torch.as_tensor(self.data.sizes, dtype=torch.float32)

    def global_flat(self) -> torch.Tensor:
        return flatten_state(self.model.state_dict()).to(self.data.device)

    def load_flat(self, flat: torch.Tensor) -> None:
        self.model.load_state_dict(unflatten_state(flat[: self.dim], self._template))

    def _shuffle_key(self) -> int:
        from .prng import shuffle_key
        return shuffle_key(self.seed, 0) if (self.shuffle and self.shuffle_mode == "feistel") else 0

    def upload(self, host_X: torch.Tensor, host_y: torch.Tensor) -> None:
        """Refresh the resident dataset from (pinned) host memory - the per-round host->device copy
        the reference performs at /root/reference/src/CFed/Classical_FL.py:47-48."""
        self.X.copy_(host_X, non_blocking=True)
        self.labels.copy_(host_y, non_blocking=True)

    # ------------------------------------------------------------------------------------
    def _state_for(self, sel):
        """Per-selection-size state: workspace, parameter rows, (graph engine) the captured round.  Keyed by the number
        of selected clients only - the selection itself lives in a device tensor that is rewritten in place, so changing
        subsets (client sampling, dropouts) reuse the same buffers and the same captured graph."""
        K = len(sel)
        st = self._graphs.get(K)
        if st is None:
            dev = self.data.device
            nbytes, off = ctypes.c_size_t(), ctypes.c_size_t()
            if self.engine == "persistent":
                _lib.call("qfx_tinycnn_round_workspace_bytes", _lib.c_int(K), ctypes.byref(nbytes), ctypes.byref(off), launches=0)
            else:
                _lib.call("qfx_tinycnn_workspace_bytes", _lib.c_int(K), ctypes.byref(nbytes), launches=0)
                _lib.call("qfx_tinycnn_stats_offset", _lib.c_int(K), ctypes.byref(off), launches=0)
            st = {
                "K": K,
                "sel": torch.empty(K, dtype=torch.int32, device=dev),
                "params": torch.empty(K, self.nd, device=dev),
                "moment": torch.empty(K, self.nd, device=dev),
                "ws": torch.empty(nbytes.value, dtype=torch.uint8, device=dev),
                "stats_off": off.value,
                "global": torch.empty(self.dim, device=dev),
                "graph": None,
                # graph engine: the captured round walks as many minibatches as the largest client overall needs
                # (exhausted clients are skipped inside the kernels), so one graph serves every selection of size K
                "n_batches": int(np.ceil(self._sizes.max() / self.batch_size)),
            }
            st["stats"] = st["ws"][off.value: off.value + K * 8].view(torch.float32).view(K, 2)
            self._graphs[K] = st
        st["sel"].copy_(torch.as_tensor(np.asarray(sel, dtype=np.int32)), non_blocking=False)
        if self.engine == "persistent":
            st["n_batches"] = int(np.ceil(self._sizes[np.asarray(sel)].max() / self.batch_size))
        st["steps"] = self.epochs * st["n_batches"]
        return st

    def _enqueue_round(self, st, rnd=0):
        """Everything of one round's local training on the current stream."""
        K = st["K"]
        st["params"][:, : self.dim] = st["global"][None, :]
        st["moment"].zero_()
        if self.engine == "persistent":
            perm_ptr = _lib.ptr(self._perm) if self._perm is not None else _lib.ptr(None)
            _lib.call("qfx_tinycnn_train_round", _lib.ptr(self.X), _lib.ptr(self.labels), _lib.ptr(self.data.offsets),
                      perm_ptr, _lib.c_int(self.X.shape[0]), _lib.ptr(st["sel"]), _lib.c_int(K), _lib.c_int(self.epochs),
                      _lib.c_int(st["n_batches"]), _lib.c_int(self.batch_size), _lib.ptr(st["params"]),
                      _lib.ptr(st["moment"]), _lib.c_int(self.nd), _lib.c_int(self.n_classes), _lib.c_float(self.lr),
                      _lib.c_float(self.momentum), _lib.c_float(self.p_drop), _lib.c_u64(self.seed), _lib.c_int(rnd),
                      _lib.ptr(self._round_dev), _lib.ptr(st["ws"]), _lib.c_u64(self._shuffle_key()),
                      _lib.c_int(self.ctas_per_client), _lib.stream_ptr(), launches=2)
            return
        _lib.call("qfx_tinycnn_begin_round", _lib.ptr(st["params"]), _lib.c_int(self.nd), _lib.c_int(K),
                  _lib.ptr(st["ws"]), _lib.stream_ptr())
        n_batches = st["n_batches"]
        perm_ptr = _lib.ptr(self._perm) if self._perm is not None else _lib.ptr(None)
        for e in range(self.epochs):
            st["stats"].zero_()
            for b in range(n_batches):
                _lib.call("qfx_tinycnn_train_step_overlapped" if self.overlap_steps else "qfx_tinycnn_train_step",
                          _lib.ptr(self.X), _lib.ptr(self.labels),
                          _lib.ptr(self.data.offsets), perm_ptr, _lib.c_int(self.X.shape[0]), _lib.ptr(st["sel"]),
                          _lib.c_int(K), _lib.c_int(e), _lib.c_int(b), _lib.c_int(self.batch_size),
                          _lib.ptr(st["params"]), _lib.ptr(st["moment"]), _lib.c_int(self.nd),
                          _lib.c_int(self.n_classes), _lib.c_float(self.lr), _lib.c_float(self.momentum),
                          _lib.c_float(self.p_drop), _lib.c_u64(self.seed), _lib.c_int(rnd), _lib.ptr(self._round_dev),
                          _lib.ptr(st["ws"]),
                          _lib.c_int(self.conv2_mode), _lib.c_u64(self._shuffle_key()), _lib.stream_ptr(),
                          launches=self.kernels_per_step)
        if self.overlap_steps:   # join the side stream's last conv2 update (also required before a capture ends)
            _lib.call("qfx_tinycnn_finish_steps", _lib.stream_ptr(), launches=0)

    def train(self, global_flat: torch.Tensor, selected, rnd: int):
        sel = [int(k) for k in selected]
        st = self._state_for(sel)
        if self.shuffle and self.shuffle_mode == "array":
            new_perm = self.data.epoch_permutations(self.epochs, self.generator)
            if self._perm is None:
                self._perm = new_perm
            else:
                self._perm.copy_(new_perm)  # same storage: the captured graph keeps pointing at it
        st["global"].copy_(global_flat[: self.dim])
        self._round_dev.fill_(int(rnd))
        if self.engine == "persistent" or not self.use_graph:
            self._enqueue_round(st, rnd=rnd)
        else:
            if st["graph"] is None:
                # dropout masks are keyed by (seed, round, client, epoch, batch); the kernels read the round
                # from device memory (self._round_dev), so one captured graph serves every round
                torch.cuda.synchronize()
                before = _lib.LAUNCH_COUNT
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=self._stream):
                    self._enqueue_round(st, rnd=0)
                st["graph"] = g
                st["graph_launches"] = _lib.LAUNCH_COUNT - before
                _lib.LAUNCH_COUNT = before
            st["graph"].replay()
            _lib.LAUNCH_COUNT += st["graph_launches"]
        self.launches += 2 if self.engine == "persistent" else st["steps"] * self.kernels_per_step
        out = st["params"][:, : self.dim]
        return out, {"train_stats": st["stats"]}

    def check_round(self, K: int) -> None:
        """Persistent engine: raise if a group barrier of the last round on the K-client workspace gave up waiting
        (synchronises the stream; for tests and debugging - a healthy run never needs it)."""
        st = self._graphs.get(K)
        if st is None or self.engine != "persistent":
            return
        flag = ctypes.c_int(0)
        _lib.call("qfx_tinycnn_round_status", _lib.c_int(K), _lib.ptr(st["ws"]), _lib.stream_ptr(), ctypes.byref(flag), launches=0)
        if flag.value:
            raise _lib.QfxError("tinycnn_round_kernel: a group barrier timed out")

    def evaluate(self, global_flat: torch.Tensor, X: torch.Tensor, y: torch.Tensor, batch_size: int = 0):
        n = X.shape[0]
        check_labels(y, self.n_classes, "evaluation data")
        if self.engine == "persistent":
            need = n * 7840 + 102400
            if self._eval_scratch is None or self._eval_scratch.numel() < need:
                self._eval_scratch = torch.empty(need, dtype=torch.uint8, device=X.device)
            out2 = torch.empty(2, device=X.device)
            prm = torch.zeros(self.nd, device=X.device)
            prm[: self.dim] = global_flat[: self.dim]
            _lib.call("qfx_tinycnn_eval_tc", _lib.ptr(X.contiguous().float()), _lib.ptr(y.to(torch.int32).contiguous()),
                      _lib.c_int(n), _lib.ptr(prm), _lib.c_int(self.n_classes), _lib.ptr(self._eval_scratch),
                      _lib.ptr(out2), _lib.stream_ptr(), launches=3)
            res = out2.cpu()
            return float(res[0]) / n, float(res[1]) / n
        need = n * 23520 + 102400
        if self._eval_scratch is None or self._eval_scratch.numel() < need:
            self._eval_scratch = torch.empty(need, dtype=torch.uint8, device=X.device)
        out2 = torch.empty(2, device=X.device)
        prm = torch.zeros(self.nd, device=X.device)
        prm[: self.dim] = global_flat[: self.dim]
        _lib.call("qfx_tinycnn_eval", _lib.ptr(X.contiguous().float()), _lib.ptr(y.to(torch.int32).contiguous()),
                  _lib.c_int(n), _lib.ptr(prm), _lib.c_int(self.n_classes), _lib.ptr(self._eval_scratch),
                  _lib.ptr(out2), _lib.stream_ptr(), launches=4)
        res = out2.cpu()
        return float(res[0]) / n, float(res[1]) / n


def check_labels(y: torch.Tensor, n_classes: int, what: str) -> None:
    """The kernels index logits by label: reject anything outside [0, n_classes) up front (torch's CrossEntropyLoss
    raises for the same input; digits like (3, 5, 8) must be remapped to 0..C-1 first)."""
    if y.numel() == 0:
        return
    lo, hi = int(y.min()), int(y.max())
    if lo < 0 or hi >= n_classes:
        raise _lib.QfxError(f"{what}: labels must lie in [0, {n_classes}), found [{lo}, {hi}]")


def smoke_tinycnn(dev) -> None:
    from ..data.synthetic import synthetic_mnist
    X, y = synthetic_mnist(96, 3, seed=0)
    data = FederatedDataset.from_clients([(X[:48], y[:48]), (X[48:], y[48:])], dev)
    for engine in ("persistent", "graph"):
        tr = TinyCNNBatchedTrainer(TinyCNN(3).to(dev), data, epochs=1, batch_size=32, use_graph=False, engine=engine)
        out, _ = tr.train(tr.global_flat(), [0, 1], 0)
        tr.check_round(2)
        loss, acc = tr.evaluate(out[0], X.to(dev), y.to(dev))
        assert np.isfinite(loss) and torch.isfinite(out).all()
