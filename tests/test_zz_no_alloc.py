"""The allocation-free aggregation entry point, called the way its intended users call it: straight
through the C ABI with raw device pointers (no workspace argument, no allocator behind it).

Those users are programs that cannot hand the library a workspace - a federated server built on the C
ABI without a tensor library, a process whose GPU memory belongs to another framework's arena, a
serving loop that must not trigger a device-synchronising cudaMalloc between client launches.  The
test allocates its inputs and outputs with torch only because a test has to put data somewhere; the
call under test receives nothing but pointers to them.

It runs last in the suite (file name) and must hold both in a fresh process and after every other GPU
test has used the library: the static workspace keeps old contents between calls, so the second call
below deliberately uses fewer clients and different data than the first.
"""
import ctypes

import pytest
import torch

from qfedx_b200.ops import _lib

pytestmark = pytest.mark.gpu
TOL = 2e-5          # the tolerance tests/test_gpu_train_agg.py uses for the aggregation kernel
MAX_CLIENTS = 4096  # QFX_FEDAVG_STATIC_MAX_CLIENTS in include/qfedx_b200.h


def _args(cp, glob, out, w, part, K, D, clip, seed, rnd, secagg, norms):
    return [_lib.ptr(cp), _lib.ptr(glob), _lib.ptr(out), _lib.ptr(w), _lib.ptr(part), _lib.c_int(K), _lib.c_int(D),
            _lib.c_int(1), _lib.c_int(D // 2), _lib.c_float(clip), _lib.c_float(0.3), _lib.c_float(0.9),
            _lib.c_u64(seed), _lib.c_int(rnd), _lib.c_int(1 if secagg else 0), _lib.ptr(norms), _lib.ptr(None)]


def _both_entry_points(K, D, clip, seed, rnd, secagg):
    g = torch.Generator().manual_seed(seed * 7919 + K)
    glob = torch.randn(D, generator=g).cuda()
    cp = (glob.cpu() + 0.05 * torch.randn(K, D, generator=g) * (1 + torch.arange(K)[:, None])).cuda().contiguous()
    w = torch.randint(1, 400, (K,), generator=g).float().cuda()
    part = (torch.rand(K, generator=g) > 0.2).int()
    part[0] = 1
    part = part.cuda()
    lib = _lib.lib()
    stream = _lib.stream_ptr()

    out_ref, norms_ref = torch.empty(D, device="cuda"), torch.empty(K, device="cuda")
    scratch = torch.empty(K, device="cuda")
    lib.qfx_fedavg_fused.restype = ctypes.c_int
    rc = lib.qfx_fedavg_fused(*_args(cp, glob, out_ref, w, part, K, D, clip, seed, rnd, secagg, norms_ref),
                              _lib.ptr(scratch), stream)
    assert rc == 0, lib.qfx_last_error().decode()

    out_new, norms_new = torch.empty(D, device="cuda"), torch.empty(K, device="cuda")
    lib.qfx_fedavg_fused_static.restype = ctypes.c_int
    rc = lib.qfx_fedavg_fused_static(*_args(cp, glob, out_new, w, part, K, D, clip, seed, rnd, secagg, norms_new),
                                     stream)
    assert rc == 0, lib.qfx_last_error().decode()
    torch.cuda.synchronize()
    return out_ref, out_new, norms_ref, norms_new


def test_static_workspace_entry_point_matches_existing_one_twice_in_a_row():
    # first call: many clients, clipping active, DP noise, masks
    a_ref, a_new, n_ref, n_new = _both_entry_points(K=37, D=5000, clip=0.4, seed=11, rnd=3, secagg=True)
    assert torch.allclose(a_new, a_ref, atol=TOL, rtol=0), float((a_new - a_ref).abs().max())
    assert torch.allclose(n_new, n_ref, rtol=1e-5)
    # second call right after it: fewer clients, other data - stale accumulators from the first call
    # (entries 0..36 of the shared array) must not leak into the clip factors
    b_ref, b_new, m_ref, m_new = _both_entry_points(K=5, D=113859, clip=0.7, seed=12, rnd=4, secagg=False)
    assert torch.allclose(b_new, b_ref, atol=TOL, rtol=0), float((b_new - b_ref).abs().max())
    assert torch.allclose(m_new, m_ref, rtol=1e-5)
    assert float((b_new - b_ref).abs().max()) < TOL and torch.isfinite(b_new).all()


def test_more_clients_than_the_static_workspace_holds_is_refused():
    lib = _lib.lib()
    lib.qfx_fedavg_fused_static.restype = ctypes.c_int
    K, D = MAX_CLIENTS + 1, 8
    cp = torch.zeros(K, D, device="cuda")
    glob = torch.zeros(D, device="cuda")
    out = torch.full((D,), 7.0, device="cuda")
    w = torch.ones(K, device="cuda")
    rc = lib.qfx_fedavg_fused_static(*_args(cp, glob, out, w, None, K, D, 0.0, 1, 0, False, None), _lib.stream_ptr())
    assert rc == -3                                              # QFX_ERR_TOO_LARGE
    msg = lib.qfx_last_error().decode()
    assert str(MAX_CLIENTS) in msg and "at most" in msg
    torch.cuda.synchronize()
    assert torch.all(out == 7.0)                                 # nothing was launched
