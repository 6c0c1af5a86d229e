"""tcgen05 implicit-GEMM conv2 kernels (single-bf16-operand variants of the graph engine) vs float64 torch ON THE CPU on
the same bf16-rounded operands - not vs cuDNN, whose fp32 convolutions may run in TF32."""
import ctypes

import pytest
import torch
import torch.nn.functional as F

from qfedx_b200.ops import _lib

pytestmark = pytest.mark.gpu


def _bf16(x):
    return x.to(torch.bfloat16).float()


def test_conv2_forward_umma_matches_torch():
    g = torch.Generator().manual_seed(0)
    n = 5
    a1 = torch.relu(torch.randn(n, 16, 14, 14, generator=g)).cuda()
    w = (0.1 * torch.randn(32, 16, 5, 5, generator=g)).cuda()
    b = (0.1 * torch.randn(32, generator=g)).cuda()
    a2 = torch.empty(n, 32, 7, 7, device="cuda")
    i2 = torch.empty(n, 32, 7, 7, device="cuda", dtype=torch.uint8)
    scratch = torch.empty(51200, device="cuda", dtype=torch.uint8)
    _lib.call("qfx_tinycnn_conv2_fwd_umma", _lib.ptr(a1), _lib.ptr(w), _lib.ptr(b), _lib.c_int(n), _lib.ptr(a2),
              _lib.ptr(i2), _lib.ptr(scratch), _lib.stream_ptr())
    torch.cuda.synchronize()
    conv = F.conv2d(_bf16(a1).cpu().double(), _bf16(w).cpu().double(), None, padding=2)
    ref, idx = F.max_pool2d(conv, 2, return_indices=True)
    ref = torch.relu(ref + b.cpu().double()[None, :, None, None])
    a2, i2 = a2.cpu().double(), i2.cpu()
    err = (a2 - ref).abs().max()
    assert float(err) < 2e-3, float(err)
    # argmax within the 2x2 window (0..3 = dy*2+dx); near-ties may legitimately differ
    yy, xx = idx // 14, idx % 14
    ref_arg = ((yy % 2) * 2 + (xx % 2)).to(torch.uint8)
    agree = (ref_arg == i2).float().mean()
    assert float(agree) > 0.995, float(agree)


def test_conv2_backward_umma_matches_torch():
    g = torch.Generator().manual_seed(1)
    n = 4
    a1 = torch.relu(torch.randn(n, 16, 14, 14, generator=g)).cuda()
    w = (0.1 * torch.randn(32, 16, 5, 5, generator=g)).cuda()
    # sparse-ish gradient like the one max-pool backward produces
    g2 = (torch.randn(n, 32, 14, 14, generator=g) * (torch.rand(n, 32, 14, 14, generator=g) < 0.25)).cuda()
    da1 = torch.empty(n, 16, 14, 14, device="cuda")
    part = torch.zeros(n, 13248, device="cuda")
    scratch = torch.empty(51200, device="cuda", dtype=torch.uint8)
    _lib.call("qfx_tinycnn_conv2_dgrad_umma", _lib.ptr(g2), _lib.ptr(w), _lib.c_int(n), _lib.ptr(da1), _lib.ptr(scratch),
              _lib.stream_ptr())
    _lib.call("qfx_tinycnn_conv2_wgrad_umma", _lib.ptr(g2), _lib.ptr(a1), _lib.c_int(n), _lib.ptr(part), _lib.stream_ptr())
    torch.cuda.synchronize()
    a1b = _bf16(a1).cpu().double().requires_grad_(True)        # float64 on the CPU: the reference has no TF32 in it
    wb = _bf16(w).cpu().double().requires_grad_(True)
    g2b = _bf16(g2).cpu().double()
    da1, part, g2c = da1.cpu().double(), part.cpu().double(), g2.cpu().double()
    out = F.conv2d(a1b, wb, None, padding=2)
    # data gradient uses bf16(g2) x bf16(w); weight gradient bf16(g2) x bf16(a1)
    (da_ref,) = torch.autograd.grad(out, a1b, g2b, retain_graph=True)
    assert float((da1 - da_ref).abs().max()) < 5e-3 * float(da_ref.abs().max())
    for i in range(n):
        oi = F.conv2d(a1b[i:i + 1], wb, None, padding=2)
        (dw_ref,) = torch.autograd.grad(oi, wb, g2b[i:i + 1])
        got = part[i, 416:416 + 12800].reshape(32, 16, 5, 5)
        assert float((got - dw_ref).abs().max()) < 5e-3 * float(dw_ref.abs().max()), i
        assert torch.allclose(part[i, 13216:13248], g2c[i].sum(dim=(1, 2)), atol=1e-4)
