"""VQC statevector ops with autograd: forward <Z> readout, adjoint / parameter-shift backward.

CUDA tensors run on the register-tiled sm_100a kernels (csrc/sv_small.cu: K1, K3, K4a, K4b of
SURVEY.md 2.5); CPU tensors run on the torch reference (ops/ref.py).  There is no silent
fallback on a GPU: a CUDA tensor with a missing library raises.
"""
from __future__ import annotations

import torch

from . import _lib, ref

BASIS = {"rx": 0, "ry": 1, "rz": 2}
ENTANGLER = {"cnot": 0, "cz": 1, "none": 2}
MAX_REG_QUBITS = 10


def _prep(angles, init_state, theta, gid):
    theta = theta.contiguous().float()
    if theta.dim() == 3:
        theta = theta[None]
    angles = None if angles is None else angles.contiguous().float()
    init_state = None if init_state is None else init_state.contiguous().float()
    gid = None if gid is None else gid.contiguous().to(torch.int32)
    return angles, init_state, theta, gid


def sv_forward(angles, theta, gid=None, basis="ry", entangler="cnot", init_state=None, want_state=False):
    """<Z_q> [B,n] (and optionally the final state [B,2^n] complex64) - no autograd."""
    angles, init_state, theta, gid = _prep(angles, init_state, theta, gid)
    G, L, n, _ = theta.shape
    src = angles if angles is not None else init_state
    B = src.shape[0]
    if not _lib.use_cuda(src, theta):
        psi = ref.vqc_state_ref(angles, theta, None if gid is None else gid.long(), basis, entangler,
                                None if init_state is None else init_state.to(torch.complex64))
        E = ref.expval_z_all(psi)
        return (E, psi) if want_state else E
    E = torch.empty(B, n, device=src.device, dtype=torch.float32)
    state = torch.empty(B, 1 << n, 2, device=src.device, dtype=torch.float32) if want_state else None
    _lib.call("qfx_sv_forward", _lib.c_int(n), _lib.c_int(L), _lib.c_int(BASIS[basis]),
              _lib.c_int(ENTANGLER[entangler]), _lib.ptr(angles), _lib.ptr(init_state), _lib.ptr(theta),
              _lib.ptr(gid), _lib.c_int(B), _lib.ptr(E), _lib.ptr(state), _lib.stream_ptr())
    return (E, torch.view_as_complex(state)) if want_state else E


def sv_backward(angles, theta, cot, gid=None, basis="ry", entangler="cnot", init_state=None,
                method="adjoint", want_angle_grad=False):
    """Gradient of sum_b sum_q cot[b,q] <Z_q> w.r.t. theta [G,L,n,2] (and angles [B,n])."""
    angles, init_state, theta, gid = _prep(angles, init_state, theta, gid)
    cot = cot.contiguous().float()
    G, L, n, _ = theta.shape
    src = angles if angles is not None else init_state
    B = src.shape[0]
    if not _lib.use_cuda(src, theta):
        th = theta.detach().clone().requires_grad_(True)
        ang = None if angles is None else angles.detach().clone().requires_grad_(want_angle_grad)
        E = ref.vqc_expvals_ref(ang, th, None if gid is None else gid.long(), basis, entangler,
                                None if init_state is None else init_state.to(torch.complex64))
        wanted = [th] + ([ang] if want_angle_grad and ang is not None else [])
        grads = torch.autograd.grad((E * cot).sum(), wanted)
        return grads[0], (grads[1] if len(grads) > 1 else None)
    gth = torch.zeros_like(theta)
    if method == "adjoint":
        gang = torch.empty(B, n, device=src.device, dtype=torch.float32) if (want_angle_grad and angles is not None) else None
        _lib.call("qfx_sv_adjoint_backward", _lib.c_int(n), _lib.c_int(L), _lib.c_int(BASIS[basis]),
                  _lib.c_int(ENTANGLER[entangler]), _lib.ptr(angles), _lib.ptr(init_state), _lib.ptr(theta),
                  _lib.ptr(gid), _lib.c_int(B), _lib.ptr(cot), _lib.ptr(gth), _lib.ptr(gang), _lib.stream_ptr())
        return gth, gang
    if method == "param_shift":
        _lib.call("qfx_sv_param_shift", _lib.c_int(n), _lib.c_int(L), _lib.c_int(BASIS[basis]),
                  _lib.c_int(ENTANGLER[entangler]), _lib.ptr(angles), _lib.ptr(init_state), _lib.ptr(theta),
                  _lib.ptr(gid), _lib.c_int(B), _lib.ptr(cot), _lib.ptr(gth), _lib.stream_ptr())
        return gth, None
    raise ValueError(f"unknown gradient method {method!r}")


class _VQCExpvals(torch.autograd.Function):
    @staticmethod
    def forward(ctx, angles, theta, gid, basis, entangler, init_state, method):
        ctx.save_for_backward(angles, theta, gid, init_state)
        ctx.cfg = (basis, entangler, method)
        return sv_forward(angles, theta, gid, basis, entangler, init_state)

    @staticmethod
    def backward(ctx, cot):
        angles, theta, gid, init_state = ctx.saved_tensors
        basis, entangler, method = ctx.cfg
        want_ang = angles is not None and ctx.needs_input_grad[0] and method == "adjoint"
        gth, gang = sv_backward(angles, theta, cot, gid, basis, entangler, init_state, method, want_ang)
        return gang, gth.reshape(theta.shape), None, None, None, None, None


def vqc_expvals(angles, theta, gid=None, basis="ry", entangler="cnot", init_state=None, method="adjoint"):
    """Differentiable <Z_q> [B,n].  ``method`` picks the backward kernel: 'adjoint' (K4a) or
    'param_shift' (K4b).  On CPU tensors torch autograd differentiates the reference."""
    src = angles if angles is not None else init_state
    if not _lib.use_cuda(src, theta):
        th = theta if theta.dim() == 4 else theta[None]
        return ref.vqc_expvals_ref(angles, th, None if gid is None else gid.long(), basis, entangler,
                                   None if init_state is None else init_state.to(torch.complex64))
    n = theta.shape[-2]
    if n > MAX_REG_QUBITS:
        raise _lib.QfxError(f"{n} qubits: use qfedx_b200.ops.large (register kernels stop at {MAX_REG_QUBITS})")
    return _VQCExpvals.apply(angles, theta, gid, basis, entangler, init_state, method)
