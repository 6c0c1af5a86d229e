"""sm_100a statevector kernels vs the float64 oracle / the torch fp32 reference of the same op."""
import math

import numpy as np
import pytest
import torch

from qfedx_b200.ops import _lib, ref
from qfedx_b200.ops import statevector as svop
from qfedx_b200.oracle import statevector as osv
from qfedx_b200.oracle import vqc as ovqc

pytestmark = pytest.mark.gpu
TOL = 2e-5


def _rand(n, L, B, G=1, seed=0):
    g = torch.Generator().manual_seed(seed)
    ang = (torch.rand(B, n, generator=g) * math.pi).cuda()
    theta = torch.randn(G, L, n, 2, generator=g).cuda()
    cot = torch.randn(B, n, generator=g).cuda()
    gid = torch.randint(0, G, (B,), generator=g).cuda() if G > 1 else None
    return ang, theta, cot, gid


def test_library_is_loaded():
    assert _lib.available(), "libqfedx_b200.so missing on a GPU box"
    assert _lib.lib().qfx_version() >= 100


@pytest.mark.parametrize("n", list(range(1, 11)))
@pytest.mark.parametrize("ent", ["cnot", "cz"])
def test_forward_matches_oracle(n, ent):
    L, B = 3, 37
    ang, theta, _, _ = _rand(n, L, B, seed=n)
    E, psi = svop.sv_forward(ang, theta, None, "ry", ent, want_state=True)
    Eo = np.stack([osv.vqc_expvals(ang[b].double().cpu().numpy(), theta[0].double().cpu().numpy(), "ry", ent)
                   for b in range(B)])
    np.testing.assert_allclose(E.cpu().numpy(), Eo, atol=TOL)
    psio = osv.run_ansatz(osv.product_state_from_angles(ang[5].double().cpu().numpy()),
                          theta[0].double().cpu().numpy(), ent)
    np.testing.assert_allclose(psi[5].cpu().numpy(), psio, atol=TOL)


@pytest.mark.parametrize("basis", ["rx", "ry", "rz"])
def test_feature_map_bases(basis):
    n, L, B = 6, 2, 19
    ang, theta, _, _ = _rand(n, L, B, seed=3)
    E = svop.sv_forward(ang, theta, None, basis, "cnot")
    Er = ref.vqc_expvals_ref(ang.cpu().double(), theta.cpu().double(), None, basis, "cnot")
    np.testing.assert_allclose(E.cpu().numpy(), Er.numpy(), atol=TOL)


def test_amplitude_encoded_input():
    n, L, B = 5, 2, 11
    g = torch.Generator().manual_seed(0)
    amp = torch.rand(B, 1 << n, generator=g)
    amp = (amp / amp.norm(dim=1, keepdim=True)).cuda()
    theta = torch.randn(1, L, n, 2, generator=g).cuda()
    E = svop.sv_forward(None, theta, None, "ry", "cnot", init_state=amp)
    Er = ref.vqc_expvals_ref(None, theta.cpu().double(), None, "ry", "cnot", amp.cpu().to(torch.complex128))
    np.testing.assert_allclose(E.cpu().numpy(), Er.numpy(), atol=TOL)


@pytest.mark.parametrize("n", [1, 2, 4, 5, 7, 8, 10])
def test_adjoint_matches_oracle_and_param_shift(n):
    L, B = 2, 21
    ang, theta, cot, _ = _rand(n, L, B, seed=10 + n)
    gth, gang = svop.sv_backward(ang, theta, cot, method="adjoint", want_angle_grad=True)
    gps, _ = svop.sv_backward(ang, theta, cot, method="param_shift")
    go = np.zeros((L, n, 2))
    gao = np.zeros((B, n))
    for b in range(B):
        g1, g2 = ovqc.grad_adjoint(ang[b].double().cpu().numpy(), theta[0].double().cpu().numpy(),
                                   cot[b].double().cpu().numpy(), want_angle_grad=True)
        go += g1
        gao[b] = g2
    np.testing.assert_allclose(gth[0].cpu().numpy(), go, atol=2e-4)
    np.testing.assert_allclose(gps[0].cpu().numpy(), go, atol=2e-4)          # ROADMAP: shift == adjoint
    np.testing.assert_allclose(gang.cpu().numpy(), gao, atol=5e-5)


def test_groups_and_autograd():
    n, L, B, G = 8, 6, 64, 4
    ang, theta, cot, gid = _rand(n, L, B, G, seed=5)
    theta.requires_grad_(True)
    E = svop.vqc_expvals(ang, theta, gid)
    (E * cot).sum().backward()
    th = theta.detach().cpu().double().requires_grad_(True)
    Er = ref.vqc_expvals_ref(ang.cpu().double(), th, gid.cpu().long())
    (Er * cot.cpu().double()).sum().backward()
    np.testing.assert_allclose(E.detach().cpu().numpy(), Er.detach().numpy(), atol=TOL)
    np.testing.assert_allclose(theta.grad.cpu().numpy(), th.grad.numpy(), atol=3e-4)


def test_bad_arguments_are_rejected_loudly():
    ang, theta, _, _ = _rand(4, 1, 2)
    with pytest.raises(_lib.QfxError):
        svop.sv_forward(torch.zeros(2, 11).cuda(), torch.zeros(1, 1, 11, 2).cuda())
