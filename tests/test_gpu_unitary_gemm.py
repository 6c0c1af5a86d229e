"""K2 tcgen05 complex GEMM vs the float64 torch reference of the same op."""
import math

import numpy as np
import pytest
import torch

from qfedx_b200.ops import ref
from qfedx_b200.ops import statevector as svop

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,L,B", [(8, 6, 200), (6, 2, 64), (9, 3, 130)])
def test_unitary_gemm_states_and_expvals(n, L, B):
    from qfedx_b200.ops.unitary_gemm import UnitaryGemm, apply_unitary_reference
    g = torch.Generator().manual_seed(n)
    theta = torch.randn(1, L, n, 2, generator=g).cuda()
    ang = (torch.rand(B, n, generator=g) * math.pi).cuda()
    ug = UnitaryGemm(theta)
    E, psi = ug.expvals(ang, want_states=True)
    psi_ref = ref.vqc_state_ref(ang.cpu().double(), theta.cpu().double())
    E_ref = ref.expval_z_all(psi_ref)
    np.testing.assert_allclose(psi.cpu().numpy(), psi_ref.numpy(), atol=2e-5)
    np.testing.assert_allclose(E.cpu().numpy(), E_ref.numpy(), atol=5e-5)
    # same numbers as the register-tiled kernel
    np.testing.assert_allclose(E.cpu().numpy(), svop.sv_forward(ang, theta).cpu().numpy(), atol=5e-5)
    # arbitrary complex input states
    s = torch.randn(B, 1 << n, dtype=torch.complex64, generator=g)
    s = (s / s.norm(dim=1, keepdim=True)).cuda()
    out = ug.apply(s)
    np.testing.assert_allclose(out.cpu().numpy(), apply_unitary_reference(theta, s).numpy(), atol=2e-5)


def test_plain_tf32_is_less_accurate_but_close():
    from qfedx_b200.ops.unitary_gemm import UnitaryGemm
    g = torch.Generator().manual_seed(0)
    theta = torch.randn(1, 4, 8, 2, generator=g).cuda()
    ang = (torch.rand(256, 8, generator=g) * math.pi).cuda()
    e3 = UnitaryGemm(theta, precision=3).expvals(ang)
    e1 = UnitaryGemm(theta, precision=1).expvals(ang)
    E_ref = ref.vqc_expvals_ref(ang.cpu().double(), theta.cpu().double()).numpy()
    err3 = np.abs(e3.cpu().numpy() - E_ref).max()
    err1 = np.abs(e1.cpu().numpy() - E_ref).max()
    assert err3 < 5e-5 and err1 < 5e-3 and err3 < err1
