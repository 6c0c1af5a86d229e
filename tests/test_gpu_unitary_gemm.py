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


@pytest.mark.parametrize("n,L,B", [(11, 3, 300), (12, 2, 256)])
def test_model_api_uses_the_unitary_gemm_for_11_and_12_qubits(n, L, B):
    """No register kernel exists beyond 10 qubits: VQC.predict / inference_expvals run on K2 there (and agree with the
    pass engine and with the float64 reference)."""
    from qfedx_b200.models import VQC
    from qfedx_b200.ops.unitary_gemm import prefers_unitary_gemm
    torch.manual_seed(n)
    m = VQC(n, L, 3, init_scale=0.7).cuda()
    g = torch.Generator().manual_seed(1)
    x = (torch.rand(B, n, generator=g) * math.pi).cuda()
    assert prefers_unitary_gemm(n, 16, B) and not prefers_unitary_gemm(n, 2, B)     # deep: GEMM; shallow: pass engine
    E = m.inference_expvals(x, "gemm")
    assert getattr(m, "_ug", None) is not None
    E_pass = m.inference_expvals(x, "passes")
    E_ref = ref.vqc_expvals_ref(x[:48].cpu().double(), m.theta.detach().cpu().double()[None]).numpy()
    np.testing.assert_allclose(E[:48].cpu().numpy(), E_ref, atol=5e-5)
    np.testing.assert_allclose(E.cpu().numpy(), E_pass.cpu().numpy(), atol=5e-5)
    pred = m.predict(x, "gemm")
    z_ref = (m.readout_scale.detach().cpu().double() * torch.from_numpy(E_ref)[:, :3] + m.readout_bias.detach().cpu().double())
    assert torch.equal(pred[:48].cpu(), z_ref.argmax(1))
    # the packed unitary is reused until theta changes, then rebuilt
    ug = m._ug
    m.inference_expvals(x, "gemm")
    assert m._ug is ug
    with torch.no_grad():
        m.theta.add_(0.1)
    E2 = m.inference_expvals(x, "gemm")
    assert m._ug is not ug and float((E2 - E).abs().max()) > 1e-3


def test_deep_circuit_over_a_large_batch_goes_to_the_gemm_and_matches_the_register_kernel():
    from qfedx_b200.models import VQC
    from qfedx_b200.ops.unitary_gemm import prefers_unitary_gemm
    assert prefers_unitary_gemm(10, 24, 8192) and not prefers_unitary_gemm(10, 4, 8192) and not prefers_unitary_gemm(8, 24, 512)
    torch.manual_seed(0)
    m = VQC(10, 24, 2, init_scale=0.5).cuda()
    x = torch.rand(8192, 10, device="cuda") * math.pi
    a, b = m.inference_expvals(x), m.inference_expvals(x, "registers")
    assert getattr(m, "_ug", None) is not None
    np.testing.assert_allclose(a.cpu().numpy(), b.cpu().numpy(), atol=1e-4)
