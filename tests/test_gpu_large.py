"""K5 pass engine and K6 fused swap (loopback on one GPU) vs the torch reference."""
import math

import numpy as np
import pytest
import torch

from qfedx_b200.ops import ref
from qfedx_b200.ops import statevector as svop

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,L,ent,tile", [(13, 2, "cnot", 12), (11, 3, "cz", 8), (16, 1, "cnot", 12), (5, 2, "cnot", 4)])
def test_pass_engine_matches_reference(n, L, ent, tile):
    from qfedx_b200.ops.large import LargeVQC
    g = torch.Generator().manual_seed(n)
    B = 3
    ang = (torch.rand(B, n, generator=g) * math.pi).cuda()
    theta = torch.randn(L * n * 2, generator=g).cuda()
    vqc = LargeVQC(n, L, ent, tile_bits=tile)
    st = vqc.states(ang, theta.reshape(1, -1))
    st_ref = ref.vqc_state_ref(ang.cpu().double(), theta.cpu().double().reshape(1, L, n, 2), None, "ry", ent)
    np.testing.assert_allclose(st.cpu().numpy(), st_ref.numpy(), atol=3e-5)
    np.testing.assert_allclose(vqc.expvals(ang, theta.reshape(1, -1)).cpu().numpy(),
                               ref.expval_z_all(st_ref).numpy(), atol=3e-5)


def test_param_shift_gradient_matches_adjoint_kernel():
    from qfedx_b200.ops.large import LargeVQC
    n, L, S = 10, 2, 4
    g = torch.Generator().manual_seed(1)
    ang = (torch.rand(S, n, generator=g) * math.pi).cuda()
    theta = torch.randn(L * n * 2, generator=g).cuda()
    cot = torch.randn(S, n, generator=g).cuda()
    grad = LargeVQC(n, L).param_shift_grad(ang, theta, cot, max_states=64)
    gadj, _ = svop.sv_backward(ang, theta.reshape(1, L, n, 2), cot)            # register-tiled adjoint (K4a)
    np.testing.assert_allclose(grad.cpu().numpy(), gadj.reshape(-1).cpu().numpy(), atol=3e-4)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_swap_fused_loopback(world):
    from qfedx_b200.ops.large import ansatz_gates
    from qfedx_b200.parallel.sharded import ShardedState, plan_sharded
    n, L = 15, 2
    g = torch.Generator().manual_seed(world)
    ang = (torch.rand(1, n, generator=g) * math.pi).cuda()
    theta = torch.randn(L * n * 2, generator=g).cuda()
    gates = ansatz_gates(n, L)
    gbits = int(math.log2(world))
    plan = plan_sharded(n, gbits, gates, tile_bits=10)
    assert plan.n_swaps >= 1
    st = ShardedState(n, world, loopback=True)
    st.reset_zero_state()
    st.run(plan, theta, ang)
    full = st.gather_logical(plan)
    st_ref = ref.vqc_state_ref(ang.cpu().double(), theta.cpu().double().reshape(1, L, n, 2))[0]
    np.testing.assert_allclose(full.cpu().numpy(), st_ref.numpy(), atol=3e-5)
    np.testing.assert_allclose(st.expvals(plan).cpu().numpy(), ref.expval_z_all(st_ref[None])[0].numpy(), atol=3e-5)
