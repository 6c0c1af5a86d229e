"""Convergence parity on a task that is NOT trivially separable: the kernel path must land where eager torch lands.

The default synthetic task is solved within a round, which would hide any numerical damage.  Here the class
templates are pulled together (contrast 0.8), the pixel noise is doubled and 10 % of the labels are re-drawn, so a
TinyCNN climbs from about 74 % to about 93 % test accuracy over five FedAvg rounds with a loss that stays near 0.32
(reference semantics: /root/reference/src/CFed/Classical_FL.py:40-81,104-157)."""
import pytest
import torch

from qfedx_b200.data import synthetic_mnist
from qfedx_b200.fl import FederatedDataset, FederatedRunner, FLConfig
from qfedx_b200.models import TinyCNN

pytestmark = pytest.mark.gpu

HARD = dict(noise=0.5, contrast=0.8, label_noise=0.1)
K, PER, ROUNDS, EPOCHS, LR = 4, 256, 5, 3, 0.02


def _eager_fedavg_float64(clients, test, state0):
    """The reference's loop restated in float64 on the CPU (no dropout, fixed sample order)."""
    Xt, yt = test
    glob = TinyCNN(3, p_drop=0.0).double()
    glob.load_state_dict({k: v.double() for k, v in state0.items()})
    curve = []
    for _ in range(ROUNDS):
        states = []
        for X, y in clients:
            m = TinyCNN(3, p_drop=0.0).double()
            m.load_state_dict(glob.state_dict())
            m.train()
            opt = torch.optim.SGD(m.parameters(), lr=LR, momentum=0.9)
            for _e in range(EPOCHS):
                for lo in range(0, len(X), 32):
                    opt.zero_grad()
                    torch.nn.functional.cross_entropy(m(X[lo:lo + 32].double()), y[lo:lo + 32]).backward()
                    opt.step()
            states.append(m.state_dict())
        n = sum(len(c[0]) for c in clients)
        glob.load_state_dict({k: sum(s[k] * (len(c[0]) / n) for s, c in zip(states, clients)) for k in states[0]})
        glob.eval()
        with torch.no_grad():
            z = glob(Xt.double())
        curve.append((float(torch.nn.functional.cross_entropy(z, yt)), float((z.argmax(1) == yt).double().mean())))
    return curve


@pytest.mark.parametrize("engine,conv2", [("persistent", "tcgen05"), ("graph", "fp32")])
def test_fedavg_curve_matches_eager_float64_on_a_hard_task(engine, conv2):
    from qfedx_b200.fl.trainers_cnn import TinyCNNBatchedTrainer
    X, y = synthetic_mnist(K * PER, 3, seed=0, **HARD)
    Xt, yt = synthetic_mnist(1000, 3, seed=1, **HARD)
    clients = [(X[k * PER:(k + 1) * PER], y[k * PER:(k + 1) * PER]) for k in range(K)]
    torch.manual_seed(0)
    model = TinyCNN(3)
    state0 = {k: v.clone() for k, v in model.state_dict().items()}
    ref = _eager_fedavg_float64(clients, (Xt, yt), state0)
    # the task is what it claims to be: not solved, not stuck
    assert 0.85 <= ref[-1][1] <= 0.95 and ref[0][1] < ref[-1][1] - 0.1, ref

    data = FederatedDataset.from_clients(clients, "cuda")
    tr = TinyCNNBatchedTrainer(model.cuda(), data, epochs=EPOCHS, batch_size=32, lr=LR, momentum=0.9, p_drop=0.0,
                               shuffle=False, engine=engine, conv2=conv2)
    runner = FederatedRunner(tr, FLConfig(num_rounds=ROUNDS, mode="weights", quiet=True), (Xt.cuda(), yt.cuda()))
    out = runner.run()
    got = list(zip(out["history"]["loss"][1:], out["accuracies"][1:]))
    for (l_ref, a_ref), (l_got, a_got) in zip(ref, got):
        assert abs(a_got - a_ref) <= 0.005, (ref, got)            # within 0.5 % accuracy every round
        assert abs(l_got - l_ref) <= 0.02 * l_ref, (ref, got)     # within 2 % loss every round
