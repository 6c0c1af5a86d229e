"""Renyi-DP accountant for the (sub)sampled Gaussian mechanism.

The reference plans to "track epsilon with Opacus" for sampling rate q, noise multiplier sigma,
T rounds and delta (/root/reference/ROADMAP.md:56-58,140-141); Opacus is not available offline,
so this is a self-contained accountant: RDP of the sampled Gaussian mechanism at integer orders
(Mironov, Talwar, Zhang 2019, eq. for integer alpha), composed additively over rounds, converted
to (epsilon, delta) with the bound of Balle et al. 2020 (the conversion Opacus uses) or the
classic one.  For q = 1 it reduces to alpha / (2 sigma^2) per round, which the tests check in
closed form.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import List, Sequence, Tuple

DEFAULT_ORDERS = tuple(range(2, 129)) + (160, 192, 256, 384, 512)


def _log_add(a: float, b: float) -> float:
    if a == -math.inf:
        return b
    if b == -math.inf:
        return a
    hi, lo = (a, b) if a > b else (b, a)
    return hi + math.log1p(math.exp(lo - hi))


def rdp_sampled_gaussian(q: float, sigma: float, alpha: int) -> float:
    """RDP epsilon at integer order alpha of one step of the sampled Gaussian mechanism."""
    if q == 0:
        return 0.0
    if sigma == 0:
        return math.inf
    if q == 1.0:
        return alpha / (2.0 * sigma ** 2)
    log_a = -math.inf
    for i in range(alpha + 1):
        term = (math.lgamma(alpha + 1) - math.lgamma(i + 1) - math.lgamma(alpha - i + 1)
                + i * math.log(q) + (alpha - i) * math.log1p(-q) + (i * i - i) / (2.0 * sigma ** 2))
        log_a = _log_add(log_a, term)
    return log_a / (alpha - 1)


def eps_from_rdp(orders: Sequence[int], rdp: Sequence[float], delta: float, improved: bool = True) -> Tuple[float, int]:
    best, best_order = math.inf, orders[0]
    for a, r in zip(orders, rdp):
        if not math.isfinite(r):
            continue
        if improved:
            e = r + math.log((a - 1) / a) - (math.log(delta) + math.log(a)) / (a - 1)
        else:
            e = r + math.log(1.0 / delta) / (a - 1)
        if e < best:
            best, best_order = e, a
    return max(best, 0.0), best_order


@dataclass
class RDPAccountant:
    orders: Sequence[int] = DEFAULT_ORDERS
    history: List[Tuple[float, float, int]] = field(default_factory=list)  # (q, sigma, steps)

    def step(self, q: float, sigma: float, steps: int = 1) -> None:
        if self.history and self.history[-1][:2] == (q, sigma):
            q0, s0, n0 = self.history[-1]
            self.history[-1] = (q0, s0, n0 + steps)
        else:
            self.history.append((q, sigma, steps))

    def rdp(self) -> List[float]:
        tot = [0.0] * len(self.orders)
        for q, sigma, steps in self.history:
            for i, a in enumerate(self.orders):
                tot[i] += steps * rdp_sampled_gaussian(q, sigma, a)
        return tot

    def epsilon(self, delta: float = 1e-5, improved: bool = True) -> float:
        if not self.history:
            return 0.0
        return eps_from_rdp(self.orders, self.rdp(), delta, improved)[0]

    def state_dict(self):
        return {"orders": list(self.orders), "history": list(self.history)}

    def load_state_dict(self, sd):
        self.orders = tuple(sd["orders"])
        self.history = [tuple(h) for h in sd["history"]]


def gaussian_epsilon_closed_form(sigma: float, steps: int, delta: float) -> float:
    """min over real alpha>1 of steps*alpha/(2 sigma^2) + log(1/delta)/(alpha-1) (classic
    conversion, q = 1): the analytic value the accountant is tested against."""
    c = steps / (2.0 * sigma ** 2)
    l = math.log(1.0 / delta)
    alpha = 1.0 + math.sqrt(l / c)
    return c * alpha + l / (alpha - 1.0)
