"""CPU: the spline oracle against golden outputs of the unmodified nsf code (tests/golden/rqs_nsf.pt,
made by oracle/gen_rqs_golden.py), and the host-side knot tables against the oracle."""
import os

import pytest
import torch

import rqs_oracle as R
from cmf_b200 import splines

GOLD = torch.load(os.path.join(os.path.dirname(__file__), "golden", "rqs_nsf.pt"), weights_only=False)


@pytest.mark.parametrize("tag", list(GOLD))
def test_oracle_matches_nsf_golden(tag):
    g = GOLD[tag]
    for inverse in (False, True):
        ref = g["inv" if inverse else "fwd"]
        y, lad, idx = R.unconstrained_rqs(g["x"], g["params"], g["K"], g["tail_bound"], inverse)
        assert torch.allclose(y, ref["y"], rtol=1e-6, atol=1e-6)
        assert torch.allclose(lad, ref["logabsdet"], rtol=1e-5, atol=1e-5)
        inside = (g["x"] >= -g["tail_bound"]) & (g["x"] <= g["tail_bound"])
        assert bool(((idx >= 0) == inside).all()) and int(idx.max()) <= g["K"] - 1
    s = g["search"]
    x = g["x"] / (g["tail_bound"] * 1.3)
    assert torch.equal(R.searchsorted(s["knots"], x), s["idx"])


@pytest.mark.parametrize("tag", list(GOLD))
def test_host_knot_tables_bit_equal_to_reference_arithmetic(tag):
    g = GOLD[tag]
    K, tb = g["K"], g["tail_bound"]
    p = g["params"].unsqueeze(0)                                       # (1, N, 3K-1)
    cw, ch, dv, *_ = splines.knot_tables(p, None, K, tb, hidden_features=37)
    uw, uh, ud = R.split_params(p, K, hidden_features=37)
    rcw, _, rch, _, rdv = R.knots(uw, uh, ud, -tb, tb, -tb, tb)
    assert torch.equal(cw, rcw) and torch.equal(ch, rch) and torch.equal(dv, rdv)


def test_host_knot_tangents_match_autograd_jvp():
    torch.manual_seed(0)
    B, d, M, K, tb = 3, 4, 5, 6, 2.5
    p = torch.randn(B, M, 3 * K - 1, dtype=torch.float64)
    P = torch.randn(B, d, M, 3 * K - 1, dtype=torch.float64)
    cw, ch, dv, CW, CH, DV = splines.knot_tables(p, P, K, tb, hidden_features=16)

    def tables(q):
        uw, uh, ud = R.split_params(q, K, hidden_features=16)
        c, _, h, _, v = R.knots(uw, uh, ud, -tb, tb, -tb, tb)
        return c, h, v
    for k in range(d):
        _, (jc, jh, jv) = torch.autograd.functional.jvp(tables, (p,), (P[:, k],))
        assert torch.allclose(CW[:, k], jc, atol=1e-12) and torch.allclose(CH[:, k], jh, atol=1e-12)
        assert torch.allclose(DV[:, k], jv, atol=1e-12)
