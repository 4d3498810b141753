"""Rational-quadratic spline coupling kernel (cmf_rqs_coupling_jvp_fwd/bwd through the C ABI)
against the CPU oracle (oracle/rqs_oracle.py, pinned to the nsf submodule's own outputs):
bin indices bit-exact, values / tangents / adjoints within 1e-4 relative (fp64 oracle)."""
import pytest
import torch

import rqs_oracle as R
from cmf_b200 import ops, splines

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _case(B, d, M, K, tb, seed=0, scale=1.5):
    g = torch.Generator().manual_seed(seed + B + 7 * M + 13 * K)
    z = (torch.rand(B, M, generator=g, dtype=torch.float64) * 2 - 1) * tb * 1.25
    z.view(-1)[:3] = torch.tensor([tb, -tb, 0.0], dtype=torch.float64)[: min(3, z.numel())]
    p = torch.randn(B, M, 3 * K - 1, generator=g, dtype=torch.float64) * scale
    V = torch.randn(B, d, M, generator=g, dtype=torch.float64)
    P = torch.randn(B, d, M, 3 * K - 1, generator=g, dtype=torch.float64)
    return z, p, V, P


SHAPES = [(3, 2, 5, 8, 3.0), (64, 6, 3, 10, 1.0), (7, 10, 11, 31, 4.0), (2, 0, 9, 4, 2.0), (50, 64, 33, 8, 3.0),
          (5, 3, 1, 1, 2.0)]


@pytest.mark.parametrize("B,d,M,K,tb", SHAPES)
def test_rqs_forward_bins_bit_exact_values_and_tangents(B, d, M, K, tb):
    z, p, V, P = _case(B, d, M, K, tb)
    zf, pf = z.float().to(DEV), p.float().to(DEV)
    Vf, Pf = (V.float().to(DEV), P.float().to(DEV)) if d else (None, None)
    cw, ch, dv, CW, CH, DV = splines.knot_tables(pf, Pf, K, tb, hidden_features=32)
    x, Vout, bins = ops.rqs_coupling_jvp(zf, Vf, cw, ch, dv, CW, CH, DV, tb)
    # (1) bins: the reference's searchsorted over the very knots the kernel saw, evaluated on the CPU
    inside = (zf.cpu() >= -tb) & (zf.cpu() <= tb)
    ref_bins = torch.where(inside, R.searchsorted(ch.cpu(), zf.cpu()), torch.full((B, M), -1, dtype=torch.long))
    assert torch.equal(bins.cpu().long(), ref_bins), "spline bin indices must be bit-exact"
    # (2) values against the fp64 oracle on the fp32-rounded inputs
    z64, p64 = zf.double().cpu(), pf.double().cpu()
    f = lambda a, q: R.unconstrained_rqs(a, q, K, tb, True, hidden_features=32)[0]
    x_ref = f(z64, p64)
    assert float((x.double().cpu() - x_ref).abs().max()) <= 1e-4 * max(1.0, float(x_ref.abs().max()))
    for k in range(d):
        _, jv = torch.autograd.functional.jvp(f, (z64, p64), (Vf[:, k].double().cpu(), Pf[:, k].double().cpu()))
        err = float((Vout[:, k].double().cpu() - jv).abs().max())
        assert err <= 1e-4 * max(1.0, float(jv.abs().max())), f"tangent column {k}: {err}"


@pytest.mark.parametrize("B,d,M,K,tb", [(3, 2, 5, 8, 3.0), (16, 6, 3, 10, 1.0), (4, 5, 7, 31, 4.0), (6, 0, 4, 5, 2.0)])
def test_rqs_adjoint_matches_fp64_autograd(B, d, M, K, tb):
    z, p, V, P = _case(B, d, M, K, tb, seed=3, scale=1.0)
    gx = torch.randn(B, M, dtype=torch.float64)
    gV = torch.randn(B, d, M, dtype=torch.float64)
    leaves = [t.float().to(DEV).requires_grad_(True) for t in ((z, p, V, P) if d else (z, p))]
    zf, pf = leaves[0], leaves[1]
    Vf, Pf = (leaves[2], leaves[3]) if d else (None, None)
    x, Vout, _ = ops.rqs_coupling_jvp(zf, Vf, *splines.knot_tables(pf, Pf, K, tb, hidden_features=32), tb)
    loss = (x * gx.float().to(DEV)).sum()
    if d:
        loss = loss + (Vout * gV.float().to(DEV)).sum()
    loss.backward()
    # fp64 oracle: forward-mode via double backward, then reverse
    ref = [t.detach().double().cpu().requires_grad_(True) for t in leaves]
    f = lambda a, q: R.unconstrained_rqs(a, q, K, tb, True, hidden_features=32)[0]
    loss_r = (f(ref[0], ref[1]) * gx).sum()
    for k in range(d):
        _, jv = torch.autograd.functional.jvp(f, (ref[0], ref[1]), (ref[2][:, k], ref[3][:, k]), create_graph=True)
        loss_r = loss_r + (jv * gV[:, k]).sum()
    loss_r.backward()
    for name, a, b in zip(("g_z", "g_params", "g_V", "g_P"), leaves, ref):
        scale = max(1.0, float(b.grad.abs().max()))
        err = float((a.grad.double().cpu() - b.grad).abs().max())
        assert err <= 2e-4 * scale, f"{name}: err {err:.3e} scale {scale:.3e}"


def test_rqs_rejects_too_many_bins():
    from cmf_b200._lib import CmfB200Error
    z = torch.zeros(2, 3, device=DEV)
    t = torch.zeros(2, 3, 34, device=DEV)
    with pytest.raises(CmfB200Error, match="num_bins"):
        ops.rqs_coupling_jvp(z, None, t, t, t, None, None, None, 1.0)
