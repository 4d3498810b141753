"""Encode-pass / loss-epilogue kernels through the C ABI against fp64 statements of the
reference code they replace: dequantise + scalar-mult + scalar-add + logit (`wrapper.py:27-30`,
`math.py:41-105`), the affine coupling in the encode direction (`acl.py:43-46,101-111`), the
reconstruction error (`non_square.py:111-114`) and the standard-normal log-prob (`gaussian.py:9-22`)."""
import math

import pytest
import torch

from conftest import load_golden
from cases import CASES, make_input
from helpers import build_product, strip_wrappers
from cmf_b200 import kernels as K, ops

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _close(a, b, rtol=2e-6, atol=1e-6, what=""):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    err, scale = float((a - b).abs().max()), float(b.abs().max())
    assert err <= rtol * scale + atol, f"{what}: err {err:.3e} scale {scale:.3e}"


@pytest.mark.parametrize("B,shape,lam", [(50, (1, 28, 28), 1e-6), (3, (3, 32, 32), 0.05), (1, (1, 2, 2), 1e-6), (7, (5,), 0.05)])
def test_preproc_logit(B, shape, lam):
    g = torch.Generator().manual_seed(B)
    pixels = torch.randint(0, 256, (B, *shape), generator=g).float()
    pixels.view(-1)[:2] = torch.tensor([0.0, 255.0])                     # the extremes of the range
    noise = torch.rand(B, *shape, generator=g)
    mult, add = (1 - 2 * lam) / 256, lam
    x = pixels.clone().to(DEV)
    y, lj = K.preproc_logit_fwd(x, noise.to(DEV), mult, add)
    assert torch.equal(x.cpu(), pixels + noise), "the noise is added in place to the caller's batch"
    u = mult * (pixels.double() + noise.double()) + add
    uc = u.clamp(1e-7, 1 - 1e-7)
    # logit is ill-conditioned at the ends of (0, 1): d logit / du = 1 / (u (1 - u)), and u itself carries
    # fp32 rounding (~1.2e-7 relative for mult * x + add) -- bound the error element by element
    ref_y = torch.log(u) - torch.log(1 - u)
    bound = 2.4e-7 * u / torch.minimum(u, 1 - u).clamp_min(1e-12) / u.clamp_min(1e-12) * u + 2e-6
    excess = ((y.double().cpu() - ref_y).abs() - bound).max()
    assert float(excess) <= 0, f"logit: {float(excess):.3e} over the conditioning bound"
    ref_lj = u[0].numel() * math.log(abs(mult)) + (-torch.log(uc) - torch.log(1 - uc)).flatten(1).sum(1, keepdim=True)
    # the log-Jacobian is n log|mult| (~ -4.3e3 for 28x28) plus a sum of n terms of the opposite sign: the
    # result is small, its fp32 rounding is that of the two large parts
    # ... plus, element by element, the conditioning of log(1 - u) / log(u) under the fp32 rounding of u
    cond = (2.4e-7 * u / torch.minimum(u, 1 - u).clamp_min(1e-12)).flatten(1).sum(1, keepdim=True)
    tol = cond + 2e-7 * u[0].numel() * abs(math.log(abs(mult)))
    assert float(((lj.double().cpu() - ref_lj).abs() - tol).max()) <= 0, "log-jac beyond its conditioning bound"
    y2, lj2 = K.preproc_logit_fwd(x, None, mult, add)                    # no noise: x untouched
    assert torch.equal(y2, y) and torch.equal(lj2, lj)


@pytest.mark.parametrize("name", ["sphere_gij_d3", "hepmass_gij_d10_warm", "img8_gij_d8", "img12_c64_gij_d16"])
def test_fused_encode_coupling_matches_the_op_string(name):
    """every ACL of the model: the fused encode-direction kernel (forward and adjoint) against the
    reference's op string evaluated in fp64 on the same conditioner output"""
    from cmf_b200.bijections import AffineCouplingBijection
    g = load_golden(name)
    density = build_product(g, DEV)
    layers = [m for m in density.modules() if isinstance(m, AffineCouplingBijection)]
    assert layers
    B = 5
    for bij in layers[:6]:
        torch.manual_seed(len(bij.x_shape))
        x = torch.randn(B, *bij.x_shape, device=DEV, requires_grad=True)
        out = bij.x_to_z(x)
        gz, glj = torch.randn_like(out["z"]), torch.randn(B, 1, device=DEV)
        ((out["z"] * gz).sum() + (out["log-jac"] * glj).sum()).backward()
        # reference op string in fp64 (acl.py:43-46 / :101-111)
        x64 = x.detach().double().requires_grad_(True)
        keep_pos, mod_pos, in_shape, dense = bij.partition()
        b64 = bij.double()
        flat = x64.flatten(1)
        cond = (b64.mask * x64) if dense else flat[:, keep_pos].view(B, *in_shape)
        res = b64.coupler(cond)
        t, s = res["shift"].flatten(1), res["log-scale"].flatten(1)
        if dense:
            t, s = t[:, mod_pos], s[:, mod_pos]
        z64 = flat.clone()
        z64[:, mod_pos] = (flat[:, mod_pos] + t) * torch.exp(s)
        ((z64 * gz.double().flatten(1)).sum() + (s.sum(1, keepdim=True) * glj.double()).sum()).backward()
        _close(out["z"].flatten(1), z64, rtol=1e-5, what="z")
        _close(out["log-jac"], s.sum(1, keepdim=True), rtol=1e-5, what="log-jac")
        _close(x.grad.flatten(1), x64.grad.flatten(1), rtol=1e-4, atol=1e-5, what="g_x")
        bij.float()


@pytest.mark.parametrize("B,n,with_c", [(50, 784, True), (3, 5, False), (1, 1, True), (6, 3072, True), (17, 64, False)])
def test_sumsq_diff(B, n, with_c):
    g = torch.Generator().manual_seed(n)
    a = torch.randn(B, n, generator=g).to(DEV).requires_grad_(True)
    c = torch.randn(B, n, generator=g).to(DEV).requires_grad_(True) if with_c else None
    out = ops.sumsq_diff(a, c)
    gout = torch.randn(B, 1, generator=g).to(DEV)
    (out * gout).sum().backward()
    d = a.detach().double() - (c.detach().double() if with_c else 0)
    _close(out, (d * d).sum(1, keepdim=True), what="sum of squares")
    _close(a.grad, 2 * gout.double() * d, what="g_a")
    if with_c:
        _close(c.grad, -2 * gout.double() * d, what="g_c")


def test_dequantiser_takes_the_fused_preprocessing_path(monkeypatch):
    """public API with the wrappers on: DequantizationDensity + the three elementwise nodes in one
    launch, same elbo as the node-by-node path on the same noise"""
    g = load_golden("img12_c64_gij_d16")
    density = build_product(g, DEV).eval()
    x = make_input(CASES["img12_c64_gij_d16"]).floor().to(DEV)           # integer pixel values
    noise = torch.rand(x.shape, generator=torch.Generator().manual_seed(0)).to(DEV)
    monkeypatch.setattr(torch, "rand_like", lambda t, **kw: noise.clone())
    calls = []
    real = K.preproc_logit_fwd
    monkeypatch.setattr(K, "preproc_logit_fwd", lambda *a: (calls.append(1), real(*a))[1])
    with torch.no_grad():
        x1 = x.clone()
        fused = density.elbo(x1)["elbo"]
        assert calls and torch.equal(x1, x + noise)
        deq = density.module
        x2 = x.clone()
        plain = deq.density.elbo(x2.add_(noise))["elbo"]                 # the unfused node chain below the dequantiser
    _close(fused, plain, rtol=1e-6, what="elbo through the fused preprocessing")
