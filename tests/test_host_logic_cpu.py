"""Host logic on CPU: the layer program, autograd wiring and fused head, with the CUDA launchers
replaced by torch emulations of their C-ABI contracts (tests/emu_kernels.py).  What is checked
here is everything ABOVE the kernels; the kernels themselves are checked on the GPU box."""
import pytest
import torch

import emu_kernels
from conftest import load_golden
from cases import CASES, make_input
from helpers import build_product, run_loss, check_against_golden, strip_wrappers, find_head, assert_close

FAST = [n for n, c in CASES.items() if not c.get("slow")]


@pytest.fixture
def emu(monkeypatch):
    import cmf_b200.kernels as K
    for name in emu_kernels.ALL:
        monkeypatch.setattr(K, name, getattr(emu_kernels, name))
    return K


@pytest.fixture
def emu_raw(monkeypatch):
    """The emulations made as opaque to autograd as the real launchers (raw C-ABI calls record nothing
    on the tape, even inside a backward pass that is itself being recorded): a graph through them can
    only come from the ops' own explicit adjoints."""
    import functools
    import cmf_b200.kernels as K

    def opaque(fn):
        @functools.wraps(fn)
        def inner(*a, **kw):
            with torch.no_grad():
                return fn(*a, **kw)
        return inner
    for name in emu_kernels.ALL:
        monkeypatch.setattr(K, name, opaque(getattr(emu_kernels, name)))
    return K


@pytest.mark.parametrize("name", [n for n in FAST if CASES[n].get("hutch")])
def test_hutch_with_cg_gradients_with_opaque_kernels(emu_raw, name):
    """hutch_with_cg differentiates a VJP (`non_square.py:199`, create_graph=True): the log-det
    surrogate's parameter gradient must survive kernels that autograd cannot see into."""
    g = load_golden(name)
    density = build_product(g)
    loss, grads = run_loss(g, density)
    check_against_golden(g, loss, grads, rtol=1e-4)


@pytest.mark.parametrize("name", FAST)
def test_fused_head_matches_reference_golden(emu, name):
    g = load_golden(name)
    density = build_product(g)
    loss, grads = run_loss(g, density)
    check_against_golden(g, loss, grads, rtol=1e-4)


@pytest.mark.parametrize("name", ["sphere_gij_d3", "img8_gij_d8", "hepmass_gij_d10_warm"])
def test_intermediates(emu, name):
    g = load_golden(name)
    density = build_product(g)
    head, _ = find_head(strip_wrappers(density))
    with torch.no_grad():
        z = g["z"]
        gram, recon = head._get_full_jac_transpose_jac(z)
        assert_close(gram, g["gram"], 2e-5, what="gram")
        logdet, recon, gram2, l1 = head._exact_log_det_jac_and_reconstruction(z, "offdiag")
        assert_close(logdet, g["logdet"], 2e-5, atol=1e-5, what="logdet")
        assert_close(l1, g["l1_offdiag"], 2e-5, what="l1")
        if g["x_pre"] is not None:
            rl = ((recon - g["x_pre"]).flatten(1) ** 2).sum(-1, keepdim=True)
            assert_close(rl, g["recon_loss"], 2e-5, atol=1e-9, what="recon")   # d = D: the error is rounding noise (~1e-14)
        # primal-only program == tangent program's primal
        assert_close(head.flow_forward(z), recon, 1e-6, what="flow_forward")


def test_program_structure_mnist(emu):
    g = load_golden("mnist_gij_d64")
    density = build_product(g)
    head, _ = find_head(strip_wrappers(density))
    prog = head.program
    assert prog.modified_elements() == 3136          # sum_l M_l
    assert prog.D == 784 and prog.d == 64
    kinds = [l.split()[0] for l in prog.describe()]
    assert kinds.count("acl") == 10


def test_jitter_retry(emu, capsys):
    g = load_golden("sphere_gij_d3")
    density = build_product(g)
    head, _ = find_head(strip_wrappers(density))
    calls = {"n": 0}
    real = head._get_full_jac_transpose_jac

    def singular(latent, create_graph=True):
        gram, recon = real(latent, create_graph)
        calls["n"] += 1
        return gram * 0.0, recon                 # rank-0 metric: Cholesky must fail until jitter lands
    head._get_full_jac_transpose_jac = singular
    logdet, _, gram, _ = head._exact_log_det_jac_and_reconstruction(g["z"], "none")
    assert head.last_cholesky_attempts == 2
    assert "WARNING: Numerical non-invertibility" in capsys.readouterr().out
    assert torch.allclose(torch.diagonal(gram, dim1=-2, dim2=-1), torch.full((64, 3), 1e-6))


def test_invalid_method_raises():
    from cmf_b200.head import FusedNonSquareHeadDensity
    with pytest.raises(ValueError, match="not a valid Jacobian calculation method"):
        FusedNonSquareHeadDensity(None, 1, "lu", (3,), "normal")
    h = FusedNonSquareHeadDensity(None, 1, "hutchinson", (3,), "normal")
    assert h.log_jacobian_method == "hutch_with_cg"


def test_no_cpu_fallback():
    """Without the emulation the product must refuse CPU tensors loudly."""
    from cmf_b200._lib import CmfB200Error
    g = load_golden("sphere_gij_d3")
    density = build_product(g)
    with pytest.raises(CmfB200Error):
        run_loss(g, density)


def test_fused_preprocessing_equals_the_node_chain(emu, monkeypatch):
    """DequantizationDensity's one-pass preprocessing (dequantise, scalar-mult, scalar-add, logit) against the
    three BijectionDensity nodes it stands in for, on the same noise"""
    g = load_golden("img8_gij_d8")
    density = build_product(g).eval()
    x = make_input(CASES["img8_gij_d8"]).floor()
    noise = torch.rand(x.shape, generator=torch.Generator().manual_seed(0))
    monkeypatch.setattr(torch, "rand_like", lambda t, **kw: noise.clone())
    with torch.no_grad():
        x1, x2 = x.clone(), x.clone()
        fused = density.elbo(x1)
        plain = density.module.density.elbo(x2.add_(noise))
    assert torch.equal(x1, x + noise)
    assert_close(fused["elbo"], plain["elbo"], 1e-6, what="elbo")
    depth = 0
    while "low-dim-x" not in fused:
        fused, depth = fused["prior-dict"], depth + 1
    assert depth >= 4          # mult, add, logit, head ... nested like the reference's result dicts


def test_hutch_cg_on_explicit_gram_cpu(emu, monkeypatch):
    """host logic of the explicit-G CG path (one d-column pass, then mat-vecs) against the reference fixture"""
    import cmf_b200.head as H
    monkeypatch.setattr(H, "_CG_EXPLICIT", "1")
    g = load_golden("hepmass_hutch_d10")
    density = build_product(g)
    loss, grads = run_loss(g, density)
    check_against_golden(g, loss, grads, rtol=1e-4)


def test_gradient_bucket_gathers_fresh_gradients():
    """GradientBucket.gather_fresh (the CUDA-graph capture's way of filling the bucket): gradients that autograd
    allocated itself are copied into their slices, parameters without a gradient get zeros, views are restored."""
    from cmf_b200.parallel import GradientBucket
    torch.manual_seed(0)
    lin, unused = torch.nn.Linear(5, 3), torch.nn.Linear(2, 2)
    params = list(lin.parameters()) + list(unused.parameters())
    bucket = GradientBucket(params)
    bucket.flat.fill_(7.0)                                    # stale content must not survive
    for p in params:
        p.grad = None
    x = torch.randn(4, 5)
    lin(x).square().sum().backward()
    want = torch.cat([lin.weight.grad.flatten(), lin.bias.grad.flatten(), torch.zeros(6)])
    bucket.gather_fresh()
    assert torch.equal(bucket.flat, want)
    off = 0
    for p in params:
        assert p.grad.data_ptr() == bucket.flat[off:].data_ptr() and p.grad.shape == p.shape
        off += p.numel()
    bucket.gather_fresh()                                     # already aliased: nothing to copy, nothing changes
    assert torch.equal(bucket.flat, want)
