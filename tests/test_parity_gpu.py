"""Parity proper: the CUDA path (through the C ABI) against (1) golden outputs of the unmodified
reference on identical inputs and weights, (2) the CPU oracle on fresh seeded inputs, and
(3) size-independent properties at the BASELINE.json MNIST-shape size.

Tolerance: BASELINE.json north_star -- log-det, metric penalty and parameter gradients within
1e-4 relative in FP32 (TF32 disabled on both sides)."""
import pytest
import torch

from conftest import load_golden
from cases import CASES, make_input
from detfill import fill_state_dict
from helpers import build_product, run_loss, check_against_golden, strip_wrappers, find_head, assert_close, record
import cmf_oracle as O

pytestmark = pytest.mark.gpu
RTOL = 1e-4


@pytest.fixture(autouse=True)
def fp32_exact():
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    yield


@pytest.mark.parametrize("name", list(CASES))
def test_cuda_path_matches_reference_golden(name):
    g = load_golden(name)
    density = build_product(g, "cuda")
    loss, grads = run_loss(g, density, "cuda")
    check_against_golden(g, loss, grads, RTOL)


@pytest.mark.parametrize("name", ["sphere_gij_d3", "sinusoid_gij_d6", "power_gij_d6", "img8_gij_d8",
                                  "img8c3_gij_d16", "mnist_gij_d64"])
def test_cuda_intermediates_match_reference_golden(name):
    g = load_golden(name)
    density = build_product(g, "cuda")
    head, _ = find_head(strip_wrappers(density))
    with torch.no_grad():
        z = g["z"].cuda()
        logdet, recon, gram, l1 = head._exact_log_det_jac_and_reconstruction(z, "offdiag")
        assert head.last_cholesky_attempts == 1
        assert_close(gram, g["gram"], RTOL, what="G")
        assert_close(logdet, g["logdet"], RTOL, atol=1e-4, what="logdet")
        assert_close(l1, g["l1_offdiag"], RTOL, what="l1 offdiag")
        _, _, _, l1d = head._exact_log_det_jac_and_reconstruction(z, "diag")
        assert_close(l1d, g["l1_diag"], RTOL, what="l1 diag")


@pytest.mark.parametrize("name,batch", [("sphere_gij_d3", 257), ("hepmass_gij_d10_warm", 33), ("img8_gij_d8", 5),
                                        # the reference's full per-GPU batches of the flat configs (two_d.py:77, tabular.py:290-296):
                                        # 1000 samples through the whole-stack kernel, 5000 / 750 through the one-launch MLP conditioner
                                        # (278 / 69 row tiles over persistent CTAs)
                                        ("sphere_gij_d3", 1000), ("power_gij_d6", 5000), ("hepmass_gij_d10_warm", 750)])
def test_cuda_path_matches_oracle_on_fresh_inputs(name, batch):
    """Different weights (salt) and inputs than the fixtures; oracle in fp64 as the tighter checker."""
    g = load_golden(name)
    case = dict(CASES[name], batch=batch)
    density = build_product(g, "cpu")
    filled = fill_state_dict(density.state_dict(), salt=11, gain=1.5)
    density.load_state_dict(filled)
    density = density.cuda().train()
    x = make_input(case) * 0.9
    from cmf_b200.non_square_helpers import get_non_square_train_metrics
    loss = get_non_square_train_metrics(g["config"])[0](strip_wrappers(density), x.cuda(), case["epoch"])["loss"]
    loss.backward()
    params = {(k[7:] if k.startswith("module.") else k): (v.double() if v.is_floating_point() else v) for k, v in filled.items()}
    params = {k: (v.requires_grad_(True) if v.is_floating_point() else v) for k, v in params.items()}
    flow = O.OracleFlow(g["schema"], params, case["x_shape"])
    ref = O.train_loss(flow, g["config"], x.double(), case["epoch"])
    ref.backward()
    assert_close(loss, ref, RTOL, what="loss")
    for pname, p in density.named_parameters():
        key = pname[7:] if pname.startswith("module.") else pname
        if params[key].grad is None:
            assert p.grad is None or float(p.grad.abs().max()) == 0
            continue
        scale = float(params[key].grad.norm()) / max(params[key].grad.numel() ** 0.5, 1)
        assert_close(p.grad, params[key].grad, RTOL, atol=RTOL * scale, what=key)


@pytest.mark.parametrize("name,batch", [("sphere_gij_d3", 64), ("hepmass_gij_d10_warm", 17), ("img8_gij_d8", 4)])
def test_eval_ood_entry_matches_oracle(name, batch):
    """The eval / OOD entry (`non_square.py:52-53,119-124`: ``elbo(x, ood=True)`` -> per-sample
    ``likelihood`` and ``reconstruction-error``) and ``extract_latent`` (`:55-62`), in eval mode,
    against the fp64 oracle on the same weights and inputs."""
    g = load_golden(name)
    case = dict(CASES[name], batch=batch)
    density = build_product(g, "cpu")
    filled = fill_state_dict(density.state_dict(), salt=5, gain=1.2)
    density.load_state_dict(filled)
    density = density.cuda().eval()
    x = make_input(case) * 0.8
    top = strip_wrappers(density)
    head, depth = find_head(top)
    xh = x.cuda()
    node = top
    for _ in range(depth):                       # push x through the elementwise nodes above the head
        xh = node.bijection.x_to_z(xh)["z"]
        node = node.prior
    with torch.no_grad():
        out = head("elbo", xh, ood=True)
        z = head("extract-latent", xh)
    params = {(k[7:] if k.startswith("module.") else k): (v.double() if v.is_floating_point() else v) for k, v in filled.items()}
    flow = O.OracleFlow(g["schema"], params, case["x_shape"])
    with torch.no_grad():
        ref = flow.elbo(x.double(), ood=True, training=False)
    assert set(out) >= {"likelihood", "reconstruction-error"}
    assert_close(out["likelihood"], ref["likelihood"], RTOL, atol=1e-4, what="ood likelihood")
    assert_close(out["reconstruction-error"], ref["reconstruction-error"], RTOL, atol=1e-6, what="ood reconstruction error")
    assert_close(z, ref["z"], RTOL, atol=1e-6, what="extract_latent")


def test_mnist_shape_properties_full_batch():
    """B = 50 (the reference's image batch), D = 784, d = 64: properties that need no oracle run.
    (i) G symmetric PSD with log det == 2 sum log diag chol (torch), (ii) J columns reproduce a
    finite-difference directional derivative of flow_forward, (iii) batch sharding is exact:
    per-sample results do not depend on which shard a sample is in."""
    g = load_golden("mnist_gij_d64")
    density = build_product(g, "cuda")
    head, _ = find_head(strip_wrappers(density))
    torch.manual_seed(0)
    z = torch.randn(50, 64, device="cuda") * 0.5
    with torch.no_grad():
        logdet, recon, gram, l1 = head._exact_log_det_jac_and_reconstruction(z, "offdiag")
        assert torch.equal(gram, gram.transpose(1, 2))
        ref_ld = 2 * torch.log(torch.diagonal(torch.linalg.cholesky(gram.double()), dim1=-2, dim2=-1)).sum(1, keepdim=True)
        assert_close(logdet, ref_ld, 1e-5, atol=1e-4, what="logdet vs fp64 cholesky of the same G")
        off = gram.abs().sum((1, 2)) - torch.diagonal(gram, dim1=-2, dim2=-1).abs().sum(1)
        assert_close(l1.flatten(), off, 1e-5, what="l1")
        # (ii) directional derivative
        v = torch.randn_like(z)
        x0, jv = head.jvp_forward(z, v)
        assert_close(x0, recon, 1e-6, what="primal of 1-column program == primal of 64-column program")
        h = 1e-3
        fd = (head.flow_forward((z.double() + h * v.double()).float()) - head.flow_forward((z.double() - h * v.double()).float())) / (2 * h)
        assert_close(jv, fd, 2e-2, what="J v vs central difference (fp32 FD noise)")
        # (iii) shard independence
        lo, _, g_lo, _ = head._exact_log_det_jac_and_reconstruction(z[:25].contiguous(), "offdiag")
        hi, _, g_hi, _ = head._exact_log_det_jac_and_reconstruction(z[25:].contiguous(), "offdiag")
        # The 64 -> 64 convolutions, the coupling, Gram and Cholesky kernels have a fixed accumulation
        # order per output element, independent of the batch; what is left of the conditioners on library
        # calls (first convolution, 1x1 tail) may pick another algorithm for another batch size.
        record("mnist_gij_d64 B=50", "sharded G vs unsharded", float((torch.cat((g_lo, g_hi)) - gram).abs().max() / gram.abs().max()), 1e-5)
        assert_close(torch.cat((g_lo, g_hi)), gram, 1e-5, what="sharded G")
        assert_close(torch.cat((lo, hi)), logdet, 1e-5, atol=1e-4, what="sharded logdet")


def _oracle_on_device(g, filled, x, epoch, chunk, dtype=torch.float64, **kw):
    """The oracle (plain torch) on the GPU in ``dtype``, on the product's weights and inputs, ``chunk``
    samples at a time (its one-column-at-a-time graph keeps every conditioner activation alive).  The
    loss is a batch mean of per-sample terms (`non_square_helpers.py:120`), so chunk losses weighted
    by chunk / B add up to the batch loss and their gradients accumulate into the batch gradient.
    Returns (flow, params with .grad, loss, per-sample elbo (B,1))."""
    params = {(k[7:] if k.startswith("module.") else k): (v.to(dtype) if v.is_floating_point() else v).cuda()
              for k, v in filled.items()}
    params = {k: (v.requires_grad_(True) if v.is_floating_point() else v) for k, v in params.items()}
    flow = O.OracleFlow(g["schema"], params, tuple(x.shape[1:]))
    B = x.shape[0]
    total, per_sample = 0.0, []
    for i in range(0, B, chunk):
        xc = x[i:i + chunk].to(dtype).cuda()
        ckw = {k: v[i:i + chunk].to(dtype).cuda() for k, v in kw.items()}
        elbo = O.elbo_per_sample(flow, g["config"], xc, epoch, **ckw)
        part = -elbo.sum() / B
        part.backward()
        total += float(part.detach())
        per_sample.append(elbo.detach())
    return flow, params, total, torch.cat(per_sample)


def _grad_deviation(grads, params):
    """max over parameters of |g - g_ref| (max-norm) relative to the reference gradient's 2-norm, and
    of the 2-norms themselves.  ``grads``: {key: tensor}, ``params``: {key: leaf with fp64 .grad}"""
    worst_entry = worst_norm = 0.0
    for key, got in grads.items():
        ref = params[key].grad
        if ref is None:
            assert got is None or float(got.abs().max()) == 0, key
            continue
        n = float(ref.norm())
        worst_entry = max(worst_entry, float((got.double() - ref.double()).abs().max()) / max(n, 1e-30))
        worst_norm = max(worst_norm, abs(float(got.double().norm()) - n) / max(n, 1e-30))
    return worst_norm, worst_entry


def _grad_l2_errors(grads, params):
    """per-parameter relative L2 error ||g - g_ref|| / ||g_ref||, and the same over the whole gradient"""
    per, num, den = {}, 0.0, 0.0
    for key, got in grads.items():
        ref = params[key].grad
        if ref is None:
            continue
        diff = float((got.double() - ref.double()).norm())
        per[key] = diff / max(float(ref.norm()), 1e-30)
        num += diff ** 2
        den += float(ref.double().norm()) ** 2
    return per, (num / den) ** 0.5


def _named_grads(density):
    return {(n[7:] if n.startswith("module.") else n): p.grad for n, p in density.named_parameters()}


def test_mnist_shape_benchmarked_batch_matches_fp64_oracle_on_device():
    """The benchmarked configuration itself -- MNIST-shape 1x28x28 -> 64, 64 hidden channels, B = 50,
    so every tcgen05 convolution launch runs on N = 50 (primal) / 3200 (tangent) images and the
    persistent CTAs walk ~66 tile pairs each -- against the oracle evaluated in fp64 on the same device.
    The oracle costs ~4.5 s per sample in fp64 (it pushes one tangent column at a time), so the loss
    whose gradient is compared is the mean elbo of 8 samples spread over the batch (first, last and
    positions straddling tile / chunk boundaries); the product still pushes all 50 samples through
    every launch, and G / log det / penalty / per-sample elbo are compared on those 8 as well.
    The same oracle in fp32 on the device (TF32 off) is the like-for-like reference run: its own
    deviation from fp64 on these inputs is what the product's deviation is held against."""
    g = load_golden("mnist_gij_d64")
    case = dict(CASES["mnist_gij_d64"], batch=50)
    epoch = case["epoch"]
    density = build_product(g, "cpu")
    filled = fill_state_dict(density.state_dict())
    density = density.cuda().train()
    x = make_input(case)
    idx = torch.tensor([0, 7, 13, 24, 25, 31, 42, 49])
    top = strip_wrappers(density)
    elbo = top.elbo(x.cuda(), **O.elbo_kwargs(g["config"], epoch))["elbo"]           # (50, 1), all samples
    loss = -elbo[idx.cuda()].mean()
    loss.backward()
    flow, params, ref_loss, ref_elbo = _oracle_on_device(g, filled, x[idx], epoch, chunk=4)
    _, params32, ref_loss32, _ = _oracle_on_device(g, filled, x[idx], epoch, chunk=8, dtype=torch.float32)
    dev_loss = abs(float(loss.detach()) - ref_loss) / abs(ref_loss)
    dev_elbo = float((elbo.detach()[idx.cuda()].double() - ref_elbo).abs().max() / ref_elbo.abs().max())
    grads32 = {k: v.grad for k, v in params32.items() if v.is_floating_point()}
    worst_norm, _ = _grad_deviation(_named_grads(density), params)
    ref32_norm, _ = _grad_deviation(grads32, params)
    per, whole = _grad_l2_errors(_named_grads(density), params)
    per32, whole32 = _grad_l2_errors(grads32, params)
    median = sorted(per.values())[len(per) // 2]
    worst_key = max(per, key=per.get)
    # intermediates
    head, depth = find_head(top)
    with torch.no_grad():
        xin, z64, _ = flow.encode(x[idx].double().cuda())
        logdet64, _, gram64, attempts = flow.exact_log_det(z64)
        zall = head("extract-latent", _preprocess(top, depth, x.cuda()))
        ld, _, gram, l1 = head._exact_log_det_jac_and_reconstruction(zall, "offdiag")
    ld, gram, l1 = ld[idx.cuda()], gram[idx.cuda()], l1[idx.cuda()]
    off = ~torch.eye(64, dtype=torch.bool, device="cuda")
    l164 = gram64.masked_select(off).view(len(idx), -1).abs().sum(1, keepdim=True)
    dev_G = float((gram.double() - gram64).abs().max() / gram64.abs().max())
    dev_ld = float((ld.double() - logdet64).abs().max() / logdet64.abs().max())
    dev_l1 = float((l1.double() - l164).abs().max() / l164.abs().max())
    name = "mnist_gij_d64 B=50"
    record(name, "loss (8 samples of the batch) vs fp64 oracle on device", dev_loss, 1e-4)
    record(name, "per-sample elbo vs fp64 oracle", dev_elbo, 1e-4)
    record(name, "G vs fp64 oracle", dev_G, 1e-4)
    record(name, "log det vs fp64 oracle", dev_ld, 1e-4)
    record(name, "l1 penalty vs fp64 oracle", dev_l1, 1e-4)
    record(name, f"grad norms vs fp64 oracle (fp32 oracle on device: {ref32_norm:.2e})", worst_norm, 1e-4 + ref32_norm)
    record(name, f"whole gradient rel-L2 vs fp64 oracle (fp32 oracle on device: {whole32:.2e})", whole, 1e-4 + whole32)
    record(name, f"per-parameter rel-L2, median of {len(per)} (fp32 oracle: {sorted(per32.values())[len(per32) // 2]:.2e})", median, 1e-4)
    record(name, f"per-parameter rel-L2, worst (fp32 oracle's worst: {max(per32.values()):.2e}) ...{worst_key[-44:]}", per[worst_key], 5e-3)
    assert attempts == 1 and head.last_cholesky_attempts == 1
    assert max(dev_loss, dev_elbo, dev_G, dev_ld, dev_l1) <= 1e-4, (dev_loss, dev_elbo, dev_G, dev_ld, dev_l1)
    assert worst_norm <= 1e-4 + ref32_norm and whole <= 1e-4 + whole32 and median <= 1e-4, (worst_norm, whole, median)
    # Single parameters can be off by ~1e-3 in ANY fp32 evaluation (the fp32 oracle shows the same on
    # its own worst parameter): a primal pre-activation within rounding of zero flips its ReLU mask
    # against fp64, which changes the gradient of everything upstream in that conditioner -- measured
    # with scripts/parity_b50.py: the outliers are all parameters of ONE coupler up to one block.
    assert per[worst_key] <= 5e-3, (worst_key, per[worst_key])


def _preprocess(top, depth, x):
    """push x through the elementwise nodes above the head (scalar-mult / scalar-add / logit)"""
    node = top
    for _ in range(depth):
        x = node.bijection.x_to_z(x)["z"]
        node = node.prior
    return x


def test_cholesky_jitter_retry_end_to_end_on_device(capsys):
    """`non_square.py:280-291` through the real kernels: a batch whose metric is singular for one
    sample fails the device factorisation (info word), the host adds eps*I to the WHOLE batch, retries
    with eps x 10 and prints the reference's warning."""
    g = load_golden("hepmass_gij_d10_warm")
    density = build_product(g, "cuda")
    head, _ = find_head(strip_wrappers(density))
    real = head._get_full_jac_transpose_jac

    def singular(latent, create_graph=True):
        gram, recon = real(latent, create_graph)
        gram = gram.clone()
        gram[3] = 0.0                        # rank-0 metric for one sample of the batch
        return gram, recon
    head._get_full_jac_transpose_jac = singular
    z = g["z"].cuda()
    logdet, _, gram, l1 = head._exact_log_det_jac_and_reconstruction(z, "offdiag")
    assert head.last_cholesky_attempts >= 2
    assert f"WARNING: Numerical non-invertibility in JtJ observed - {head.last_cholesky_attempts} attempts needed to fix" in capsys.readouterr().out
    assert torch.isfinite(logdet).all()
    # what came back is the jittered metric (the reference returns it too, `:296`), and log det is its log det
    jitter = sum(1e-6 * 10 ** k for k in range(head.last_cholesky_attempts - 1))
    assert_close(torch.diagonal(gram[3]), torch.full((10,), jitter), 1e-5, what="jittered diagonal")
    ref = torch.linalg.slogdet(gram.double())[1].unsqueeze(1)
    assert_close(logdet, ref, 1e-4, atol=1e-3, what="log det of the jittered metric")
    # the same loop on the CPU oracle takes the same number of attempts
    flow_gram = gram.double().cpu() - jitter * torch.eye(10, dtype=torch.float64)
    eps, attempts = 1e-6, 1
    while True:
        if torch.linalg.cholesky_ex(flow_gram.float())[1].eq(0).all():
            break
        flow_gram = flow_gram + eps * torch.eye(10, dtype=torch.float64)
        attempts += 1
        eps *= 10
    assert attempts == head.last_cholesky_attempts


@pytest.mark.parametrize("name,batch,m", [("hepmass_hutch_d10", 9, 1), ("hepmass_hutch_m2_plain", 5, 2), ("img8_hutch_d8", 3, 1)])
def test_hutch_with_cg_gradients_match_fp64_oracle(name, batch, m):
    """hutch_with_cg in TRAINING: the estimator is a gradient surrogate (`non_square.py:249-256`), so
    what matters is d(loss)/d(parameters) through stopgrad((J^T J)^-1 eps) . (J^T J eps) -- a VJP that
    is differentiated again.  Fresh weights and inputs, probes injected on both sides, fp64 oracle."""
    from cases import hutch_eps
    g = load_golden(name)
    case = dict(CASES[name], batch=batch)
    density = build_product(g, "cpu")
    filled = fill_state_dict(density.state_dict(), salt=3, gain=1.3)
    density.load_state_dict(filled)
    density = density.cuda().train()
    x = make_input(case) * 0.95
    top = strip_wrappers(density)
    head, _ = find_head(top)
    d = g["config"]["latent_dimension"]
    eps = hutch_eps(name + "/fresh", (batch, d, m))
    head._draw_hutchinson = lambda shape, dev, dtype: eps.to(device=dev, dtype=dtype)
    from cmf_b200.non_square_helpers import get_non_square_train_metrics
    loss = get_non_square_train_metrics(g["config"])[0](top, x.cuda(), case["epoch"])["loss"]
    loss.backward()
    _, params, ref_loss, _ = _oracle_on_device(g, filled, x, case["epoch"], chunk=batch, hutchinson_eps=eps)
    worst_norm, worst_entry = _grad_deviation(_named_grads(density), params)
    loss = loss.detach()
    record(name, "hutch_with_cg training loss vs fp64 oracle", abs(float(loss) - ref_loss) / abs(ref_loss), 1e-4)
    record(name, "hutch_with_cg grad norms vs fp64 oracle", worst_norm, 2e-3)
    record(name, "hutch_with_cg grad entries vs fp64 oracle (of the norm)", worst_entry, 2e-3)
    assert abs(float(loss) - ref_loss) <= 1e-4 * abs(ref_loss)
    # the fp32 CG weights are reproducible to ~1e-3 only (tests/test_oracle_golden.py: gpytorch's own
    # linear_cg and its restatement drift apart ~30x per iteration in fp32), and they scale this gradient
    assert worst_norm <= 2e-3 and worst_entry <= 2e-3, (worst_norm, worst_entry)
    # the log-det surrogate must actually reach the parameters of the injective flow
    flow_params = [p for n_, p in density.named_parameters() if p.grad is not None]
    assert len(flow_params) == len([k for k, v in params.items() if v.is_floating_point() and v.grad is not None])


@pytest.mark.parametrize("name", ["power_nsfc_gij_d4", "hepmass_nsfc_gij_d10"])
def test_nsfc_model_spline_bins_bit_exact_and_matches_fp64_oracle(name):
    """``nsf-c`` layers on the JVP stack, model level: (1) every bin index the spline
    kernel selects inside the model equals the reference's ``searchsorted`` over the very knots it was
    given (`gitmodules/nsf/utils/torchutils.py:133-138`); (2) G, log det, penalty against the fp64
    oracle (autograd JVP over the restated nsf transform) on fresh weights and inputs."""
    import rqs_oracle as R
    from cmf_b200 import ops
    g = load_golden(name)
    case = dict(CASES[name], batch=37)
    density = build_product(g, "cpu")
    filled = fill_state_dict(density.state_dict(), salt=4, gain=2.0)
    density.load_state_dict(filled)
    density = density.cuda().train()
    x = make_input(case) * 1.7                      # some features beyond the tail bound after the flow
    head, depth = find_head(strip_wrappers(density))
    calls, real = [], ops.rqs_coupling_jvp

    def spy(z_m, V_m, cw, ch, dv, CW, CH, DV, tb):
        out = real(z_m, V_m, cw, ch, dv, CW, CH, DV, tb)
        calls.append((z_m.detach().cpu(), ch.detach().cpu(), out[2].cpu(), tb))
        return out
    ops.rqs_coupling_jvp = spy
    try:
        with torch.no_grad():
            z = head("extract-latent", x.cuda())
            logdet, recon, gram, l1 = head._exact_log_det_jac_and_reconstruction(z, "offdiag")
    finally:
        ops.rqs_coupling_jvp = real
    n_layers = sum(1 for l in g["schema"] if l["type"] == "nsf-c")
    assert len(calls) == 2 * n_layers               # identity half + transformed half per layer
    outside = 0
    for z_m, ch, bins, tb in calls:
        inside = (z_m >= -tb) & (z_m <= tb)
        ref = torch.where(inside, R.searchsorted(ch, z_m), torch.full_like(bins, -1, dtype=torch.long))
        assert torch.equal(bins.long(), ref), "spline bin indices must be bit-exact"
        outside += int((~inside).sum())
    params = {(k[7:] if k.startswith("module.") else k): (v.double() if v.is_floating_point() else v) for k, v in filled.items()}
    flow = O.OracleFlow(g["schema"], params, case["x_shape"])
    with torch.no_grad():
        _, z64, _ = flow.encode(x.double())
        ld64, _, gram64, _ = flow.exact_log_det(z64)
    assert_close(z, z64, RTOL, atol=1e-6, what="latent")
    assert_close(gram, gram64, RTOL, what="G")
    assert_close(logdet, ld64, RTOL, atol=1e-4, what="log det")
    record(name, "nsf-c model G vs fp64 oracle", float((gram.double().cpu() - gram64).abs().max() / gram64.abs().max()), RTOL)


def test_hutch_with_cg_matches_oracle():
    """log_jacobian_method=hutch_with_cg with injected probes (the image-config default)."""
    g = load_golden("hepmass_gij_d10_warm")
    density = build_product(g, "cuda")
    head, _ = find_head(strip_wrappers(density))
    head.log_jacobian_method = "hutch_with_cg"
    head.num_hutchinson_samples = 2
    torch.manual_seed(3)
    z = (torch.randn(6, 10) * 0.3)
    eps = torch.randn(6, 10, 2)
    params = {(k[7:] if k.startswith("module.") else k): (v.double() if v.is_floating_point() else v) for k, v in fill_state_dict(density.state_dict()).items()}
    flow = O.OracleFlow(g["schema"], {k: v.cpu() for k, v in params.items()}, (21,))
    flow.head_cfg["max_cg_iterations"] = head.max_cg_iterations
    est_ref, recon_ref, jtj_ref = flow.approx_log_det(z.double(), eps.double(), training=False)
    with torch.no_grad():
        head.train()
        est, recon, jtj = head._approx_log_det_jac_and_reconstruction(z.cuda(), eps.cuda())
    assert_close(jtj, jtj_ref, RTOL, what="J^T J eps")
    assert_close(est, est_ref, 5e-4, what="hutchinson estimate")
    assert_close(recon, recon_ref, RTOL, what="reconstruction")


def test_checkpoint_roundtrip_keeps_reference_keys(tmp_path):
    g = load_golden("img8_gij_d8")
    density = build_product(g, "cuda")
    keys = list(density.state_dict())
    assert all(k.startswith("module.density.") for k in keys)          # DataParallel + dequantiser prefixes
    assert "module.density.prior.prior.prior.prior.bijection.mask" in keys
    tail_keys = [k for k in keys if k.endswith("inverse_permutation")]
    assert len(tail_keys) == 1
    torch.save(density.state_dict(), tmp_path / "ckpt.pt")
    fresh = build_product(g, "cpu")
    fresh.load_state_dict(torch.load(tmp_path / "ckpt.pt"))
    loss, grads = run_loss(g, fresh.cuda(), "cuda")
    check_against_golden(g, loss, grads, RTOL)


@pytest.mark.parametrize("name", ["sphere_gij_d3", "hepmass_gij_d10_warm", "img8_gij_d8"])
def test_cuda_graph_step_is_bit_identical_to_eager(name):
    """cmf_b200.graphed.GraphedTrainStep replays the same kernels: loss and every gradient must be
    bitwise equal to the eager step, on the captured batch and on a second batch."""
    from cmf_b200.graphed import GraphedTrainStep
    from cmf_b200.non_square_helpers import get_non_square_train_metrics
    from cmf_b200.parallel import GradientBucket
    g = load_golden(name)
    density = strip_wrappers(build_product(g, "cuda")).train()
    bucket = GradientBucket(density.parameters())
    tm = get_non_square_train_metrics(g["config"])[0]
    epoch = g["case"]["epoch"]
    step = GraphedTrainStep(density, tm, bucket)
    x0 = make_input(CASES[name]).cuda()
    for x in (x0, (x0 * 0.97).contiguous()):
        bucket.zero()
        loss_e = tm(density, x, epoch)["loss"]
        loss_e.backward()
        ref, loss_e = bucket.flat.clone(), loss_e.detach().clone()   # no eager autograd graph may stay alive:
        # its AccumulateGrad nodes are bound to the legacy stream and would be reused inside the capture
        loss_g = step(x, epoch).clone()
        if name.startswith("img"):      # cuDNN may pick another conv algorithm on the capture stream
            assert_close(loss_g, loss_e, 1e-6, what="loss")
            assert_close(bucket.flat, ref, 1e-5, what="grads")
        else:
            assert torch.equal(loss_g, loss_e)
            assert torch.equal(bucket.flat, ref)
    assert step.eager_fallbacks == 0 and len(step._graphs) == 1


def test_hutch_cg_product_graph_replay_matches_eager():
    """The CG loop of hutch_with_cg replays its J^T J product as a CUDA graph: same kernels as eager launches, same
    loss and parameter gradients; and the graph survives an in-place parameter update."""
    import copy
    import cmf_b200.head as H
    g = load_golden("img8_hutch_d8")
    results = []
    for use_graph in (True, False):
        H._CG_GRAPH, H._CG_EXPLICIT = use_graph, "0"          # the implicit products (the explicit-G path has its own test)
        try:
            density = build_product(g, "cuda")
            head, _ = find_head(strip_wrappers(density))
            loss, grads = run_loss(g, density, "cuda")
            if use_graph:
                assert head.cg_graph_replays > 0, "the CG product was not replayed from a graph"
                with torch.no_grad():
                    for p in density.parameters():
                        p.mul_(1.01)
                n0 = head.cg_graph_replays
                loss_b, _ = run_loss(g, density, "cuda")
                assert head.cg_graph_replays > n0 and torch.isfinite(loss_b)
                copy.deepcopy(density)                      # the captured graph must not get in the way
            results.append((loss.detach().clone(), {k: v.clone() for k, v in grads.items()}))
        finally:
            H._CG_GRAPH, H._CG_EXPLICIT = True, "auto"
    (l1, g1), (l0, g0) = results
    # same kernels either way; the fixture's 8-channel conditioners run on cuDNN, whose data-gradient algorithms are
    # not bitwise reproducible run to run, and CG amplifies that: 1e-5, not torch.equal
    assert abs(float(l1) - float(l0)) <= 1e-5 * abs(float(l0))
    assert set(g1) == set(g0)
    for k in g0:                                             # (library weight gradients are not bitwise reproducible run to run)
        assert_close(g1[k], g0[k], rtol=1e-4, atol=1e-7, what=k)


@pytest.mark.parametrize("name", ["hepmass_hutch_d10", "img8_hutch_d8"])
def test_hutch_cg_on_explicit_gram_matches_implicit_products(name):
    """For small latents the CG loop runs on G = J^T J built by one d-column pass instead of ~12 network passes
    (head._CG_EXPLICIT): same operator, so the same loss and gradients as the implicit products, and the same
    reference fixture within the contract tolerance."""
    import cmf_b200.head as H
    g = load_golden(name)
    results = []
    for mode in ("1", "0"):
        H._CG_EXPLICIT = mode
        try:
            density = build_product(g, "cuda")
            loss, grads = run_loss(g, density, "cuda")
            if mode == "1":
                check_against_golden(g, loss, grads, rtol=1e-4)
            results.append((loss, grads))
        finally:
            H._CG_EXPLICIT = "auto"
    (l1, g1), (l0, g0) = results
    assert abs(float(l1) - float(l0)) <= 1e-5 * abs(float(l0))
    for k in g0:
        assert_close(g1[k], g0[k], rtol=1e-4, atol=1e-7, what=k)
