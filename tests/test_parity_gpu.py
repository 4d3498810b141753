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
from helpers import build_product, run_loss, check_against_golden, strip_wrappers, find_head, assert_close
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


@pytest.mark.parametrize("name,batch", [("sphere_gij_d3", 257), ("hepmass_gij_d10_warm", 33), ("img8_gij_d8", 5)])
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
        # not bitwise: cuDNN picks its conv algorithm per batch size (B*d = 1600 vs 3200 images), and
        # fp32 rounding of different algorithms differs at the 1e-4 level in G for this ill-conditioned flow
        assert_close(torch.cat((g_lo, g_hi)), gram, 1e-3, what="sharded G")
        assert_close(torch.cat((lo, hi)), logdet, 1e-3, atol=1e-3, what="sharded logdet")


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
