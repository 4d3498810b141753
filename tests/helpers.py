"""Shared helpers for the parity tests (product vs golden fixtures vs oracle)."""
import torch

from cases import CASES, make_input, grad_probe
from detfill import fill_state_dict
from cmf_b200.factory import get_density
from cmf_b200.non_square_helpers import get_non_square_train_metrics


def build_product(golden, device="cpu"):
    case = golden["case"]
    density = get_density(golden["schema"], case["x_shape"])
    filled = fill_state_dict(density.state_dict())
    density.load_state_dict(filled)
    density.train()
    return density.to(device)


def strip_wrappers(density):
    """below DataParallelDensity and the dequantiser (its noise is folded into the fixture)"""
    top = density.module if hasattr(density, "module") else density
    if type(top).__name__ == "DequantizationDensity":
        top = top.density
    return top


def find_head(top):
    node, depth = top, 0
    while type(node).__name__ != "FusedNonSquareHeadDensity":
        node = node.prior
        depth += 1
    return node, depth


def run_loss(golden, density, device="cpu"):
    name = golden["name"]
    x = make_input(CASES[name]).to(device)
    train_metrics = get_non_square_train_metrics(golden["config"])[0]
    loss = train_metrics(strip_wrappers(density), x, golden["case"]["epoch"])["loss"]
    loss.backward()
    grads = {}
    for pname, p in density.named_parameters():
        if p.grad is not None:
            key = pname[len("module."):] if pname.startswith("module.") else pname
            grads[key] = p.grad.detach().cpu()
    return loss.detach().cpu(), grads


def assert_close(a, b, rtol, atol=0.0, what=""):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    scale = max(float(b.abs().max()), 1e-30)
    err = float((a - b).abs().max())
    assert err <= rtol * scale + atol, f"{what}: max err {err:.3e} vs scale {scale:.3e} (rtol {rtol})"


def reference_rounding_noise(golden):
    """Worst relative deviation of the reference's own fp32 gradients from an fp64 evaluation of
    the same algorithm (oracle/add_fp64_probes.py).  ~1e-7 for the well-conditioned flat configs,
    1.6e-4 for MNIST-shape d=64 where cond(G) ~ 5e3..8e3 amplifies fp32 rounding in G^{-1}."""
    if "grads_fp64" not in golden:
        return 0.0
    worst = 0.0
    for k, p64 in golden["grads_fp64"].items():
        if k in golden["grads"]:
            worst = max(worst, abs(golden["grads"][k]["norm"] - p64["norm"]) / max(p64["norm"], 1e-30))
    return worst


def check_against_golden(golden, loss, grads, rtol):
    """loss and every parameter gradient within ``rtol`` relative of the reference's -- or within
    3x the reference's own fp32 rounding error against fp64 when that is larger (an fp32 result
    cannot be asked to agree with another fp32 result more closely than either agrees with the
    exact value)."""
    rtol = max(rtol, 3.0 * reference_rounding_noise(golden))
    assert_close(loss, golden["loss"], rtol, what="loss")
    assert set(golden["grads"]) == set(grads), "parameters receiving gradients differ from the reference"
    for k, probe in golden["grads"].items():
        mine = grad_probe(k, grads[k])
        tol = rtol * max(probe["norm"], 1e-12)
        assert abs(mine["norm"] - probe["norm"]) <= tol, f"{k}: norm {mine['norm']} vs {probe['norm']}"
        assert abs(mine["dot"] - probe["dot"]) <= 8 * tol, f"{k}: dot {mine['dot']} vs {probe['dot']}"
        assert_close(mine["head"], probe["head"], rtol, atol=tol, what=k)
