"""The ``--model non-square`` config surface (same keys, defaults and schema format as the
reference's `config/two_d.py:263-309`, `config/tabular.py:280-354`, `config/images.py:120-178`,
`config/config.py:39-81`, `config/schemas.py:1-105,381-531`), restricted to what that model needs.

``get_config(dataset, "non-square")`` returns the merged dict; ``get_schema(config)`` returns the
list of layer dicts the factory consumes.  Unknown keys are accepted and ignored exactly like the
reference's ``--config k=v`` overlay (`main.py:32-45,79`) -- in particular ``g_ij_global_loss`` and
``hutchinson_samples`` under ``cholesky`` are silent no-ops there too.
"""
from __future__ import annotations

from typing import Dict, List

TWO_D = ["sphere", "sinusoid-1-6", "sinusoid-1-3", "hemisphere-2-6", "torus", "swissroll", "moebius",
         "hyperboloid", "river", "null6d", "s2inr6", "s4inr6", "randomized-s2inr6", "pure-line",
         "2uniforms", "8gaussians", "checkerboard", "2spirals", "rings", "annulus"]
TABULAR = {"power": (2, 5000), "gas": (2, 2500), "hepmass": (10, 750), "miniboone": (21, 400), "bsds300": (30, 250)}
IMAGES = ["mnist", "fashion-mnist", "cifar10", "svhn", "celeba"]


def config_group(dataset: str) -> str:
    if dataset in TABULAR:
        return "tabular"
    if dataset in IMAGES:
        return "images"
    return "2d"   # the reference registers 50+ simulated manifolds here (config/two_d.py:4-60)


_HEAD_COMMON = {
    "non_square": True, "m_flow": False, "num_u_channels": 0,
    "use_cond_affine": False, "pure_cond_affine": False,
    "g_kk_loss": False, "g_ij_loss": False,
    "elbo_regularization_param": 1, "metric_regularization_param": 1,
    "num_valid_elbo_samples": 1, "num_test_elbo_samples": 1,
    "underlying_flow": "realnvp", "opt": "adam", "lr_schedule": "none", "weight_decay": 0.0,
    "max_grad_norm": None, "early_stopping": True,
}


def get_config(dataset: str, model: str = "non-square", use_baseline: bool = False) -> Dict:
    if model != "non-square":
        raise ValueError("cmf_b200 implements the --model non-square path only")
    if use_baseline:
        raise ValueError("--baseline selects the m_flow head, which is outside the fused hot path")
    group = config_group(dataset)
    if group == "2d":
        cfg = {
            "dequantize": False, "batch_norm": False,
            "train_batch_size": 1000, "valid_batch_size": 1000, "test_batch_size": 10000,
            "schema_type": "flat-realnvp", "num_density_layers": 5,
            "coupler_shared_nets": True, "coupler_hidden_channels": [10, 10],
            "lr": 3e-4, "max_epochs": 1000, "epochs_per_test": 50, "max_bad_valid_epochs": 50,
            "regularization_param": 1, "log_jacobian_method": "cholesky", "latent_dimension": 2,
            "likelihood_warmup": False, "likelihood_warmup_start": 500, "likelihood_warmup_end": 1000,
            "prior": "affine",
            "vis_log_prob_min": -3, "vis_log_prob_max": -1,      # plotting only (config/two_d.py); carried for key parity
        }
    elif group == "tabular":
        d, bs = TABULAR[dataset]
        cfg = {
            "dequantize": False, "batch_norm": False, "resnet_batchnorm": False, "ignore_batch_effects": False,
            "train_batch_size": bs, "valid_batch_size": 500, "test_batch_size": 500,
            "schema_type": "flat-realnvp", "num_density_layers": 10,
            "coupler_shared_nets": True, "coupler_hidden_channels": [128] * 4,
            "lr": 1e-4, "max_epochs": 1000, "epochs_per_test": 5, "max_bad_valid_epochs": 20,
            "regularization_param": 50, "log_jacobian_method": "cholesky",
            "hutchinson_distribution": "normal", "hutchinson_samples": 1, "latent_dimension": d,
            "likelihood_warmup": True, "likelihood_warmup_start": 25, "likelihood_warmup_end": 50,
            "prior": "realnvp", "prior_num_density_layers": 5, "prior_hidden_channels": [32] * 2,
            # not on the hot path (batch norm is off, FID is evaluation); carried so the dict equals the reference's
            "prior_batch_norm": False, "batch_norm_use_running_averages": False, "batch_norm_apply_affine": True,
            "smaller_realnvp": False, "use_fid": True, "num_fid_samples": 10000,
        }
    else:
        cfg = {
            "dequantize": True, "batch_norm": False, "resnet_batchnorm": False, "ignore_batch_effects": False,
            "train_batch_size": 50, "valid_batch_size": 50, "test_batch_size": 50,
            "schema_type": "multiscale-realnvp", "g_hidden_channels": [64] * 8, "smaller_realnvp": False,
            "num_density_layers": 10,
            "lr": 1e-4, "max_epochs": 1000, "epochs_per_test": 10, "max_bad_valid_epochs": 20,
            "regularization_param": 50, "log_jacobian_method": "hutch_with_cg",
            "hutchinson_distribution": "normal", "hutchinson_samples": 1, "latent_dimension": 20,
            "likelihood_warmup": True, "likelihood_warmup_start": 25, "likelihood_warmup_end": 50,
            "cg_tolerance": 1,
            "prior": "realnvp", "prior_num_density_layers": 10, "prior_hidden_channels": [32] * 4,
            "logit_tf_lambda": 1e-6 if dataset in ("mnist", "fashion-mnist") else 0.05,
            "logit_tf_scale": 256,
            # not on the hot path; carried so the dict equals the reference's (config/images.py:120-178)
            "prior_batch_norm": False, "batch_norm_use_running_averages": True, "batch_norm_momentum": 0.1,
            "batch_norm_apply_affine": True, "test_metric": False, "use_fid": True, "num_fid_samples": 10000,
            "fid_dims": 2048,
        }
    return {"model": model, "dataset": dataset, **_HEAD_COMMON, **cfg}


# ---------------------------------------------------------------------------------------------
# schema (config/schemas.py)
# ---------------------------------------------------------------------------------------------
def _flat_realnvp(num_layers: int, hidden: List[int]) -> List[dict]:
    """`schemas.py:476-531` with shared tanh MLP couplers; the ``normalise`` entries it interleaves
    are removed again when ``batch_norm`` is False (`schemas.py:19-22`), so they are not emitted."""
    coupler = {"independent_nets": False,
               "shift_log_scale_net": {"type": "mlp", "hidden_channels": list(hidden), "activation": "tanh"}}
    out = [{"type": "flatten"}]
    for i in range(num_layers):
        out.append({"type": "acl", "mask_type": "alternating-channel", "reverse_mask": i % 2 != 0,
                    "coupler": coupler, "num_u_channels": 0})
    return out


def _multiscale_realnvp(hidden: List[int], resnet_batchnorm: bool, ignore_batch_effects: bool,
                        smaller: bool) -> List[dict]:
    """`schemas.py:381-441`."""
    cb = lambda r: {"type": "acl", "mask_type": "checkerboard", "reverse_mask": r}
    sc = lambda r: {"type": "acl", "mask_type": "split-channel", "reverse_mask": r}
    if smaller:
        base = [cb(False), cb(True), {"type": "squeeze", "factor": 2}, sc(False), sc(True),
                {"type": "split", "non_square": True}, cb(False), cb(True)]
    else:
        base = [cb(False), cb(True), cb(False), {"type": "squeeze", "factor": 2}, sc(True), sc(False), sc(True),
                {"type": "split", "non_square": True}, cb(False), cb(True), cb(False), cb(True)]
    out = []
    for layer in base:
        if layer["type"] == "acl":
            layer = {**layer, "num_u_channels": 0,
                     "coupler": {"independent_nets": False,
                                 "shift_log_scale_net": {"type": "resnet", "hidden_channels": list(hidden),
                                                         "batchnorm": resnet_batchnorm,
                                                         "ignore_batch_effects": ignore_batch_effects}}}
        out.append(layer)
    return out


def get_schema(config: Dict) -> List[dict]:
    if not config.get("non_square", False):
        raise ValueError("cmf_b200.get_schema builds non-square schemas only")
    if config.get("batch_norm", False):
        raise ValueError("flow-level batch norm needs cross-sample statistics and is off in every "
                         "non-square config (config/two_d.py:71, tabular.py:309, images.py:127)")
    ty = config["schema_type"]
    if ty == "flat-realnvp":
        if not config.get("coupler_shared_nets", True):
            raise ValueError("independent coupler nets are not used by any non-square config")
        body = _flat_realnvp(config["num_density_layers"], config["coupler_hidden_channels"])
    elif ty == "multiscale-realnvp":
        body = _multiscale_realnvp(config["g_hidden_channels"], config.get("resnet_batchnorm", True),
                                   config.get("ignore_batch_effects", False), config.get("smaller_realnvp", False))
    elif ty == "nsf":
        # config/schemas.py:152-162,586-626 with the `normalise` entries already removed (:22).  Only the
        # coupling variant can carry tangents; the reference itself has no spline _jvp at all.
        if config["autoregressive"] or config.get("use_linear", True):
            raise ValueError("schema_type nsf on the non-square JVP stack needs autoregressive=False and "
                             "use_linear=False (nsf-ar and LU-linear layers have no JVP rule)")
        body = [{"type": "flatten"}]
        for i in range(config["num_density_layers"]):
            body += [{"type": "rand-channel-perm"},
                     {"type": "nsf-c", "num_hidden_channels": config["num_hidden_channels"],
                      "num_hidden_layers": config["num_hidden_layers"], "num_bins": config["num_bins"],
                      "tail_bound": config["tail_bound"], "activation": "relu",
                      "dropout_probability": config["dropout_probability"], "reverse_mask": i % 2 == 0}]
        body.append({"type": "rand-channel-perm"})
    else:
        raise ValueError(f"schema_type {ty} has no JVP path in the reference")

    head = {"type": "non-square-head",
            "regularization_param": config["regularization_param"],
            "log_jacobian_method": config["log_jacobian_method"],
            "hutchinson_distribution": config.get("hutchinson_distribution", "normal"),
            "hutchinson_samples": config.get("hutchinson_samples", 1),
            "m_flow": config["m_flow"],
            "max_cg_iterations": config.get("max_cg_iterations", None),
            "cg_tolerance": config.get("cg_tolerance", 1),
            "latent_dimension": config["latent_dimension"],
            "metric_regularization_param": config["metric_regularization_param"]}
    tail = [{"type": "non-square-base", "latent_dimension": config["latent_dimension"], "m_flow": config["m_flow"]}]
    if config["prior"] == "affine":
        tail.append({"type": "affine", "per_channel": False})
    elif config["prior"] == "realnvp":
        tail += _flat_realnvp(config["prior_num_density_layers"], config["prior_hidden_channels"])
    else:
        raise ValueError(f"prior {config['prior']} is not built here (nsf prior: SURVEY 8f N2)")

    pre = [{"type": "dequantization"}] if config["dequantize"] else []
    if config.get("logit_tf_lambda") is not None and config.get("logit_tf_scale") is not None:
        lam, scale = config["logit_tf_lambda"], config["logit_tf_scale"]
        pre += [{"type": "scalar-mult", "value": (1 - 2 * lam) / scale},
                {"type": "scalar-add", "value": lam}, {"type": "logit"}]
    elif config.get("centering_tf_scale") is not None:
        pre += [{"type": "scalar-mult", "value": 1 / config["centering_tf_scale"]},
                {"type": "scalar-add", "value": -.5}]
    return pre + [head] + body + tail
