"""Parity cases shared by ``oracle/gen_golden.py`` (drives the real reference, build container
only), ``tests/test_oracle_golden.py`` (oracle vs golden, CPU) and ``tests/test_parity_gpu.py``
(CUDA path vs golden + oracle, GPU box).  TEST INFRASTRUCTURE ONLY.

Each case names a reference dataset key (which selects the config group, `config/two_d.py`,
`config/tabular.py`, `config/images.py`), the data shape actually fed (the reference builds its
model from ``x_train.shape[1:]``, `factory.py:55-56`), ``--config`` style overrides, the batch
size and the epoch passed to the loss closure (`non_square_helpers.py:99-123`).
"""
from __future__ import annotations

import math

import torch

from detfill import det_tensor

COMMON = dict(log_jacobian_method="cholesky")

CASES = {
    # BASELINE.json configs[0]: sphere, cholesky, g_ij_loss, d=3
    "sphere_gij_d3": dict(dataset="sphere", x_shape=(3,), batch=64, epoch=0, data="sphere",
                          overrides=dict(g_ij_loss=True, latent_dimension=3, **COMMON)),
    # same geometry, d=2 (true manifold dimension) and the diagonal penalty variant
    "sphere_gkk_d2": dict(dataset="sphere", x_shape=(3,), batch=48, epoch=0, data="sphere",
                          overrides=dict(g_kk_loss=True, latent_dimension=2, **COMMON)),
    # BASELINE.json configs[1]: sinusoid-1-6, d=6, hutchinson_samples=1 and g_ij_global_loss (both no-ops)
    "sinusoid_gij_d6": dict(dataset="sinusoid-1-6", x_shape=(6,), batch=64, epoch=0, data="uniform",
                            overrides=dict(g_ij_loss=True, latent_dimension=6, hutchinson_samples=1,
                                           g_ij_global_loss=True, **COMMON)),
    # plain ELBO (no metric term), odd latent dim
    "sinusoid_plain_d3": dict(dataset="sinusoid-1-6", x_shape=(6,), batch=32, epoch=0, data="uniform",
                              overrides=dict(latent_dimension=3, **COMMON)),
    # BASELINE.json configs[2]: power-shape D=6 (d=10 is not runnable in the reference: SURVEY D5)
    "power_gij_d2": dict(dataset="power", x_shape=(6,), batch=32, epoch=60, data="uniform",
                         overrides=dict(g_ij_loss=True, latent_dimension=2, **COMMON)),
    "power_gij_d6": dict(dataset="power", x_shape=(6,), batch=24, epoch=60, data="uniform",
                         overrides=dict(g_ij_loss=True, latent_dimension=6, **COMMON)),
    # warm-up ramp: epoch 40 of [25, 50] -> likelihood weight 0.6, metric weight 0.6*5
    "hepmass_gij_d10_warm": dict(dataset="hepmass", x_shape=(21,), batch=16, epoch=40, data="uniform",
                                 overrides=dict(g_ij_loss=True, latent_dimension=10,
                                                metric_regularization_param=5, **COMMON)),
    # warm-up epoch 0: likelihood weight 0 -> reconstruction only, no log-det (non_square.py:105-109)
    "hepmass_recon_only": dict(dataset="hepmass", x_shape=(21,), batch=16, epoch=0, data="uniform",
                               overrides=dict(g_ij_loss=True, latent_dimension=10, **COMMON)),
    # image architecture at reduced size: 1x8x8, two residual blocks of 8 channels, d=8
    "img8_gij_d8": dict(dataset="mnist", x_shape=(1, 8, 8), batch=6, epoch=60, data="pixels",
                        overrides=dict(g_ij_loss=True, latent_dimension=8, g_hidden_channels=[8, 8],
                                       prior_num_density_layers=2, **COMMON)),
    # 3-channel image architecture (CIFAR-like), 3x8x8, d=16
    "img8c3_gij_d16": dict(dataset="cifar10", x_shape=(3, 8, 8), batch=4, epoch=60, data="pixels",
                           overrides=dict(g_ij_loss=True, latent_dimension=16, g_hidden_channels=[8, 8],
                                          prior_num_density_layers=2, **COMMON)),
    # image architecture with the SHIPPED hidden width of 64 (config/images.py:137) at reduced size 1x12x12:
    # a second reference-pinned case (besides the slow full MNIST shape) that reaches the tcgen05
    # conditioner convolutions (64 -> 64 channels) -- 12x12 checkerboard layers, 6x6 split-channel layers
    "img12_c64_gij_d16": dict(dataset="mnist", x_shape=(1, 12, 12), batch=3, epoch=60, data="pixels",
                              overrides=dict(g_ij_loss=True, latent_dimension=16, g_hidden_channels=[64, 64],
                                             prior_num_density_layers=2, **COMMON)),
    # the same for the 3-channel (CIFAR-shape) architecture: first convolutions with 3 / 6 / 12 input channels, conditioner
    # tails with 6 / 12 / 24 outputs, 64 -> 64 convolutions at 8x8 / 4x4 / 2x2
    "img8c3_c64_gij_d16": dict(dataset="cifar10", x_shape=(3, 8, 8), batch=3, epoch=60, data="pixels",
                               overrides=dict(g_ij_loss=True, latent_dimension=16, g_hidden_channels=[64, 64],
                                              prior_num_density_layers=2, **COMMON)),
    # log_jacobian_method=hutch_with_cg (the image-config default, config/images.py:148) with the Hutchinson
    # probes injected (``hutch=True``: gen_golden.py replaces torch.randn inside the reference's draw,
    # non_square.py:206, by the deterministic tensor of ``hutch_eps``).  With g_ij_loss the reference
    # takes the penalty over a broadcast of J^T J eps (non_square.py:98-99,258) -- reproduced as is.
    "hepmass_hutch_d10": dict(dataset="hepmass", x_shape=(21,), batch=12, epoch=60, data="uniform", hutch=True,
                              overrides=dict(g_ij_loss=True, latent_dimension=10, log_jacobian_method="hutch_with_cg",
                                             hutchinson_samples=1)),
    # two probes per sample, plain ELBO (the penalty's masked_select cannot broadcast for m = 2)
    "hepmass_hutch_m2_plain": dict(dataset="hepmass", x_shape=(21,), batch=8, epoch=60, data="uniform", hutch=True,
                                   overrides=dict(latent_dimension=10, log_jacobian_method="hutch_with_cg",
                                                  hutchinson_samples=2)),
    # image architecture, method left at the image default (hutch_with_cg)
    "img8_hutch_d8": dict(dataset="mnist", x_shape=(1, 8, 8), batch=4, epoch=60, data="pixels", hutch=True,
                          overrides=dict(g_ij_loss=True, latent_dimension=8, g_hidden_channels=[8, 8],
                                         prior_num_density_layers=2, log_jacobian_method="hutch_with_cg")),
    # the image configs' default method at the shipped hidden width: the m-column JVP and the explicit G of the CG loop run
    # on the tcgen05 / first-conv / tail kernels, the doubly differentiated VJP on library convolutions
    "img12_c64_hutch_d16": dict(dataset="mnist", x_shape=(1, 12, 12), batch=3, epoch=60, data="pixels", hutch=True,
                                overrides=dict(g_ij_loss=True, latent_dimension=16, g_hidden_channels=[64, 64],
                                               prior_num_density_layers=2, log_jacobian_method="hutch_with_cg")),
    # north_star's "rational-quadratic-spline coupling" layers ON the JVP stack (nsf-c):
    # schema_type nsf with the coupling variant and no LU-linear layers.  The reference has no spline
    # _jvp; oracle/gen_golden.py gives its bijection class the autograd JVP of the nsf transform
    # (torch.autograd.functional.jvp over flow.inverse) and then runs the reference unchanged.
    "power_nsfc_gij_d4": dict(dataset="power", x_shape=(6,), batch=10, epoch=60, data="uniform",
                              overrides=dict(g_ij_loss=True, latent_dimension=4, schema_type="nsf", autoregressive=False,
                                             use_linear=False, num_density_layers=3, num_hidden_layers=2,
                                             num_hidden_channels=16, num_bins=6, tail_bound=3, dropout_probability=0.0,
                                             **COMMON)),
    "hepmass_nsfc_gij_d10": dict(dataset="hepmass", x_shape=(21,), batch=6, epoch=60, data="uniform",
                                 overrides=dict(g_ij_loss=True, latent_dimension=10, schema_type="nsf", autoregressive=False,
                                                use_linear=False, num_density_layers=4, num_hidden_layers=2,
                                                num_hidden_channels=32, num_bins=8, tail_bound=3, dropout_probability=0.0,
                                                **COMMON)),
    # BASELINE.json configs[4] (CIFAR-shape latent sweep) at reduced size: 3x16x16 -> d=160 > 128,
    # the regime of the panel-blocked Cholesky and the multi-tile tcgen05 Gram
    "img16c3_gij_d160": dict(dataset="cifar10", x_shape=(3, 16, 16), batch=2, epoch=60, data="pixels",
                             overrides=dict(g_ij_loss=True, latent_dimension=160, g_hidden_channels=[8, 8],
                                            prior_num_density_layers=2, **COMMON), slow=True),
    # BASELINE.json configs[3] at full size: MNIST-shape 1x28x28 -> 64, B=2 (slow on CPU)
    "mnist_gij_d64": dict(dataset="mnist", x_shape=(1, 28, 28), batch=2, epoch=60, data="pixels",
                          overrides=dict(g_ij_loss=True, latent_dimension=64, **COMMON), slow=True),
}


def make_input(case: dict) -> torch.Tensor:
    shape = (case["batch"], *case["x_shape"])
    u = det_tensor("x/" + case["dataset"], shape, 1.0, salt=7)
    if case["data"] == "sphere":      # unit sphere, cf. cmf/datasets/two_d.py:394-400
        return u / u.norm(dim=1, keepdim=True)
    if case["data"] == "pixels":      # de-quantised pixel values in [0, 256): the noise of
        return (u + 1.0) * 127.9 + 0.05   # wrapper.py:27-30 is folded into the fixture
    return u


def hutch_eps(name: str, shape) -> torch.Tensor:
    """The injected Hutchinson probes of a ``hutch=True`` case: (B, d, m), roughly unit variance."""
    return det_tensor("hutch/" + name, shape, 1.7, salt=9)


def grad_probe(name: str, grad: torch.Tensor) -> dict:
    """Small, order-sensitive summary of one parameter gradient."""
    g = grad.detach().double().flatten()
    w = det_tensor("probe/" + name, g.shape, 1.0, salt=3, dtype=torch.float64)
    return dict(norm=float(g.norm()), dot=float((g * w).sum()), head=g[:16].clone().float())
