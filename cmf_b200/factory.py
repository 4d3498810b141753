"""Schema -> density tree (`cmf/models/factory.py:55-162,383-464`) for the layer types
``--model non-square`` produces, with :class:`FusedNonSquareHeadDensity` at the head position
(`factory.py:121-144`).  Module attribute names match the reference, so ``state_dict`` keys are
identical, including the leading ``module.`` the reference gets from ``DataParallelDensity``
whenever ``x_shape[0] != 2`` (`factory.py:76-81`).
"""
from __future__ import annotations

import math
from typing import Sequence

from . import bijections as bj
from . import densities as dn
from .head import FusedNonSquareHeadDensity
from .networks import make_coupler
from .parallel import DataParallelDensity


def get_density(schema: Sequence[dict], x_train_or_shape):
    """``x_train_or_shape``: a tensor (N, *x_shape) like the reference takes, or the shape itself."""
    x_shape = tuple(x_train_or_shape.shape[1:]) if hasattr(x_train_or_shape, "shape") else tuple(x_train_or_shape)
    density = get_density_recursive(list(schema), x_shape)
    if x_shape[0] != 2:
        density = DataParallelDensity(density)
    return density


def get_density_recursive(schema, x_shape):
    if not schema:
        return dn.standard_gaussian(x_shape)
    layer, tail = schema[0], schema[1:]
    ty = layer["type"]

    if ty == "dequantization":
        return dn.DequantizationDensity(get_density_recursive(tail, x_shape))

    if ty == "split":
        half = (x_shape[0] // 2, *x_shape[1:])
        return dn.SplitDensity(density_1=get_density_recursive(tail, half),
                               density_2=dn.standard_gaussian(half), dim=1, non_square=layer["non_square"])

    if ty == "non-square-head":
        if layer["m_flow"]:
            raise ValueError("m_flow selects ManifoldFlowHeadDensity (no log-det): outside the fused hot path")
        d = layer["latent_dimension"]
        max_cg = min(layer["max_cg_iterations"], d) if layer["max_cg_iterations"] else d
        return FusedNonSquareHeadDensity(
            prior=get_density_recursive(tail, x_shape),
            regularization_param=layer["regularization_param"],
            log_jacobian_method=layer["log_jacobian_method"],
            x_shape=x_shape,
            hutchinson_distribution=layer["hutchinson_distribution"],
            num_hutchinson_samples=layer["hutchinson_samples"],
            max_cg_iterations=max_cg,
            cg_tolerance=layer["cg_tolerance"],
        )

    if ty == "non-square-base":
        d = layer["latent_dimension"]
        return dn.NonSquareTailDensity(prior=get_density_recursive(tail, (d,)), x_shape=x_shape,
                                       latent_dimension=d, detach_before_prior=layer["m_flow"])

    bijection = get_bijection(layer, x_shape)
    if layer.get("num_u_channels", 0) != 0:
        raise ValueError("CIF u-channels are not part of the non-square configs (num_u_channels: 0)")
    return dn.BijectionDensity(bijection=bijection, prior=get_density_recursive(tail, bijection.z_shape))


def get_bijection(layer: dict, x_shape):
    ty = layer["type"]
    if ty == "acl":
        return get_acl_bijection(layer, x_shape)
    if ty == "squeeze":
        return bj.Squeeze2dBijection(x_shape, layer["factor"])
    if ty == "flatten":
        return bj.ViewBijection(x_shape, (int(math.prod(x_shape)),))
    if ty == "logit":
        return bj.LogitBijection(x_shape)
    if ty == "scalar-mult":
        return bj.ScalarMultiplicationBijection(x_shape, layer["value"])
    if ty == "scalar-add":
        return bj.ScalarAdditionBijection(x_shape, layer["value"])
    if ty == "affine":
        return bj.AffineBijection(x_shape, layer["per_channel"])
    if ty == "rand-channel-perm":                       # factory.py:291-292
        return bj.RandomChannelwisePermutationBijection(x_shape)
    if ty == "nsf-c":                                   # factory.py:316-327
        assert len(x_shape) == 1
        if layer["activation"] != "relu":
            raise ValueError("nsf-c: only the relu conditioner (config/schemas.py:608) is built")
        return bj.CoupledRationalQuadraticSplineBijection(
            num_input_channels=x_shape[0], num_hidden_layers=layer["num_hidden_layers"],
            num_hidden_channels=layer["num_hidden_channels"], num_bins=layer["num_bins"], tail_bound=layer["tail_bound"],
            dropout_probability=layer["dropout_probability"], reverse_mask=layer["reverse_mask"])
    if ty in ("nsf-ar", "linear"):
        raise ValueError(f"layer type {ty} has no JVP rule (autoregressive splines / LU linear layers cannot sit on "
                         "the non-square JVP stack; use autoregressive=False, use_linear=False)")
    raise ValueError(f"Invalid layer type {ty} for the non-square path")


def get_acl_bijection(cfg: dict, x_shape):
    c = x_shape[0]
    if cfg["mask_type"] == "checkerboard":
        coupler = make_coupler((c, *x_shape[1:]), c, cfg["coupler"])
        return bj.Checkerboard2dAffineCouplingBijection(x_shape, coupler, cfg["reverse_mask"])

    def coupler_factory(num_passthrough):
        return make_coupler((num_passthrough, *x_shape[1:]), c - num_passthrough, cfg["coupler"])

    if cfg["mask_type"] == "alternating-channel":
        return bj.AlternatingChannelwiseAffineCouplingBijection(x_shape, coupler_factory, cfg["reverse_mask"])
    if cfg["mask_type"] == "split-channel":
        return bj.SplitChannelwiseAffineCouplingBijection(x_shape, coupler_factory, cfg["reverse_mask"])
    raise ValueError(f"Invalid mask type {cfg['mask_type']}")
