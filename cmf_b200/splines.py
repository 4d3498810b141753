"""Host side of the rational-quadratic spline coupling: knot tables and their tangents.

The tables are built with the SAME torch ops, in the same order, as
`gitmodules/nsf/nde/transforms/splines/rational_quadratic.py:13-52,69-92` (softmax -> affine ->
cumsum -> pad -> rescale -> end pins; softplus derivatives with the linear-tail constant), so the
fp32 knots -- and hence the bin indices the kernel finds -- are bit-identical to the reference's.
Their tangents are the JVPs of those ops applied to the conditioner tangents (tangent index next
to the batch).  Everything here is differentiable torch code; the device work is
``cmf_b200.ops.rqs_coupling_jvp``.
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import numpy as np
import torch
import torch.nn.functional as F

Tensor = torch.Tensor

DEFAULT_MIN_BIN_WIDTH = 1e-3
DEFAULT_MIN_BIN_HEIGHT = 1e-3
DEFAULT_MIN_DERIVATIVE = 1e-3


def _cum_knots(unnormalized: Tensor, min_size: float, lo: float, hi: float) -> Tuple[Tensor, Tensor]:
    """(knots (..., K+1), normalised bin sizes (..., K)): rational_quadratic.py:74-81 / :85-92."""
    K = unnormalized.shape[-1]
    sizes = F.softmax(unnormalized, dim=-1)
    sizes = min_size + (1 - min_size * K) * sizes
    cum = torch.cumsum(sizes, dim=-1)
    cum = F.pad(cum, pad=(1, 0), mode="constant", value=0.0)
    cum = (hi - lo) * cum + lo
    cum = torch.cat((torch.full_like(cum[..., :1], lo), cum[..., 1:-1], torch.full_like(cum[..., :1], hi)), dim=-1)
    return cum, sizes


def _cum_knots_tangent(sizes: Tensor, unnormalized_dot: Tensor, min_size: float, lo: float, hi: float) -> Tensor:
    """JVP of :func:`_cum_knots`: sizes (B,M,K) (post-affine), unnormalized_dot (B,d,M,K) -> (B,d,M,K+1)."""
    K = sizes.shape[-1]
    soft = (sizes - min_size) / (1 - min_size * K)                       # the softmax itself
    s = soft.unsqueeze(1)
    soft_dot = s * (unnormalized_dot - (s * unnormalized_dot).sum(-1, keepdim=True))
    cum_dot = (hi - lo) * (1 - min_size * K) * torch.cumsum(soft_dot, dim=-1)
    zero = torch.zeros_like(cum_dot[..., :1])
    return torch.cat((zero, cum_dot[..., :-1], zero), dim=-1)            # pinned end knots do not move


def knot_tables(params: Tensor, params_dot: Optional[Tensor], num_bins: int, tail_bound: float,
                hidden_features: Optional[int] = None,
                min_bin_width: float = DEFAULT_MIN_BIN_WIDTH, min_bin_height: float = DEFAULT_MIN_BIN_HEIGHT,
                min_derivative: float = DEFAULT_MIN_DERIVATIVE):
    """params (B,M,3K-1) = conditioner output per transformed element (`coupling.py:437-440`),
    params_dot (B,d,M,3K-1) its tangents or None.  Returns (cw, ch, dv, CW, CH, DV)."""
    K = num_bins
    assert params.shape[-1] == 3 * K - 1
    uw, uh, ud = params[..., :K], params[..., K:2 * K], params[..., 2 * K:]
    scale = 1.0
    if hidden_features is not None:            # coupling.py:442-447
        scale = 1.0 / np.sqrt(hidden_features)
        uw, uh = uw / np.sqrt(hidden_features), uh / np.sqrt(hidden_features)
    # linear tails: rational_quadratic.py:29-33
    ud = F.pad(ud, pad=(1, 1))
    constant = math.log(math.exp(1 - min_derivative) - 1)
    ud = torch.cat((torch.full_like(ud[..., :1], constant), ud[..., 1:-1], torch.full_like(ud[..., :1], constant)), dim=-1)
    cw, w_sizes = _cum_knots(uw, min_bin_width, -tail_bound, tail_bound)
    ch, h_sizes = _cum_knots(uh, min_bin_height, -tail_bound, tail_bound)
    dv = min_derivative + F.softplus(ud)
    if params_dot is None:
        return cw, ch, dv, None, None, None
    uw_dot, uh_dot, ud_dot = params_dot[..., :K] * scale, params_dot[..., K:2 * K] * scale, params_dot[..., 2 * K:]
    CW = _cum_knots_tangent(w_sizes, uw_dot, min_bin_width, -tail_bound, tail_bound)
    CH = _cum_knots_tangent(h_sizes, uh_dot, min_bin_height, -tail_bound, tail_bound)
    zero = ud_dot.new_zeros(*ud_dot.shape[:-1], 1)
    DV = torch.sigmoid(ud).unsqueeze(1) * torch.cat((zero, ud_dot, zero), dim=-1)
    return cw, ch, dv, CW, CH, DV
