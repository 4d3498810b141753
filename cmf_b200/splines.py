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


# ---------------------------------------------------------------------------------------------
# Plain-torch evaluation from knot tables (encode pass x -> z and sampling; the z -> x tangent
# push runs on the kernel).  Arithmetic of rational_quadratic.py:94-134 on the tables above.
# ---------------------------------------------------------------------------------------------
def _bin_of(knots: Tensor, inputs: Tensor) -> Tensor:
    """`utils/torchutils.py:133-138`: number of knots <= input, minus one, with the last knot nudged up."""
    bumped = torch.cat((knots[..., :-1], knots[..., -1:] + 1e-6), dim=-1)
    return (inputs.unsqueeze(-1) >= bumped).sum(-1) - 1


def rq_spline_from_tables(inputs: Tensor, cw: Tensor, ch: Tensor, dv: Tensor, inverse: bool, tail_bound: float):
    """inputs (B,M); cw, ch, dv (B,M,K+1) from :func:`knot_tables` -> (outputs (B,M), logabsdet (B,M)).
    Linear tails: identity with zero log-derivative outside [-tail_bound, tail_bound]
    (rational_quadratic.py:21-52)."""
    inside = (inputs >= -tail_bound) & (inputs <= tail_bound)
    u = torch.where(inside, inputs, torch.zeros_like(inputs))
    k = _bin_of(ch if inverse else cw, u).clamp_(0, cw.shape[-1] - 2).unsqueeze(-1)
    pick = lambda t: t.gather(-1, k).squeeze(-1)
    x0, y0 = pick(cw), pick(ch)
    w, h = pick(cw[..., 1:]) - x0, pick(ch[..., 1:]) - y0
    delta = h / w
    d0, d1 = pick(dv), pick(dv[..., 1:])
    s = d0 + d1 - 2 * delta
    if inverse:
        a = (u - y0) * s + h * (delta - d0)
        b = h * d0 - (u - y0) * s
        c = -delta * (u - y0)
        root = (2 * c) / (-b - torch.sqrt(b * b - 4 * a * c))
        out = root * w + x0
        tt = root * (1 - root)
        den = delta + s * tt
        num = delta * delta * (d1 * root * root + 2 * delta * tt + d0 * (1 - root) * (1 - root))
        lad = -(torch.log(num) - 2 * torch.log(den))
    else:
        theta = (u - x0) / w
        tt = theta * (1 - theta)
        den = delta + s * tt
        out = y0 + h * (delta * theta * theta + d0 * tt) / den
        num = delta * delta * (d1 * theta * theta + 2 * delta * tt + d0 * (1 - theta) * (1 - theta))
        lad = torch.log(num) - 2 * torch.log(den)
    return torch.where(inside, out, inputs), torch.where(inside, lad, torch.zeros_like(lad))


class ResidualNet(torch.nn.Module):
    """The conditioner of an ``nsf-c`` layer: `gitmodules/nsf/nn/resnet.py:62-105` with
    ``context_features=None``, no batch norm, relu -- same attribute names, hence the same
    state-dict keys (``initial_layer``, ``blocks.{i}.linear_layers.{0,1}``, ``final_layer``).
    ``jvp_batched`` pushes all d tangents at once with the reference's JVP rules (`jvp_layers.py:38-64`:
    linear drops the bias, relu -> (x > 0) v; dropout is the identity at p = 0 / in eval)."""

    class Block(torch.nn.Module):
        def __init__(self, features: int, dropout_probability: float):
            super().__init__()
            self.linear_layers = torch.nn.ModuleList([torch.nn.Linear(features, features) for _ in range(2)])
            self.dropout = torch.nn.Dropout(p=dropout_probability)
            torch.nn.init.uniform_(self.linear_layers[-1].weight, -1e-3, 1e-3)
            torch.nn.init.uniform_(self.linear_layers[-1].bias, -1e-3, 1e-3)

    def __init__(self, in_features: int, out_features: int, hidden_features: int, num_blocks: int,
                 dropout_probability: float = 0.0):
        super().__init__()
        self.hidden_features = hidden_features
        self.initial_layer = torch.nn.Linear(in_features, hidden_features)
        self.blocks = torch.nn.ModuleList([ResidualNet.Block(hidden_features, dropout_probability)
                                           for _ in range(num_blocks)])
        self.final_layer = torch.nn.Linear(hidden_features, out_features)

    def forward(self, x: Tensor) -> Tensor:
        h = self.initial_layer(x)
        for blk in self.blocks:
            t = blk.linear_layers[0](torch.relu(h))
            t = blk.linear_layers[1](blk.dropout(torch.relu(t)))
            h = h + t
        return self.final_layer(h)

    def jvp_batched(self, x: Tensor, V: Optional[Tensor]):
        """x (B,p), V (B,d,p) | None -> y (B,o), Y (B,d,o) | None"""
        if V is None:
            return self.forward(x), None
        if self.training and any(blk.dropout.p > 0 for blk in self.blocks):
            raise ValueError("dropout > 0 has no JVP rule in the reference (jvp_layers.py:6-35)")
        lin = lambda layer, T: F.linear(T, layer.weight)
        h, H = self.initial_layer(x), lin(self.initial_layer, V)
        for blk in self.blocks:
            T = H * (h > 0).to(H.dtype).unsqueeze(1)
            t = blk.linear_layers[0](torch.relu(h))
            T = lin(blk.linear_layers[0], T) * (t > 0).to(H.dtype).unsqueeze(1)
            t = blk.linear_layers[1](torch.relu(t))
            h, H = h + t, H + lin(blk.linear_layers[1], T)
        return self.final_layer(h), lin(self.final_layer, H)
