"""CPU oracle for the rational-quadratic spline coupling (TEST INFRASTRUCTURE ONLY: imported by
tests/ and nothing else).  Plain-torch restatement of

  * `gitmodules/nsf/nde/transforms/splines/rational_quadratic.py:13-52`
        (unconstrained_rational_quadratic_spline, linear tails),
  * `:54-134` (rational_quadratic_spline, forward and inverse branches),
  * `gitmodules/nsf/utils/torchutils.py:133-138` (searchsorted),
  * the parameter split / scaling of `gitmodules/nsf/nde/transforms/coupling.py:437-447`.

Pinned by `oracle/gen_rqs_golden.py`: outputs of the UNMODIFIED nsf functions imported from
/root/reference in the build container, stored in tests/golden/rqs_nsf.pt
(tests/test_rqs_oracle.py checks this file against them, bins bit-exact).
Works in any dtype; tangents and gradients come from torch autograd over this code.
"""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn.functional as F

MIN_W = MIN_H = MIN_D = 1e-3


def searchsorted(bin_locations, inputs, eps=1e-6):
    """torchutils.py:133-138 (the in-place `+= eps` on the last knot is applied to a copy)."""
    locs = bin_locations.clone()
    locs[..., -1] += eps
    return torch.sum(inputs[..., None] >= locs, dim=-1) - 1


def knots(unnormalized_widths, unnormalized_heights, unnormalized_derivatives, left, right, bottom, top):
    """rational_quadratic.py:69-92: (cumwidths, widths, cumheights, heights, derivatives)."""
    num_bins = unnormalized_widths.shape[-1]
    widths = F.softmax(unnormalized_widths, dim=-1)
    widths = MIN_W + (1 - MIN_W * num_bins) * widths
    cumwidths = torch.cumsum(widths, dim=-1)
    cumwidths = F.pad(cumwidths, pad=(1, 0), mode="constant", value=0.0)
    cumwidths = (right - left) * cumwidths + left
    cumwidths = torch.cat((torch.full_like(cumwidths[..., :1], left), cumwidths[..., 1:-1],
                           torch.full_like(cumwidths[..., :1], right)), -1)
    widths = cumwidths[..., 1:] - cumwidths[..., :-1]
    derivatives = MIN_D + F.softplus(unnormalized_derivatives)
    heights = F.softmax(unnormalized_heights, dim=-1)
    heights = MIN_H + (1 - MIN_H * num_bins) * heights
    cumheights = torch.cumsum(heights, dim=-1)
    cumheights = F.pad(cumheights, pad=(1, 0), mode="constant", value=0.0)
    cumheights = (top - bottom) * cumheights + bottom
    cumheights = torch.cat((torch.full_like(cumheights[..., :1], bottom), cumheights[..., 1:-1],
                            torch.full_like(cumheights[..., :1], top)), -1)
    heights = cumheights[..., 1:] - cumheights[..., :-1]
    return cumwidths, widths, cumheights, heights, derivatives


def spline_from_tables(inputs, cumwidths, cumheights, derivatives, inverse):
    """rational_quadratic.py:94-134 given the knot tables; returns (outputs, logabsdet, bin_idx)."""
    widths = cumwidths[..., 1:] - cumwidths[..., :-1]
    heights = cumheights[..., 1:] - cumheights[..., :-1]
    bin_idx = searchsorted(cumheights if inverse else cumwidths, inputs)[..., None]
    g = lambda t: t.gather(-1, bin_idx)[..., 0]
    input_cumwidths, input_bin_widths = g(cumwidths), g(widths)
    input_cumheights, input_heights = g(cumheights), g(heights)
    input_delta = g(heights / widths)
    input_derivatives, input_derivatives_plus_one = g(derivatives), g(derivatives[..., 1:])
    if inverse:
        s = input_derivatives + input_derivatives_plus_one - 2 * input_delta
        a = (inputs - input_cumheights) * s + input_heights * (input_delta - input_derivatives)
        b = input_heights * input_derivatives - (inputs - input_cumheights) * s
        c = -input_delta * (inputs - input_cumheights)
        discriminant = b.pow(2) - 4 * a * c
        root = (2 * c) / (-b - torch.sqrt(discriminant))
        outputs = root * input_bin_widths + input_cumwidths
        tt = root * (1 - root)
        denominator = input_delta + s * tt
        dnum = input_delta.pow(2) * (input_derivatives_plus_one * root.pow(2) + 2 * input_delta * tt
                                     + input_derivatives * (1 - root).pow(2))
        return outputs, -(torch.log(dnum) - 2 * torch.log(denominator)), bin_idx[..., 0]
    theta = (inputs - input_cumwidths) / input_bin_widths
    tt = theta * (1 - theta)
    numerator = input_heights * (input_delta * theta.pow(2) + input_derivatives * tt)
    denominator = input_delta + (input_derivatives + input_derivatives_plus_one - 2 * input_delta) * tt
    outputs = input_cumheights + numerator / denominator
    dnum = input_delta.pow(2) * (input_derivatives_plus_one * theta.pow(2) + 2 * input_delta * tt
                                 + input_derivatives * (1 - theta).pow(2))
    return outputs, torch.log(dnum) - 2 * torch.log(denominator), bin_idx[..., 0]


def split_params(transform_params, num_bins, hidden_features=None):
    """coupling.py:437-447 + the linear-tail padding of rational_quadratic.py:29-33."""
    uw = transform_params[..., :num_bins]
    uh = transform_params[..., num_bins:2 * num_bins]
    ud = transform_params[..., 2 * num_bins:]
    if hidden_features is not None:
        uw, uh = uw / np.sqrt(hidden_features), uh / np.sqrt(hidden_features)
    ud = F.pad(ud, pad=(1, 1))
    constant = math.log(math.exp(1 - MIN_D) - 1)
    ud = torch.cat((torch.full_like(ud[..., :1], constant), ud[..., 1:-1], torch.full_like(ud[..., :1], constant)), -1)
    return uw, uh, ud


def unconstrained_rqs(inputs, transform_params, num_bins, tail_bound, inverse, hidden_features=None):
    """rational_quadratic.py:13-52 (+ parameter split): identity outside [-tail_bound, tail_bound].
    Returns (outputs, logabsdet, bin_idx with -1 outside)."""
    uw, uh, ud = split_params(transform_params, num_bins, hidden_features)
    inside = (inputs >= -tail_bound) & (inputs <= tail_bound)
    cw, _, ch, _, dv = knots(uw, uh, ud, -tail_bound, tail_bound, -tail_bound, tail_bound)
    safe = torch.where(inside, inputs, torch.zeros_like(inputs))
    out, lad, idx = spline_from_tables(safe, cw, ch, dv, inverse)
    return (torch.where(inside, out, inputs), torch.where(inside, lad, torch.zeros_like(lad)),
            torch.where(inside, idx, torch.full_like(idx, -1)))
