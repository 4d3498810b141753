"""Bijection nodes of the non-square density tree: parameter containers with the reference's
names (`cmf/models/components/bijections/*.py`) plus the plain-PyTorch ``x_to_z`` / ``z_to_x``
maps used by the encode pass (`non_square.py:65`) and by sampling.  The JVP direction is NOT
implemented here: the fused head compiles these nodes into a layer program that runs on the
sm_100a kernels (``cmf_b200.program``).

Dispatch contract kept from `bijection.py:13-38`: ``x_to_z(x) -> {"z", "log-jac"}``,
``z_to_x(z) -> {"x", "log-jac"}``, with the same shape asserts.
"""
from __future__ import annotations

import math
from typing import Tuple

import torch
import torch.nn as nn

Tensor = torch.Tensor


class Bijection(nn.Module):
    def __init__(self, x_shape, z_shape):
        super().__init__()
        self.x_shape, self.z_shape = tuple(x_shape), tuple(z_shape)

    def x_to_z(self, x: Tensor, **kw):
        assert tuple(x.shape[1:]) == self.x_shape, (x.shape, self.x_shape)
        out = self._x_to_z(x, **kw)
        assert tuple(out["z"].shape[1:]) == self.z_shape
        return out

    def z_to_x(self, z: Tensor, **kw):
        assert tuple(z.shape[1:]) == self.z_shape, (z.shape, self.z_shape)
        out = self._z_to_x(z, **kw)
        assert tuple(out["x"].shape[1:]) == self.x_shape
        return out

    def _x_to_z(self, x, **kw): raise NotImplementedError
    def _z_to_x(self, z, **kw): raise NotImplementedError


def _zeros_log_jac(t: Tensor) -> Tensor:
    return torch.zeros(t.shape[0], 1, dtype=t.dtype, device=t.device)


# ---- index maps -----------------------------------------------------------------------------
class ReshapingBijection(Bijection):
    """`reshaping.py:8-29`.  ``reshape_z`` is also what the layer program probes with an
    ``arange`` to turn the bijection into a static position map."""

    def _x_to_z(self, x, **kw): return {"z": self.reshape_x(x), "log-jac": _zeros_log_jac(x)}
    def _z_to_x(self, z, **kw): return {"x": self.reshape_z(z), "log-jac": _zeros_log_jac(z)}
    def reshape_x(self, x): raise NotImplementedError
    def reshape_z(self, z): raise NotImplementedError


class ViewBijection(ReshapingBijection):
    """`reshaping.py:61-66` (the schema's ``flatten``)."""
    def reshape_x(self, x): return x.reshape(x.shape[0], *self.z_shape)
    def reshape_z(self, z): return z.reshape(z.shape[0], *self.x_shape)


class Squeeze2dBijection(ReshapingBijection):
    """`reshaping.py:69-114`: x (c,h,w) <-> z (c f^2, h/f, w/f), z channel = c f^2 + a f + b for the
    pixel at (f i + a, f j + b)."""

    def __init__(self, x_shape, factor: int):
        c, h, w = x_shape
        assert h % factor == 0 and w % factor == 0
        super().__init__(x_shape, (c * factor * factor, h // factor, w // factor))
        self.factor = factor

    def reshape_x(self, x):
        c, h, w = self.x_shape
        f = self.factor
        return x.reshape(-1, c, h // f, f, w // f, f).permute(0, 1, 3, 5, 2, 4).reshape(-1, *self.z_shape)

    def reshape_z(self, z):
        c, h, w = self.x_shape
        f = self.factor
        return z.reshape(-1, c, f, f, h // f, w // f).permute(0, 1, 4, 2, 5, 3).reshape(-1, c, h, w)


class RandomChannelwisePermutationBijection(ReshapingBijection):
    """`reshaping.py:32-44` (the schema's ``rand-channel-perm`` between ``nsf-c`` layers)."""

    def __init__(self, x_shape):
        super().__init__(x_shape, x_shape)
        self.register_buffer("permutation", torch.randperm(x_shape[0]))
        self.register_buffer("inverse_permutation", torch.argsort(self.permutation))

    def reshape_x(self, x): return x[:, self.permutation]
    def reshape_z(self, z): return z[:, self.inverse_permutation]


# ---- elementwise preprocessing above the head (`math.py:9-105`) -------------------------------
class ScalarMultiplicationBijection(Bijection):
    def __init__(self, x_shape, value):
        assert value != 0.0, "Scalar multiplication by zero is not a bijection"
        super().__init__(x_shape, x_shape)
        self.value = value

    def _log_jac(self, t):
        n = int(math.prod(self.x_shape))
        return torch.full((t.shape[0], 1), n * math.log(abs(self.value)), dtype=t.dtype, device=t.device)

    def _x_to_z(self, x, **kw): return {"z": self.value * x, "log-jac": self._log_jac(x)}
    def _z_to_x(self, z, **kw): return {"x": z / self.value, "log-jac": -self._log_jac(z)}


class ScalarAdditionBijection(Bijection):
    def __init__(self, x_shape, value):
        super().__init__(x_shape, x_shape)
        self.value = value

    def _x_to_z(self, x, **kw): return {"z": x + self.value, "log-jac": _zeros_log_jac(x)}
    def _z_to_x(self, z, **kw): return {"x": z - self.value, "log-jac": _zeros_log_jac(z)}


class LogitBijection(Bijection):
    _EPS = 1e-7

    def __init__(self, x_shape):
        super().__init__(x_shape, x_shape)

    def _log_dF(self, x):
        xc = x.clamp(self._EPS, 1 - self._EPS)
        return (-torch.log(xc) - torch.log(1 - xc)).flatten(1).sum(1, keepdim=True)

    def _x_to_z(self, x, **kw):
        return {"z": torch.log(x) - torch.log(1 - x), "log-jac": self._log_dF(x)}

    def _z_to_x(self, z, **kw):
        x = torch.sigmoid(z)
        return {"x": x, "log-jac": -self._log_dF(x)}


class AffineBijection(Bijection):
    """`affine.py:11-41`: learned elementwise affine, the d-dim prior flow of the 2-D configs."""

    def __init__(self, x_shape, per_channel: bool):
        super().__init__(x_shape, x_shape)
        if per_channel:
            shape = (x_shape[0], *[1] * (len(x_shape) - 1))
            self.log_jac_factor = int(math.prod(x_shape[1:]))
        else:
            shape, self.log_jac_factor = tuple(x_shape), 1
        self.shift = nn.Parameter(torch.zeros(shape))
        self.log_scale = nn.Parameter(torch.zeros(shape))

    def _lj(self, b): return (self.log_jac_factor * self.log_scale.sum()).view(1, 1).expand(b, 1)

    def _x_to_z(self, x, **kw):
        return {"z": x * torch.exp(self.log_scale) + self.shift, "log-jac": self._lj(x.shape[0])}

    def _z_to_x(self, z, **kw):
        return {"x": (z - self.shift) * torch.exp(-self.log_scale), "log-jac": -self._lj(z.shape[0])}


# ---- affine coupling layers (`acl.py`) -------------------------------------------------------
class AffineCouplingBijection(Bijection):
    """Common part of `acl.py:7-26`.  Subclasses define the passthrough/modified partition both
    as tensor ops (encode pass) and as flat position lists (``partition()``, layer program)."""

    def __init__(self, x_shape, coupler: nn.Module):
        super().__init__(x_shape, x_shape)
        self.coupler = coupler

    def _couple(self, inputs):
        out = self.coupler(inputs)
        return out["shift"], out["log-scale"]

    def _encode_maps(self, device):
        """int32 index maps of ``cmf_coupling_x_to_z_fwd/bwd`` for this layer's partition (cached per device)"""
        cache = self.__dict__.setdefault("_enc_maps", {})
        key = str(device)
        if key not in cache:
            _, mod_pos, _, dense = self.partition()
            mod_pos = mod_pos.to(device=device, dtype=torch.int64)
            D, M = int(math.prod(self.x_shape)), int(mod_pos.numel())
            pos2j = torch.full((D,), -1, dtype=torch.int32, device=device)
            pos2j[mod_pos] = torch.arange(M, dtype=torch.int32, device=device)
            if dense:       # the conditioner emits a full-resolution output; element at position i reads q = i
                qmap = mod_pos.to(torch.int32)
                q2pos = torch.full((D,), -1, dtype=torch.int32, device=device)
                q2pos[mod_pos] = qmap
            else:           # compact conditioner output in the order of mod_pos
                qmap, q2pos = None, mod_pos.to(torch.int32)
            cache[key] = (pos2j, qmap, q2pos)
        return cache[key]

    def _x_to_z_fused(self, x, conditioner_input):
        """encode direction of `acl.py:43-46,101-111` on the fused kernel: one launch instead of the
        split / exp / add / mul / combine / masked-sum string"""
        from . import ops
        st, _ = self.coupler.st_batched(conditioner_input, None)
        z, lj = ops.coupling_x_to_z(x.flatten(1), st, *self._encode_maps(x.device))
        return {"z": z.view_as(x), "log-jac": lj}

    def partition(self) -> Tuple[Tensor, Tensor, tuple, bool]:
        """(passthrough positions, modified positions, conditioner input shape, dense_input).

        Positions index the flattened ``x_shape``; the modified list is in conditioner-output
        order.  ``dense_input``: the conditioner sees the full ``x_shape`` tensor with the modified
        positions zeroed (checkerboard); otherwise it sees the compact passthrough tensor."""
        raise NotImplementedError


class Checkerboard2dAffineCouplingBijection(AffineCouplingBijection):
    """`acl.py:29-78`: mask = 1 (passthrough) where row+col is odd, flipped by ``reverse_mask``."""

    def __init__(self, x_shape, coupler, reverse_mask: bool):
        super().__init__(x_shape, coupler)
        c, h, w = x_shape
        rows = torch.arange(h).view(h, 1)
        cols = torch.arange(w).view(1, w)
        mask = ((rows + cols) % 2 == 1).to(torch.get_default_dtype()).expand(c, h, w)
        self.register_buffer("mask", (1 - mask if reverse_mask else mask).contiguous())

    def _x_to_z(self, x, **kw):
        return self._x_to_z_fused(x, self.mask * x)

    def _z_to_x(self, z, **kw):
        m = self.mask
        t, s = self._couple(m * z)
        return {"x": m * z + (1 - m) * (z * torch.exp(-s) - t),
                "log-jac": -((1 - m) * s).flatten(1).sum(1, keepdim=True)}

    def partition(self):
        flat = self.mask.flatten() > 0.5
        idx = torch.arange(flat.numel(), device=flat.device)
        return idx[flat], idx[~flat], self.x_shape, True


class ChannelwiseAffineCouplingBijection(AffineCouplingBijection):
    """`acl.py:81-166`."""

    def __init__(self, x_shape, coupler_factory, num_passthrough_channels: int, reverse_mask: bool):
        if reverse_mask:
            num_passthrough_channels = x_shape[0] - num_passthrough_channels
        assert num_passthrough_channels > 0, "Not a bijection without passthrough"
        super().__init__(x_shape, coupler_factory(num_passthrough_channels))
        self.reverse_mask = reverse_mask

    def _split(self, t):
        a, b = self._do_split(t)
        return (b, a) if self.reverse_mask else (a, b)

    def _combine(self, keep, moved):
        a, b = (moved, keep) if self.reverse_mask else (keep, moved)
        return self._do_combine(a, b)

    def _x_to_z(self, x, **kw):
        return self._x_to_z_fused(x, self._split(x)[0])

    def _z_to_x(self, z, **kw):
        keep, moved = self._split(z)
        t, s = self._couple(keep)
        return {"x": self._combine(keep, moved * torch.exp(-s) - t),
                "log-jac": -s.flatten(1).sum(1, keepdim=True)}

    def partition(self):
        n = int(math.prod(self.x_shape))
        probe = torch.arange(n).view(1, *self.x_shape)
        keep, moved = self._split(probe)
        return keep.flatten(), moved.flatten(), tuple(keep.shape[1:]), False


class SplitChannelwiseAffineCouplingBijection(ChannelwiseAffineCouplingBijection):
    """`acl.py:169-189`."""

    def __init__(self, x_shape, coupler_factory, reverse_mask: bool):
        self.num_passthrough_channels = x_shape[0] // 2
        super().__init__(x_shape, coupler_factory, self.num_passthrough_channels, reverse_mask)

    def _do_split(self, t):
        return t[:, : self.num_passthrough_channels], t[:, self.num_passthrough_channels:]

    def _do_combine(self, a, b):
        return torch.cat((a, b), dim=1)


class AlternatingChannelwiseAffineCouplingBijection(ChannelwiseAffineCouplingBijection):
    """`acl.py:192-214`: even channels / odd channels."""

    def __init__(self, x_shape, coupler_factory, reverse_mask: bool):
        super().__init__(x_shape, coupler_factory, (x_shape[0] + 1) // 2, reverse_mask)

    def _do_split(self, t):
        return t[:, ::2], t[:, 1::2]

    def _do_combine(self, a, b):
        out = a.new_empty(a.shape[0], *self.x_shape)
        out[:, ::2] = a
        out[:, 1::2] = b
        return out


# ---- rational-quadratic spline coupling (`bijections/nsf.py:45-84`) -----------------------------
class _UnconditionalRQ(nn.Module):
    """Parameters of nsf's ``PiecewiseRationalQuadraticCDF`` on the identity features
    (`gitmodules/nsf/nde/transforms/nonlinearities.py:337-369`, linear tails, random init)."""

    def __init__(self, features: int, num_bins: int):
        super().__init__()
        self.unnormalized_widths = nn.Parameter(torch.rand(features, num_bins))
        self.unnormalized_heights = nn.Parameter(torch.rand(features, num_bins))
        self.unnormalized_derivatives = nn.Parameter(torch.rand(features, num_bins - 1))


class _RQCouplingFlow(nn.Module):
    """Container with the attribute names of nsf's ``PiecewiseRationalQuadraticCouplingTransform``
    (`gitmodules/nsf/nde/transforms/coupling.py:14-57,399-431`) so state-dict keys match."""

    def __init__(self, features: int, even: bool, net_factory, num_bins: int):
        super().__init__()
        mask = torch.zeros(features, dtype=torch.uint8)
        mask[(0 if even else 1)::2] += 1                      # utils/torchutils.py:87-98
        idx = torch.arange(features)
        self.register_buffer("identity_features", idx[mask <= 0])
        self.register_buffer("transform_features", idx[mask > 0])
        n_id, n_tr = int(self.identity_features.numel()), int(self.transform_features.numel())
        self.transform_net = net_factory(n_id, n_tr * (3 * num_bins - 1))
        self.unconditional_transform = _UnconditionalRQ(n_id, num_bins)


class CoupledRationalQuadraticSplineBijection(Bijection):
    """``nsf-c`` layer (`bijections/nsf.py:45-84`): alternating binary mask, ResidualNet conditioner,
    linear tails, ``apply_unconditional_transform=True`` (the identity features go through a spline
    with free parameters).  ``x_to_z`` / ``z_to_x`` here are plain torch on the reference's knot
    arithmetic; the z -> x push WITH tangents -- which the reference does not have at all
    -- is compiled into the layer program (``cmf_b200.program.RqsStep``) and runs on
    ``cmf_rqs_coupling_jvp_fwd/bwd``."""

    def __init__(self, num_input_channels: int, num_hidden_layers: int, num_hidden_channels: int, num_bins: int,
                 tail_bound: float, dropout_probability: float, reverse_mask: bool):
        from .splines import ResidualNet
        super().__init__((num_input_channels,), (num_input_channels,))
        self.num_bins, self.tail_bound = num_bins, float(tail_bound)
        self.flow = _RQCouplingFlow(
            num_input_channels, reverse_mask,
            lambda n_in, n_out: ResidualNet(n_in, n_out, num_hidden_channels, num_hidden_layers, dropout_probability),
            num_bins)

    # -- knot tables ---------------------------------------------------------------------------
    def unconditional_tables(self, batch: int):
        """(cw, ch, dv), each (batch, n_id, K+1): the free-parameter spline of the identity half"""
        from .splines import knot_tables
        u = self.flow.unconditional_transform
        params = torch.cat((u.unnormalized_widths, u.unnormalized_heights, u.unnormalized_derivatives), dim=-1)
        cw, ch, dv, _, _, _ = knot_tables(params.unsqueeze(0), None, self.num_bins, self.tail_bound)
        exp = lambda t: t.expand(batch, *t.shape[1:]).contiguous()
        return exp(cw), exp(ch), exp(dv)

    def conditional_tables(self, x_id: Tensor, V_id):
        """knot tables of the transformed half (and their tangents) from the conditioner"""
        from .splines import knot_tables
        net = self.flow.transform_net
        p, P = net.jvp_batched(x_id, V_id)
        B, M = x_id.shape[0], int(self.flow.transform_features.numel())
        p = p.view(B, M, 3 * self.num_bins - 1)
        if P is not None:
            P = P.view(B, P.shape[1], M, 3 * self.num_bins - 1)
        return knot_tables(p, P, self.num_bins, self.tail_bound, hidden_features=net.hidden_features)

    # -- encode pass / sampling (plain torch) ---------------------------------------------------
    def _x_to_z(self, x, **kw):
        from .splines import rq_spline_from_tables
        f = self.flow
        x_id, x_tr = x[:, f.identity_features], x[:, f.transform_features]
        cw, ch, dv, _, _, _ = self.conditional_tables(x_id, None)              # coupling.py:74-81
        z_tr, lad = rq_spline_from_tables(x_tr, cw, ch, dv, False, self.tail_bound)
        z_id, lad_id = rq_spline_from_tables(x_id, *self.unconditional_tables(x.shape[0]), False, self.tail_bound)
        z = torch.empty_like(x)
        z[:, f.identity_features] = z_id
        z[:, f.transform_features] = z_tr
        return {"z": z, "log-jac": (lad.sum(1) + lad_id.sum(1)).view(-1, 1)}

    def _z_to_x(self, z, **kw):
        from .splines import rq_spline_from_tables
        f = self.flow
        z_id, z_tr = z[:, f.identity_features], z[:, f.transform_features]
        x_id, lad_id = rq_spline_from_tables(z_id, *self.unconditional_tables(z.shape[0]), True, self.tail_bound)
        cw, ch, dv, _, _, _ = self.conditional_tables(x_id, None)              # coupling.py:103-114
        x_tr, lad = rq_spline_from_tables(z_tr, cw, ch, dv, True, self.tail_bound)
        x = torch.empty_like(z)
        x[:, f.identity_features] = x_id
        x[:, f.transform_features] = x_tr
        return {"x": x, "log-jac": (lad.sum(1) + lad_id.sum(1)).view(-1, 1)}

    def partition(self):
        f = self.flow
        return f.identity_features, f.transform_features
