"""``torch.autograd.Function`` ops over the sm_100a kernels (through ``cmf_b200.kernels`` ->
C ABI).  These are the only places the hot path touches the GPU outside the conditioner
networks' library calls.

The exact (cholesky) path needs first-order adjoints only: the reference's "double backward" is one
reverse pass over a graph whose nodes are JVPs (`trainer.py:216`, `non_square.py:298-311`), and
each op here is one such node with an explicit adjoint w.r.t. primal *and* tangent inputs.
``hutch_with_cg`` additionally differentiates a VJP (`non_square.py:199`,
``autograd_F.vjp(..., create_graph=True)`` over a PRIMAL-only run): the primal coupling and the
gather ops therefore issue their adjoints as differentiable nodes when the backward pass itself is
being recorded (``torch.is_grad_enabled()`` inside ``backward``).
"""
from __future__ import annotations

from typing import Optional, Sequence, Tuple

import torch
from torch.autograd import Function

from . import kernels

Tensor = torch.Tensor


class NotPositiveDefiniteError(RuntimeError):
    """Raised where ``torch.linalg.cholesky`` raises in the reference (`non_square.py:282-288`)."""


class _DeferredInfo:
    """While a training step is being captured into / replayed from a CUDA graph the host cannot
    read the Cholesky ``info`` word (a device->host sync is illegal during capture).  In this mode
    :func:`chol_logdet_l1` appends the device word to ``pending`` instead; the owner of the graph
    reads all of them after the replay and falls back to the eager path (which runs the
    reference's ``+eps*I`` retry loop) if any is non-zero."""
    active = False
    pending: list = []


class defer_pd_check:
    def __enter__(self):
        _DeferredInfo.active, _DeferredInfo.pending = True, []
        return _DeferredInfo.pending

    def __exit__(self, *exc):
        _DeferredInfo.active = False
        return False


class _CouplingJVP(Function):
    @staticmethod
    def forward(ctx, z_m, V_m, st, ST, map_, inv):
        x_m, V_out = kernels.coupling_jvp_fwd(z_m, V_m, st, ST, map_)
        ctx.save_for_backward(z_m, V_m, st, ST)
        ctx.maps = (map_, inv)
        ctx.set_materialize_grads(False)
        if V_out is None:
            ctx.mark_non_differentiable()
            return x_m, None
        return x_m, V_out

    @staticmethod
    def backward(ctx, g_x, g_V):
        z_m, V_m, st, ST = ctx.saved_tensors
        map_, inv = ctx.maps
        if V_m is None and torch.is_grad_enabled():
            # autograd.grad(..., create_graph=True) over a primal-only run: the VJP J^T u of
            # `non_square.py:190-201` (hutch_with_cg) must itself carry a graph
            g_z, g_st = _CouplingPrimalVJP.apply(z_m, st, g_x, map_, inv)
            return g_z, None, g_st, None, None, None
        if V_m is not None and g_V is None:
            g_V = torch.zeros_like(V_m)
        g_z, g_Vin, g_st, g_ST = kernels.coupling_jvp_bwd(z_m, V_m, st, ST, map_, inv, g_x, g_V)
        return g_z, g_Vin, g_st, g_ST, None, None


class _CouplingPrimalVJP(Function):
    """The primal coupling's VJP as a differentiable node: (z_m, st, g_x) -> (g_z, g_st) by the
    adjoint kernel with d = 0,

        e = exp(-s);   g_z = e g_x;   g_t = -g_x;   g_s = -z e g_x,

    and ITS adjoint (the "double backward" the reference gets from autograd when
    `non_square.py:199` calls ``autograd_F.vjp(..., create_graph=True)``): with cotangents
    (a, b_t, b_s) of (g_z, g_t, g_s),

        d/dg_x = a e - b_t - b_s z e;   d/dz = -b_s e g_x;   d/ds = e g_x (b_s z - a);   d/dt = 0.

    Primal-sized elementwise work (B x M, no tangent factor), issued as torch ops on the device."""

    @staticmethod
    def forward(ctx, z_m, st, g_x, map_, inv):
        if g_x is None:
            g_x = torch.zeros_like(z_m)
        g_z, _, g_st, _ = kernels.coupling_jvp_bwd(z_m, None, st, None, map_, inv, g_x, None)
        ctx.save_for_backward(z_m, st, g_x)
        ctx.map = map_
        return g_z, g_st

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, a, b):
        z, st, gx = ctx.saved_tensors
        Q = st.shape[1] // 2
        q = None if ctx.map is None else ctx.map.long()
        pick = (lambda t: t) if q is None else (lambda t: t.index_select(1, q))
        e = torch.exp(-pick(st[:, Q:]))
        a = torch.zeros_like(z) if a is None else a
        b = torch.zeros_like(st) if b is None else b
        bt, bs = pick(b[:, :Q]), pick(b[:, Q:])
        g_gx = a * e - bt - bs * z * e
        g_z = -bs * e * gx
        g_s = e * gx * (bs * z - a)
        g_st = torch.zeros_like(st)
        if q is None:
            g_st[:, Q:] = g_s
        else:
            g_st[:, Q:].index_copy_(1, q, g_s)
        return g_z, g_st, g_gx, None, None


def coupling_jvp(z_m: Tensor, V_m: Optional[Tensor], st: Tensor, ST: Optional[Tensor],
                 map_: Optional[Tensor] = None, inv: Optional[Tensor] = None) -> Tuple[Tensor, Optional[Tensor]]:
    """Affine coupling on the compact transformed half, primal + d tangents (acl.py:53-66,125-146).

    z_m (B,M), V_m (B,d,M) or None (primal only, acl.py:48-51), st (B,2Q) = [shift | log_scale],
    ST (B,d,2Q) their tangents; ``map_``/``inv`` (int32) relate compact index j to conditioner
    position q when the conditioner emits more positions than are transformed (checkerboard)."""
    return _CouplingJVP.apply(z_m, V_m, st, ST, map_, inv)


class _GatherRows(Function):
    @staticmethod
    def forward(ctx, in0, in1, idx, inv0, inv1):
        ctx.inv = (inv0, inv1)
        ctx.has1 = in1 is not None
        ctx.idx, ctx.n0 = idx, in0.shape[-1]
        return kernels.gather_rows(in0, in1, idx)

    @staticmethod
    def backward(ctx, g):
        inv0, inv1 = ctx.inv
        if torch.is_grad_enabled():
            # create_graph=True (the VJP of hutch_with_cg, `non_square.py:199`): a gather's adjoint is a
            # gather, so it is issued through the differentiable op with the roles of the maps swapped
            idx, n0 = ctx.idx, ctx.n0
            neg = torch.full_like(idx, -1)
            g0 = g1 = None
            if ctx.needs_input_grad[0]:
                g0 = _GatherRows.apply(g.contiguous(), None, inv0, torch.where((idx >= 0) & (idx < n0), idx, neg), None)
            if ctx.has1 and ctx.needs_input_grad[1]:
                g1 = _GatherRows.apply(g.contiguous(), None, inv1, torch.where(idx >= n0, idx - n0, neg), None)
            return g0, g1, None, None, None
        g0 = kernels.gather_rows(g, None, inv0) if ctx.needs_input_grad[0] else None
        g1 = kernels.gather_rows(g, None, inv1) if (ctx.has1 and ctx.needs_input_grad[1]) else None
        return g0, g1, None, None, None


def gather_rows(in0: Tensor, in1: Optional[Tensor], idx: Tensor, inv0: Tensor,
                inv1: Optional[Tensor] = None) -> Tensor:
    """out[..., i] = idx[i] < 0 ? 0 : cat(in0, in1)[..., idx[i]].  ``inv0``/``inv1`` are the
    adjoint maps: for source element s of in0 (in1), the output slot that reads it, or -1."""
    return _GatherRows.apply(in0, in1, idx, inv0, inv1)


class _Gram(Function):
    @staticmethod
    def forward(ctx, *blocks):
        G = None
        i = 0
        while i < len(blocks):                      # pairs of blocks share one launch
            if i + 1 < len(blocks):
                G = kernels.gram_fwd2(blocks[i], blocks[i + 1], G)
                i += 2
            else:
                G = kernels.gram_fwd(blocks[i], G)
                i += 1
        ctx.save_for_backward(*blocks)
        return G

    @staticmethod
    def backward(ctx, g_G):
        g_G = g_G.contiguous()
        return tuple(kernels.gram_bwd(V, g_G) if need else None
                     for V, need in zip(ctx.saved_tensors, ctx.needs_input_grad))


def gram(blocks: Sequence[Tensor]) -> Tensor:
    """G = sum_i V_i V_i^T over the compact blocks V_i (B,d,n_i) of J^T (non_square.py:307-308)."""
    return _Gram.apply(*blocks)


class _CholLogdetL1(Function):
    @staticmethod
    def forward(ctx, G, penalty):
        L, logdet, l1_off, l1_diag, info = kernels.chol_logdet_fwd(G)
        ctx.save_for_backward(L, G)
        ctx.penalty = penalty
        l1 = l1_off if penalty == "offdiag" else l1_diag
        ctx.mark_non_differentiable(info)
        ctx.set_materialize_grads(False)
        return logdet.unsqueeze(1), l1.unsqueeze(1), info

    @staticmethod
    def backward(ctx, g_logdet, g_l1, _g_info):
        L, G = ctx.saved_tensors
        B = G.shape[0]
        if g_logdet is None:
            g_logdet = torch.zeros(B, dtype=G.dtype, device=G.device)
        penalty = ctx.penalty if g_l1 is not None else "none"
        g_G = kernels.chol_logdet_bwd(L, G, g_logdet.reshape(B).contiguous(),
                                      None if g_l1 is None else g_l1.reshape(B).contiguous(), penalty)
        return g_G, None


def chol_logdet_l1(G: Tensor, penalty: str = "none") -> Tuple[Tensor, Tensor]:
    """(log det G (B,1), metric penalty (B,1)) for penalty in {"none","offdiag","diag"}.

    Raises :class:`NotPositiveDefiniteError` exactly where the reference's
    ``torch.linalg.cholesky`` raises, so the caller can replay the ``+eps*I`` retry loop of
    `non_square.py:280-288`.  Reading ``info`` is the one host sync of the path (the reference
    syncs at the same place: cuSOLVER's potrf info check)."""
    if penalty not in kernels.PENALTY:
        raise ValueError(f"unknown penalty {penalty}")
    logdet, l1, info = _CholLogdetL1.apply(G, penalty)
    if _DeferredInfo.active:
        _DeferredInfo.pending.append(info)
        bad = 0
    else:
        bad = int(info.item())
    if bad:
        raise NotPositiveDefiniteError(
            f"cmf_chol_logdet_fwd: leading minor not positive for sample {bad - 1}")
    if penalty == "none":
        l1 = torch.zeros_like(logdet)
    return logdet, l1


class _RqsCouplingJVP(Function):
    @staticmethod
    def forward(ctx, z_m, V_m, cw, ch, dv, CW, CH, DV, tail_bound):
        x_m, V_out, bins = kernels.rqs_coupling_jvp_fwd(z_m, V_m, cw, ch, dv, CW, CH, DV, tail_bound)
        ctx.save_for_backward(z_m, V_m, cw, ch, dv, CW, CH, DV)
        ctx.tail_bound = tail_bound
        ctx.set_materialize_grads(False)
        ctx.mark_non_differentiable(bins)
        if V_out is None:
            return x_m, None, bins
        return x_m, V_out, bins

    @staticmethod
    def backward(ctx, g_x, g_V, _g_bins):
        z_m, V_m, cw, ch, dv, CW, CH, DV = ctx.saved_tensors
        if V_m is not None and g_V is None:
            g_V = torch.zeros_like(V_m)
        out = kernels.rqs_coupling_jvp_bwd(z_m, V_m, cw, ch, dv, CW, CH, DV, ctx.tail_bound, g_x, g_V)
        return (*out, None)


def rqs_coupling_jvp(z_m: Tensor, V_m: Optional[Tensor], cw: Tensor, ch: Tensor, dv: Tensor,
                     CW: Optional[Tensor], CH: Optional[Tensor], DV: Optional[Tensor], tail_bound: float):
    """Inverse rational-quadratic spline (z -> x) on the transformed half, primal + d tangents.

    Knot tables from :func:`cmf_b200.splines.knot_tables`.  Returns ``(x_m, V_out, bin)``; ``bin``
    (int32, -1 outside the interval) is bit-identical to the reference's
    ``utils.searchsorted(cumheights, inputs)`` (`gitmodules/nsf/utils/torchutils.py:133-138`)."""
    return _RqsCouplingJVP.apply(z_m, V_m, cw, ch, dv, CW, CH, DV, tail_bound)


class _Conv3x3C64(Function):
    """3x3 / 64 -> 64 / pad 1 convolution on the tcgen05 kernels.  The data gradient is the forward
    kernel on the transposed, flipped weights; the weight gradient is the pixels-as-K kernel
    (library ``aten::convolution_backward`` only for image widths it does not cover)."""

    @staticmethod
    def forward(ctx, x, weight, bias, wsplit):
        if wsplit is None:
            wsplit = kernels.conv3x3_c64_prep(weight, False)
        ctx.save_for_backward(x, weight)
        ctx.has_bias = bias is not None
        return kernels.conv3x3_c64(x, wsplit, bias)

    @staticmethod
    def backward(ctx, g_y):
        x, weight = ctx.saved_tensors
        g_y = g_y.contiguous()
        need_x, need_w = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        g_x = kernels.conv3x3_c64(g_y, kernels.conv3x3_c64_prep(weight, True), None, drain=2) if need_x else None   # a gradient: see networks._BWD_DRAIN
        g_w = None
        if need_w and kernels.conv3x3_c64_wgrad_usable(x.shape[0], x.shape[2], x.shape[3]):
            g_w = kernels.conv3x3_c64_wgrad(x, g_y)
        elif need_w:
            g_w = torch.ops.aten.convolution_backward(g_y, x, weight, None, (1, 1), (1, 1), (1, 1), False, (0, 0), 1,
                                                      (False, True, False))[1]
        g_b = kernels.channel_sum(g_y) if (ctx.has_bias and ctx.needs_input_grad[2]) else None
        return g_x, g_w, g_b, None


def conv3x3_c64(x: Tensor, weight: Tensor, bias: Optional[Tensor] = None, wsplit: Optional[Tensor] = None) -> Tensor:
    """``F.conv2d(x, weight, bias, padding=1)`` for weight (64,64,3,3) (`networks.py:35-94`: the
    residual-block convolutions of the coupler ResNet).  ``wsplit`` = a cached
    ``kernels.conv3x3_c64_prep(weight, False)`` shared by the primal and the tangent call."""
    return _Conv3x3C64.apply(x, weight, bias, wsplit)


def conv3x3_c64_usable(x: Tensor, weight: Tensor) -> bool:
    return (x.dim() == 4 and tuple(weight.shape) == (64, 64, 3, 3) and x.shape[1] == 64
            and x.dtype == torch.float32 and kernels.conv3x3_c64_usable(x.shape[0], x.shape[2], x.shape[3]))


# ------------------------------------------------------------------ encode pass / loss epilogue (N4)
class _CouplingXToZ(Function):
    @staticmethod
    def forward(ctx, x, st, pos2j, qmap, q2pos):
        z, lj = kernels.coupling_x_to_z_fwd(x, st, pos2j, qmap)
        ctx.save_for_backward(z, st)
        ctx.maps = (pos2j, qmap, q2pos)
        ctx.set_materialize_grads(False)
        return z, lj

    @staticmethod
    def backward(ctx, g_z, g_lj):
        z, st = ctx.saved_tensors
        pos2j, qmap, q2pos = ctx.maps
        g_x, g_st = kernels.coupling_x_to_z_bwd(z, st, pos2j, qmap, q2pos, g_z, None if g_lj is None else g_lj.reshape(-1))
        return g_x, g_st, None, None, None


def coupling_x_to_z(x: Tensor, st: Tensor, pos2j: Tensor, qmap: Optional[Tensor], q2pos: Tensor) -> Tuple[Tensor, Tensor]:
    """Affine coupling in the encode direction on the flattened layer input (`acl.py:43-46,101-111`):
    z = x at passthrough positions, (x + t) e^s at modified ones; log-jac (B,1) = sum s.  st (B,2Q) =
    [shift | log-scale] as the coupler emits it; the int32 maps relate flat positions, compact indices and
    conditioner positions (see include/cmf_b200.h)."""
    return _CouplingXToZ.apply(x, st, pos2j, qmap, q2pos)


class _SumSqDiff(Function):
    @staticmethod
    def forward(ctx, a, c):
        ctx.save_for_backward(a, c)
        return kernels.sumsq_diff_fwd(a, c)

    @staticmethod
    def backward(ctx, g):
        a, c = ctx.saved_tensors
        g_a = kernels.sumsq_diff_bwd(a, c, g.reshape(-1).contiguous())
        return (g_a if ctx.needs_input_grad[0] else None), (-g_a if (c is not None and ctx.needs_input_grad[1]) else None)


def sumsq_diff(a: Tensor, c: Optional[Tensor] = None) -> Tensor:
    """(B,1): sum over everything but the batch of (a - c)^2 -- the reconstruction error of
    `non_square.py:111-114` and the quadratic form of `gaussian.py:9-22`."""
    a2 = a.flatten(1)
    return _SumSqDiff.apply(a2, None if c is None else c.flatten(1))


# ------------------------------------------------------------------ whole-stack kernel, small flat configs (N1-ii)
class _FlatStackGram(Function):
    @staticmethod
    def forward(ctx, z, params, desc):
        G, x = kernels.flat_stack_fwd(desc, z, params)
        ctx.save_for_backward(z, params)
        ctx.desc = desc
        ctx.set_materialize_grads(False)
        return G, x

    @staticmethod
    def backward(ctx, g_G, g_x):
        z, params = ctx.saved_tensors
        g_z, g_params = kernels.flat_stack_bwd(ctx.desc, z, params, None if g_G is None else g_G.contiguous(),
                                               None if g_x is None else g_x.contiguous())
        return g_z, g_params, None


def flat_stack_gram(z: Tensor, params: Tensor, desc) -> Tuple[Tensor, Tensor]:
    """(G (B,d,d), x (B,D)) of a small flat flow in one launch: tail pad, every coupling layer with its
    tanh-MLP conditioner (primal + d tangents) and J^T J (`non_square.py:298-311,397-410`,
    `acl.py:125-146`, `networks.py:206-224`).  ``desc`` from :class:`cmf_b200.program.FlatStackPlan`."""
    return _FlatStackGram.apply(z, params, desc)


# ------------------------------------------------------------------ whole tanh-MLP conditioner, flat couplers (N1-ii)
class _MlpJvp(Function):
    @staticmethod
    def forward(ctx, x, V, nlin, *wb):
        weights, biases = wb[:nlin], wb[nlin:]
        wpack, widths = kernels.mlp_pack(weights, biases)
        need = any(ctx.needs_input_grad)
        st, ST, saved = kernels.mlp_jvp_fwd(x, V, wpack, widths, save=need)
        if need:
            tensors = [x, wpack, *saved] + ([V] if V is not None else [])
            ctx.save_for_backward(*tensors)
        ctx.widths, ctx.has_V = widths, V is not None
        ctx.set_materialize_grads(False)
        return st, ST

    @staticmethod
    def backward(ctx, g_st, g_ST):
        x, wpack, save_T, save_U, *rest = ctx.saved_tensors
        V = rest[0] if ctx.has_V else None
        if g_st is None:
            g_st = torch.zeros(x.shape[0], ctx.widths[-1], dtype=x.dtype, device=x.device)
        g_x, g_V, g_W, g_b = kernels.mlp_jvp_bwd(x, V, wpack, ctx.widths, (save_T, save_U), g_st.contiguous(),
                                                 None if g_ST is None else g_ST.contiguous())
        return (g_x, g_V, None, *g_W, *g_b)


def mlp_jvp(x: Tensor, V: Optional[Tensor], weights, biases) -> Tuple[Tensor, Optional[Tensor]]:
    """A tanh MLP on the primal x (B, n_in) and, if given, on the tangents V (B, d, n_in) by the
    reference's JVP rules (`networks.py:24-32,206-224`, `jvp_layers.py:38-47`), the whole stack in
    one launch; its adjoint (inputs, weights, biases) in one launch plus a fixed-order reduction."""
    return _MlpJvp.apply(x, V, len(weights), *weights, *biases)


# ------------------------------------------------------------------ tail of the image couplers' conditioner (N1-i)
class _ResnetTail(Function):
    @staticmethod
    def forward(ctx, h, Hv, inv, P, w2, b2, sw, sb):
        y, Y = kernels.resnet_tail_fwd(h, Hv, inv, P, w2, b2, sw, sb)
        tensors = [h, w2, b2, sw, sb] + ([Hv] if Hv is not None else []) + ([inv] if inv is not None else [])
        ctx.save_for_backward(*tensors)
        ctx.has_Hv, ctx.has_inv, ctx.P = Hv is not None, inv is not None, P
        ctx.set_materialize_grads(False)
        return y, Y

    @staticmethod
    def backward(ctx, g_y, g_Y):
        h, w2, b2, sw, sb, *rest = ctx.saved_tensors
        Hv = rest.pop(0) if ctx.has_Hv else None
        inv = rest.pop(0) if ctx.has_inv else None
        g_h, g_Hv, g_w2, g_b2, g_sw, g_sb = kernels.resnet_tail_bwd(
            h, Hv, inv, ctx.P, w2, b2, sw, sb, None if g_y is None else g_y.contiguous(), None if g_Y is None else g_Y.contiguous())
        return g_h, g_Hv, None, None, g_w2, g_b2, g_sw, g_sb


def resnet_tail(h: Tensor, Hv: Optional[Tensor], inv: Optional[Tensor], P: int, w2: Tensor, b2: Tensor, sw: Tensor, sb: Tensor):
    """ReLU -> 1x1 convolution -> scaled tanh of the coupler ResNet (`networks.py:96-113,156-161`) on the primal
    hidden activations h (B, C, HW) and the tangents Hv (B d, C, HW), at the positions ``inv`` selects:
    y (B, O, P), Y (B, d, O, P)."""
    return _ResnetTail.apply(h, Hv, inv, P, w2, b2, sw, sb)


# ------------------------------------------------------------------ first convolution of the image couplers' conditioner (N1-i)
class _ConvIn(Function):
    @staticmethod
    def forward(ctx, x, weight):
        ctx.save_for_backward(x, weight)
        return kernels.convin_fwd(x, weight)

    @staticmethod
    def backward(ctx, g_y):
        x, weight = ctx.saved_tensors
        g_y = g_y.contiguous()
        g_x = kernels.convin_dgrad(g_y, weight) if ctx.needs_input_grad[0] else None
        g_w = kernels.convin_wgrad(x, g_y) if ctx.needs_input_grad[1] else None
        return g_x, g_w


def conv_in(x: Tensor, weight: Tensor) -> Tensor:
    """3x3 convolution C_in -> 64, padding 1, no bias (`networks.py:131-139`): x (N, C_in, H, W) -> (N, 64, H, W)."""
    return _ConvIn.apply(x, weight)
