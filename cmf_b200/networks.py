"""Conditioner networks of the affine coupling layers, with *tangent-batched* JVPs.

State-dict names and module indices match the reference (`networks.py:116-161,206-224`,
`couplers.py:27-59`) so reference checkpoints load unchanged.  The JVP rules are the
reference's (`jvp_layers.py:38-64`, `networks.py:62-79,108-113`): linear/conv tangents drop the
bias, relu -> (x>0)*v, tanh -> (1-y^2)*v, scaled tanh -> w*v.  The difference is the schedule:
the reference pushes ONE tangent column per call and recomputes the primal every time
(`non_square.py:302-306`, d calls per layer); here the primal runs once on (B, ...) and all d
tangents run as one library call on a (B*d, ...) tensor.  These are cuBLAS / cuDNN library calls
 and ordinary autograd nodes; the hand-written kernels start where their outputs
are consumed (``cmf_b200.ops.coupling_jvp``).
"""
from __future__ import annotations

from typing import Sequence, Tuple

import os

import torch
import torch.nn as nn
import torch.nn.functional as F

Tensor = torch.Tensor

# CMF_CONV=cudnn keeps every conditioner convolution on the library path (A/B measurements)
_TC_CONV = os.environ.get("CMF_CONV", "tc") != "cudnn"
# K = 32 stages per FP32 drain in the data-gradient launches of the backward pass.  Forward values
# pass through G^-1 (errors amplified by cond(G) ~ 5e3) and keep the per-stage drain; a gradient is
# consumed as is, so 4e-7 instead of 2e-7 relative error buys 7 % of those launches.
_BWD_DRAIN = int(os.environ.get("CMF_BWD_DRAIN", "2"))
# drain interval of the forward launches (primal and tangent push); see DESIGN.md 3a for the measurements
_FWD_DRAIN = int(os.environ.get("CMF_FWD_DRAIN", "1"))


# CMF_CONVIN=cudnn keeps the first (C_in -> 64) convolution of the image couplers on F.conv2d
_CONVIN_OWN = os.environ.get("CMF_CONVIN", "own") != "cudnn"
# CMF_TAIL=torch keeps the ReLU -> 1x1 conv -> scaled tanh tail of the image couplers on library ops
_TAIL_FUSED = os.environ.get("CMF_TAIL", "fused") != "torch"
# CMF_MLP=torch keeps the tanh-MLP conditioners of the flat couplers on F.linear + elementwise ops
_MLP_FUSED = os.environ.get("CMF_MLP", "fused") != "torch"


class recompute_activations:
    """Activation recomputation for the residual-stack node: inside this context (or with
    ``CMF_RECOMPUTE=1`` in the environment) ``_ResidualStackJoint`` keeps only the stack's INPUT (primal
    and tangents) for the backward pass and re-runs its forward launches there, instead of keeping two
    (B d + B) x 64 x H x W buffers per residual block (51 GiB at the MNIST-shape batch of 50; the
    CIFAR-shape d = 512 sweep fits B = 6 without this).  Costs one extra forward of the stack in the
    backward pass; the gradients are bit-identical (same launches on the same data)."""
    depth = 1 if os.environ.get("CMF_RECOMPUTE", "0") == "1" else 0

    def __enter__(self):
        recompute_activations.depth += 1

    def __exit__(self, *exc):
        recompute_activations.depth -= 1
        return False


class library_convs:
    """Inside this context every conditioner convolution is the library call ``F.conv2d`` and every
    MLP conditioner the ``F.linear`` stack (which autograd can differentiate twice).  Used for the primal run whose VJP is itself differentiated
    (`non_square.py:199`, ``create_graph=True`` of hutch_with_cg): the tcgen05 convolution ops carry
    first-order adjoints only."""
    depth = 0

    def __enter__(self):
        library_convs.depth += 1

    def __exit__(self, *exc):
        library_convs.depth -= 1
        return False


def _tc_conv_enabled() -> bool:
    return _TC_CONV and library_convs.depth == 0


def _conv3x3_pair(conv: nn.Conv2d, h: Tensor, Hv):
    """``conv(h)`` and the bias-free tangent push ``conv.weight * Hv`` (`jvp_layers.py:49-64`).
    64 -> 64 channel 3x3 convolutions run on the tcgen05 kernel (weights pre-split once for both
    calls); other shapes are library calls."""
    from . import kernels, ops
    if _tc_conv_enabled() and ops.conv3x3_c64_usable(h, conv.weight):
        ws = kernels.conv3x3_c64_prep(conv.weight.detach(), False)
        y = ops.conv3x3_c64(h, conv.weight, conv.bias, ws)
        if Hv is not None and ops.conv3x3_c64_usable(Hv, conv.weight):
            return y, ops.conv3x3_c64(Hv, conv.weight, None, ws)
        Y = None if Hv is None else F.conv2d(Hv, conv.weight, None, padding=1)
        return y, Y
    return conv(h), (None if Hv is None else F.conv2d(Hv, conv.weight, None, padding=1))


def _bcast(t: Tensor) -> Tensor:
    """(B, ...) -> (B, 1, ...) so a primal-derived factor broadcasts over the tangent axis."""
    return t.unsqueeze(1)


class TangentMLP(nn.Sequential):
    """Linear / activation stack (`networks.py:206-224`); keys ``{2i}.weight``, ``{2i}.bias``."""

    def __init__(self, num_in: int, hidden: Sequence[int], num_out: int, activation: str):
        act = {"tanh": nn.Tanh, "relu": nn.ReLU}[activation]
        layers, prev = [], num_in
        for h in hidden:
            layers += [nn.Linear(prev, h), act()]
            prev = h
        layers.append(nn.Linear(prev, num_out))
        super().__init__(*layers)
        self.activation = activation
        self.widths = [num_in, *hidden, num_out]

    def _fused(self, x: Tensor, d: int) -> bool:
        """the whole stack as one launch of the hand-written kernel (``ops.mlp_jvp``)"""
        from . import kernels
        return (_MLP_FUSED and library_convs.depth == 0 and self.activation == "tanh"
                and kernels.mlp_usable(self.widths, d, x))

    def _linears(self):
        lin = [m for m in self if isinstance(m, nn.Linear)]
        return [m.weight for m in lin], [m.bias for m in lin]

    def forward(self, x: Tensor) -> Tensor:
        if self._fused(x, 0):
            from . import ops
            return ops.mlp_jvp(x, None, *self._linears())[0]
        return super().forward(x)

    def jvp_batched(self, x: Tensor, V: Tensor) -> Tuple[Tensor, Tensor]:
        """x (B,p), V (B,d,p) -> y (B,o), Y (B,d,o)."""
        if self._fused(x, V.shape[1]):
            from . import ops
            return ops.mlp_jvp(x, V, *self._linears())
        for mod in self:
            if isinstance(mod, nn.Linear):
                x, V = mod(x), F.linear(V, mod.weight)
            elif isinstance(mod, nn.Tanh):
                x = torch.tanh(x)
                V = V * _bcast(1 - x * x)
            elif isinstance(mod, nn.ReLU):
                V = V * _bcast((x > 0).to(V.dtype))
                x = torch.relu(x)
            else:
                raise ValueError(f"jvp unsupported for module {type(mod).__name__}")
        return x, V


class ResidualBlock(nn.Module):
    """relu-conv3x3-relu-conv3x3 + skip, no batch norm (`networks.py:35-94` with
    ``use_batchnorm=False``, the only variant the non-square image config builds)."""

    def __init__(self, channels: int):
        super().__init__()
        self.conv1 = nn.Conv2d(channels, channels, 3, padding=1, bias=True)
        self.conv2 = nn.Conv2d(channels, channels, 3, padding=1, bias=True)

    def forward(self, x: Tensor) -> Tensor:
        h = x
        for conv in (self.conv1, self.conv2):
            h, _ = _conv3x3_pair(conv, torch.relu(h), None)
        return h + x

    def jvp_batched(self, x: Tensor, V: Tensor, d: int) -> Tuple[Tensor, Tensor]:
        """x (B,C,H,W), V (B*d,C,H,W)."""
        h, Hv = x, V
        for conv in (self.conv1, self.conv2):
            Hv = _mask_rows(Hv, h > 0, d)
            h, Hv = _conv3x3_pair(conv, torch.relu(h), Hv)
        return h + x, Hv + V


class _ResidualStackTangent(torch.autograd.Function):
    """Tangent push through a run of residual blocks (`networks.py:62-79` applied per block) with
    every elementwise neighbour of the convolutions folded into a convolution epilogue:

        A0 = m0 * V;  A1 = m1 * conv1(A0);  S = conv2(A1) + V;  next A0 = m0' * S      (m = primal > 0)

    forward : conv1 epilogue applies m1; conv2 epilogue adds the skip and also emits the masked copy
              the next block consumes (two outputs, one pass).
    backward: g_C1 = m1 * dgrad2(g) (epilogue mask), g = m0 * dgrad1(g_C1) + g (mask, then skip add);
              weight gradients from the saved masked activations.
    The masks come from the PRIMAL activations (not differentiated: the reference's relu JVP
    `jvp_layers.py:38-46` is piecewise constant in them)."""

    @staticmethod
    def forward(ctx, V, d, *tensors):
        from . import kernels
        K = len(tensors) // 4
        w1s, w2s = tensors[2::4], tensors[3::4]
        S = V.contiguous()
        A0 = _mask_rows(S, tensors[0] > 0, d)
        # the relu-JVP masks as packed sign bits (one 8-byte word per pixel): read by the epilogues of
        # the forward and the backward launches, and 32x smaller to keep for the backward
        xs = [kernels.relu_mask_pack(t) for t in tensors[0::4]]
        h1s = [kernels.relu_mask_pack(t) for t in tensors[1::4]]
        saved = []
        for k in range(K):
            A1 = kernels.conv3x3_c64(A0, kernels.conv3x3_c64_prep(w1s[k], False), None, mask_src=h1s[k], mask_div=d,
                                     mask_first=True)
            ws2 = kernels.conv3x3_c64_prep(w2s[k], False)
            saved += [A0, A1]
            if k + 1 < K:
                S, A0 = kernels.conv3x3_c64(A1, ws2, None, mask_src=xs[k + 1], mask_div=d, res=S, second=True)
            else:
                S = kernels.conv3x3_c64(A1, ws2, None, res=S)
        ctx.d, ctx.K = d, K
        ctx.save_for_backward(*xs, *h1s, *w1s, *w2s, *saved)
        return S

    @staticmethod
    def backward(ctx, g):
        from . import kernels
        K, d = ctx.K, ctx.d
        t = ctx.saved_tensors
        xs, h1s, w1s, w2s, saved = t[:K], t[K:2 * K], t[2 * K:3 * K], t[3 * K:4 * K], t[4 * K:]
        g = g.contiguous()
        grads = [None] * (4 * K)
        for k in reversed(range(K)):
            A0, A1 = saved[2 * k], saved[2 * k + 1]
            grads[4 * k + 3] = kernels.conv3x3_c64_wgrad(A1, g)
            gC1 = kernels.conv3x3_c64(g, kernels.conv3x3_c64_prep(w2s[k], True), None, mask_src=h1s[k], mask_div=d,
                                      mask_first=True, drain=_BWD_DRAIN)
            grads[4 * k + 2] = kernels.conv3x3_c64_wgrad(A0, gC1)
            g = kernels.conv3x3_c64(gC1, kernels.conv3x3_c64_prep(w1s[k], True), None, mask_src=xs[k], mask_div=d,
                                    mask_first=True, res=g, drain=_BWD_DRAIN)
        return (g, None, *grads)


class _ResidualStackJoint(torch.autograd.Function):
    """Primal AND tangents through a run of residual blocks as ONE autograd node on the tcgen05
    convolution kernels (`networks.py:62-79` of the reference applied per block, for the primal and
    for every tangent column).  Per block, with m0 = (x > 0), m1 = (h1 > 0) taken from the primal:

        primal   :  h1 = conv1(m0 x) + b1;      x' = conv2(m1 h1) + b2 + x
        tangents :  C1 = conv1(m0 V);           V' = conv2(m1 C1) + V

    Both rows have the same structure, so every saved activation is ONE buffer of B*d + B images
    (tangent images first, the B primal images last) and the backward pass runs each data-gradient
    and each weight-gradient launch once on that joint batch (`cmf_conv3x3_c64_joint`): the weight
    gradient of a convolution is the sum of its primal and tangent uses, which is exactly what one
    launch over the concatenated images accumulates.  That removes the 2/3 of the N = B launches that
    the separate primal autograd convolutions used to issue (they fill a third of the SMs).
    In the forward pass the tangent launch of a convolution needs the primal result of the SAME
    convolution (its relu-JVP mask), so the primal still runs one small launch ahead.
    The masks come from the primal activations and are not differentiated (the reference's relu JVP,
    `jvp_layers.py:38-46`, is piecewise constant in them)."""

    @staticmethod
    def _run(h, V, d, tensors):
        """the forward launches; returns (x_out, S_out, xbits, hbits, [A0_0, A1_0, A0_1, ...])"""
        from . import kernels
        K = len(tensors) // 4
        B, Nt = h.shape[0], V.shape[0]
        shape = (Nt + B, *h.shape[1:])
        x, S = h.contiguous(), V.contiguous()
        new = lambda: torch.empty(shape, dtype=h.dtype, device=h.device)
        A0 = new()
        xb = kernels.relu_mask_pack(x, relu_out=A0[Nt:])
        torch.mul(S.view(B, d, *S.shape[1:]), _bcast((x > 0).to(S.dtype)), out=A0[:Nt].view(B, d, *S.shape[1:]))
        xbits, hbits, saved = [], [], []
        for k in range(K):
            w1, b1, w2, b2 = tensors[4 * k:4 * k + 4]
            ws1, ws2 = kernels.conv3x3_c64_prep(w1, False), kernels.conv3x3_c64_prep(w2, False)
            h1 = kernels.conv3x3_c64(A0[Nt:], ws1, b1, drain=_FWD_DRAIN)                 # primal, one launch ahead
            A1 = new()
            hb = kernels.relu_mask_pack(h1, relu_out=A1[Nt:])
            kernels.conv3x3_c64(A0[:Nt], ws1, None, mask_src=hb, mask_div=d, mask_first=True, out=A1[:Nt], drain=_FWD_DRAIN)
            x = kernels.conv3x3_c64(A1[Nt:], ws2, b2, res=x, drain=_FWD_DRAIN)
            xbits.append(xb); hbits.append(hb); saved += [A0, A1]
            if k + 1 < K:
                A0 = new()
                xb = kernels.relu_mask_pack(x, relu_out=A0[Nt:])
                S, _ = kernels.conv3x3_c64(A1[:Nt], ws2, None, mask_src=xb, mask_div=d, res=S, second=True, out2=A0[:Nt],
                                           drain=_FWD_DRAIN)
            else:
                S = kernels.conv3x3_c64(A1[:Nt], ws2, None, res=S, drain=_FWD_DRAIN)
        return x, S, xbits, hbits, saved

    @staticmethod
    def forward(ctx, h, V, d, *tensors):
        K = len(tensors) // 4
        x, S, xbits, hbits, saved = _ResidualStackJoint._run(h, V, d, tensors)
        ctx.d, ctx.K, ctx.Nt = d, K, V.shape[0]
        ctx.recompute = recompute_activations.depth > 0
        if ctx.recompute:
            ctx.save_for_backward(h, V, *tensors)
        else:
            ctx.save_for_backward(*xbits, *hbits, *tensors[0::4], *tensors[2::4], *saved)
        return x, S

    @staticmethod
    def backward(ctx, g_h, g_S):
        from . import kernels
        K, d, Nt = ctx.K, ctx.d, ctx.Nt
        t = ctx.saved_tensors
        if ctx.recompute:
            tensors = t[2:]
            _, _, xbits, hbits, saved = _ResidualStackJoint._run(t[0], t[1], d, tensors)
            w1s, w2s = tensors[0::4], tensors[2::4]
        else:
            xbits, hbits, w1s, w2s, saved = t[:K], t[K:2 * K], t[2 * K:3 * K], t[3 * K:4 * K], t[4 * K:]
        g = torch.empty_like(saved[0])
        for part, src in ((g[:Nt], g_S), (g[Nt:], g_h)):
            if src is None:
                part.zero_()
            else:
                part.copy_(src)
        grads = [None] * (4 * K)
        for k in reversed(range(K)):
            A0, A1 = saved[2 * k], saved[2 * k + 1]
            grads[4 * k + 2] = kernels.conv3x3_c64_wgrad(A1, g)
            grads[4 * k + 3] = kernels.channel_sum(g[Nt:])
            gC1 = kernels.conv3x3_c64(g, kernels.conv3x3_c64_prep(w2s[k], True), None, mask_src=hbits[k], mask_div=d,
                                      mask_first=True, drain=_BWD_DRAIN, tail=Nt)
            grads[4 * k + 0] = kernels.conv3x3_c64_wgrad(A0, gC1)
            grads[4 * k + 1] = kernels.channel_sum(gC1[Nt:])
            g = kernels.conv3x3_c64(gC1, kernels.conv3x3_c64_prep(w1s[k], True), None, mask_src=xbits[k], mask_div=d,
                                    mask_first=True, res=g, drain=_BWD_DRAIN, tail=Nt)
        return (g[Nt:], g[:Nt], None, *grads)


def _residual_stack_jvp(blocks, h: Tensor, Hv: Tensor, d: int):
    """primal + tangents through consecutive residual blocks: one joint autograd node on the tcgen05
    kernels when the shapes qualify, the per-block library path otherwise"""
    from . import kernels
    N, _, H, W = Hv.shape
    fused = (_tc_conv_enabled() and all(b.conv1.weight.shape == (64, 64, 3, 3) for b in blocks) and Hv.shape[1] == 64
             and h.dtype == torch.float32 and N == h.shape[0] * d
             and kernels.conv3x3_c64_usable(N + h.shape[0], H, W) and kernels.conv3x3_c64_wgrad_usable(N + h.shape[0], H, W))
    if not fused:
        for b in blocks:
            h, Hv = b.jvp_batched(h, Hv, d)
        return h, Hv
    args = []
    for b in blocks:
        args += [b.conv1.weight, b.conv1.bias, b.conv2.weight, b.conv2.bias]
    return _ResidualStackJoint.apply(h, Hv, d, *args)


def _mask_rows(V: Tensor, keep: Tensor, d: int) -> Tensor:
    """V (B*d, ...) * keep (B, ...) broadcast over the d tangents of each sample."""
    B = keep.shape[0]
    return (V.view(B, d, *V.shape[1:]) * _bcast(keep.to(V.dtype))).view_as(V)


class ScaledTanhResNet(nn.Module):
    """`networks.py:116-161` (get_resnet, no BN) inside `networks.py:96-113` (ScaledTanh2dModule).
    Keys: ``weights``, ``bias``, ``module.0.weight``, ``module.{i}.conv{1,2}.*``, ``module.{n+2}.*``."""

    def __init__(self, num_in: int, hidden: Sequence[int], num_out: int):
        super().__init__()
        hidden = list(hidden)
        width = hidden[0] if hidden else num_out
        layers = [nn.Conv2d(num_in, width, 3, padding=1, bias=False)]
        for width in hidden:
            layers.append(ResidualBlock(width))
        layers += [nn.ReLU(), nn.Conv2d(width, num_out, 1, bias=True)]
        self.module = nn.Sequential(*layers)
        self.weights = nn.Parameter(torch.ones(num_out, 1, 1))
        self.bias = nn.Parameter(torch.zeros(num_out, 1, 1))

    def _convin_own(self, mod, x: Tensor) -> bool:
        """the first convolution on the hand-written kernels (``ops.conv_in``)"""
        from . import kernels
        return (_CONVIN_OWN and library_convs.depth == 0 and isinstance(mod, nn.Conv2d) and mod.bias is None
                and mod.padding == (1, 1) and mod.stride == (1, 1) and mod.dilation == (1, 1) and mod.groups == 1
                and kernels.convin_usable(x, mod.weight))

    def _body(self, x: Tensor) -> Tensor:
        """everything before the last ReLU, on the primal alone"""
        mods = list(self.module[:-2])
        if mods and self._convin_own(mods[0], x):
            from . import ops
            x, mods = ops.conv_in(x, mods[0].weight), mods[1:]
        for mod in mods:
            x = mod(x)
        return x

    def forward(self, x: Tensor) -> Tensor:
        if self._tail_fused(x):
            return self._tail_apply(self._body(x), None, None)[0]
        return self.weights * torch.tanh(self.module(x)) + self.bias

    # -- compact tail ------------------------------------------------------------------------
    # A checkerboard coupling layer uses the conditioner output at half of the spatial positions
    # only (`acl.py:53-66`: everything is multiplied by 1 - mask).  The tail of the net
    # (ReLU -> 1x1 conv -> scaled tanh, `networks.py:156-161,96-113`) acts per position, so it is
    # evaluated at those positions alone: the hidden activations are gathered once (cmf_gather_rows,
    # fused with the ReLU mask product the JVP rule needs anyway) and the coupling kernel receives
    # COMPACT (shift, log-scale) tensors -- no half-used sectors, no zero-filled adjoints.
    supports_compact_output = True

    def _tail(self):
        relu, last = self.module[-2], self.module[-1]
        assert isinstance(relu, nn.ReLU) and isinstance(last, nn.Conv2d) and last.kernel_size == (1, 1)
        return last

    def _tail_fused(self, x: Tensor) -> bool:
        """the tail as one launch of the hand-written kernel (``ops.resnet_tail``)"""
        from . import kernels
        last = self.module[-1]
        return (_TAIL_FUSED and library_convs.depth == 0 and isinstance(last, nn.Conv2d) and last.kernel_size == (1, 1)
                and last.bias is not None and kernels.resnet_tail_usable(x, last.weight.shape[1], last.weight.shape[0]))

    def _tail_apply(self, h: Tensor, Hv, pos):
        """h (B,C,H,W) hidden activations before the last ReLU, Hv (B*d,C,H,W) their tangents or None, pos as in
        :meth:`jvp_batched` -> (y, Y): compact (B,O,P) / (B,d,O,P) with ``pos``, else (B,O,H,W) / (B,d,O,H,W)"""
        from . import ops
        last = self._tail()
        B, _, H, W = h.shape
        O = last.weight.shape[0]
        inv, P = (None, H * W) if pos is None else (pos[1], pos[0].numel())
        y, Y = ops.resnet_tail(h.flatten(2), None if Hv is None else Hv.flatten(2), inv, P, last.weight.flatten(1), last.bias,
                               self.weights.flatten(), self.bias.flatten())
        if pos is None:
            y = y.view(B, O, H, W)
            Y = None if Y is None else Y.view(B, -1, O, H, W)
        return y, Y

    def forward_compact(self, x: Tensor, pos) -> Tensor:
        """x (B,C,H,W), pos = (idx int32 (P,), inv int32 (H*W,), pos int64 (P,)) -> y (B,O,P)."""
        if self._tail_fused(x):
            return self._tail_apply(self._body(x), None, pos)[0]
        pos = pos[2]
        last = self._tail()
        h = self.module[:-2](x).flatten(2).index_select(2, pos)
        y = torch.matmul(last.weight.flatten(1), torch.relu(h)) + last.bias.view(1, -1, 1)
        return self.weights.view(1, -1, 1) * torch.tanh(y) + self.bias.view(1, -1, 1)

    def jvp_batched(self, x: Tensor, V: Tensor, pos=None) -> Tuple[Tensor, Tensor]:
        """x (B,C,H,W), V (B,d,C,H,W) -> y (B,O,H,W), Y (B,d,O,H,W); with ``pos`` = (idx int32 (P,),
        inv int32 (H*W,), pos int64 (P,)) the outputs are compact: y (B,O,P), Y (B,d,O,P)."""
        B, d = V.shape[:2]
        Hv = V.reshape(B * d, *V.shape[2:])
        h = x
        fused_tail = self._tail_fused(x)
        body = list(self.module if (pos is None and not fused_tail) else self.module[:-2])
        i = 0
        while i < len(body):
            mod = body[i]
            i += 1
            if isinstance(mod, ResidualBlock):
                run = [mod]
                while i < len(body) and isinstance(body[i], ResidualBlock):
                    run.append(body[i])
                    i += 1
                h, Hv = _residual_stack_jvp(run, h, Hv, d)
            elif isinstance(mod, nn.Conv2d) and self._convin_own(mod, h) and Hv.shape[0] + h.shape[0] <= 65535:
                from . import ops
                both = ops.conv_in(torch.cat((Hv, h), 0), mod.weight)       # one launch for tangents + primal
                Hv, h = both[: Hv.shape[0]], both[Hv.shape[0]:]
            elif isinstance(mod, nn.Conv2d):
                Hv = F.conv2d(Hv, mod.weight, None, stride=mod.stride, padding=mod.padding)
                h = mod(h)
            elif isinstance(mod, nn.ReLU):
                Hv = _mask_rows(Hv, h > 0, d)
                h = torch.relu(h)
            else:
                raise ValueError(f"jvp unsupported for module {type(mod).__name__}")
        if fused_tail:
            return self._tail_apply(h, Hv, pos)
        if pos is not None:
            from . import ops
            idx, inv, pos64 = pos
            last = self._tail()
            w2 = last.weight.flatten(1)                                  # (O, C)
            hc = h.flatten(2).index_select(2, pos64)                     # (B, C, P)
            Hc = ops.gather_rows(Hv.flatten(2), None, idx, inv)          # (B*d, C, P)
            Hc = Hc.view(B, d, *Hc.shape[1:]) * _bcast((hc > 0).to(Hc.dtype))
            h = torch.matmul(w2, torch.relu(hc)) + last.bias.view(1, -1, 1)
            Hv = torch.matmul(w2, Hc)                                    # (B, d, O, P)
            y = torch.tanh(h)
            w = self.weights.view(1, -1, 1)
            return w * y + self.bias.view(1, -1, 1), Hv * _bcast((1 - y * y) * w)
        y = torch.tanh(h)
        Hv = Hv.view(B, d, *Hv.shape[1:]) * _bcast((1 - y * y) * self.weights)
        return self.weights * y + self.bias, Hv


def make_coupler_net(in_shape, num_out: int, cfg: dict) -> nn.Module:
    """`factory.py:467-507` for the net types reachable from ``--model non-square``."""
    if cfg["type"] == "mlp":
        assert len(in_shape) == 1
        return TangentMLP(in_shape[0], cfg["hidden_channels"], num_out, cfg["activation"])
    if cfg["type"] == "resnet":
        assert len(in_shape) == 3
        if cfg.get("batchnorm", True):
            raise ValueError("resnet batchnorm is not on the non-square path "
                             "(config/images.py:128 sets resnet_batchnorm=False)")
        return ScaledTanhResNet(in_shape[0], cfg["hidden_channels"], num_out)
    raise ValueError(f"Invalid net type {cfg['type']} for the non-square JVP path")


class ChunkedSharedCoupler(nn.Module):
    """One net; channel halves of its output are (shift, log-scale) (`couplers.py:27-59`)."""

    def __init__(self, net: nn.Module):
        super().__init__()
        self.shift_log_scale_net = net

    def forward(self, x: Tensor):
        y = self.shift_log_scale_net(x)
        c = y.shape[1] // 2
        return {"shift": y[:, :c], "log-scale": y[:, c:]}

    @property
    def supports_compact_output(self) -> bool:
        return bool(getattr(self.shift_log_scale_net, "supports_compact_output", False))

    def st_batched(self, x: Tensor, V, pos=None):
        """Flat conditioner outputs for the coupling kernel: st (B,2Q) = [shift | log_scale]
        and, if V (B,d,...) is given, ST (B,d,2Q).  ``pos``: evaluate the net's tail at these
        spatial positions only (compact outputs, Q = channels * len(pos))."""
        net = self.shift_log_scale_net
        if V is None:
            y = net(x) if pos is None else net.forward_compact(x, pos)
            return y.flatten(1), None
        y, Y = net.jvp_batched(x, V) if pos is None else net.jvp_batched(x, V, pos)
        return y.flatten(1), Y.flatten(2)


class IndependentCoupler(nn.Module):
    """Separate shift / log-scale nets (`couplers.py:6-24`)."""

    def __init__(self, shift_net: nn.Module, log_scale_net: nn.Module):
        super().__init__()
        self.shift_net, self.log_scale_net = shift_net, log_scale_net

    def forward(self, x: Tensor):
        return {"shift": self.shift_net(x), "log-scale": self.log_scale_net(x)}

    supports_compact_output = False

    def st_batched(self, x: Tensor, V, pos=None):
        assert pos is None
        if V is None:
            return torch.cat((self.shift_net(x).flatten(1), self.log_scale_net(x).flatten(1)), 1), None
        t, T = self.shift_net.jvp_batched(x, V)
        s, S = self.log_scale_net.jvp_batched(x, V)
        return torch.cat((t.flatten(1), s.flatten(1)), 1), torch.cat((T.flatten(2), S.flatten(2)), 2)


def make_coupler(in_shape, channels_per_output: int, cfg: dict) -> nn.Module:
    """`factory.py:413-464`."""
    if cfg["independent_nets"]:
        return IndependentCoupler(make_coupler_net(in_shape, channels_per_output, cfg["shift_net"]),
                                  make_coupler_net(in_shape, channels_per_output, cfg["log_scale_net"]))
    return ChunkedSharedCoupler(make_coupler_net(in_shape, 2 * channels_per_output, cfg["shift_log_scale_net"]))
