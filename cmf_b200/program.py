"""Layer program: the z -> x half of the non-square flow compiled once into a list of kernel
steps (replaces the Python walk of `non_square.py:146-184` that rebuilds ``jvp_stack`` /
``transform_stack`` every call, and the d-fold loop of `non_square.py:298-311`).

Data layout.  At every point of the flow the logical per-sample tensor (flattened size D) is held
as a few COMPACT blocks: block i stores ``n_i`` elements whose logical flat positions are the
static vector ``pos_i``; a block may be a *zero block* (no storage: zero-padding of
`non_square.py:397-404` and `split.py:50-52`).  The primal of a block is ``x (B, n_i)``, its
tangents ``V (B, d, n_i)`` -- tangent index next to the batch, so the conditioner call is a free
``view(B*d, C, H, W)`` and ``G = sum_i V_i V_i^T`` reads both operands K-major from one buffer.

Consequences:
  * reshapes / squeeze / flatten / split-pad / tail-pad never move data: they only rewrite
    ``pos_i`` at compile time (`reshaping.py:28-29`, `split.py:26-30`, `non_square.py:406-410`);
  * a coupling layer reads and writes ONLY the half it transforms (kernel (a) traffic equals the
    algorithmic 16 M (1+d) bytes/sample forward); the passthrough half is passed on by reference;
  * when consecutive layers use different partitions (tail permutation -> checkerboard ->
    split-channel -> checkerboard) blocks are re-grouped by one gather kernel; in the steady state
    (alternating masks) nothing is copied.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import ops
from .bijections import AffineCouplingBijection, CoupledRationalQuadraticSplineBijection, ReshapingBijection
from .densities import BijectionDensity, NonSquareTailDensity, SplitDensity

Tensor = torch.Tensor


@dataclass
class SymBlock:
    slot: Optional[int]          # runtime slot holding (x, V); None for a zero block
    pos: np.ndarray              # (n,) int64 logical flat positions, in storage order

    @property
    def zero(self) -> bool:
        return self.slot is None


@dataclass
class GatherStep:
    out: int
    srcs: Tuple[int, ...]        # 1 or 2 source slots
    idx: np.ndarray              # (n_out,) int32 into cat(srcs) or -1
    invs: Tuple[np.ndarray, ...]  # per source: (n_src,) int32 output index or -1


@dataclass
class ZerosStep:
    out: int
    n: int


@dataclass
class AclStep:
    bijection: AffineCouplingBijection
    cond: int                    # slot holding the conditioner input (B, n_in)
    in_shape: tuple
    mod: int                     # slot holding the compact half to transform
    out: int
    map: Optional[np.ndarray]    # (M,) int32 conditioner position of compact element j (None = identity)
    inv: Optional[np.ndarray]    # (Q,) int32 compact index of conditioner position q or -1
    M: int = 0
    Q: int = 0
    pos: Optional[np.ndarray] = None   # (P,) spatial positions at which the conditioner tail is evaluated
    hw: int = 0                        # H*W of the conditioner output (compact tail only)


@dataclass
class RqsStep:
    """``nsf-c`` layer (`bijections/nsf.py:45-84`, inverse direction of nsf `coupling.py:90-122`): the
    identity half goes through the free-parameter spline, the conditioner (with all tangents) reads the
    result, the transformed half goes through the conditional spline -- both on the spline kernel."""
    bijection: CoupledRationalQuadraticSplineBijection
    ident: int                   # slot holding the identity features (B, n_id), natural order
    mod: int                     # slot holding the transformed features (B, M), natural order
    out_ident: int
    out_mod: int
    n_id: int = 0
    M: int = 0


class FlatStackPlan:
    """Descriptor of a small flat flow for the whole-stack kernel (``cmf_flat_stack_fwd/bwd``): built when
    every node between head and tail is an identity view or a channelwise affine coupling whose
    conditioner is one shared tanh MLP, and the sizes fit the kernel (the 2-D configs of
    `config/two_d.py`: D <= 8, d <= 8, widths <= 16).  ``params()`` packs the model's parameters into
    the flat buffer the kernel reads (a differentiable ``torch.cat``: autograd routes the kernel's
    packed gradient back to every parameter)."""

    def __init__(self, nodes, tail):
        from .bijections import ChannelwiseAffineCouplingBijection, ViewBijection
        from .networks import ChunkedSharedCoupler, TangentMLP
        from . import kernels
        self.ok = False
        D, d = tail.flattened_dims, tail.latent_dimension
        layers, tensors, off, nh, widest = [], [], 0, None, 0
        for node in reversed(nodes):                           # z -> x order
            if isinstance(node, ViewBijection):
                if int(np.prod(node.x_shape)) != D or len(node.z_shape) != 1:
                    return
                continue
            if not isinstance(node, ChannelwiseAffineCouplingBijection) or tuple(node.x_shape) != (D,):
                return
            net = getattr(node.coupler, "shift_log_scale_net", None)
            if not isinstance(node.coupler, ChunkedSharedCoupler) or not isinstance(net, TangentMLP) or net.activation != "tanh":
                return
            keep, mod, _, dense = node.partition()
            linears = [m for m in net if isinstance(m, torch.nn.Linear)]
            if dense or (nh is not None and len(linears) - 1 != nh):
                return
            nh = len(linears) - 1
            widths = [linears[0].in_features] + [m.out_features for m in linears]
            w_off, b_off = [], []
            for m in linears:
                w_off.append(off); off += m.weight.numel(); tensors.append(m.weight)
                b_off.append(off); off += m.bias.numel(); tensors.append(m.bias)
            widest = max(widest, *widths)
            layers.append((keep.tolist(), mod.tolist(), widths, w_off, b_off))
        if not layers or not kernels.flat_stack_usable(D, d, len(layers), nh, widest, off):
            return
        perm = tail.permutation.detach().cpu().numpy().astype(np.int64)
        desc = [D, d, len(layers), nh, off] + [int(p) for p in perm[:d]]
        for keep, mod, widths, w_off, b_off in layers:
            desc += [len(keep), len(mod)] + [int(i) for i in keep] + [int(i) for i in mod] + widths + w_off + b_off
        self.desc = np.asarray(desc, dtype=np.int32)
        self.tensors, self.D, self.d, self.P = tensors, D, d, off
        self.ok = True

    def params(self) -> Tensor:
        return torch.cat([t.reshape(-1) for t in self.tensors])


class LayerProgram:
    def __init__(self, head_prior):
        """``head_prior``: the density subtree below the head (``head.prior``)."""
        self.nodes, self.tail = self._collect(head_prior)
        self.d = self.tail.latent_dimension
        self.steps: List[object] = []
        self._n_slots = 0
        self._dev_cache: Dict[Tuple[int, str], Tensor] = {}
        self._compile()
        plan = FlatStackPlan(self.nodes, self.tail)
        self.flat_plan = plan if plan.ok else None

    # ------------------------------------------------------------------ tree walk
    @staticmethod
    def _collect(node):
        """x -> z list of nodes between the head and the tail; ValueError exactly where
        `non_square.py:160-168` raises."""
        nodes = []
        while not isinstance(node, NonSquareTailDensity):
            if isinstance(node, BijectionDensity):
                bij = node.bijection
                if not isinstance(bij, (AffineCouplingBijection, ReshapingBijection, CoupledRationalQuadraticSplineBijection)):
                    raise ValueError(f"Cannot handle bijection of class {type(bij).__name__} on the JVP path")
                nodes.append(bij)
                node = node.prior
            elif isinstance(node, SplitDensity):
                if not node.non_square:
                    raise ValueError("Cannot handle a square SplitDensity on the non-square JVP path")
                nodes.append(node)
                node = node.density_1
            else:
                raise ValueError(f"Cannot handle density of class {type(node).__name__}")
        return nodes, node

    # ------------------------------------------------------------------ compile
    def _new_slot(self) -> int:
        self._n_slots += 1
        return self._n_slots - 1

    def _compile(self):
        tail = self.tail
        D, d = tail.flattened_dims, self.d
        perm = tail.permutation.detach().cpu().numpy().astype(np.int64)
        # z_k sits at logical position perm[k]; everything else is zero padding (non_square.py:397-404)
        blocks = [SymBlock(self._new_slot(), perm[:d].copy())]
        if D > d:
            blocks.append(SymBlock(None, np.sort(perm[d:])))
        for node in reversed(self.nodes):
            if isinstance(node, ReshapingBijection):
                bufs = list(node.buffers())                        # a permutation layer indexes with its buffer
                probe = torch.arange(D, device=bufs[0].device if bufs else None).view(1, *node.z_shape)
                src = node.reshape_z(probe).flatten().cpu().numpy()      # x[p] = z[src[p]]
                new_pos = np.empty(D, dtype=np.int64)
                new_pos[src] = np.arange(D)
                for b in blocks:
                    b.pos = new_pos[b.pos]
            elif isinstance(node, SplitDensity):
                assert node.dim == 1
                blocks.append(SymBlock(None, np.arange(D, 2 * D, dtype=np.int64)))
                D *= 2
            elif isinstance(node, CoupledRationalQuadraticSplineBijection):
                blocks = self._compile_rqs(node, blocks, D)
            else:
                blocks = self._compile_acl(node, blocks, D)
        self.final_blocks = blocks
        self.D = D

    def _compile_acl(self, bij, blocks: List[SymBlock], D: int) -> List[SymBlock]:
        keep_pos, mod_pos, in_shape, dense = bij.partition()
        keep_pos = keep_pos.cpu().numpy().astype(np.int64)
        mod_pos = mod_pos.cpu().numpy().astype(np.int64)
        is_mod = np.zeros(D, dtype=bool)
        is_mod[mod_pos] = True
        # 1. no block may straddle the partition
        conformed: List[SymBlock] = []
        for b in blocks:
            m = is_mod[b.pos]
            if m.all() or not m.any():
                conformed.append(b)
            elif b.zero:
                conformed += [SymBlock(None, b.pos[~m]), SymBlock(None, b.pos[m])]
            else:
                for sel in (~m, m):
                    sub = np.nonzero(sel)[0]
                    inv = np.full(b.pos.size, -1, dtype=np.int32)
                    inv[sub] = np.arange(sub.size, dtype=np.int32)
                    out = self._new_slot()
                    self.steps.append(GatherStep(out, (b.slot,), sub.astype(np.int32), (inv,)))
                    conformed.append(SymBlock(out, b.pos[sel]))
        keep_blocks = [b for b in conformed if not is_mod[b.pos[0]]]
        mod_blocks = [b for b in conformed if is_mod[b.pos[0]]]
        # 2. conditioner input: dense x_shape tensor with modified positions zeroed, or the compact
        #    passthrough tensor in natural order
        if dense:
            target = np.arange(D, dtype=np.int64)
            target[is_mod] = -1
        else:
            target = keep_pos
        cond = self._assemble(keep_blocks, target, D)
        # 3. the compact half to transform, in conditioner-output order
        mod = self._assemble(mod_blocks, mod_pos, D)
        # 4. relation between compact index j and conditioner position q
        pos, hw = None, 0
        if dense and getattr(bij.coupler, "supports_compact_output", False) and len(in_shape) == 3:
            # checkerboard with a per-position net tail: the conditioner emits compact outputs at
            # the modified spatial positions (same set for every channel), in (channel, position)
            # order == the order of mod_pos
            C, hw = in_shape[0], in_shape[1] * in_shape[2]
            sp = np.unique(mod_pos % hw)
            if sp.size * C == mod_pos.size and np.array_equal(
                    mod_pos, (np.arange(C)[:, None] * hw + sp[None, :]).reshape(-1)):
                pos = sp
        if pos is not None:
            cmap, cinv, Q = None, None, mod_pos.size
        elif dense:
            cmap = mod_pos.astype(np.int32)
            cinv = np.full(D, -1, dtype=np.int32)
            cinv[mod_pos] = np.arange(mod_pos.size, dtype=np.int32)
            Q = D
        else:
            cmap, cinv, Q = None, None, mod_pos.size
        out = self._new_slot()
        self.steps.append(AclStep(bij, cond, tuple(in_shape), mod, out, cmap, cinv, mod_pos.size, Q, pos, hw))
        return keep_blocks + [SymBlock(out, mod_pos)]

    def _compile_rqs(self, bij, blocks: List[SymBlock], D: int) -> List[SymBlock]:
        """Both halves change (a spline does not fix 0, so zero padding is materialised too): the
        identity features and the transformed features are each assembled into one compact slot in
        natural order and replaced by the step's two outputs."""
        ident_pos, mod_pos = (t.cpu().numpy().astype(np.int64) for t in bij.partition())
        assert ident_pos.size + mod_pos.size == D
        ident = self._assemble_any(blocks, ident_pos, D)
        mod = self._assemble_any(blocks, mod_pos, D)
        out_i, out_m = self._new_slot(), self._new_slot()
        self.steps.append(RqsStep(bij, ident, mod, out_i, out_m, int(ident_pos.size), int(mod_pos.size)))
        return [SymBlock(out_i, ident_pos), SymBlock(out_m, mod_pos)]

    def _assemble_any(self, blocks: Sequence[SymBlock], target: np.ndarray, D: int) -> int:
        """:meth:`_assemble` over whatever blocks hold the target positions (blocks may straddle them)"""
        want = np.zeros(D, dtype=bool)
        want[target] = True
        return self._assemble([b for b in blocks if want[b.pos].any()], target, D)

    def _assemble(self, blocks: Sequence[SymBlock], target: np.ndarray, D: int) -> int:
        """Slot whose element i is the value at logical position target[i] (0 where target[i] < 0
        or the position belongs to a zero block)."""
        real = [b for b in blocks if not b.zero]
        if len(real) == 1 and real[0].pos.size == target.size and np.array_equal(real[0].pos, target):
            return real[0].slot
        if not real:
            out = self._new_slot()
            self.steps.append(ZerosStep(out, int(target.size)))
            return out
        while len(real) > 2:          # concatenate pairwise until two sources remain
            a, b = real.pop(), real.pop()
            out = self._new_slot()
            na, nb = a.pos.size, b.pos.size
            self.steps.append(GatherStep(out, (a.slot, b.slot), np.arange(na + nb, dtype=np.int32),
                                         (np.arange(na, dtype=np.int32), np.arange(na, na + nb, dtype=np.int32))))
            real.append(SymBlock(out, np.concatenate([a.pos, b.pos])))
        where = np.full(D, -1, dtype=np.int64)   # logical position -> index into cat(real)
        base = 0
        for b in real:
            where[b.pos] = base + np.arange(b.pos.size)
            base += b.pos.size
        idx = np.where(target >= 0, where[np.maximum(target, 0)], -1).astype(np.int32)
        invs, base = [], 0
        for b in real:
            inv = np.full(b.pos.size, -1, dtype=np.int32)
            hit = (idx >= base) & (idx < base + b.pos.size)
            inv[idx[hit] - base] = np.nonzero(hit)[0].astype(np.int32)
            invs.append(inv)
            base += b.pos.size
        out = self._new_slot()
        self.steps.append(GatherStep(out, tuple(b.slot for b in real), idx, tuple(invs)))
        return out

    # ------------------------------------------------------------------ run
    def _dev(self, arr: Optional[np.ndarray], device, key, dtype=torch.int32) -> Optional[Tensor]:
        if arr is None:
            return None
        k = (key, str(device))
        t = self._dev_cache.get(k)
        if t is None:
            t = torch.as_tensor(arr, dtype=dtype, device=device)
            self._dev_cache[k] = t
        return t

    def run(self, z: Tensor, V0: Optional[Tensor]):
        """Push z (B,d) and tangents V0 (B,m,d) (None = primal only, `non_square.py:313-320`)
        through the flow.  Returns ``[(pos (n,) int64 device tensor, x (B,n), V (B,m,n)|None)]`` for
        the non-zero final blocks and the positions of the final zero blocks."""
        dev = z.device
        B = z.shape[0]
        m = None if V0 is None else V0.shape[1]
        slots: Dict[int, Tuple[Tensor, Optional[Tensor]]] = {0: (z.contiguous(), V0)}
        for si, step in enumerate(self.steps):
            if isinstance(step, GatherStep):
                xs = [slots[s][0] for s in step.srcs]
                vs = [slots[s][1] for s in step.srcs]
                idx = self._dev(step.idx, dev, (si, "idx"))
                invs = [self._dev(a, dev, (si, "inv", i)) for i, a in enumerate(step.invs)]
                x1 = xs[1] if len(xs) > 1 else None
                i1 = invs[1] if len(invs) > 1 else None
                x = ops.gather_rows(xs[0], x1, idx, invs[0], i1)
                V = None
                if m is not None:
                    V = ops.gather_rows(vs[0], vs[1] if len(vs) > 1 else None, idx, invs[0], i1)
                slots[step.out] = (x, V)
            elif isinstance(step, ZerosStep):
                x = torch.zeros(B, step.n, dtype=z.dtype, device=dev)
                V = None if m is None else torch.zeros(B, m, step.n, dtype=z.dtype, device=dev)
                slots[step.out] = (x, V)
            elif isinstance(step, RqsStep):
                bij, tb = step.bijection, step.bijection.tail_bound
                x_id, V_id = slots[step.ident]
                ucw, uch, udv = bij.unconditional_tables(B)
                zero_t = None if m is None else torch.zeros(B, m, *ucw.shape[1:], dtype=z.dtype, device=dev)
                x_id, V_id, _ = ops.rqs_coupling_jvp(x_id, V_id, ucw, uch, udv, zero_t, zero_t, zero_t, tb)
                z_m, V_m = slots[step.mod]
                x_m, V_m, _ = ops.rqs_coupling_jvp(z_m, V_m, *bij.conditional_tables(x_id, V_id), tb)
                slots[step.out_ident], slots[step.out_mod] = (x_id, V_id), (x_m, V_m)
            else:
                xc, Vc = slots[step.cond]
                x_in = xc.view(B, *step.in_shape)
                V_in = None if m is None else Vc.view(B, m, *step.in_shape)
                pos = None
                if step.pos is not None:
                    pinv = np.full(step.hw, -1, dtype=np.int32)
                    pinv[step.pos] = np.arange(step.pos.size, dtype=np.int32)
                    pos = (self._dev(step.pos.astype(np.int32), dev, (si, "pidx")),
                           self._dev(pinv, dev, (si, "pinv")),
                           self._dev(step.pos, dev, (si, "p64"), torch.int64))
                st, ST = step.bijection.coupler.st_batched(x_in, V_in, pos)
                z_m, V_m = slots[step.mod]
                slots[step.out] = ops.coupling_jvp(z_m, V_m, st, ST,
                                                   self._dev(step.map, dev, (si, "map")),
                                                   self._dev(step.inv, dev, (si, "cinv")))
        out, zero_pos = [], []
        for bi, b in enumerate(self.final_blocks):
            pos = self._dev(b.pos, dev, ("final", bi), torch.int64)
            if b.zero:
                zero_pos.append(pos)
            else:
                out.append((pos, *slots[b.slot]))
        return out, zero_pos

    # ------------------------------------------------------------------ introspection
    def describe(self) -> List[str]:
        lines = []
        for s in self.steps:
            if isinstance(s, GatherStep):
                lines.append(f"gather  {s.srcs} -> {s.out}  n_out={s.idx.size}")
            elif isinstance(s, ZerosStep):
                lines.append(f"zeros   -> {s.out}  n={s.n}")
            elif isinstance(s, RqsStep):
                lines.append(f"rqs     ident={s.ident} mod={s.mod} -> {s.out_ident},{s.out_mod}  n_id={s.n_id} M={s.M}")
            else:
                lines.append(f"acl     {type(s.bijection).__name__}: cond={s.cond} mod={s.mod} -> {s.out}  "
                             f"M={s.M} Q={s.Q} {'mapped' if s.map is not None else ('compact-tail' if s.pos is not None else 'identity')}")
        return lines

    def modified_elements(self) -> int:
        """sum_l M_l, the factor of the kernel-(a) byte model."""
        return sum(s.M for s in self.steps if isinstance(s, AclStep))

    def spline_elements(self) -> int:
        """elements pushed through the spline kernel per sample (identity + transformed halves)"""
        return sum(s.M + s.n_id for s in self.steps if isinstance(s, RqsStep))
