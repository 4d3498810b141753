"""CPU oracle for the non-square log-det + g_ij hot path of k-flouris/cmf.

TEST INFRASTRUCTURE ONLY.  Nothing under ``cmf_b200/`` may import this file; it is
used by ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` as the *checker* / CPU baseline.

What it is: a plain-PyTorch (CPU, fp32 or fp64) restatement of the reference's
algorithm for ``--model non-square`` -- the same sequence of operations, in the same
order, one tangent column at a time, recomputing the conditioner primal for every
column exactly as the reference does (`non_square.py:298-311`).  It is *functional*:
it consumes the reference's schema (a list of layer dicts, `config/schemas.py:1-27`)
and a flat ``params`` dict keyed by the reference's ``state_dict`` names, so the same
weights can be pushed through the real reference (in the build container, see
``oracle/gen_golden.py``), through this oracle, and through the CUDA path.

Parity pin: ``tests/golden/*.pt`` hold outputs of the UNMODIFIED reference
(imported from /root/reference by ``oracle/gen_golden.py``) on seeded inputs and
weights; ``tests/test_oracle_golden.py`` checks this file against them.  The reference
ships no tests / golden vectors of its own for this path, so the
pin is "outputs of the reference itself run here".

Reference citations are ``path:line`` relative to the reference checkout.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor
Params = Dict[str, Tensor]


# --------------------------------------------------------------------------------------
# conditioner networks: primal call and the reference's hand-written JVP rules
# --------------------------------------------------------------------------------------
class _Mlp:
    """`networks.py:206-224` (get_mlp) + `networks.py:24-32` (NN_Sequential_JVP.jvp).

    Sequential indices: Linear at 0,2,4,...; activation at odd indices (tanh for the
    shared-coupler nets used by every non-square config, `schemas.py:492-500`).
    """

    def __init__(self, params: Params, prefix: str, num_hidden: int, activation: str):
        self.w = [params[f"{prefix}{2 * i}.weight"] for i in range(num_hidden + 1)]
        self.b = [params[f"{prefix}{2 * i}.bias"] for i in range(num_hidden + 1)]
        assert activation in ("tanh", "relu")
        self.activation = activation

    def __call__(self, x: Tensor) -> Tensor:
        n = len(self.w)
        for i in range(n):
            x = F.linear(x, self.w[i], self.b[i])
            if i + 1 < n:
                x = torch.tanh(x) if self.activation == "tanh" else torch.relu(x)
        return x

    def jvp(self, x: Tensor, v: Tensor) -> Tuple[Tensor, Tensor]:
        n = len(self.w)
        for i in range(n):
            # jvp_layers.py:49-52: primal with bias, tangent through the weight only
            x, v = F.linear(x, self.w[i], self.b[i]), F.linear(v, self.w[i])
            if i + 1 < n:
                x, v = _act_jvp(self.activation, x, v)
        return x, v


def _act_jvp(name: str, x: Tensor, v: Tensor) -> Tuple[Tensor, Tensor]:
    """`jvp_layers.py:38-46`."""
    if name == "relu":
        return torch.relu(x), (x > 0) * v
    y = torch.tanh(x)
    return y, (1 - y ** 2) * v


class _ResNet:
    """`networks.py:116-161` (get_resnet, no batch norm) wrapped in
    `networks.py:96-113` (ScaledTanh2dModule); residual block `networks.py:35-94`.

    State-dict layout under ``prefix``: ``weights``, ``bias`` (scaled tanh),
    ``module.0.weight`` (3x3 in-conv, no bias), ``module.{1..n}.conv{1,2}.{weight,bias}``,
    ``module.{n+2}.{weight,bias}`` (1x1 out-conv; index n+1 is the ReLU).
    """

    def __init__(self, params: Params, prefix: str, num_blocks: int):
        p = params
        self.out_w, self.out_b = p[prefix + "weights"], p[prefix + "bias"]
        self.conv_in = p[prefix + "module.0.weight"]
        self.blocks = []
        for i in range(1, num_blocks + 1):
            self.blocks.append(tuple(p[f"{prefix}module.{i}.{n}"] for n in
                                     ("conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias")))
        self.conv_out_w = p[f"{prefix}module.{num_blocks + 2}.weight"]
        self.conv_out_b = p[f"{prefix}module.{num_blocks + 2}.bias"]

    def __call__(self, x: Tensor) -> Tensor:
        h = F.conv2d(x, self.conv_in, None, padding=1)
        for w1, b1, w2, b2 in self.blocks:
            r = F.conv2d(torch.relu(h), w1, b1, padding=1)
            r = F.conv2d(torch.relu(r), w2, b2, padding=1)
            h = r + h
        h = F.conv2d(torch.relu(h), self.conv_out_w, self.conv_out_b)
        return self.out_w * torch.tanh(h) + self.out_b

    def jvp(self, x: Tensor, v: Tensor) -> Tuple[Tensor, Tensor]:
        # jvp_layers.py:55-64: tangent conv has no bias
        h, hv = F.conv2d(x, self.conv_in, None, padding=1), F.conv2d(v, self.conv_in, None, padding=1)
        for w1, b1, w2, b2 in self.blocks:  # networks.py:62-79
            r, rv = _act_jvp("relu", h, hv)
            r, rv = F.conv2d(r, w1, b1, padding=1), F.conv2d(rv, w1, None, padding=1)
            r, rv = _act_jvp("relu", r, rv)
            r, rv = F.conv2d(r, w2, b2, padding=1), F.conv2d(rv, w2, None, padding=1)
            h, hv = r + h, rv + hv
        h, hv = _act_jvp("relu", h, hv)
        h, hv = F.conv2d(h, self.conv_out_w, self.conv_out_b), F.conv2d(hv, self.conv_out_w, None)
        h, hv = _act_jvp("tanh", h, hv)  # networks.py:108-113
        return self.out_w * h + self.out_b, self.out_w * hv


def _make_net(params: Params, prefix: str, net_cfg: dict):
    """`factory.py:467-507` (get_coupler_net), the two net types the non-square configs use."""
    if net_cfg["type"] == "mlp":
        return _Mlp(params, prefix, len(net_cfg["hidden_channels"]), net_cfg["activation"])
    if net_cfg["type"] == "resnet":
        if net_cfg.get("batchnorm", True):
            raise ValueError("oracle: resnet batchnorm is outside the non-square hot path (SURVEY 2)")
        return _ResNet(params, prefix, len(net_cfg["hidden_channels"]))
    raise ValueError(f"oracle: unsupported coupler net {net_cfg['type']}")


class _Coupler:
    """`couplers.py:27-59`: ChunkedSharedCoupler (one net, channel halves = shift, log-scale)
    or IndependentCoupler (`couplers.py:6-24`)."""

    def __init__(self, params: Params, prefix: str, cfg: dict):
        self.shared = not cfg["independent_nets"]
        if self.shared:
            self.net = _make_net(params, prefix + "shift_log_scale_net.", cfg["shift_log_scale_net"])
        else:
            self.shift_net = _make_net(params, prefix + "shift_net.", cfg["shift_net"])
            self.log_scale_net = _make_net(params, prefix + "log_scale_net.", cfg["log_scale_net"])

    @staticmethod
    def _halves(t: Tensor) -> Tuple[Tensor, Tensor]:
        c = t.shape[1]
        assert c % 2 == 0
        return t[:, : c // 2], t[:, c // 2:]

    def __call__(self, x: Tensor) -> Tuple[Tensor, Tensor]:
        if self.shared:
            return self._halves(self.net(x))
        return self.shift_net(x), self.log_scale_net(x)

    def jvp(self, x: Tensor, v: Tensor):
        """returns (shift, shift_dot, log_scale, log_scale_dot)"""
        if self.shared:
            y, yd = self.net.jvp(x, v)
            (t, s), (td, sd) = self._halves(y), self._halves(yd)
            return t, td, s, sd
        t, td = self.shift_net.jvp(x, v)
        s, sd = self.log_scale_net.jvp(x, v)
        return t, td, s, sd


# --------------------------------------------------------------------------------------
# flow layers (z <-> x maps and their JVPs)
# --------------------------------------------------------------------------------------
class _Layer:
    """A node of the density chain below the head.  ``x_to_z`` returns (z, log_jac or None)."""
    x_shape: Tuple[int, ...]
    z_shape: Tuple[int, ...]

    def x_to_z(self, x): raise NotImplementedError
    def z_to_x(self, z): raise NotImplementedError
    def jvp(self, z, v): raise NotImplementedError


def _checkerboard_mask(c: int, h: int, w: int, reverse: bool, like: Tensor) -> Tensor:
    """`acl.py:68-78`: 1 where (row+col) is odd, flipped by reverse_mask."""
    ii = torch.arange(h, device=like.device).view(h, 1)
    jj = torch.arange(w, device=like.device).view(1, w)
    m = ((ii + jj) % 2 == 1).to(like.dtype).expand(c, h, w)
    return 1 - m if reverse else m


class _CheckerboardAcl(_Layer):
    """`acl.py:29-66` Checkerboard2dAffineCouplingBijection."""

    def __init__(self, shape, reverse, coupler: _Coupler, like: Tensor):
        self.x_shape = self.z_shape = tuple(shape)
        self.mask = _checkerboard_mask(*shape, reverse, like)
        self.coupler = coupler

    def x_to_z(self, x):
        m = self.mask
        t, s = self.coupler(m * x)
        z = m * x + (1 - m) * ((x + t) * torch.exp(s))
        return z, ((1 - m) * s).flatten(1).sum(1, keepdim=True)

    def z_to_x(self, z):
        m = self.mask
        t, s = self.coupler(m * z)
        return m * z + (1 - m) * (z * torch.exp(-s) - t)

    def jvp(self, z, v):
        m = self.mask
        t, td, s, sd = self.coupler.jvp(m * z, m * v)
        x = m * z + (1 - m) * (z * torch.exp(-s) - t)
        moved = torch.exp(-s) * ((1 - m) * v - (1 - m) * z * sd) - td
        return x, m * v + (1 - m) * moved


class _ChannelAcl(_Layer):
    """`acl.py:81-214`: split-channel (`:169-189`) and alternating-channel (`:192-214`) ACLs."""

    def __init__(self, shape, kind, reverse, coupler: _Coupler):
        self.x_shape = self.z_shape = tuple(shape)
        assert kind in ("split-channel", "alternating-channel")
        self.kind, self.reverse, self.coupler = kind, reverse, coupler
        self.n_first = shape[0] // 2 if kind == "split-channel" else (shape[0] + 1) // 2

    def _split(self, t):
        if self.kind == "split-channel":
            a, b = t[:, : self.n_first], t[:, self.n_first:]
        else:
            a, b = t[:, ::2], t[:, 1::2]
        return (b, a) if self.reverse else (a, b)  # (passthrough, modified)

    def _combine(self, keep, moved):
        a, b = (moved, keep) if self.reverse else (keep, moved)
        if self.kind == "split-channel":
            return torch.cat((a, b), dim=1)
        out = a.new_empty(a.shape[0], *self.x_shape)
        out[:, ::2] = a
        out[:, 1::2] = b
        return out

    def x_to_z(self, x):
        keep, moved = self._split(x)
        t, s = self.coupler(keep)
        return self._combine(keep, (moved + t) * torch.exp(s)), s.flatten(1).sum(1, keepdim=True)

    def z_to_x(self, z):
        keep, moved = self._split(z)
        t, s = self.coupler(keep)
        return self._combine(keep, moved * torch.exp(-s) - t)

    def jvp(self, z, v):
        keep, moved = self._split(z)
        vkeep, vmoved = self._split(v)
        t, td, s, sd = self.coupler.jvp(keep, vkeep)
        x = self._combine(keep, moved * torch.exp(-s) - t)
        xd = self._combine(vkeep, torch.exp(-s) * (vmoved - moved * sd) - td)
        return x, xd


class _Reshape(_Layer):
    """`reshaping.py:8-29`: JVP of an index map is the same index map on the tangent."""

    def __init__(self, x_shape, z_shape, fwd, inv):
        self.x_shape, self.z_shape, self._fwd, self._inv = tuple(x_shape), tuple(z_shape), fwd, inv

    def x_to_z(self, x): return self._fwd(x), None
    def z_to_x(self, z): return self._inv(z)
    def jvp(self, z, v): return self._inv(z), self._inv(v)


def _flatten_layer(shape) -> _Reshape:
    """`reshaping.py:61-66` ViewBijection used by the `flatten` schema entry (`factory.py:243-244`)."""
    n = int(math.prod(shape))
    return _Reshape(shape, (n,), lambda x: x.reshape(x.shape[0], n), lambda z: z.reshape(z.shape[0], *shape))


def _squeeze_layer(shape, f) -> _Reshape:
    """`reshaping.py:69-114` Squeeze2dBijection: x (c,h,w) <-> z (c f^2, h/f, w/f)."""
    c, h, w = shape
    zs = (c * f * f, h // f, w // f)

    def fwd(x):
        b = x.shape[0]
        return x.reshape(b, c, h // f, f, w // f, f).permute(0, 1, 3, 5, 2, 4).reshape(b, *zs)

    def inv(z):
        b = z.shape[0]
        return z.reshape(b, c, f, f, h // f, w // f).permute(0, 1, 4, 2, 5, 3).reshape(b, c, h, w)

    return _Reshape(shape, zs, fwd, inv)


class _RandPerm(_Layer):
    """`reshaping.py:32-44` RandomChannelwisePermutationBijection (``rand-channel-perm``)."""

    def __init__(self, shape, params: Params, prefix: str):
        self.x_shape = self.z_shape = tuple(shape)
        self.perm, self.inv = params[prefix + "permutation"], params[prefix + "inverse_permutation"]

    def x_to_z(self, x): return x[:, self.perm], None
    def z_to_x(self, z): return z[:, self.inv]
    def jvp(self, z, v): return z[:, self.inv], v[:, self.inv]


class _NsfCoupling(_Layer):
    """``nsf-c``: `bijections/nsf.py:45-84` = nsf ``PiecewiseRationalQuadraticCouplingTransform``
    (`gitmodules/nsf/nde/transforms/coupling.py:60-122,399-473`) with an alternating binary mask
    (`utils/torchutils.py:87-98`), a ResidualNet conditioner (`nn/resnet.py:10-105`, relu, no batch
    norm, no context), linear tails and the unconditional spline on the identity features
    (`nonlinearities.py:337-406`).  The spline arithmetic is ``oracle/rqs_oracle.py``.
    The reference has no ``_jvp`` for this bijection: the oracle's is the autograd JVP of
    the z -> x map, ``torch.autograd.functional.jvp`` -- the same definition ``oracle/gen_golden.py``
    gives the reference class when it generates the fixture."""

    def __init__(self, shape, layer: dict, params: Params, prefix: str):
        import rqs_oracle as R
        self.R = R
        self.x_shape = self.z_shape = tuple(shape)
        n = shape[0]
        start = 0 if layer["reverse_mask"] else 1                 # even=reverse_mask: nsf.py:74-77
        mask = torch.zeros(n, dtype=torch.bool)
        mask[start::2] = True
        idx = torch.arange(n)
        self.tr, self.ident = idx[mask], idx[~mask]
        self.K, self.tb = layer["num_bins"], float(layer["tail_bound"])
        self.hidden, self.blocks = layer["num_hidden_channels"], layer["num_hidden_layers"]
        self.p, self.prefix = params, prefix + "flow."

    def _net(self, h):
        p, pre = self.p, self.prefix + "transform_net."
        h = F.linear(h, p[pre + "initial_layer.weight"], p[pre + "initial_layer.bias"])
        for i in range(self.blocks):
            t = F.linear(torch.relu(h), p[f"{pre}blocks.{i}.linear_layers.0.weight"], p[f"{pre}blocks.{i}.linear_layers.0.bias"])
            t = F.linear(torch.relu(t), p[f"{pre}blocks.{i}.linear_layers.1.weight"], p[f"{pre}blocks.{i}.linear_layers.1.bias"])
            h = h + t
        return F.linear(h, p[pre + "final_layer.weight"], p[pre + "final_layer.bias"])

    def _uncond(self, t, inverse):
        u = self.prefix + "unconditional_transform."
        prm = torch.cat((self.p[u + "unnormalized_widths"], self.p[u + "unnormalized_heights"],
                         self.p[u + "unnormalized_derivatives"]), dim=-1)
        prm = prm.unsqueeze(0).expand(t.shape[0], *prm.shape)
        out, lad, _ = self.R.unconstrained_rqs(t, prm, self.K, self.tb, inverse)
        return out, lad.sum(1, keepdim=True)

    def _cond(self, t, ident, inverse):
        prm = self._net(ident).view(t.shape[0], self.tr.numel(), 3 * self.K - 1)
        out, lad, _ = self.R.unconstrained_rqs(t, prm, self.K, self.tb, inverse, hidden_features=self.hidden)
        return out, lad.sum(1, keepdim=True)

    def x_to_z(self, x):                                           # coupling.py:60-88
        ident, tr = x[:, self.ident.to(x.device)], x[:, self.tr.to(x.device)]
        z_tr, lad = self._cond(tr, ident, False)
        z_id, lad_id = self._uncond(ident, False)
        z = torch.empty_like(x)
        z[:, self.ident.to(x.device)] = z_id
        z[:, self.tr.to(x.device)] = z_tr
        return z, lad + lad_id

    def z_to_x(self, z):                                           # coupling.py:90-122
        ident, tr = z[:, self.ident.to(z.device)], z[:, self.tr.to(z.device)]
        x_id, _ = self._uncond(ident, True)
        x_tr, _ = self._cond(tr, x_id, True)
        x = torch.empty_like(z)
        x[:, self.ident.to(z.device)] = x_id
        x[:, self.tr.to(z.device)] = x_tr
        return x

    def jvp(self, z, v):
        return torch.autograd.functional.jvp(self.z_to_x, z, v, create_graph=torch.is_grad_enabled())


class _Elementwise(_Layer):
    """`math.py:9-105` preprocessing bijections above the head (scalar-mult, scalar-add, logit).
    Only ``x_to_z`` is on the path.  They sit ABOVE the head as BijectionDensity nodes, so their
    log-jacobians are added to the head's elbo on the way out (`exact.py:23-30`); they depend on
    the data only, never on a parameter."""

    def __init__(self, shape, kind, value=None):
        self.x_shape = self.z_shape = tuple(shape)
        self.kind, self.value = kind, value

    def x_to_z(self, x):
        n = x[0].numel()
        if self.kind == "scalar-mult":        # math.py:70-88
            return self.value * x, torch.full((x.shape[0], 1), n * math.log(abs(self.value)), dtype=x.dtype, device=x.device)
        if self.kind == "scalar-add":         # math.py:91-105
            return x + self.value, torch.zeros((x.shape[0], 1), dtype=x.dtype, device=x.device)
        xc = x.clamp(1e-7, 1 - 1e-7)          # logit, math.py:41-53
        lj = (-torch.log(xc) - torch.log(1 - xc)).flatten(1).sum(1, keepdim=True)
        return torch.log(x) - torch.log(1 - x), lj


class _SplitPad(_Layer):
    """`split.py:15-30,50-52` in non-square mode: encode keeps the first channel half
    (density_1); z->x and its JVP zero-pad the dropped half."""

    def __init__(self, shape):
        self.x_shape = tuple(shape)
        self.z_shape = (shape[0] // 2, *shape[1:])

    def x_to_z(self, x): return torch.chunk(x, 2, dim=1)[0], None
    def z_to_x(self, z): return torch.cat((z, torch.zeros_like(z)), dim=1)
    def jvp(self, z, v): return self.z_to_x(z), self.z_to_x(v)


class _Tail(_Layer):
    """`non_square.py:365-410` NonSquareTailDensity: permute + keep first d; zero-pad + un-permute."""

    def __init__(self, shape, d, params: Params, prefix: str):
        self.x_shape, self.z_shape, self.d = tuple(shape), (d,), d
        self.n = int(math.prod(shape))
        self.mask = params[prefix + "mask"]
        self.perm = params[prefix + "permutation"]
        self.inv_perm = params[prefix + "inverse_permutation"]

    def x_to_z(self, x):
        return x.flatten(1)[:, self.perm][:, self.mask], None

    def z_to_x(self, z):
        padded = torch.zeros(z.shape[0], self.n, dtype=z.dtype, device=z.device)
        padded[:, : self.d] = z
        return padded[:, self.inv_perm].view(z.shape[0], *self.x_shape)

    def jvp(self, z, v): return self.z_to_x(z), self.z_to_x(v)


class _Affine(_Layer):
    """`affine.py:11-41` AffineBijection (per_channel=False): the 2-D configs' d-dim prior flow."""

    def __init__(self, shape, params: Params, prefix: str):
        self.x_shape = self.z_shape = tuple(shape)
        self.shift, self.log_scale = params[prefix + "shift"], params[prefix + "log_scale"]

    def x_to_z(self, x):
        lj = torch.sum(self.log_scale).view(1, 1).expand(x.shape[0], 1)
        return x * torch.exp(self.log_scale) + self.shift, lj


def _std_normal_log_prob(z: Tensor) -> Tensor:
    """`gaussian.py:9-22` with mean 0, stddev 1 (`factory.py:200-205`)."""
    f = z.flatten(1)
    return -0.5 * f.shape[1] * math.log(2 * math.pi) - 0.5 * torch.sum(f ** 2, dim=1, keepdim=True)


# --------------------------------------------------------------------------------------
# schema walk (restates `factory.py:86-162` for the layer types reachable from non-square)
# --------------------------------------------------------------------------------------
def strip_data_parallel_prefix(state: Params) -> Params:
    """`factory.py:76-81` wraps everything but 2-channel data in DataParallelDensity => 'module.' keys."""
    if state and all(k.startswith("module.") for k in state):
        return {k[len("module."):]: v for k, v in state.items()}
    return dict(state)


class OracleFlow:
    """The reference density tree for ``--model non-square``, flattened into three chains.

    ``pre``   elementwise preprocessing above the head (images only)
    ``flow``  head -> tail: the injective part whose Jacobian is J  (x -> z order)
    ``prior`` the d-dimensional square flow below the tail, ending in N(0, I)
    """

    def __init__(self, schema: Sequence[dict], params: Params, x_shape: Sequence[int]):
        params = strip_data_parallel_prefix(params)
        like = next(v for v in params.values() if v.is_floating_point())
        self.params = params
        self.pre: List[_Layer] = []
        self.flow: List[_Layer] = []
        self.prior: List[_Layer] = []
        self.head_cfg: Optional[dict] = None
        self.tail: Optional[_Tail] = None
        self.x_shape = tuple(x_shape)

        prefix, shape = "", tuple(x_shape)
        chain = self.pre
        for layer in schema:
            ty = layer["type"]
            if ty == "dequantization":            # wrapper.py:27-30; noise is pre-added by the caller
                prefix += "density."
                continue
            if ty == "non-square-head":           # factory.py:121-144
                if layer["m_flow"]:
                    raise ValueError("oracle: m_flow baseline head is out of scope")
                if layer["log_jacobian_method"] not in ("cholesky", "hutch_with_cg"):
                    raise ValueError(f"{layer['log_jacobian_method']} not a valid Jacobian calculation method")
                self.head_cfg, self.head_shape = dict(layer), shape
                chain = self.flow
                prefix += "prior."
                continue
            if ty == "non-square-base":           # factory.py:146-155
                self.tail = _Tail(shape, layer["latent_dimension"], params, prefix)
                self.flow.append(self.tail)
                chain, shape = self.prior, (layer["latent_dimension"],)
                prefix += "prior."
                continue
            if ty == "split":                     # factory.py:104-116
                if not layer["non_square"]:
                    raise ValueError("oracle: square split is not on the non-square path")
                node = _SplitPad(shape)
                prefix += "density_1."
            elif ty == "acl":
                coupler = _Coupler(params, prefix + "bijection.coupler.", layer["coupler"])
                if layer.get("num_u_channels", 0):
                    raise ValueError("oracle: CIF u-channels are out of scope")
                if layer["mask_type"] == "checkerboard":
                    node = _CheckerboardAcl(shape, layer["reverse_mask"], coupler, like)
                else:
                    node = _ChannelAcl(shape, layer["mask_type"], layer["reverse_mask"], coupler)
                prefix += "prior."
            elif ty == "flatten":
                node = _flatten_layer(shape); prefix += "prior."
            elif ty == "rand-channel-perm":
                node = _RandPerm(shape, params, prefix + "bijection."); prefix += "prior."
            elif ty == "nsf-c":
                node = _NsfCoupling(shape, layer, params, prefix + "bijection."); prefix += "prior."
            elif ty == "squeeze":
                node = _squeeze_layer(shape, layer["factor"]); prefix += "prior."
            elif ty in ("scalar-mult", "scalar-add", "logit"):
                node = _Elementwise(shape, ty, layer.get("value")); prefix += "prior."
            elif ty == "affine":
                assert not layer["per_channel"]
                node = _Affine(shape, params, prefix + "bijection."); prefix += "prior."
            else:
                raise ValueError(f"Cannot handle layer of type {ty}")   # cf. non_square.py:167-168
            chain.append(node)
            shape = node.z_shape
        assert self.head_cfg is not None and self.tail is not None
        self.d = self.tail.d
        self.regularization_param = self.head_cfg["regularization_param"]

    # ---- encode pass: `non_square.py:65` prior.elbo(x) -----------------------------------
    def encode(self, x: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
        """returns (x after preprocessing, low-dim latent z (B,d), low_dim_elbo (B,1))."""
        self.pre_log_jac = 0.0
        for node in self.pre:
            x, lj = node.x_to_z(x)
            self.pre_log_jac = self.pre_log_jac + lj
        h = x
        for node in self.flow:
            h, _ = node.x_to_z(h)     # log-jacs above the tail are dropped: non_square.py:176-177
        z = h
        u, log_jac = z, 0.0
        for node in self.prior:
            u, lj = node.x_to_z(u)
            if lj is not None:
                log_jac = log_jac + lj
        return x, z, _std_normal_log_prob(u) + log_jac

    # ---- `non_square.py:313-329` ----------------------------------------------------------
    def flow_forward(self, z: Tensor) -> Tensor:
        for node in reversed(self.flow):
            z = node.z_to_x(z)
        return z

    def jvp_forward(self, z: Tensor, v: Tensor) -> Tuple[Tensor, Tensor]:
        for node in reversed(self.flow):
            z, v = node.jvp(z, v)
        return z, v

    # ---- `non_square.py:298-311` ----------------------------------------------------------
    def full_jac_transpose_jac(self, z: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
        cols = []
        recon = None
        for i in range(self.d):
            e = torch.zeros_like(z)
            e[:, i] = 1
            recon, jv = self.jvp_forward(z, e)
            cols.append(jv.flatten(1))
        jac = torch.stack(cols, dim=2)                       # (B, D, d)
        return torch.bmm(jac.transpose(1, 2), jac), recon, jac

    # ---- `non_square.py:262-296` ----------------------------------------------------------
    def exact_log_det(self, z: Tensor):
        gram, recon, jac = self.full_jac_transpose_jac(z)
        eps, attempts = 1e-6, 1
        eye = torch.eye(self.d, dtype=gram.dtype, device=gram.device).expand(z.shape[0], self.d, self.d)
        while True:
            try:
                chol = torch.linalg.cholesky(gram)
                break
            except RuntimeError:
                gram = gram + eps * eye
                attempts += 1
                eps *= 10
        log_det = 2 * torch.sum(torch.log(torch.diagonal(chol, dim1=-2, dim2=-1)), dim=1, keepdim=True)
        return log_det, recon, gram, attempts

    # ---- `non_square.py:190-258` (hutch_with_cg; eps injected for reproducibility) ---------
    def approx_log_det(self, z: Tensor, eps_samples: Tensor, training: bool = True):
        """``eps_samples``: (B, d, m) Hutchinson probes (the reference draws them inside,
        `non_square.py:204-213`).  CG = gpytorch.utils.linear_cg semantics restated in
        :func:`batched_cg` (`gitmodules/gpytorch/gpytorch/utils/linear_cg.py:89`)."""
        b, d, m = eps_samples.shape
        zr = z.repeat_interleave(m, dim=0)

        def to_vec(t): return t.transpose(1, 2).reshape(b * m, d)
        def to_ten(v): return v.reshape(b, m, d).transpose(1, 2)

        def jtj_vec(lat, vec, graph):
            if not graph:
                lat = lat.detach()
                with torch.no_grad():
                    recon, jv = self.jvp_forward(lat, vec)
            else:
                recon, jv = self.jvp_forward(lat, vec)
            fwd = lambda t: self.flow_forward(t).flatten(1)
            _, out = torch.autograd.functional.vjp(fwd, lat, jv.flatten(1), create_graph=graph)
            return out, recon

        max_iter = self.head_cfg["max_cg_iterations"]
        max_iter = min(max_iter, d) if max_iter else d          # factory.py:123-129
        inv = batched_cg(lambda t: to_ten(jtj_vec(zr, to_vec(t), False)[0]), eps_samples,
                         max_iter=max_iter, tolerance=self.head_cfg["cg_tolerance"]).detach()
        jtj_eps_vec, recon_rep = jtj_vec(zr, to_vec(eps_samples), training)
        jtj_eps = to_ten(jtj_eps_vec)
        est = torch.mean(torch.sum(inv * jtj_eps, dim=1, keepdim=True), dim=2)
        return est, recon_rep[::m], jtj_eps

    # ---- `non_square.py:64-129` -------------------------------------------------------------
    def elbo(self, x: Tensor, add_reconstruction=True, add_diagonal_metric_reg=False,
             add_offdiagonal_metric_reg=False, likelihood_wt=1.0, metric_wt=1.0, ood=False,
             hutchinson_eps: Optional[Tensor] = None, training: bool = True) -> Dict[str, Tensor]:
        x, z, low_dim_elbo = self.encode(x)
        out: Dict[str, Tensor] = {"z": z, "low_dim_elbo": low_dim_elbo}
        if not math.isclose(likelihood_wt, 0.0, abs_tol=1e-8):     # np.isclose(likelihood_wt, 0.)
            if self.head_cfg["log_jacobian_method"] == "cholesky" or not training:
                log_det, recon, gram, attempts = self.exact_log_det(z)
                out["attempts"] = attempts
            else:
                log_det, recon, gram = self.approx_log_det(z, hutchinson_eps, training)
            likelihood = low_dim_elbo - log_det / 2.0
            if add_diagonal_metric_reg:                               # non_square.py:87-92
                l1 = torch.diagonal(gram, dim1=-2, dim2=-1).abs().sum(1, keepdim=True)
            elif add_offdiagonal_metric_reg:                          # non_square.py:95-100
                off = ~torch.eye(gram.shape[1], dtype=torch.bool, device=gram.device)
                l1 = gram.masked_select(off).view(gram.shape[0], -1).abs().sum(1, keepdim=True)
            else:
                l1 = 0
            out.update(log_det=log_det, gram=gram)
        else:                                                         # non_square.py:105-109
            l1, likelihood = 0, 0
            recon = self.flow_forward(z)
        if add_reconstruction:
            recon_loss = ((recon - x).flatten(1) ** 2).sum(-1, keepdim=True)
        else:
            recon_loss = 0
        out.update(l1=l1, likelihood=likelihood, recon=recon, recon_loss=recon_loss)
        if ood:
            return {"likelihood": likelihood, "reconstruction-error": recon_loss, **out}
        out["head_elbo"] = likelihood_wt * likelihood - self.regularization_param * recon_loss - metric_wt * l1
        out["elbo"] = out["head_elbo"] + self.pre_log_jac        # exact.py:23-30 for the nodes above the head
        return out


def batched_cg(matmul, rhs: Tensor, max_iter: int, tolerance: float,
               stop_updating_after: float = 1e-10) -> Tensor:
    """Preconditioner-free batched conjugate gradients with the exact control flow of
    gpytorch's ``linear_cg`` (`gitmodules/gpytorch/gpytorch/utils/linear_cg.py:89-330`,
    updates `:42-86`): the rhs is normalised by its column 2-norm; ``alpha``/``beta`` use
    the "safe division" (denominator < 1e-10 -> result 0); columns whose residual norm
    fell below ``stop_updating_after`` stop moving; the loop runs at most
    ``min(max_iter, rows)`` times and breaks once ``k >= min(10, max_iter-1)`` and the
    mean residual norm is below ``tolerance``; the solution is un-normalised at the end.
    rhs: (B, d, m); ``matmul`` maps (B, d, m) -> (B, d, m)."""
    eps = 1e-10
    rhs_norm = rhs.norm(2, dim=-2, keepdim=True)
    rhs_is_zero = rhs_norm < eps
    rhs_norm = rhs_norm.masked_fill(rhs_is_zero, 1)
    rhs = rhs / rhs_norm
    x = torch.zeros_like(rhs)
    r = rhs - matmul(x)
    converged = r.norm(2, dim=-2, keepdim=True) < stop_updating_after
    if bool(converged.all()):
        return x * rhs_norm
    p = r.clone()
    rz = (r * r).sum(-2, keepdim=True)
    for k in range(min(max_iter, rhs.shape[-2])):
        ap = matmul(p)
        pap = (p * ap).sum(-2, keepdim=True)
        small = pap < eps
        alpha = (rz / pap.masked_fill(small, 1)).masked_fill(small, 0).masked_fill(converged, 0)
        r = r - alpha * ap
        x = x + alpha * p
        rz_old, rz = rz, (r * r).sum(-2, keepdim=True)
        small = rz_old < eps
        beta = (rz / rz_old.masked_fill(small, 1)).masked_fill(small, 0)
        p = p * beta + r
        res = r.norm(2, dim=-2, keepdim=True).masked_fill(rhs_is_zero, 0)
        converged = res < stop_updating_after
        if k >= min(10, max_iter - 1) and bool(res.mean() < tolerance):
            break
    return x * rhs_norm


# --------------------------------------------------------------------------------------
# loss closure: `non_square_helpers.py:31-135`
# --------------------------------------------------------------------------------------
def loss_weights(config: dict, epoch: int):
    """(likelihood_weight, add_reconstruction) for one epoch; m_flow=False => num_objectives=1."""
    n_obj = 2 if config["m_flow"] else 1
    if config["likelihood_warmup"]:
        lo, hi = n_obj * config["likelihood_warmup_start"], n_obj * config["likelihood_warmup_end"]
        if (epoch + 1) % n_obj == 0:
            lw = 0.0 if epoch <= lo else (1.0 if epoch >= hi else (epoch - lo) / (hi - lo))  # np.interp
        else:
            lw = 0.0
    else:
        lw = float((epoch + 1) % n_obj == 0)
    return lw, epoch % n_obj == 0


def train_loss(flow: OracleFlow, config: dict, x: Tensor, epoch: int = 0, **kw) -> Tensor:
    return -elbo_per_sample(flow, config, x, epoch, **kw).mean()


def elbo_kwargs(config: dict, epoch: int) -> dict:
    """the keyword arguments the loss closure passes to ``density.elbo`` (`non_square_helpers.py:99-123`)"""
    lw, add_rec = loss_weights(config, epoch)
    if config["g_kk_loss"]:
        assert not config["g_ij_loss"]
        return dict(likelihood_wt=lw * config["elbo_regularization_param"],
                    metric_wt=lw * config["metric_regularization_param"], add_reconstruction=add_rec,
                    add_diagonal_metric_reg=add_rec, add_offdiagonal_metric_reg=False)
    if config["g_ij_loss"]:
        assert config["latent_dimension"] != 1
        return dict(likelihood_wt=lw * config["elbo_regularization_param"],
                    metric_wt=lw * config["metric_regularization_param"], add_reconstruction=add_rec,
                    add_diagonal_metric_reg=False, add_offdiagonal_metric_reg=add_rec)
    return dict(likelihood_wt=lw, add_reconstruction=add_rec)


def elbo_per_sample(flow: OracleFlow, config: dict, x: Tensor, epoch: int = 0, **kw) -> Tensor:
    """(B, 1): what the loss closure averages"""
    return flow.elbo(x, **elbo_kwargs(config, epoch), **kw)["elbo"]


# --------------------------------------------------------------------------------------
# parameter inventory: names and shapes of the reference state dict for a schema
# --------------------------------------------------------------------------------------
def _net_spec(prefix: str, net_cfg: dict, in_shape, out_channels: int) -> Dict[str, tuple]:
    spec = {}
    hidden = list(net_cfg["hidden_channels"])
    if net_cfg["type"] == "mlp":                       # networks.py:206-224
        widths = [in_shape[0]] + hidden + [out_channels]
        for i in range(len(widths) - 1):
            spec[f"{prefix}{2 * i}.weight"] = (widths[i + 1], widths[i])
            spec[f"{prefix}{2 * i}.bias"] = (widths[i + 1],)
    elif net_cfg["type"] == "resnet":                  # networks.py:116-161
        c = hidden[0] if hidden else out_channels
        spec[prefix + "weights"] = (out_channels, 1, 1)
        spec[prefix + "bias"] = (out_channels, 1, 1)
        spec[prefix + "module.0.weight"] = (c, in_shape[0], 3, 3)
        for i, c in enumerate(hidden, start=1):
            for conv in ("conv1", "conv2"):
                spec[f"{prefix}module.{i}.{conv}.weight"] = (c, c, 3, 3)
                spec[f"{prefix}module.{i}.{conv}.bias"] = (c,)
        spec[f"{prefix}module.{len(hidden) + 2}.weight"] = (out_channels, c, 1, 1)
        spec[f"{prefix}module.{len(hidden) + 2}.bias"] = (out_channels,)
    else:
        raise ValueError(net_cfg["type"])
    return spec


def param_spec(schema: Sequence[dict], x_shape: Sequence[int]) -> Dict[str, tuple]:
    """{reference state-dict key (no 'module.' prefix): shape} for every learnable tensor and
    tail buffer the path touches (`factory.py:86-162, 383-507`)."""
    spec: Dict[str, tuple] = {}
    prefix, shape = "", tuple(x_shape)
    for layer in schema:
        ty = layer["type"]
        if ty == "dequantization":
            prefix += "density."
        elif ty == "split":
            shape = (shape[0] // 2, *shape[1:]); prefix += "density_1."
        elif ty == "non-square-base":
            n = int(math.prod(shape))
            for b in ("mask", "permutation", "inverse_permutation"):
                spec[prefix + b] = (n,)
            shape = (layer["latent_dimension"],); prefix += "prior."
        elif ty == "acl":
            c = shape[0]
            if layer["mask_type"] == "checkerboard":
                n_in, n_out = c, c
            else:
                first = c // 2 if layer["mask_type"] == "split-channel" else (c + 1) // 2
                n_in = c - first if layer["reverse_mask"] else first
                n_out = c - n_in
            cp = prefix + "bijection.coupler."
            cc = layer["coupler"]
            if cc["independent_nets"]:
                spec.update(_net_spec(cp + "shift_net.", cc["shift_net"], (n_in, *shape[1:]), n_out))
                spec.update(_net_spec(cp + "log_scale_net.", cc["log_scale_net"], (n_in, *shape[1:]), n_out))
            else:
                spec.update(_net_spec(cp + "shift_log_scale_net.", cc["shift_log_scale_net"],
                                      (n_in, *shape[1:]), 2 * n_out))
            prefix += "prior."
        elif ty == "affine":
            spec[prefix + "bijection.shift"] = shape
            spec[prefix + "bijection.log_scale"] = shape
            prefix += "prior."
        elif ty == "flatten":
            shape = (int(math.prod(shape)),); prefix += "prior."
        elif ty == "rand-channel-perm":
            spec[prefix + "bijection.permutation"] = (shape[0],)
            spec[prefix + "bijection.inverse_permutation"] = (shape[0],)
            prefix += "prior."
        elif ty == "nsf-c":
            n, K, hdn = shape[0], layer["num_bins"], layer["num_hidden_channels"]
            n_tr = len(range(0 if layer["reverse_mask"] else 1, n, 2))
            n_id = n - n_tr
            fp = prefix + "bijection.flow."
            spec[fp + "transform_net.initial_layer.weight"] = (hdn, n_id)
            spec[fp + "transform_net.initial_layer.bias"] = (hdn,)
            for i in range(layer["num_hidden_layers"]):
                for j in (0, 1):
                    spec[f"{fp}transform_net.blocks.{i}.linear_layers.{j}.weight"] = (hdn, hdn)
                    spec[f"{fp}transform_net.blocks.{i}.linear_layers.{j}.bias"] = (hdn,)
            spec[fp + "transform_net.final_layer.weight"] = (n_tr * (3 * K - 1), hdn)
            spec[fp + "transform_net.final_layer.bias"] = (n_tr * (3 * K - 1),)
            spec[fp + "unconditional_transform.unnormalized_widths"] = (n_id, K)
            spec[fp + "unconditional_transform.unnormalized_heights"] = (n_id, K)
            spec[fp + "unconditional_transform.unnormalized_derivatives"] = (n_id, K - 1)
            prefix += "prior."
        elif ty == "squeeze":
            f = layer["factor"]
            shape = (shape[0] * f * f, shape[1] // f, shape[2] // f); prefix += "prior."
        else:  # head, elementwise preprocessing
            prefix += "prior."
    return spec


def init_params(schema: Sequence[dict], x_shape: Sequence[int], dtype=torch.float32) -> Params:
    """Zero-valued params with the right names/shapes/dtypes (fill with ``detfill.fill_state_dict``)."""
    out: Params = {}
    for key, shape in param_spec(schema, x_shape).items():
        leaf = key.rsplit(".", 1)[-1]
        if leaf == "mask":
            d = next(l["latent_dimension"] for l in schema if l["type"] == "non-square-base")
            out[key] = torch.arange(shape[0]) < d            # non_square.py:377
        elif leaf in ("permutation", "inverse_permutation"):
            out[key] = torch.arange(shape[0])
        else:
            out[key] = torch.zeros(shape, dtype=dtype)
    return out
