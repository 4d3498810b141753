"""Tensor-level launchers over the C ABI (``include/cmf_b200.h``): validate, allocate outputs with
torch (device memory + stream plumbing only), pass raw pointers.  Every function requires CUDA
fp32 contiguous tensors and raises otherwise -- there is no CPU path.
"""
from __future__ import annotations

from typing import Optional, Sequence, Tuple

import torch

from . import _lib

Tensor = torch.Tensor

PENALTY = {"none": 0, "offdiag": 1, "diag": 2}


def _ptr(t: Optional[Tensor]):
    return None if t is None else t.data_ptr()


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _req(t: Tensor, name: str, dtype=torch.float32) -> Tensor:
    if not t.is_cuda:
        raise _lib.CmfB200Error(f"{name} must be a CUDA tensor (cmf_b200 has no CPU fallback)")
    if t.dtype != dtype:
        raise _lib.CmfB200Error(f"{name} must be {dtype}, got {t.dtype}")
    return t if t.is_contiguous() else t.contiguous()


def coupling_jvp_fwd(z_m: Tensor, V_m: Optional[Tensor], st: Tensor, ST: Optional[Tensor],
                     map_: Optional[Tensor]) -> Tuple[Tensor, Optional[Tensor]]:
    """z_m (B,M), V_m (B,d,M)|None, st (B,2Q), ST (B,d,2Q)|None, map (M,) int32|None."""
    lib = _lib.load()
    z_m, st = _req(z_m, "z_m"), _req(st, "st")
    B, M = z_m.shape
    Q = st.shape[1] // 2
    d = 0 if V_m is None else V_m.shape[1]
    if d:
        V_m, ST = _req(V_m, "V_m"), _req(ST, "ST")
        assert V_m.shape == (B, d, M) and ST.shape == (B, d, 2 * Q), (V_m.shape, ST.shape)
    if map_ is not None:
        map_ = _req(map_, "map", torch.int32)
        assert map_.numel() == M
    x_m = torch.empty_like(z_m)
    V_out = torch.empty_like(V_m) if d else None
    with torch.cuda.device(z_m.device):
        _lib.check(lib.cmf_coupling_jvp_fwd(B, d, M, Q, _ptr(z_m), _ptr(V_m), _ptr(st), _ptr(ST),
                                            _ptr(map_), _ptr(x_m), _ptr(V_out), _stream()),
                   "cmf_coupling_jvp_fwd")
    return x_m, V_out


def coupling_jvp_bwd(z_m, V_m, st, ST, map_, inv, g_x, g_V):
    """returns (g_z (B,M), g_Vin (B,d,M)|None, g_st (B,2Q), g_ST (B,d,2Q)|None)"""
    lib = _lib.load()
    z_m, st = _req(z_m, "z_m"), _req(st, "st")
    B, M = z_m.shape
    Q = st.shape[1] // 2
    d = 0 if V_m is None else V_m.shape[1]
    if d:
        V_m, ST, g_V = _req(V_m, "V_m"), _req(ST, "ST"), _req(g_V, "g_V")
    if g_x is not None:
        g_x = _req(g_x, "g_x")
    g_z, g_st = torch.empty_like(z_m), torch.empty_like(st)
    g_Vin = torch.empty_like(V_m) if d else None
    g_ST = torch.empty_like(ST) if d else None
    nbytes = lib.cmf_coupling_jvp_bwd_workspace(B, d, M, Q)
    ws = torch.empty(max(nbytes // 4, 1), dtype=torch.float32, device=z_m.device) if nbytes else None
    with torch.cuda.device(z_m.device):
        _lib.check(lib.cmf_coupling_jvp_bwd(B, d, M, Q, _ptr(z_m), _ptr(V_m), _ptr(st), _ptr(ST),
                                            _ptr(map_), _ptr(inv), _ptr(g_x), _ptr(g_V),
                                            _ptr(g_z), _ptr(g_Vin), _ptr(g_st), _ptr(g_ST),
                                            _ptr(ws), nbytes, _stream()),
                   "cmf_coupling_jvp_bwd")
    return g_z, g_Vin, g_st, g_ST


def gather_rows(in0: Tensor, in1: Optional[Tensor], idx: Tensor) -> Tensor:
    """out[..., i] = idx[i] < 0 ? 0 : cat(in0, in1, -1)[..., idx[i]]; leading dims are rows."""
    lib = _lib.load()
    in0 = _req(in0, "in0")
    idx = _req(idx, "idx", torch.int32)
    lead = in0.shape[:-1]
    n0 = in0.shape[-1]
    n1 = 0
    if in1 is not None:
        in1 = _req(in1, "in1")
        assert in1.shape[:-1] == lead
        n1 = in1.shape[-1]
    rows = 1
    for s in lead:
        rows *= s
    out = torch.empty(*lead, idx.numel(), dtype=torch.float32, device=in0.device)
    with torch.cuda.device(in0.device):
        _lib.check(lib.cmf_gather_rows(rows, n0, n1, idx.numel(), _ptr(in0), _ptr(in1), _ptr(idx),
                                       _ptr(out), _stream()), "cmf_gather_rows")
    return out


GRAM_PATH = {"auto": 0, "fma": 1, "tc": 2}


def gram_fwd(V: Tensor, G: Optional[Tensor] = None, path: str = "auto") -> Tensor:
    """G (B,d,d) (+)= V V^T for V (B,d,n); accumulates into G when given.  ``path``: "auto" |
    "fma" (FP32 FMA SYRK) | "tc" (tcgen05 3xTF32, raises when the shape does not qualify)."""
    lib = _lib.load()
    V = _req(V, "V")
    B, d, n = V.shape
    acc = G is not None
    if G is None:
        G = torch.empty(B, d, d, dtype=torch.float32, device=V.device)
    with torch.cuda.device(V.device):
        _lib.check(lib.cmf_gram_fwd_path(B, d, n, _ptr(V), _ptr(G), int(acc), GRAM_PATH[path], _stream()),
                   "cmf_gram_fwd")
    return G


def gram_fwd2(V1: Tensor, V2: Tensor, G: Optional[Tensor] = None) -> Tensor:
    """G (B,d,d) (+)= V1 V1^T + V2 V2^T for two blocks (B,d,n1), (B,d,n2) of the same samples: one launch,
    one accumulation over n1 + n2 columns when both qualify for the tcgen05 kernel."""
    lib = _lib.load()
    V1, V2 = _req(V1, "V1"), _req(V2, "V2")
    B, d, n1 = V1.shape
    if V2.shape[:2] != (B, d):
        raise ValueError(f"gram_fwd2: blocks disagree: {tuple(V1.shape)} vs {tuple(V2.shape)}")
    acc = G is not None
    if G is None:
        G = torch.empty(B, d, d, dtype=torch.float32, device=V1.device)
    with torch.cuda.device(V1.device):
        _lib.check(lib.cmf_gram_fwd2(B, d, n1, _ptr(V1), V2.shape[2], _ptr(V2), _ptr(G), int(acc), _stream()),
                   "cmf_gram_fwd2")
    return G


def gram_bwd(V: Tensor, g_G: Tensor, path: str = "auto") -> Tensor:
    lib = _lib.load()
    V, g_G = _req(V, "V"), _req(g_G, "g_G")
    B, d, n = V.shape
    g_V = torch.empty_like(V)
    with torch.cuda.device(V.device):
        _lib.check(lib.cmf_gram_bwd_path(B, d, n, _ptr(V), _ptr(g_G), _ptr(g_V), GRAM_PATH[path], _stream()),
                   "cmf_gram_bwd")
    return g_V


def chol_logdet_fwd(G: Tensor):
    """returns (L (B,d,d), logdet (B,), l1_off (B,), l1_diag (B,), info int32[1])"""
    lib = _lib.load()
    G = _req(G, "G")
    B, d, _ = G.shape
    L = torch.empty_like(G)
    logdet = torch.empty(B, dtype=torch.float32, device=G.device)
    l1_off, l1_diag = torch.empty_like(logdet), torch.empty_like(logdet)
    info = torch.zeros(1, dtype=torch.int32, device=G.device)
    with torch.cuda.device(G.device):
        _lib.check(lib.cmf_chol_logdet_fwd(B, d, _ptr(G), _ptr(L), _ptr(logdet), _ptr(l1_off),
                                           _ptr(l1_diag), _ptr(info), _stream()), "cmf_chol_logdet_fwd")
    return L, logdet, l1_off, l1_diag, info


def chol_logdet_bwd(L: Tensor, G: Tensor, g_logdet: Tensor, g_l1: Optional[Tensor], penalty: str) -> Tensor:
    lib = _lib.load()
    L, G, g_logdet = _req(L, "L"), _req(G, "G"), _req(g_logdet, "g_logdet")
    if g_l1 is not None:
        g_l1 = _req(g_l1, "g_l1")
    B, d, _ = G.shape
    g_G = torch.empty_like(G)
    with torch.cuda.device(G.device):
        _lib.check(lib.cmf_chol_logdet_bwd(B, d, _ptr(L), _ptr(G), _ptr(g_logdet), _ptr(g_l1),
                                           PENALTY[penalty], _ptr(g_G), _stream()), "cmf_chol_logdet_bwd")
    return g_G


def rqs_coupling_jvp_fwd(z_m: Tensor, V_m: Optional[Tensor], cw: Tensor, ch: Tensor, dv: Tensor,
                         CW: Optional[Tensor], CH: Optional[Tensor], DV: Optional[Tensor], tail_bound: float):
    """z_m (B,M), V_m (B,d,M)|None, cw/ch/dv (B,M,K+1), CW/CH/DV (B,d,M,K+1)|None ->
    (x_m (B,M), V_out (B,d,M)|None, bin (B,M) int32)."""
    lib = _lib.load()
    z_m, cw, ch, dv = _req(z_m, "z_m"), _req(cw, "cw"), _req(ch, "ch"), _req(dv, "dv")
    B, M = z_m.shape
    K = cw.shape[-1] - 1
    assert cw.shape == ch.shape == dv.shape == (B, M, K + 1), (cw.shape, ch.shape, dv.shape)
    d = 0 if V_m is None else V_m.shape[1]
    if d:
        V_m, CW, CH, DV = _req(V_m, "V_m"), _req(CW, "CW"), _req(CH, "CH"), _req(DV, "DV")
        assert V_m.shape == (B, d, M) and CW.shape == CH.shape == DV.shape == (B, d, M, K + 1)
    x_m = torch.empty_like(z_m)
    V_out = torch.empty_like(V_m) if d else None
    bins = torch.empty(B, M, dtype=torch.int32, device=z_m.device)
    with torch.cuda.device(z_m.device):
        _lib.check(lib.cmf_rqs_coupling_jvp_fwd(B, d, M, K, float(tail_bound), _ptr(z_m), _ptr(V_m), _ptr(cw), _ptr(ch),
                                                _ptr(dv), _ptr(CW), _ptr(CH), _ptr(DV), _ptr(x_m), _ptr(V_out),
                                                _ptr(bins), _stream()), "cmf_rqs_coupling_jvp_fwd")
    return x_m, V_out, bins


def rqs_coupling_jvp_bwd(z_m, V_m, cw, ch, dv, CW, CH, DV, tail_bound, g_x, g_V):
    """returns (g_z, g_Vin|None, g_cw, g_ch, g_dv, g_CW|None, g_CH|None, g_DV|None)"""
    lib = _lib.load()
    z_m, cw, ch, dv = _req(z_m, "z_m"), _req(cw, "cw"), _req(ch, "ch"), _req(dv, "dv")
    B, M = z_m.shape
    K = cw.shape[-1] - 1
    d = 0 if V_m is None else V_m.shape[1]
    if d:
        V_m, CW, CH, DV, g_V = _req(V_m, "V_m"), _req(CW, "CW"), _req(CH, "CH"), _req(DV, "DV"), _req(g_V, "g_V")
    if g_x is not None:
        g_x = _req(g_x, "g_x")
    g_z, g_cw, g_ch, g_dv = torch.empty_like(z_m), torch.empty_like(cw), torch.empty_like(ch), torch.empty_like(dv)
    g_Vin = torch.empty_like(V_m) if d else None
    g_CW, g_CH, g_DV = (torch.empty_like(CW), torch.empty_like(CH), torch.empty_like(DV)) if d else (None, None, None)
    with torch.cuda.device(z_m.device):
        _lib.check(lib.cmf_rqs_coupling_jvp_bwd(B, d, M, K, float(tail_bound), _ptr(z_m), _ptr(V_m), _ptr(cw), _ptr(ch),
                                                _ptr(dv), _ptr(CW), _ptr(CH), _ptr(DV), _ptr(g_x), _ptr(g_V),
                                                _ptr(g_z), _ptr(g_Vin), _ptr(g_cw), _ptr(g_ch), _ptr(g_dv),
                                                _ptr(g_CW), _ptr(g_CH), _ptr(g_DV), _stream()),
                   "cmf_rqs_coupling_jvp_bwd")
    return g_z, g_Vin, g_cw, g_ch, g_dv, g_CW, g_CH, g_DV


# ------------------------------------------------------------------ conditioner convolution (N1)
CONV_WSPLIT_FLOATS = 2 * 9 * 64 * 64


def conv3x3_c64_usable(N: int, H: int, W: int) -> bool:
    return bool(_lib.load().cmf_conv3x3_c64_usable(int(N), int(H), int(W)))


def conv3x3_c64_prep(weight: Tensor, transpose: bool) -> Tensor:
    """weight (64,64,3,3) -> pre-split (TF32 head | remainder), tap-major weights for
    ``conv3x3_c64``; ``transpose`` selects the data-gradient variant."""
    lib = _lib.load()
    weight = _req(weight, "weight")
    assert weight.shape == (64, 64, 3, 3), weight.shape
    out = torch.empty(CONV_WSPLIT_FLOATS, dtype=torch.float32, device=weight.device)
    with torch.cuda.device(weight.device):
        _lib.check(lib.cmf_conv3x3_c64_prep(_ptr(weight), int(bool(transpose)), _ptr(out), _stream()),
                   "cmf_conv3x3_c64_prep")
    return out


def relu_mask_pack(x: Tensor, relu_out: Optional[Tensor] = None) -> Tensor:
    """x (N,64,H,W) primal activations -> packed signs (N, H*W, 2) int32 for ``conv3x3_c64(mask_src=...)``:
    bit e of word 0 / 1 = (x[n, e, p] > 0) / (x[n, 32 + e, p] > 0).  With ``relu_out`` (same shape, contiguous, not
    x itself) the same pass also writes max(x, 0) there."""
    lib = _lib.load()
    x = _req(x, "x")
    N, C, H, W = x.shape
    assert C == 64
    bits = torch.empty(N, H * W, 2, dtype=torch.int32, device=x.device)
    with torch.cuda.device(x.device):
        if relu_out is None:
            _lib.check(lib.cmf_relu_mask_pack(N, H * W, _ptr(x), _ptr(bits), _stream()), "cmf_relu_mask_pack")
        else:
            assert relu_out.is_cuda and relu_out.dtype == torch.float32 and relu_out.is_contiguous() and relu_out.shape == x.shape
            assert relu_out.data_ptr() != x.data_ptr()
            _lib.check(lib.cmf_relu_and_pack(N, H * W, _ptr(x), _ptr(bits), _ptr(relu_out), _stream()), "cmf_relu_and_pack")
    return bits


def conv3x3_c64(x: Tensor, wsplit: Tensor, bias: Optional[Tensor] = None, mask_src: Optional[Tensor] = None,
                mask_div: int = 1, mask_first: bool = False, res: Optional[Tensor] = None, second: bool = False,
                drain: int = 1, tail: int = 0, out: Optional[Tensor] = None, out2: Optional[Tensor] = None):
    """x (N,64,H,W), wsplit from ``conv3x3_c64_prep`` -> y (N,64,H,W), with the fused epilogue of
    ``cmf_conv3x3_c64_fused``:  t = conv(x) + bias; if mask_first: t *= m; if res: t += res; y = t;
    ``second`` additionally returns y2 = t * m  (m = mask_src[n // mask_div] > 0).  ``mask_src`` is
    either the float activations (N // mask_div, 64, H, W) or their packed signs from
    ``relu_mask_pack`` (int32, (N // mask_div, H*W, 2)).
    ``tail`` = Nt > 0 selects the joint batch of ``cmf_conv3x3_c64_joint``: images [0, Nt) are tangents
    (mask row n // mask_div, no bias), images [Nt, N) the primal images of samples 0 .. N - Nt - 1
    (mask row n - Nt, bias).  ``out`` / ``out2``: write into these (contiguous, same shape as x)."""
    lib = _lib.load()
    x, wsplit = _req(x, "x"), _req(wsplit, "wsplit")
    N, C, H, W = x.shape
    assert C == 64 and wsplit.numel() == CONV_WSPLIT_FLOATS
    if bias is not None:
        bias = _req(bias, "bias")
        assert bias.numel() == 64
    n_rows = (N - tail) if tail else (N // mask_div if mask_src is not None else 0)
    packed = mask_src is not None and mask_src.dtype == torch.int32
    if packed:
        assert mask_src.is_cuda and mask_src.is_contiguous()
        assert (tail or N) % mask_div == 0 and tuple(mask_src.shape) == (n_rows, H * W, 2), (mask_src.shape, x.shape, mask_div, tail)
    elif mask_src is not None:
        assert not tail, "the joint batch takes packed masks"
        mask_src = _req(mask_src, "mask_src")
        assert N % mask_div == 0 and tuple(mask_src.shape) == (N // mask_div, 64, H, W), (mask_src.shape, x.shape, mask_div)
    if res is not None:
        res = _req(res, "res")
        assert res.shape == x.shape
    assert not second or mask_src is not None
    for o in (out, out2):
        if o is not None:
            assert o.is_cuda and o.dtype == torch.float32 and o.is_contiguous() and o.shape == x.shape
    y = torch.empty_like(x) if out is None else out
    y2 = (torch.empty_like(x) if out2 is None else out2) if second else None
    with torch.cuda.device(x.device):
        if tail:
            assert mask_src is None or tail // mask_div == N - tail
            _lib.check(lib.cmf_conv3x3_c64_joint(N, int(tail), H, W, _ptr(x), _ptr(wsplit), _ptr(bias), _ptr(mask_src),
                                                 int(mask_div), int(bool(mask_first)), _ptr(res), _ptr(y), _ptr(y2),
                                                 int(drain), _stream()), "cmf_conv3x3_c64_joint")
        elif packed or mask_src is None:      # the extended entry point (drain interval); mask_bits may be NULL
            _lib.check(lib.cmf_conv3x3_c64_fused_bits(N, H, W, _ptr(x), _ptr(wsplit), _ptr(bias), _ptr(mask_src),
                                                      int(mask_div), int(bool(mask_first)), _ptr(res), _ptr(y), _ptr(y2),
                                                      int(drain), _stream()), "cmf_conv3x3_c64_fused_bits")
        else:
            assert drain == 1
            _lib.check(lib.cmf_conv3x3_c64_fused(N, H, W, _ptr(x), _ptr(wsplit), _ptr(bias), _ptr(mask_src), int(mask_div),
                                                 int(bool(mask_first)), _ptr(res), _ptr(y), _ptr(y2), _stream()),
                       "cmf_conv3x3_c64_fused")
    return (y, y2) if second else y


def conv3x3_c64_wgrad_usable(N: int, H: int, W: int) -> bool:
    return bool(_lib.load().cmf_conv3x3_c64_wgrad_usable(int(N), int(H), int(W)))


def conv3x3_c64_wgrad(x: Tensor, g_y: Tensor, deterministic: Optional[bool] = None) -> Tensor:
    """x, g_y (N,64,H,W) -> g_w (64,64,3,3) = weight gradient of ``conv2d(x, w, padding=1)``.
    ``deterministic``: bitwise run-to-run reproducible accumulation order (~9 % slower); None follows
    ``torch.are_deterministic_algorithms_enabled()`` like aten's own convolution_backward."""
    lib = _lib.load()
    x, g_y = _req(x, "x"), _req(g_y, "g_y")
    N, C, H, W = x.shape
    assert C == 64 and g_y.shape == x.shape
    if deterministic is None:
        deterministic = torch.are_deterministic_algorithms_enabled()
    g_w = torch.empty(64, 64, 3, 3, dtype=torch.float32, device=x.device)
    nbytes = lib.cmf_conv3x3_c64_wgrad_workspace(N, H, W)
    ws = torch.empty(max(nbytes // 4, 1), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(lib.cmf_conv3x3_c64_wgrad_det(N, H, W, _ptr(x), _ptr(g_y), _ptr(ws), 0, _ptr(g_w),
                                                 int(bool(deterministic)), _stream()),
                   "cmf_conv3x3_c64_wgrad")
    return g_w


# ------------------------------------------------------------------ encode pass / loss epilogue (N4)
def preproc_logit_fwd(x: Tensor, noise: Optional[Tensor], mult: float, add: float) -> Tuple[Tensor, Tensor]:
    """x (B, ...) gets ``noise`` added IN PLACE (wrapper.py:27-30); returns (logit(mult * x + add) with x's
    shape, log-Jacobian (B,1))."""
    lib = _lib.load()
    if not (x.is_cuda and x.dtype == torch.float32 and x.is_contiguous()):
        raise _lib.CmfB200Error("preproc_logit: x must be a contiguous CUDA float32 tensor (it is updated in place)")
    B = x.shape[0]
    n = x[0].numel() if B else 0
    if noise is not None:
        noise = _req(noise, "noise")
        assert noise.shape == x.shape
    y = torch.empty_like(x)
    lj = torch.empty(B, 1, dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(lib.cmf_preproc_logit_fwd(B, n, _ptr(x), _ptr(noise), float(mult), float(add), _ptr(y), _ptr(lj), _stream()),
                   "cmf_preproc_logit_fwd")
    return y, lj


def coupling_x_to_z_fwd(x: Tensor, st: Tensor, pos2j: Tensor, qmap: Optional[Tensor]) -> Tuple[Tensor, Tensor]:
    """x (B,D) flat, st (B,2Q), pos2j (D) int32, qmap (M) int32|None -> z (B,D), log-jac (B,1)"""
    lib = _lib.load()
    x, st, pos2j = _req(x, "x"), _req(st, "st"), _req(pos2j, "pos2j", torch.int32)
    B, D = x.shape
    Q = st.shape[1] // 2
    assert pos2j.numel() == D
    z = torch.empty_like(x)
    lj = torch.empty(B, 1, dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(lib.cmf_coupling_x_to_z_fwd(B, D, Q, _ptr(x), _ptr(st), _ptr(pos2j), _ptr(qmap), _ptr(z), _ptr(lj), _stream()),
                   "cmf_coupling_x_to_z_fwd")
    return z, lj


def coupling_x_to_z_bwd(z: Tensor, st: Tensor, pos2j: Tensor, qmap: Optional[Tensor], q2pos: Tensor,
                        g_z: Optional[Tensor], g_lj: Optional[Tensor]) -> Tuple[Tensor, Tensor]:
    lib = _lib.load()
    z, st = _req(z, "z"), _req(st, "st")
    B, D = z.shape
    Q = st.shape[1] // 2
    g_z = None if g_z is None else _req(g_z, "g_z")
    g_lj = None if g_lj is None else _req(g_lj, "g_lj")
    g_x, g_st = torch.empty_like(z), torch.empty_like(st)
    with torch.cuda.device(z.device):
        _lib.check(lib.cmf_coupling_x_to_z_bwd(B, D, Q, _ptr(z), _ptr(st), _ptr(pos2j), _ptr(qmap), _ptr(q2pos), _ptr(g_z),
                                               _ptr(g_lj), _ptr(g_x), _ptr(g_st), _stream()), "cmf_coupling_x_to_z_bwd")
    return g_x, g_st


def sumsq_diff_fwd(a: Tensor, c: Optional[Tensor]) -> Tensor:
    """a, c (B,n) -> (B,1): sum_i (a - c)^2 (c None: sum a^2)"""
    lib = _lib.load()
    a = _req(a, "a")
    c = None if c is None else _req(c, "c")
    B, n = a.shape
    out = torch.empty(B, 1, dtype=torch.float32, device=a.device)
    with torch.cuda.device(a.device):
        _lib.check(lib.cmf_sumsq_diff_fwd(B, n, _ptr(a), _ptr(c), _ptr(out), _stream()), "cmf_sumsq_diff_fwd")
    return out


def sumsq_diff_bwd(a: Tensor, c: Optional[Tensor], g: Tensor) -> Tensor:
    lib = _lib.load()
    a, g = _req(a, "a"), _req(g, "g")
    c = None if c is None else _req(c, "c")
    B, n = a.shape
    g_a = torch.empty_like(a)
    with torch.cuda.device(a.device):
        _lib.check(lib.cmf_sumsq_diff_bwd(B, n, _ptr(a), _ptr(c), _ptr(g), _ptr(g_a), _stream()), "cmf_sumsq_diff_bwd")
    return g_a


# ------------------------------------------------------------------ whole-stack kernel, small flat configs (N1-ii)
def flat_stack_usable(D: int, d: int, L: int, nh: int, max_width: int, P: int) -> bool:
    return bool(_lib.load().cmf_flat_stack_usable(int(D), int(d), int(L), int(nh), int(max_width), int(P)))


def _desc_ptr(desc):
    import numpy as np
    assert isinstance(desc, np.ndarray) and desc.dtype == np.int32 and desc.flags["C_CONTIGUOUS"]
    return desc.ctypes.data, int(desc.size)


def flat_stack_fwd(desc, z: Tensor, params: Tensor) -> Tuple[Tensor, Tensor]:
    """desc: host int32 descriptor (see include/cmf_b200.h), z (B,d), params (P,) -> G (B,d,d), x (B,D)"""
    lib = _lib.load()
    z, params = _req(z, "z"), _req(params, "params")
    B, d = z.shape
    D = int(desc[0])
    G = torch.empty(B, d, d, dtype=torch.float32, device=z.device)
    x = torch.empty(B, D, dtype=torch.float32, device=z.device)
    ptr, n = _desc_ptr(desc)
    with torch.cuda.device(z.device):
        _lib.check(lib.cmf_flat_stack_fwd(ptr, n, B, _ptr(z), _ptr(params), _ptr(G), _ptr(x), _stream()), "cmf_flat_stack_fwd")
    return G, x


def flat_stack_bwd(desc, z: Tensor, params: Tensor, g_G: Optional[Tensor], g_x: Optional[Tensor]) -> Tuple[Tensor, Tensor]:
    """-> g_z (B,d), g_params (P,) (the per-CTA partials summed in a fixed order)"""
    lib = _lib.load()
    z, params = _req(z, "z"), _req(params, "params")
    g_G = None if g_G is None else _req(g_G, "g_G")
    g_x = None if g_x is None else _req(g_x, "g_x")
    B, d = z.shape
    P = params.numel()
    g_z = torch.empty_like(z)
    nparts = lib.cmf_flat_stack_bwd_partials(B)
    parts = torch.empty(max(nparts, 1), P, dtype=torch.float32, device=z.device)
    ptr, n = _desc_ptr(desc)
    with torch.cuda.device(z.device):
        _lib.check(lib.cmf_flat_stack_bwd(ptr, n, B, _ptr(z), _ptr(params), _ptr(g_G), _ptr(g_x), _ptr(g_z), _ptr(parts),
                                          _stream()), "cmf_flat_stack_bwd")
    return g_z, (parts.sum(0) if nparts else torch.zeros_like(params))


# ------------------------------------------------------------------------------------------------
# tangent-batched tanh-MLP conditioner (flat couplers)
# ------------------------------------------------------------------------------------------------
MLP_TILE = 128
MLP_PART_PER_LIN = MLP_TILE * MLP_TILE + MLP_TILE


def _int_array(values):
    import ctypes as C
    return (C.c_int * len(values))(*[int(v) for v in values])


def _ptr_array(tensors):
    import ctypes as C
    return (C.c_void_p * len(tensors))(*[None if t is None else t.data_ptr() for t in tensors])


def mlp_usable(widths: Sequence[int], d: int, x: Tensor) -> bool:
    """the fused MLP launches take these widths / this tangent count, and x is theirs to take (fp32 CUDA matrix)"""
    return (x.is_cuda and x.dtype == torch.float32 and x.dim() == 2
            and bool(_lib.load().cmf_mlp_usable(len(widths) - 1, _int_array(widths), int(d))))


def mlp_pack(weights: Sequence[Tensor], biases: Sequence[Tensor]) -> Tuple[Tensor, list]:
    """Linear weights (out, in) / biases (out,) -> (wpack, widths): both operand layouts, zero padded to 128 x 128."""
    lib = _lib.load()
    weights = [_req(w, "weight") for w in weights]
    biases = [_req(b, "bias") for b in biases]
    widths = [weights[0].shape[1]] + [w.shape[0] for w in weights]
    for l, w in enumerate(weights):
        if w.shape[1] != widths[l] or biases[l].shape != (widths[l + 1],):
            raise _lib.CmfB200Error("mlp_pack: Linear shapes do not chain")
    nlin = len(weights)
    wpack = torch.empty(lib.cmf_mlp_pack_floats(nlin), dtype=torch.float32, device=weights[0].device)
    with torch.cuda.device(wpack.device):
        _lib.check(lib.cmf_mlp_pack(nlin, _int_array(widths), _ptr_array(weights), _ptr_array(biases), _ptr(wpack), _stream()),
                   "cmf_mlp_pack")
    return wpack, widths


def mlp_jvp_fwd(x: Tensor, V: Optional[Tensor], wpack: Tensor, widths: Sequence[int], save: bool):
    """x (B, n_in), V (B, d, n_in) or None -> st (B, n_out), ST (B, d, n_out) or None, saved = (save_T, save_U) or None"""
    lib = _lib.load()
    x = _req(x, "x")
    V = None if V is None else _req(V, "V")
    B, d, nlin = x.shape[0], (0 if V is None else V.shape[1]), len(widths) - 1
    if x.shape[1] != widths[0] or (V is not None and (V.shape[0] != B or V.shape[2] != widths[0])):
        raise _lib.CmfB200Error("mlp_jvp_fwd: input shapes do not match the first Linear")
    st = torch.empty(B, widths[-1], dtype=torch.float32, device=x.device)
    ST = None if V is None else torch.empty(B, d, widths[-1], dtype=torch.float32, device=x.device)
    saved = None
    if save:
        n = lib.cmf_mlp_save_floats(B, d, nlin)
        saved = (torch.empty(max(n, 4), dtype=torch.float32, device=x.device), torch.empty(max(n, 4), dtype=torch.float32, device=x.device))
    with torch.cuda.device(x.device):
        _lib.check(lib.cmf_mlp_jvp_fwd(B, d, nlin, _int_array(widths), _ptr(x), _ptr(V), _ptr(wpack), _ptr(st), _ptr(ST),
                                       _ptr(saved[0]) if save else None, _ptr(saved[1]) if save else None, _stream()),
                   "cmf_mlp_jvp_fwd")
    return st, ST, saved


def mlp_jvp_bwd(x: Tensor, V: Optional[Tensor], wpack: Tensor, widths: Sequence[int], saved, g_st: Tensor, g_ST: Optional[Tensor]):
    """-> g_x, g_V (or None), [g_W[l] (out, in) views], [g_b[l]] (the per-CTA partials reduced in a fixed order)"""
    lib = _lib.load()
    x, g_st = _req(x, "x"), _req(g_st, "g_st")
    V = None if V is None else _req(V, "V")
    B, d, nlin = x.shape[0], (0 if V is None else V.shape[1]), len(widths) - 1
    if V is not None:
        g_ST = torch.zeros(B, d, widths[-1], dtype=torch.float32, device=x.device) if g_ST is None else _req(g_ST, "g_ST")
    g_x = torch.empty_like(x)
    g_V = None if V is None else torch.empty_like(V)
    partial = torch.empty(max(lib.cmf_mlp_partial_floats(B, d, nlin), 4), dtype=torch.float32, device=x.device)
    gpack = torch.empty(nlin, MLP_PART_PER_LIN, dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(lib.cmf_mlp_jvp_bwd(B, d, nlin, _int_array(widths), _ptr(x), _ptr(V), _ptr(wpack), _ptr(saved[0]), _ptr(saved[1]),
                                       _ptr(g_st), _ptr(g_ST), _ptr(g_x), _ptr(g_V), _ptr(partial), _ptr(gpack), _stream()),
                   "cmf_mlp_jvp_bwd")
    g_W = [gpack[l, :MLP_TILE * MLP_TILE].view(MLP_TILE, MLP_TILE)[:widths[l + 1], :widths[l]] for l in range(nlin)]
    g_b = [gpack[l, MLP_TILE * MLP_TILE:MLP_TILE * MLP_TILE + widths[l + 1]] for l in range(nlin)]
    return g_x, g_V, g_W, g_b


# ------------------------------------------------------------------------------------------------
# tail of the image couplers' conditioner: ReLU -> 1x1 conv -> scaled tanh, primal + tangents
# ------------------------------------------------------------------------------------------------
def resnet_tail_usable(x: Tensor, C: int, num_out: int) -> bool:
    """the fused tail takes C hidden channels -> num_out outputs, and x (the net's input) is fp32 on a CUDA device"""
    return x.is_cuda and x.dtype == torch.float32 and bool(_lib.load().cmf_resnet_tail_usable(int(C), int(num_out)))


def resnet_tail_fwd(h: Tensor, Hv: Optional[Tensor], inv: Optional[Tensor], P: int, w2: Tensor, b2: Tensor, sw: Tensor, sb: Tensor):
    """h (B, C, HW), Hv (B d, C, HW) or None, inv int32 (HW) or None, w2 (O, C), b2 / sw / sb (O) -> y (B, O, P), Y (B, d, O, P) or None"""
    lib = _lib.load()
    h, w2, b2, sw, sb = _req(h, "h"), _req(w2, "w2"), _req(b2, "b2"), _req(sw, "sw"), _req(sb, "sb")
    B, C, HW = h.shape
    O = w2.shape[0]
    d = 0
    if Hv is not None:
        Hv = _req(Hv, "Hv")
        assert Hv.shape[1:] == (C, HW) and (B == 0 or Hv.shape[0] % B == 0)
        d = Hv.shape[0] // max(B, 1)
    inv = None if inv is None else _req(inv, "inv", torch.int32)
    assert w2.shape == (O, C) and b2.numel() == O and sw.numel() == O and sb.numel() == O and (inv is None or inv.numel() == HW)
    y = torch.empty(B, O, P, dtype=torch.float32, device=h.device)
    Y = None if Hv is None else torch.empty(B, d, O, P, dtype=torch.float32, device=h.device)
    with torch.cuda.device(h.device):
        _lib.check(lib.cmf_resnet_tail_fwd(B, d, C, HW, int(P), O, _ptr(h), _ptr(Hv), _ptr(inv), _ptr(w2), _ptr(b2), _ptr(sw), _ptr(sb),
                                           _ptr(y), _ptr(Y), _stream()), "cmf_resnet_tail_fwd")
    return y, Y


def resnet_tail_bwd(h: Tensor, Hv: Optional[Tensor], inv: Optional[Tensor], P: int, w2: Tensor, b2: Tensor, sw: Tensor, sb: Tensor,
                    g_y: Optional[Tensor], g_Y: Optional[Tensor]):
    """-> g_h (B, C, HW), g_Hv (B d, C, HW) or None (dense: zeros at unused positions), g_w2 (O, C), g_b2, g_sw, g_sb (O)"""
    lib = _lib.load()
    h, w2, b2, sw, sb = _req(h, "h"), _req(w2, "w2"), _req(b2, "b2"), _req(sw, "sw"), _req(sb, "sb")
    B, C, HW = h.shape
    O = w2.shape[0]
    d = 0
    if Hv is not None:
        Hv = _req(Hv, "Hv")
        d = Hv.shape[0] // max(B, 1)
        g_Y = torch.zeros(B, d, O, P, dtype=torch.float32, device=h.device) if g_Y is None else _req(g_Y, "g_Y")
    inv = None if inv is None else _req(inv, "inv", torch.int32)
    g_y = None if g_y is None else _req(g_y, "g_y")
    g_h = torch.empty_like(h)
    g_Hv = None if Hv is None else torch.empty_like(Hv)
    partial = torch.empty(max(lib.cmf_resnet_tail_partial_floats(B, d, HW, int(P), C, O), 4), dtype=torch.float32, device=h.device)
    gp = torch.empty(O * C + 3 * O, dtype=torch.float32, device=h.device)
    with torch.cuda.device(h.device):
        _lib.check(lib.cmf_resnet_tail_bwd(B, d, C, HW, int(P), O, _ptr(h), _ptr(Hv), _ptr(inv), _ptr(w2), _ptr(b2), _ptr(sw), _ptr(sb),
                                           _ptr(g_y), _ptr(g_Y), _ptr(g_h), _ptr(g_Hv), _ptr(partial), _ptr(gp), _stream()),
                   "cmf_resnet_tail_bwd")
    n = O * C
    return g_h, g_Hv, gp[:n].view(O, C), gp[n:n + O], gp[n + O:n + 2 * O], gp[n + 2 * O:]


# ------------------------------------------------------------------------------------------------
# first convolution of the image couplers' conditioner (C_in -> 64, 3x3, padding 1, no bias)
# ------------------------------------------------------------------------------------------------
def convin_usable(x: Tensor, weight: Tensor) -> bool:
    return (x.is_cuda and x.dtype == torch.float32 and x.dim() == 4 and weight.dim() == 4 and tuple(weight.shape[2:]) == (3, 3)
            and weight.shape[1] == x.shape[1] and x.shape[0] <= 65535
            and bool(_lib.load().cmf_convin_usable(int(weight.shape[1]), int(weight.shape[0]), int(x.shape[2]), int(x.shape[3]))))


def convin_fwd(x: Tensor, weight: Tensor) -> Tensor:
    lib = _lib.load()
    x, weight = _req(x, "x"), _req(weight, "weight")
    N, Cin, H, W = x.shape
    y = torch.empty(N, weight.shape[0], H, W, dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(lib.cmf_convin_fwd(N, Cin, H, W, _ptr(x), _ptr(weight), _ptr(y), _stream()), "cmf_convin_fwd")
    return y


def convin_dgrad(g_y: Tensor, weight: Tensor) -> Tensor:
    lib = _lib.load()
    g_y, weight = _req(g_y, "g_y"), _req(weight, "weight")
    N, _, H, W = g_y.shape
    Cin = weight.shape[1]
    g_x = torch.empty(N, Cin, H, W, dtype=torch.float32, device=g_y.device)
    with torch.cuda.device(g_y.device):
        _lib.check(lib.cmf_convin_dgrad(N, Cin, H, W, _ptr(g_y), _ptr(weight), _ptr(g_x), _stream()), "cmf_convin_dgrad")
    return g_x


def convin_wgrad(x: Tensor, g_y: Tensor) -> Tensor:
    """-> g_w (64, C_in, 3, 3): per-block partials reduced in block order"""
    lib = _lib.load()
    x, g_y = _req(x, "x"), _req(g_y, "g_y")
    N, Cin, H, W = x.shape
    Co = g_y.shape[1]
    nparts = lib.cmf_convin_wgrad_partials(N, H, W)
    partial = torch.empty(max(nparts, 1) * Co * Cin * 9, dtype=torch.float32, device=x.device)
    g_w = torch.empty(Co, Cin, 3, 3, dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(lib.cmf_convin_wgrad(N, Cin, H, W, _ptr(x), _ptr(g_y), _ptr(partial), _ptr(g_w), _stream()), "cmf_convin_wgrad")
    return g_w


def channel_sum(x: Tensor) -> Tensor:
    """x (N, C, ...) -> (C,): sum over the batch and every trailing dimension (the bias gradient of a convolution)"""
    lib = _lib.load()
    x = _req(x, "x")
    N, C = x.shape[0], x.shape[1]
    HW = x[0, 0].numel() if N else 1
    buf = torch.empty(C * (1 + lib.cmf_channel_sum_splits(N, C)), dtype=torch.float32, device=x.device)
    out, partial = buf[:C], buf[C:]
    with torch.cuda.device(x.device):
        _lib.check(lib.cmf_channel_sum(N, C, HW, _ptr(x), _ptr(partial), _ptr(out), _stream()), "cmf_channel_sum")
    return out
