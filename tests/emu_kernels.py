"""Plain-torch stand-ins for ``cmf_b200.kernels`` (TEST INFRASTRUCTURE ONLY).

The product has no CPU path; these emulate each C-ABI launcher's *contract* (same arguments,
shapes and index-map semantics as include/cmf_b200.h) so the host logic -- layer program,
autograd wiring, head, sharding -- can be exercised on the GPU-less build container.  They are
installed by the ``emu`` fixture through monkeypatching and are never importable from cmf_b200.
"""
import math

import torch

PENALTY = {"none": 0, "offdiag": 1, "diag": 2}


def _maps(M, Q, map_, dev):
    return torch.arange(M, device=dev) if map_ is None else map_.long()


def coupling_jvp_fwd(z_m, V_m, st, ST, map_):
    B, M = z_m.shape
    Q = st.shape[1] // 2
    q = _maps(M, Q, map_, z_m.device)
    t, s = st[:, :Q][:, q], st[:, Q:][:, q]
    e = torch.exp(-s)
    x = z_m * e - t
    if V_m is None:
        return x, None
    T, S = ST[:, :, :Q][:, :, q], ST[:, :, Q:][:, :, q]
    return x, e.unsqueeze(1) * (V_m - z_m.unsqueeze(1) * S) - T


def coupling_jvp_bwd(z_m, V_m, st, ST, map_, inv, g_x, g_V):
    B, M = z_m.shape
    Q = st.shape[1] // 2
    q = _maps(M, Q, map_, z_m.device)
    s = st[:, Q:][:, q]
    e = torch.exp(-s)
    gx = torch.zeros_like(z_m) if g_x is None else g_x
    g_st = torch.zeros_like(st)
    if V_m is None:
        A = C = torch.zeros_like(z_m)
        g_Vin = g_ST = None
    else:
        S = ST[:, :, Q:][:, :, q]
        A, C = (S * g_V).sum(1), (V_m * g_V).sum(1)
        g_Vin = e.unsqueeze(1) * g_V
        g_ST = torch.zeros_like(ST)
        g_ST[:, :, q] = -g_V
        g_ST[:, :, Q + q] = -(e * z_m).unsqueeze(1) * g_V
    g_z = e * (gx - A)
    g_st[:, q] = -gx
    g_st[:, Q + q] = -z_m * e * gx - e * (C - z_m * A)
    return g_z, g_Vin, g_st, g_ST


def gather_rows(in0, in1, idx):
    src = in0 if in1 is None else torch.cat((in0, in1), dim=-1)
    idx = idx.long()
    out = src[..., idx.clamp(min=0)]
    return torch.where(idx >= 0, out, torch.zeros((), dtype=out.dtype))


def gram_fwd(V, G=None):
    g = V @ V.transpose(1, 2)
    if G is None:
        return g
    G += g
    return G


def gram_fwd2(V1, V2, G=None):
    return gram_fwd(V2, gram_fwd(V1, G))


def gram_bwd(V, g_G):
    return (g_G + g_G.transpose(1, 2)) @ V


def chol_logdet_fwd(G):
    B, d, _ = G.shape
    L, info_each = torch.linalg.cholesky_ex(G)
    bad = (info_each != 0).nonzero()
    info = torch.zeros(1, dtype=torch.int32)
    if bad.numel():
        info[0] = int(bad.max()) + 1
    logdet = 2 * torch.log(torch.diagonal(L, dim1=-2, dim2=-1)).sum(1)
    diag = torch.diagonal(G, dim1=-2, dim2=-1).abs().sum(1)
    off = G.abs().sum((1, 2)) - diag
    return L, logdet, off, diag, info


def chol_logdet_bwd(L, G, g_logdet, g_l1, penalty):
    d = G.shape[1]
    Ginv = torch.cholesky_inverse(L)
    out = g_logdet.view(-1, 1, 1) * Ginv
    if penalty != "none" and g_l1 is not None:
        eye = torch.eye(d, dtype=torch.bool)
        mask = ~eye if penalty == "offdiag" else eye
        out = out + g_l1.view(-1, 1, 1) * torch.sign(G) * mask
    return out


def conv3x3_c64_usable(N, H, W):
    return N > 0 and (H * W) % 4 == 0 and 128 + ((W + 4) & ~3) + W + 1 <= 256


def conv3x3_c64_prep(weight, transpose):
    """stand-in for the pre-split weights: the (possibly transposed + flipped) fp32 weights"""
    return (weight.flip(2, 3).transpose(0, 1) if transpose else weight).contiguous()


def relu_mask_pack(x, relu_out=None):
    """(N,64,H,W) -> (N, H*W, 2) int32, bit e of word h = x[n, 32 h + e, p] > 0; relu_out (optional) <- max(x, 0)"""
    N, C, H, W = x.shape
    if relu_out is not None:
        torch.clamp_min(x, 0, out=relu_out)
    pos = (x > 0).flatten(2).view(N, 2, 32, H * W).to(torch.int64)                 # (N, word, bit, p)
    words = (pos << torch.arange(32, device=x.device).view(1, 1, 32, 1)).sum(2)      # (N, 2, p), 0 .. 2^32-1
    words = torch.where(words >= 2 ** 31, words - 2 ** 32, words)
    return words.transpose(1, 2).contiguous().to(torch.int32)


def _unpack_mask(bits, H, W):
    N = bits.shape[0]
    w = bits.to(torch.int64) & 0xFFFFFFFF                                            # (N, p, 2)
    b = (w.unsqueeze(-1) >> torch.arange(32, device=bits.device)) & 1                # (N, p, 2, 32)
    return b.permute(0, 2, 3, 1).reshape(N, 64, H, W).bool()


def conv3x3_c64(x, wsplit, bias=None, mask_src=None, mask_div=1, mask_first=False, res=None, second=False, drain=1,
                tail=0, out=None, out2=None):
    N = x.shape[0]
    t = torch.nn.functional.conv2d(x, wsplit, None, padding=1)
    if bias is not None:
        if tail:
            t[tail:] = t[tail:] + bias.view(1, -1, 1, 1)
        else:
            t = t + bias.view(1, -1, 1, 1)
    m = None
    if mask_src is not None:
        keep = _unpack_mask(mask_src, *x.shape[2:]) if mask_src.dtype == torch.int32 else mask_src > 0
        keep = keep.to(t.dtype)
        m = torch.cat((keep.repeat_interleave(mask_div, dim=0), keep)) if tail else keep.repeat_interleave(mask_div, dim=0)
    if mask_first:
        t = t * m
    if res is not None:
        t = t + res
    if out is not None:
        out.copy_(t)
        t = out
    if second:
        t2 = t * m
        if out2 is not None:
            out2.copy_(t2)
            t2 = out2
        return t, t2
    return t


def conv3x3_c64_wgrad_usable(N, H, W):
    return N > 0 and (H * W) % 4 == 0 and 128 + ((W + 4) & ~3) + W + 1 <= 228


def conv3x3_c64_wgrad(x, g_y, deterministic=None):
    w = torch.zeros(64, 64, 3, 3, dtype=x.dtype, device=x.device)
    return torch.ops.aten.convolution_backward(g_y, x, w, None, (1, 1), (1, 1), (1, 1), False, (0, 0), 1,
                                               (False, True, False))[1]


def _rqs_inverse(z, cw, ch, dv, tb):
    """inverse spline from knot tables with linear tails: the restated nsf arithmetic of oracle/rqs_oracle.py"""
    import rqs_oracle as R
    inside = (z >= -tb) & (z <= tb)
    out, _, idx = R.spline_from_tables(torch.where(inside, z, torch.zeros_like(z)), cw, ch, dv, True)
    return torch.where(inside, out, z), torch.where(inside, idx, torch.full_like(idx, -1))


def _rqs_push(z, V, cw, ch, dv, CW, CH, DV, tb, graph):
    f = lambda a, b, c, e: _rqs_inverse(a, b, c, e, tb)[0]
    x, bins = _rqs_inverse(z, cw, ch, dv, tb)
    if V is None:
        return x, None, bins
    cols = [torch.autograd.functional.jvp(f, (z, cw, ch, dv), (V[:, k], CW[:, k], CH[:, k], DV[:, k]), create_graph=graph)[1]
            for k in range(V.shape[1])]
    return x, torch.stack(cols, 1), bins


def rqs_coupling_jvp_fwd(z_m, V_m, cw, ch, dv, CW, CH, DV, tail_bound):
    x, Vo, bins = _rqs_push(z_m, V_m, cw, ch, dv, CW, CH, DV, tail_bound, False)
    return x.detach(), None if Vo is None else Vo.detach(), bins.to(torch.int32)


def rqs_coupling_jvp_bwd(z_m, V_m, cw, ch, dv, CW, CH, DV, tail_bound, g_x, g_V):
    names = [z_m, V_m, cw, ch, dv, CW, CH, DV]
    with torch.enable_grad():
        leaves = [None if t is None else t.detach().clone().requires_grad_(True) for t in names]
        z, V, a, b, c, A, Bq, Cq = leaves
        x, Vo, _ = _rqs_push(z, V, a, b, c, A, Bq, Cq, tail_bound, True)
        loss = 0.0 if g_x is None else (x * g_x).sum()
        if Vo is not None and g_V is not None:
            loss = loss + (Vo * g_V).sum()
        live = [t for t in leaves if t is not None]
        grads = torch.autograd.grad(loss, live, allow_unused=True)
    it = iter(grads)
    out = []
    for t in leaves:
        if t is None:
            out.append(None)
        else:
            g = next(it)
            out.append(torch.zeros_like(t) if g is None else g.detach())
    g_z, g_Vin, g_cw, g_ch, g_dv, g_CW, g_CH, g_DV = out
    return g_z, g_Vin, g_cw, g_ch, g_dv, g_CW, g_CH, g_DV


def preproc_logit_fwd(x, noise, mult, add):
    if noise is not None:
        x.add_(noise)
    u = mult * x + add
    uc = u.clamp(1e-7, 1 - 1e-7)
    n = x[0].numel()
    lj = n * math.log(abs(mult)) + (-torch.log(uc) - torch.log(1 - uc)).flatten(1).sum(1, keepdim=True)
    return torch.log(u) - torch.log(1 - u), lj


def coupling_x_to_z_fwd(x, st, pos2j, qmap):
    B, D = x.shape
    Q = st.shape[1] // 2
    j = pos2j.long()
    mod = j >= 0
    jj = j.clamp(min=0)
    q = jj if qmap is None else qmap.long()[jj]
    t, s = st[:, :Q][:, q], st[:, Q:][:, q]
    z = torch.where(mod, (x + t) * torch.exp(s), x)
    return z, torch.where(mod, s, torch.zeros_like(s)).sum(1, keepdim=True)


def coupling_x_to_z_bwd(z, st, pos2j, qmap, q2pos, g_z, g_lj):
    B, D = z.shape
    Q = st.shape[1] // 2
    j = pos2j.long()
    mod = j >= 0
    jj = j.clamp(min=0)
    q = jj if qmap is None else qmap.long()[jj]
    gz = torch.zeros_like(z) if g_z is None else g_z
    e = torch.exp(st[:, Q:][:, q])
    g_x = torch.where(mod, gz * e, gz)
    i = q2pos.long()
    used = i >= 0
    ii = i.clamp(min=0)
    glj = torch.zeros(B, 1, dtype=z.dtype) if g_lj is None else g_lj.view(B, 1)
    gt = torch.where(used, gz[:, ii] * torch.exp(st[:, Q:]), torch.zeros_like(st[:, Q:]))
    gs = torch.where(used, gz[:, ii] * z[:, ii] + glj, torch.zeros_like(st[:, Q:]))
    return g_x, torch.cat((gt, gs), dim=1)


def sumsq_diff_fwd(a, c):
    d = a if c is None else a - c
    return (d * d).sum(1, keepdim=True)


def sumsq_diff_bwd(a, c, g):
    d = a if c is None else a - c
    return 2 * g.view(-1, 1) * d


def flat_stack_usable(D, d, L, nh, max_width, P):
    return 1 <= D <= 8 and 1 <= d <= min(8, D) and 1 <= L <= 16 and 1 <= nh <= 4 and max_width <= 16 and P <= 4096


def _flat_stack_eval(desc, z, params):
    """the whole-stack kernel's contract in batched torch ops (descriptor layout: include/cmf_b200.h)"""
    it = iter(int(v) for v in desc)
    D, d, L, nh, P = (next(it) for _ in range(5))
    pos = [next(it) for _ in range(d)]
    B = z.shape[0]
    x = z.new_zeros(B, D)
    x = x.index_copy(1, torch.tensor(pos), z)
    V = z.new_zeros(B, d, D)
    V[:, torch.arange(d), torch.tensor(pos)] = 1.0
    for _ in range(L):
        n_keep, n_mod = next(it), next(it)
        keep = torch.tensor([next(it) for _ in range(n_keep)])
        mod = torch.tensor([next(it) for _ in range(n_mod)])
        widths = [next(it) for _ in range(nh + 2)]
        w_off = [next(it) for _ in range(nh + 1)]
        b_off = [next(it) for _ in range(nh + 1)]
        h, H = x[:, keep], V[:, :, keep]
        for j in range(nh + 1):
            W = params[w_off[j]: w_off[j] + widths[j] * widths[j + 1]].view(widths[j + 1], widths[j])
            bias = params[b_off[j]: b_off[j] + widths[j + 1]]
            h, H = h @ W.t() + bias, H @ W.t()
            if j < nh:
                h = torch.tanh(h)
                H = H * (1 - h * h).unsqueeze(1)
        t, s, T, S = h[:, :n_mod], h[:, n_mod:], H[:, :, :n_mod], H[:, :, n_mod:]
        e = torch.exp(-s)
        zq, Vq = x[:, mod], V[:, :, mod]
        x = x.index_copy(1, mod, zq * e - t)
        V = V.index_copy(2, mod, e.unsqueeze(1) * (Vq - zq.unsqueeze(1) * S) - T)
    return V @ V.transpose(1, 2), x


def flat_stack_fwd(desc, z, params):
    G, x = _flat_stack_eval(desc, z, params)
    return G.detach(), x.detach()


def flat_stack_bwd(desc, z, params, g_G, g_x):
    with torch.enable_grad():
        zz, pp = z.detach().clone().requires_grad_(True), params.detach().clone().requires_grad_(True)
        G, x = _flat_stack_eval(desc, zz, pp)
        loss = (G * (torch.zeros_like(G) if g_G is None else g_G)).sum() + (x * (torch.zeros_like(x) if g_x is None else g_x)).sum()
        gz, gp = torch.autograd.grad(loss, (zz, pp), allow_unused=True)
    return (torch.zeros_like(z) if gz is None else gz), (torch.zeros_like(params) if gp is None else gp)


def channel_sum(x):
    return x.sum(dim=[0] + list(range(2, x.dim())))


def convin_usable(x, weight):
    return (x.dtype == torch.float32 and x.dim() == 4 and weight.dim() == 4 and tuple(weight.shape[2:]) == (3, 3)
            and weight.shape[0] == 64 and weight.shape[1] == x.shape[1] <= 16 and x.shape[3] <= 64)


def convin_fwd(x, weight):
    return torch.nn.functional.conv2d(x, weight, None, padding=1)


def convin_dgrad(g_y, weight):
    return torch.nn.grad.conv2d_input((g_y.shape[0], weight.shape[1], g_y.shape[2], g_y.shape[3]), weight, g_y, padding=1)


def convin_wgrad(x, g_y):
    return torch.nn.grad.conv2d_weight(x, (g_y.shape[1], x.shape[1], 3, 3), g_y, padding=1)


def resnet_tail_usable(x, C, num_out):
    return x.dtype == torch.float32 and 1 <= C <= 64 and 1 <= num_out <= 32


def _tail_eval(h, Hv, inv, P, w2, b2, sw, sb):
    """`networks.py:96-113,156-161`, `jvp_layers.py:38-64` at the selected positions"""
    B, C, HW = h.shape
    if inv is not None:
        sel = torch.nonzero(inv >= 0).flatten()
        sel = sel[torch.argsort(inv[sel])]
        assert sel.numel() == P
        h = h.index_select(2, sel)
        Hv = None if Hv is None else Hv.index_select(2, sel)
    th = torch.tanh(torch.matmul(w2, torch.relu(h)) + b2.view(1, -1, 1))
    y = sw.view(1, -1, 1) * th + sb.view(1, -1, 1)
    Y = None
    if Hv is not None:
        d = Hv.shape[0] // B
        Hm = Hv.view(B, d, C, -1) * (h > 0).to(h.dtype).unsqueeze(1)
        Y = torch.matmul(w2, Hm) * ((1 - th * th) * sw.view(1, -1, 1)).unsqueeze(1)
    return y, Y


def resnet_tail_fwd(h, Hv, inv, P, w2, b2, sw, sb):
    y, Y = _tail_eval(h, Hv, inv, P, w2, b2, sw, sb)
    return y.detach(), (None if Y is None else Y.detach())


def resnet_tail_bwd(h, Hv, inv, P, w2, b2, sw, sb, g_y, g_Y):
    with torch.enable_grad():
        ins = [t.detach().clone().requires_grad_(True) for t in (h, w2, b2, sw, sb)] + ([Hv.detach().clone().requires_grad_(True)] if Hv is not None else [])
        y, Y = _tail_eval(ins[0], ins[5] if Hv is not None else None, inv, P, *ins[1:5])
        loss = (0 if g_y is None else (y * g_y).sum()) + (0 if (Y is None or g_Y is None) else (Y * g_Y).sum())
        if not torch.is_tensor(loss):
            gr = [None] * len(ins)
        else:
            gr = list(torch.autograd.grad(loss, ins, allow_unused=True))
    gr = [torch.zeros_like(t) if g is None else g for g, t in zip(gr, ins)]
    return gr[0], (gr[5] if Hv is not None else None), gr[1], gr[2], gr[3], gr[4]


def mlp_usable(widths, d, x):
    return x.dim() == 2 and x.dtype == torch.float32 and len(widths) - 1 <= 8 and max(widths) <= 128 and 0 <= d < 128


def mlp_pack(weights, biases):
    widths = [weights[0].shape[1]] + [w.shape[0] for w in weights]
    return torch.cat([w.reshape(-1) for w in weights] + [b.reshape(-1) for b in biases]), widths


def _mlp_unpack(wpack, widths):
    ws, bs, o = [], [], 0
    for l in range(len(widths) - 1):
        n = widths[l] * widths[l + 1]
        ws.append(wpack[o:o + n].view(widths[l + 1], widths[l]))
        o += n
    for l in range(len(widths) - 1):
        bs.append(wpack[o:o + widths[l + 1]])
        o += widths[l + 1]
    return ws, bs


def _mlp_eval(x, V, ws, bs):
    """`networks.py:24-32,206-224`, `jvp_layers.py:38-47`: Linear (bias-free on the tangents), tanh with its JVP mask"""
    for l, (W, b) in enumerate(zip(ws, bs)):
        x = x @ W.t() + b
        V = None if V is None else V @ W.t()
        if l + 1 < len(ws):
            x = torch.tanh(x)
            V = None if V is None else V * (1 - x * x).unsqueeze(1)
    return x, V


def mlp_jvp_fwd(x, V, wpack, widths, save):
    st, ST = _mlp_eval(x, V, *_mlp_unpack(wpack, widths))
    saved = (x.new_zeros(4), x.new_zeros(4)) if save else None
    return st.detach(), (None if ST is None else ST.detach()), saved


def mlp_jvp_bwd(x, V, wpack, widths, saved, g_st, g_ST):
    with torch.enable_grad():
        xx = x.detach().clone().requires_grad_(True)
        VV = None if V is None else V.detach().clone().requires_grad_(True)
        ws, bs = _mlp_unpack(wpack.detach(), widths)
        ws = [w.clone().requires_grad_(True) for w in ws]
        bs = [b.clone().requires_grad_(True) for b in bs]
        st, ST = _mlp_eval(xx, VV, ws, bs)
        loss = (st * g_st).sum() + (0 if (ST is None or g_ST is None) else (ST * g_ST).sum())
        ins = [xx] + ([VV] if VV is not None else []) + ws + bs
        gr = list(torch.autograd.grad(loss, ins, allow_unused=True))
        gr = [torch.zeros_like(t) if g is None else g for g, t in zip(gr, ins)]
    g_x = gr.pop(0)
    g_V = gr.pop(0) if VV is not None else None
    n = len(ws)
    return g_x, g_V, gr[:n], gr[n:]


ALL = ["channel_sum", "convin_usable", "convin_fwd", "convin_dgrad", "convin_wgrad", "resnet_tail_usable", "resnet_tail_fwd", "resnet_tail_bwd", "mlp_usable", "mlp_pack", "mlp_jvp_fwd", "mlp_jvp_bwd", "flat_stack_usable", "flat_stack_fwd", "flat_stack_bwd", "preproc_logit_fwd", "coupling_x_to_z_fwd", "coupling_x_to_z_bwd", "sumsq_diff_fwd", "sumsq_diff_bwd",
       "rqs_coupling_jvp_fwd", "rqs_coupling_jvp_bwd", "coupling_jvp_fwd", "coupling_jvp_bwd", "gather_rows", "gram_fwd", "gram_fwd2", "gram_bwd",
       "chol_logdet_fwd", "chol_logdet_bwd", "conv3x3_c64_usable", "conv3x3_c64_prep", "conv3x3_c64", "relu_mask_pack",
       "conv3x3_c64_wgrad_usable", "conv3x3_c64_wgrad"]
