"""Each C-ABI kernel against a plain-torch statement of its contract (tests/emu_kernels.py) on the
same seeded inputs, computed in fp64 on the CPU.  fp32 kernels: rtol 1e-5 unless noted."""
import pytest
import torch

import emu_kernels as E

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _rand(*shape, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed + sum(shape))
    return (torch.rand(*shape, generator=g, dtype=torch.float64) * 2 - 1) * scale


def _close(a, b, rtol=1e-5, atol=1e-6, what=""):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    err = float((a - b).abs().max())
    scale = float(b.abs().max())
    assert err <= rtol * scale + atol, f"{what}: err {err:.3e} scale {scale:.3e}"


def _checker_maps(M, Q, seed=0):
    g = torch.Generator().manual_seed(seed)
    q = torch.randperm(Q, generator=g)[:M].sort().values
    inv = torch.full((Q,), -1, dtype=torch.int32)
    inv[q] = torch.arange(M, dtype=torch.int32)
    return q.int(), inv


COUPLING_SHAPES = [  # B, d, M, Q (Q == M -> identity map; else mapped)
    (3, 5, 8, 8), (50, 64, 392, 392), (7, 64, 196, 392), (2, 3, 1, 1), (5, 0, 12, 12),
    (4, 6, 3, 3), (9, 10, 11, 21), (2, 130, 64, 64), (1, 1, 4, 4), (6, 2, 2, 5),
    # 128-bit path with the tangent chunks on the warps of a CTA (8 and 4 chunks, some of them empty; a last block with
    # idle lanes), and its workspace + finalize form (tiny grid, d > 256)
    (130, 24, 100, 100), (40, 12, 36, 36), (1, 260, 8, 8),
]


@pytest.mark.parametrize("B,d,M,Q", COUPLING_SHAPES)
def test_coupling_fwd_bwd(B, d, M, Q):
    import cmf_b200.kernels as K
    z, st = _rand(B, M, seed=1), _rand(B, 2 * Q, seed=2, scale=0.7)
    V = _rand(B, d, M, seed=3) if d else None
    ST = _rand(B, d, 2 * Q, seed=4) if d else None
    map_, inv = (None, None) if M == Q else _checker_maps(M, Q)
    x_ref, V_ref = E.coupling_jvp_fwd(z, V, st, ST, map_)
    f = lambda t: None if t is None else t.float().to(DEV)
    x, Vo = K.coupling_jvp_fwd(f(z), f(V), f(st), f(ST), f(map_) if map_ is None else map_.to(DEV))
    _close(x, x_ref, what="x_m")
    if d:
        _close(Vo, V_ref, what="V_out")
    gx, gV = _rand(B, M, seed=5), (_rand(B, d, M, seed=6) if d else None)
    ref = E.coupling_jvp_bwd(z, V, st, ST, map_, inv, gx, gV)
    got = K.coupling_jvp_bwd(f(z), f(V), f(st), f(ST), None if map_ is None else map_.to(DEV),
                             None if inv is None else inv.to(DEV), f(gx), f(gV))
    for name, a, b in zip(("g_z", "g_Vin", "g_st", "g_ST"), got, ref):
        if b is not None:
            _close(a, b, rtol=2e-5, atol=2e-6, what=name)
    # g_x = None is treated as zero
    ref0 = E.coupling_jvp_bwd(z, V, st, ST, map_, inv, None, gV)
    got0 = K.coupling_jvp_bwd(f(z), f(V), f(st), f(ST), None if map_ is None else map_.to(DEV),
                              None if inv is None else inv.to(DEV), None, f(gV))
    _close(got0[2], ref0[2], rtol=2e-5, atol=2e-6, what="g_st (no g_x)")


def test_coupling_unaligned_falls_back_to_scalar_path():
    import cmf_b200.kernels as K
    B, d, M = 4, 8, 16
    z, st, V, ST = _rand(B, M), _rand(B, 2 * M, scale=0.5), _rand(B, d, M), _rand(B, d, 2 * M)
    x_ref, V_ref = E.coupling_jvp_fwd(z, V, st, ST, None)
    buf = torch.empty(B * M + 1, device=DEV)
    z_dev = buf[1:].view(B, M)          # 4-byte aligned only
    z_dev.copy_(z.float())
    x, Vo = K.coupling_jvp_fwd(z_dev, V.float().to(DEV), st.float().to(DEV), ST.float().to(DEV), None)
    _close(x, x_ref); _close(Vo, V_ref)


@pytest.mark.parametrize("rows,n0,n1,n_out", [((3,), 5, 0, 9), ((4, 6), 17, 13, 40), ((50, 64), 392, 0, 784), ((2,), 1, 1, 1)])
def test_gather_rows(rows, n0, n1, n_out):
    import cmf_b200.kernels as K
    g = torch.Generator().manual_seed(n_out)
    a = _rand(*rows, n0)
    b = _rand(*rows, n1, seed=9) if n1 else None
    idx = torch.randint(-1, n0 + n1, (n_out,), generator=g).int()
    ref = E.gather_rows(a, b, idx)
    got = K.gather_rows(a.float().to(DEV), None if b is None else b.float().to(DEV), idx.to(DEV))
    assert torch.equal(got.cpu(), ref.float())      # pure data movement: bit exact


@pytest.mark.parametrize("B,d,n", [(3, 3, 3), (5, 6, 6), (4, 10, 21), (6, 64, 784), (2, 64, 392), (2, 128, 300),
                                   (1, 200, 77), (3, 16, 1), (2, 65, 130)])
def test_gram_fwd_bwd(B, d, n):
    import cmf_b200.kernels as K
    V = _rand(B, d, n)
    Vd = V.float().to(DEV)
    G = K.gram_fwd(Vd)
    _close(G, E.gram_fwd(V), rtol=2e-6 * max(1, n ** 0.5), what="G")
    assert torch.equal(G, G.transpose(1, 2)), "mirror tiles must be bitwise symmetric"
    V2 = _rand(B, d, max(1, n // 2), seed=5)
    G2 = K.gram_fwd(V2.float().to(DEV), G.clone())
    _close(G2, E.gram_fwd(V) + E.gram_fwd(V2), rtol=1e-5, what="G accumulate")
    gG = _rand(B, d, d, seed=7)
    _close(K.gram_bwd(Vd, gG.float().to(DEV)), E.gram_bwd(V, gG), rtol=1e-5, what="g_V")


@pytest.mark.parametrize("B,d,n", [(3, 64, 392), (600, 64, 8), (2, 64, 132), (5, 33, 256), (1, 17, 4), (7, 50, 520), (2, 6, 12)])
def test_gram_bwd_small_d_kernel(B, d, n):
    """d <= 64, n % 4 == 0: the staged-S kernel (ragged last 128-column chunk, chunk groups split over
    CTAs when B is small, zero-padded S) against the fp64 contract"""
    import cmf_b200.kernels as K
    V, gG = _rand(B, d, n, seed=11), _rand(B, d, d, seed=12)
    got = K.gram_bwd(V.float().to(DEV), gG.float().to(DEV), path="fma")
    _close(got, E.gram_bwd(V, gG), rtol=1e-5, what="g_V small-d")


TC_SHAPES = [(5, 64, 392), (50, 64, 784), (3, 128, 96), (4, 20, 40), (2, 256, 128), (3, 200, 36),
             (7, 16, 32), (1, 512, 64), (9, 48, 100), (2, 130, 260)]


@pytest.mark.parametrize("B,d,n", TC_SHAPES)
def test_gram_tc_matches_fp64_and_fma(B, d, n):
    """tcgen05 3xTF32 Gram (cmf_gram_fwd_path CMF_GRAM_TC) is FP32-accurate: same tolerance as the
    FMA kernel against the fp64 contract, bitwise symmetric, accumulate mode, ragged tails."""
    import cmf_b200.kernels as K
    V = _rand(B, d, n)
    Vd = V.float().to(DEV)
    ref = E.gram_fwd(V)
    G = K.gram_fwd(Vd, path="tc")
    _close(G, ref, rtol=2e-6 * max(1, n ** 0.5), what="G tc")
    assert torch.equal(G, G.transpose(1, 2)), "G must be bitwise symmetric"
    G_fma = K.gram_fwd(Vd, path="fma")
    _close(G, G_fma, rtol=2e-6 * max(1, n ** 0.5), what="tc vs fma")
    V2 = _rand(B, d, n, seed=5)
    G2 = K.gram_fwd(V2.float().to(DEV), G.clone(), path="tc")
    _close(G2, ref + E.gram_fwd(V2), rtol=1e-5, what="G tc accumulate")
    gG = _rand(B, d, d, seed=7)
    g_ref = E.gram_bwd(V, gG)
    if d % 4:          # the adjoint stages g_G by TMA: row stride must be a multiple of 16 bytes
        from cmf_b200 import _lib
        with pytest.raises(_lib.CmfB200Error):
            K.gram_bwd(Vd, gG.float().to(DEV), path="tc")
        _close(K.gram_bwd(Vd, gG.float().to(DEV)), g_ref, rtol=1e-5, what="g_V auto")
        return
    g_tc = K.gram_bwd(Vd, gG.float().to(DEV), path="tc")
    _close(g_tc, g_ref, rtol=1e-5, what="g_V tc")
    _close(g_tc, K.gram_bwd(Vd, gG.float().to(DEV), path="fma"), rtol=1e-5, what="g_V tc vs fma")


@pytest.mark.parametrize("B,d,n", [(700, 64, 392), (1000, 32, 200), (500, 128, 260), (641, 48, 100), (450, 96, 36), (1201, 16, 72)])
def test_gram_fwd_panel_kernel_many_panels_per_cta(B, d, n):
    """The persistent tcgen05 forward (d <= 128) with several panels per CTA: the stage ring, the TMEM staging slots, the
    ping-pong accumulator (odd and even numbers of drain groups per panel: n = 260 -> 3, n = 36 -> 1) and the cross
    accumulator hand over from panel to panel; ragged last panel (B not a multiple of the samples per panel), panels
    of fewer than 128 rows (d = 48, 96).  Accumulate mode and the two-block launch on the same shapes."""
    import cmf_b200.kernels as K
    V = _rand(B, d, n, seed=31)
    Vd = V.float().to(DEV)
    ref = E.gram_fwd(V)
    G = K.gram_fwd(Vd, path="tc")
    _close(G, ref, rtol=2e-6 * max(1, n ** 0.5), what="G panel")
    assert torch.equal(G, G.transpose(1, 2)), "G must be bitwise symmetric"
    assert torch.equal(G, K.gram_fwd(Vd, path="tc")), "not reproducible"
    G2 = K.gram_fwd2(Vd, Vd, G.clone())
    _close(G2, 3 * ref, rtol=1e-5, what="G two blocks, accumulate")


@pytest.mark.parametrize("B,d,n", [(300, 64, 392), (200, 96, 392), (260, 128, 260), (400, 32, 200), (180, 72, 520), (700, 16, 136)])
def test_gram_bwd_persistent_kernel_many_items_per_cta(B, d, n):
    """The persistent tcgen05 adjoint (d <= 128) with several (sample, n tile) items per CTA: every ring (V stages, the
    per-sample S buffers, the TMEM staging slots of the two issuers, the accumulator pair) wraps many times; odd and even
    numbers of K stages per item (d = 96, 72: three), ragged last tiles, d not a multiple of 32.  Twice: the same bits."""
    import cmf_b200.kernels as K
    V, gG = _rand(B, d, n, seed=21), _rand(B, d, d, seed=22)
    Vd, gd = V.float().to(DEV), gG.float().to(DEV)
    got = K.gram_bwd(Vd, gd, path="tc")
    _close(got, E.gram_bwd(V, gG), rtol=1e-5, what="g_V persistent")
    assert torch.equal(got, K.gram_bwd(Vd, gd, path="tc")), "not reproducible"


@pytest.mark.parametrize("B,d,n1,n2", [(5, 64, 392, 392), (3, 64, 392, 40), (2, 200, 36, 260), (9, 48, 100, 32),
                                       (3, 64, 392, 8), (2, 6, 12, 20), (4, 130, 64, 64)])
def test_gram_two_blocks_in_one_launch(B, d, n1, n2):
    """cmf_gram_fwd2: G = V1 V1^T + V2 V2^T (non_square.py:307-308 over two layer blocks) -- one tcgen05
    launch over the n1 + n2 columns when both qualify (ragged last K chunk of either block), the
    two-call fallback otherwise; also in accumulate mode"""
    import cmf_b200.kernels as K
    V1, V2 = _rand(B, d, n1, seed=21), _rand(B, d, n2, seed=22)
    ref = E.gram_fwd(V1) + E.gram_fwd(V2)
    G = K.gram_fwd2(V1.float().to(DEV), V2.float().to(DEV))
    _close(G, ref, rtol=2e-6 * max(1, (n1 + n2) ** 0.5), what="G two blocks")
    assert torch.equal(G, G.transpose(1, 2)), "G must be bitwise symmetric"
    G0 = _rand(B, d, d, seed=23)
    G0 = G0 + G0.transpose(1, 2)
    Ga = K.gram_fwd2(V1.float().to(DEV), V2.float().to(DEV), G0.float().to(DEV))
    _close(Ga, ref + G0, rtol=1e-5, what="G two blocks, accumulate")


def test_gram_tc_badly_scaled_columns():
    """Trained CMF models drive most G_kk toward zero (SURVEY 7 hard parts): tangent rows scaled by
    10^-k must keep fp32-level relative accuracy element-wise against |v_i||v_j|."""
    import cmf_b200.kernels as K
    B, d, n = 4, 64, 392
    V = _rand(B, d, n) * (10.0 ** (-torch.arange(d, dtype=torch.float64) / 8.0)).view(1, d, 1)
    G = K.gram_fwd(V.float().to(DEV), path="tc").double().cpu()
    ref = E.gram_fwd(V)
    nrm = V.norm(dim=2)
    bound = nrm.unsqueeze(2) * nrm.unsqueeze(1)
    assert float(((G - ref).abs() / bound).max()) < 2e-6


def test_gram_tc_rejects_unaligned_shapes():
    import cmf_b200.kernels as K
    from cmf_b200 import _lib
    with pytest.raises(_lib.CmfB200Error):
        K.gram_fwd(torch.zeros(2, 6, 6, device=DEV), path="tc")
    with pytest.raises(_lib.CmfB200Error):
        K.gram_fwd(torch.zeros(2, 64, 30, device=DEV), path="tc")


def _spd(B, d, seed=0, cond=1e3):
    A = _rand(B, d, max(d, 4) * 2, seed=seed)
    G = A @ A.transpose(1, 2) / A.shape[2]
    scale = torch.logspace(0, -torch.log10(torch.tensor(cond)).item() / 2, d, dtype=torch.float64)
    return G * scale.view(1, d, 1) * scale.view(1, 1, d) + 1e-6 * torch.eye(d, dtype=torch.float64)


@pytest.mark.parametrize("B,d", [(1, 1), (7, 2), (33, 3), (100, 6), (17, 10), (5, 20), (4, 31), (6, 32), (5, 33), (3, 50),
                                 (9, 64), (3, 100), (2, 127), (2, 128),
                                 (3, 129), (2, 200), (2, 256), (2, 512), (1, 1000)])
def test_chol_logdet_fwd_bwd(B, d):
    import cmf_b200.kernels as K
    G = _spd(B, d)
    L_ref, ld_ref, off_ref, dia_ref, _ = E.chol_logdet_fwd(G)
    Gd = G.float().to(DEV)
    L, ld, off, dia, info = K.chol_logdet_fwd(Gd)
    assert int(info.item()) == 0
    _close(L, L_ref, rtol=2e-5, what="L")
    _close(ld, ld_ref, rtol=1e-5, atol=1e-4, what="logdet")
    _close(off, off_ref, what="l1_off"); _close(dia, dia_ref, what="l1_diag")
    assert float(L.triu(1).abs().max()) == 0.0
    gl, g1 = _rand(B, seed=3), _rand(B, seed=4)
    for pen in ("none", "offdiag", "diag"):
        ref = E.chol_logdet_bwd(L_ref, G, gl, g1, pen)
        got = K.chol_logdet_bwd(L, Gd, gl.float().to(DEV), g1.float().to(DEV), pen)
        # G^{-1} amplifies rounding by cond(G) ~ 1e3
        _close(got, ref, rtol=5e-4, what=f"g_G[{pen}]")


def test_chol_reports_non_positive_definite():
    import cmf_b200.kernels as K
    G = _spd(6, 8)
    G[4] = -G[4]
    G[1, 3, 3] = float("nan")
    *_, info = K.chol_logdet_fwd(G.float().to(DEV))
    assert int(info.item()) == 5          # 1 + largest failing sample index
    *_, info = K.chol_logdet_fwd(_spd(6, 8).float().to(DEV))
    assert int(info.item()) == 0
    # one-sample-per-CTA kernel (32 <= d <= 128): negative pivot late in sample 3, NaN in sample 1
    G = _spd(5, 64)
    G[3, 60, 60] = -1.0
    G[1, 20, 5] = G[1, 5, 20] = float("nan")
    _, ld, *_, info = K.chol_logdet_fwd(G.float().to(DEV))
    assert int(info.item()) == 4
    assert torch.isnan(ld[3]) and torch.isnan(ld[1]) and torch.isfinite(ld[[0, 2, 4]]).all()
    # panel-blocked kernel (d > 128): failure in a later panel of sample 2, NaN in sample 0
    G = _spd(4, 160)
    G[2, 150, 150] = -1.0
    G[0, 40, 7] = G[0, 7, 40] = float("nan")
    _, ld, *_, info = K.chol_logdet_fwd(G.float().to(DEV))
    assert int(info.item()) == 3
    assert torch.isnan(ld[2]) and torch.isnan(ld[0]) and torch.isfinite(ld[1]) and torch.isfinite(ld[3])


def test_chol_d_above_limit_is_rejected():
    import cmf_b200.kernels as K
    from cmf_b200._lib import CmfB200Error
    with pytest.raises(CmfB200Error, match="1 <= d"):
        K.chol_logdet_fwd(torch.eye(1025, device=DEV).repeat(2, 1, 1))


def test_ops_autograd_matches_torch_autograd():
    """The autograd.Function wrappers against torch autograd over the emulation (fp64)."""
    from cmf_b200 import ops
    B, d, M = 5, 6, 12
    leaves64 = [_rand(B, M, seed=1), _rand(B, d, M, seed=2), _rand(B, 2 * M, seed=3, scale=0.5), _rand(B, d, 2 * M, seed=4)]
    ref_in = [t.clone().requires_grad_(True) for t in leaves64]
    x_r, V_r = E.coupling_jvp_fwd(*ref_in, None)
    G_r = V_r @ V_r.transpose(1, 2) + 0.5 * torch.eye(d, dtype=torch.float64)
    ld_r = torch.logdet(G_r)
    off = ~torch.eye(d, dtype=torch.bool)
    loss_r = (ld_r * 0.7).sum() + (G_r.abs() * off).sum() * 0.3 + (x_r ** 2).sum()
    loss_r.backward()
    dev_in = [t.float().to(DEV).requires_grad_(True) for t in leaves64]
    x, V = ops.coupling_jvp(*dev_in)
    G = ops.gram([V[:, :, :5].contiguous(), V[:, :, 5:].contiguous()]) + 0.5 * torch.eye(d, device=DEV)
    ld, l1 = ops.chol_logdet_l1(G, "offdiag")
    loss = (ld * 0.7).sum() + l1.sum() * 0.3 + (x ** 2).sum()
    loss.backward()
    _close(loss, loss_r, rtol=1e-5)
    for a, b in zip(dev_in, ref_in):
        _close(a.grad, b.grad, rtol=2e-4, atol=1e-5)


@pytest.mark.parametrize("name,batch", [("sphere_gij_d3", 1000), ("sphere_gkk_d2", 77), ("sinusoid_gij_d6", 1000), ("sinusoid_plain_d3", 5)])
def test_flat_stack_kernel_matches_the_layer_program(name, batch):
    """Whole-stack kernel against the per-layer program it replaces (coupling kernel +
    torch MLP JVPs + Gram kernel), on the 2-D config models: G, primal output, and the adjoint w.r.t. the
    latent and EVERY parameter for random cotangents."""
    import cmf_b200.head as H
    from conftest import load_golden
    from cases import CASES
    from detfill import fill_state_dict
    from helpers import build_product, strip_wrappers, find_head
    g = load_golden(name)
    density = build_product(g, "cpu")
    density.load_state_dict(fill_state_dict(density.state_dict(), salt=2, gain=1.5))
    density = density.to(DEV).train()
    head, _ = find_head(strip_wrappers(density))
    assert head.program.flat_plan is not None, "the 2-D configs must qualify for the whole-stack kernel"
    d = head.program.d
    gen = torch.Generator().manual_seed(batch)
    z0 = torch.randn(batch, d, generator=gen).to(DEV)
    gG = torch.randn(batch, d, d, generator=gen).to(DEV)
    gX = torch.randn(batch, *head.x_shape, generator=gen).to(DEV)
    results = []
    for flat in (True, False):
        H._FLAT_STACK = flat
        try:
            density.zero_grad()
            z = z0.clone().requires_grad_(True)
            gram, recon = head._get_full_jac_transpose_jac(z)
            ((gram * gG).sum() + (recon * gX).sum()).backward()
            results.append((gram.detach(), recon.detach(), z.grad.clone(),
                            {n: p.grad.clone() for n, p in head.named_parameters() if p.grad is not None}))
        finally:
            H._FLAT_STACK = True
    (G1, x1, gz1, gp1), (G0, x0, gz0, gp0) = results
    _close(G1, G0, rtol=1e-5, what="G")
    _close(x1, x0, rtol=1e-5, what="x")
    _close(gz1, gz0, rtol=1e-4, atol=1e-5, what="g_z")
    assert set(gp1) == set(gp0) and gp1
    for n in gp0:
        _close(gp1[n], gp0[n], rtol=1e-4, atol=1e-5 * float(gp0[n].abs().max()) + 1e-7, what=n)


def test_flat_stack_rejects_what_it_cannot_hold():
    import numpy as np
    import cmf_b200.kernels as K
    assert not K.flat_stack_usable(21, 10, 10, 4, 128, 500000)         # tabular: widths of 128 stay on the layer program
    assert K.flat_stack_usable(3, 3, 5, 2, 10, 840)
    from cmf_b200._lib import CmfB200Error
    bad = np.asarray([3, 3, 5, 2, 840, 0, 1], dtype=np.int32)             # truncated descriptor
    with pytest.raises(CmfB200Error):
        K.flat_stack_fwd(bad, torch.zeros(2, 3, device=DEV), torch.zeros(840, device=DEV))


# ------------------------------------------------------------------------------------------------
# whole tanh-MLP conditioner (cmf_mlp_pack / cmf_mlp_jvp_fwd / cmf_mlp_jvp_bwd)
# ------------------------------------------------------------------------------------------------
MLP_SHAPES = [  # B, d, widths
    (50, 6, [3, 128, 128, 128, 128, 6]),        # power-shape coupler, S = 18 samples per tile, ragged last tile
    (37, 10, [11, 128, 128, 128, 128, 20]),     # hepmass-shape coupler, S = 11
    (300, 3, [2, 10, 10, 2]),                   # 2-D coupler widths (small-tile paths of the weight gradient)
    (5, 0, [4, 32, 32, 8]),                     # primal only (encode pass / prior flow)
    (260, 0, [21, 128, 128, 42]),               # primal only, three tiles
    (9, 2, [5, 7]),                             # a single Linear
    (23, 5, [40, 100, 70, 90]),                 # every width between the 32 / 64 thresholds and 128
    (3, 127, [6, 33, 64, 65, 12]),              # one sample per tile
    (1, 1, [1, 1, 1]),
]


def _mlp_case(B, d, widths, seed=0):
    g = torch.Generator().manual_seed(seed + B + d + sum(widths))
    r = lambda *s: torch.rand(*s, generator=g, dtype=torch.float64) * 2 - 1
    ws = [r(widths[l + 1], widths[l]) * (1.5 / widths[l] ** 0.5) for l in range(len(widths) - 1)]
    bs = [r(widths[l + 1]) * 0.5 for l in range(len(widths) - 1)]
    x, V = r(B, widths[0]), (r(B, d, widths[0]) if d else None)
    g_st, g_ST = r(B, widths[-1]), (r(B, d, widths[-1]) if d else None)
    return ws, bs, x, V, g_st, g_ST


@pytest.mark.parametrize("B,d,widths", MLP_SHAPES)
def test_mlp_jvp_fwd_bwd(B, d, widths):
    """One-launch MLP conditioner against the Linear / tanh / JVP-mask stack in fp64
    (`networks.py:24-32,206-224`, `jvp_layers.py:38-47`): outputs, and the adjoint w.r.t. inputs,
    tangents, every weight and bias."""
    import cmf_b200.kernels as K
    ws, bs, x, V, g_st, g_ST = _mlp_case(B, d, widths)
    f = lambda t: None if t is None else t.float().to(DEV)
    assert K.mlp_usable(widths, d, f(x))
    wpack, w2 = K.mlp_pack([f(w) for w in ws], [f(b) for b in bs])
    assert w2 == widths
    st, ST, saved = K.mlp_jvp_fwd(f(x), f(V), wpack, widths, save=True)
    st0, ST0, _ = K.mlp_jvp_fwd(f(x), f(V), wpack, widths, save=False)          # inference launch: same values
    assert torch.equal(st, st0) and (ST is None or torch.equal(ST, ST0))
    epack, _ = E.mlp_pack(ws, bs)
    rst, rST, _ = E.mlp_jvp_fwd(x, V, epack, widths, False)
    _close(st, rst, what="st")
    if d:
        _close(ST, rST, what="ST")
    g_x, g_V, g_W, g_b = K.mlp_jvp_bwd(f(x), f(V), wpack, widths, saved, f(g_st), f(g_ST))
    rg_x, rg_V, rg_W, rg_b = E.mlp_jvp_bwd(x, V, epack, widths, None, g_st, g_ST)
    _close(g_x, rg_x, rtol=2e-5, what="g_x")
    if d:
        _close(g_V, rg_V, rtol=2e-5, what="g_V")
    for l in range(len(ws)):
        assert g_W[l].shape == ws[l].shape and g_b[l].shape == bs[l].shape
        _close(g_W[l], rg_W[l], rtol=2e-5, what=f"g_W[{l}]")
        _close(g_b[l], rg_b[l], rtol=2e-5, what=f"g_b[{l}]")
    # bitwise reproducible (fixed reduction orders everywhere)
    g_x2, g_V2, g_W2, g_b2 = K.mlp_jvp_bwd(f(x), f(V), wpack, widths, saved, f(g_st), f(g_ST))
    assert torch.equal(g_x, g_x2) and all(torch.equal(a, b) for a, b in zip(g_W + g_b, g_W2 + g_b2))


def test_mlp_op_through_autograd_and_many_tiles():
    """ops.mlp_jvp at the tabular step's size (B = 5000, d = 6: 278 tiles over 148 persistent CTAs) against the
    F.linear stack it replaces, through autograd."""
    from cmf_b200 import networks, ops
    torch.manual_seed(0)
    net = networks.TangentMLP(3, [128] * 4, 6, "tanh").to(DEV)
    x = torch.randn(5000, 3, device=DEV, requires_grad=True)
    V = torch.randn(5000, 6, 3, device=DEV, requires_grad=True)
    gy, gY = torch.randn(5000, 6, device=DEV), torch.randn(5000, 6, 6, device=DEV)
    res = []
    for fused in (True, False):
        networks._MLP_FUSED = fused
        try:
            for t in (x, V, *net.parameters()):
                t.grad = None
            y, Y = net.jvp_batched(x, V)
            ((y * gy).sum() + (Y * gY).sum()).backward()
            res.append((y.detach(), Y.detach(), x.grad.clone(), V.grad.clone(), [p.grad.clone() for p in net.parameters()]))
        finally:
            networks._MLP_FUSED = True
    (y1, Y1, gx1, gV1, gp1), (y0, Y0, gx0, gV0, gp0) = res
    _close(y1, y0, what="y"); _close(Y1, Y0, what="Y")
    _close(gx1, gx0, rtol=1e-4, what="g_x"); _close(gV1, gV0, rtol=1e-4, what="g_V")
    for a, b in zip(gp1, gp0):
        _close(a, b, rtol=1e-4, what="g_param")
    # primal-only entry (encode pass) and no-grad mode
    y2 = net(x.detach())
    _close(y2, y0, what="forward")


def test_mlp_rejects_what_it_cannot_hold():
    import cmf_b200.kernels as K
    from cmf_b200._lib import CmfB200Error
    x = torch.zeros(4, 3, device=DEV)
    assert not K.mlp_usable([3, 512, 512, 6], 6, x)                   # wider than a tile: stays on F.linear
    assert not K.mlp_usable([3, 128, 6], 128, x)                      # 1 + d rows of a sample must fit a tile
    assert not K.mlp_usable([3, 128, 6], 2, x.cpu())
    wpack, widths = K.mlp_pack([torch.zeros(8, 3, device=DEV)], [torch.zeros(8, device=DEV)])
    with pytest.raises(CmfB200Error):
        K.mlp_jvp_fwd(torch.zeros(4, 5, device=DEV), None, wpack, widths, save=False)     # wrong input width


# ------------------------------------------------------------------------------------------------
# tail of the image couplers' conditioner (cmf_resnet_tail_fwd / _bwd)
# ------------------------------------------------------------------------------------------------
TAIL_SHAPES = [  # B, d, C, HW, selection ("all" | "checker" | "third"), O
    (3, 5, 64, 196, "checker", 2), (2, 4, 8, 64, "all", 2), (2, 0, 64, 49, "third", 6), (1, 3, 64, 130, "all", 12),
    (2, 2, 20, 33, "checker", 24), (4, 64, 64, 784, "checker", 2), (2, 3, 64, 256, "all", 4), (1, 0, 5, 1, "all", 1),
]


@pytest.mark.parametrize("B,d,C,HW,sel,O", TAIL_SHAPES)
def test_resnet_tail_fwd_bwd(B, d, C, HW, sel, O):
    """ReLU -> 1x1 conv -> scaled tanh with its JVP at the selected positions (`networks.py:96-113,156-161`,
    `jvp_layers.py:38-64`) against fp64: outputs, dense hidden gradients, every parameter gradient."""
    import cmf_b200.kernels as K
    g = torch.Generator().manual_seed(B + d + C + HW + O)
    r = lambda *s: torch.rand(*s, generator=g, dtype=torch.float64) * 2 - 1
    keep = {"all": torch.ones(HW, dtype=torch.bool), "checker": torch.arange(HW) % 2 == 0, "third": torch.arange(HW) % 3 == 1}[sel]
    P = int(keep.sum())
    inv = None
    if sel != "all":
        inv = torch.full((HW,), -1, dtype=torch.int32)
        inv[keep] = torch.arange(P, dtype=torch.int32)
    h, Hv = r(B, C, HW), (r(B * d, C, HW) if d else None)
    w2, b2, sw, sb = r(O, C) * (1.5 / C ** 0.5), r(O) * 0.3, r(O) + 1.5, r(O)
    g_y, g_Y = r(B, O, P), (r(B, d, O, P) if d else None)
    f = lambda t: None if t is None else t.float().to(DEV)
    dinv = None if inv is None else inv.to(DEV)
    assert K.resnet_tail_usable(f(h), C, O)
    y, Y = K.resnet_tail_fwd(f(h), f(Hv), dinv, P, f(w2), f(b2), f(sw), f(sb))
    ry, rY = E.resnet_tail_fwd(h, Hv, inv, P, w2, b2, sw, sb)
    _close(y, ry, what="y")
    if d:
        _close(Y, rY, what="Y")
    out = K.resnet_tail_bwd(f(h), f(Hv), dinv, P, f(w2), f(b2), f(sw), f(sb), f(g_y), f(g_Y))
    ref = E.resnet_tail_bwd(h, Hv, inv, P, w2, b2, sw, sb, g_y, g_Y)
    for name, a, b in zip(("g_h", "g_Hv", "g_w2", "g_b2", "g_sw", "g_sb"), out, ref):
        if b is not None:
            _close(a, b, rtol=3e-5, what=name)
    out2 = K.resnet_tail_bwd(f(h), f(Hv), dinv, P, f(w2), f(b2), f(sw), f(sb), f(g_y), f(g_Y))
    assert all(torch.equal(a, b) for a, b in zip(out, out2) if a is not None)          # fixed reduction order


def test_resnet_tail_in_the_conditioner():
    """ScaledTanhResNet with the fused tail against the library tail, through autograd, compact and full outputs."""
    from cmf_b200 import networks
    torch.backends.cuda.matmul.allow_tf32 = False          # the library side of the comparison in plain FP32
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(0)
    net = networks.ScaledTanhResNet(2, [64, 64], 4).to(DEV)
    with torch.no_grad():
        net.weights.add_(0.3 * torch.randn_like(net.weights)); net.bias.add_(0.3 * torch.randn_like(net.bias))
    B, d, H, W = 3, 5, 14, 14
    x = torch.randn(B, 2, H, W, device=DEV, requires_grad=True)
    V = torch.randn(B, d, 2, H, W, device=DEV, requires_grad=True)
    keep = (torch.arange(H * W) // W + torch.arange(H * W) % W) % 2 == 0
    idx = torch.nonzero(keep).flatten()
    inv = torch.full((H * W,), -1, dtype=torch.int32); inv[idx] = torch.arange(idx.numel(), dtype=torch.int32)
    pos = (idx.int().to(DEV), inv.to(DEV), idx.to(DEV))
    for p in (None, pos):
        res = []
        for fused in (True, False):
            networks._TAIL_FUSED = fused
            try:
                for t in (x, V, *net.parameters()):
                    t.grad = None
                y, Y = net.jvp_batched(x, V, p)
                gen = torch.Generator(device=DEV).manual_seed(1)
                ((y * torch.randn(y.shape, device=DEV, generator=gen)).sum() + (Y * torch.randn(Y.shape, device=DEV, generator=gen)).sum()).backward()
                res.append((y.detach(), Y.detach(), x.grad.clone(), V.grad.clone(), [q.grad.clone() for q in net.parameters()]))
                y1 = net(x.detach()) if p is None else net.forward_compact(x.detach(), p)
                _close(y1, y.detach(), what="primal-only entry")
            finally:
                networks._TAIL_FUSED = True
        (y1, Y1, gx1, gV1, gp1), (y0, Y0, gx0, gV0, gp0) = res
        _close(y1, y0, what="y"); _close(Y1, Y0, what="Y")
        _close(gx1, gx0, rtol=1e-4, what="g_x"); _close(gV1, gV0, rtol=1e-4, what="g_V")
        for a, b in zip(gp1, gp0):
            _close(a, b, rtol=1e-4, atol=1e-6 + 1e-5 * float(b.abs().max()), what="g_param")


@pytest.mark.parametrize("shape", [(50, 64, 28, 28), (3, 64, 14, 14), (7, 5, 3, 3), (0, 4, 2, 2), (2, 1, 1, 1), (9, 64, 7, 5)])
def test_channel_sum(shape):
    """bias gradient of a convolution (sum over images and pixels), incl. a batch slice of a larger tensor"""
    import cmf_b200.kernels as K
    x = _rand(shape[0] + 2, *shape[1:], seed=3)
    xd = x.float().to(DEV)
    out = K.channel_sum(xd[2:])
    ref = x[2:].sum(dim=(0, 2, 3))
    err = float((out.double().cpu() - ref).abs().max())
    assert err <= 1e-6 * (float(x[2:].abs().sum(dim=(0, 2, 3)).max()) + 1e-30) + 1e-12
    assert torch.equal(out, K.channel_sum(xd[2:]))


def test_relu_and_pack_in_one_pass():
    """max(x, 0) into a batch slice of a larger buffer + the packed sign bits, against the two separate kernels"""
    import cmf_b200.kernels as K
    x = _rand(5, 64, 14, 14, seed=11).float().to(DEV)
    big = torch.full((9, 64, 14, 14), 7.0, device=DEV)
    bits = K.relu_mask_pack(x, relu_out=big[4:])
    assert torch.equal(bits, K.relu_mask_pack(x))
    assert torch.equal(big[4:], torch.clamp_min(x, 0)) and bool((big[:4] == 7.0).all())


def test_empty_batches_of_the_conditioner_kernels():
    """B = 0 through the MLP, tail and first-convolution entry points: empty outputs, zero parameter gradients, no launch error"""
    import cmf_b200.kernels as K
    z = lambda *s: torch.zeros(*s, device=DEV)
    # MLP
    ws, bs = [torch.randn(16, 3, device=DEV), torch.randn(4, 16, device=DEV)], [torch.randn(16, device=DEV), torch.randn(4, device=DEV)]
    wpack, widths = K.mlp_pack(ws, bs)
    st, ST, saved = K.mlp_jvp_fwd(z(0, 3), z(0, 2, 3), wpack, widths, save=True)
    assert st.shape == (0, 4) and ST.shape == (0, 2, 4)
    g_x, g_V, g_W, g_b = K.mlp_jvp_bwd(z(0, 3), z(0, 2, 3), wpack, widths, saved, z(0, 4), z(0, 2, 4))
    assert g_x.shape == (0, 3) and all(float(g.abs().max()) == 0 for g in g_W + g_b)
    # tail
    w2, b2, sw, sb = torch.randn(2, 64, device=DEV), z(2), torch.ones(2, device=DEV), z(2)
    y, Y = K.resnet_tail_fwd(z(0, 64, 16), z(0, 64, 16), None, 16, w2, b2, sw, sb)
    assert y.shape == (0, 2, 16)
    out = K.resnet_tail_bwd(z(0, 64, 16), None, None, 16, w2, b2, sw, sb, z(0, 2, 16), None)
    assert float(out[2].abs().max()) == 0 and float(out[3].abs().max()) == 0
    # first convolution
    w = torch.randn(64, 2, 3, 3, device=DEV)
    assert K.convin_fwd(z(0, 2, 8, 8), w).shape == (0, 64, 8, 8)
    assert K.convin_dgrad(z(0, 64, 8, 8), w).shape == (0, 2, 8, 8)
    assert float(K.convin_wgrad(z(0, 2, 8, 8), z(0, 64, 8, 8)).abs().max()) == 0
