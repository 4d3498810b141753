"""tcgen05 conditioner convolution against an FP64 statement of the same op
(`networks.py:35-94` of the reference: 3x3, 64 -> 64 channels, zero padding 1), through the C ABI.
Tolerance: 2e-6 of the output scale (the north-star bar is 1e-4 relative; the kernel measures
2-6e-7, cuDNN FP32's autotuned algorithms 4-7e-6 on these shapes)."""
import pytest
import torch
import torch.nn.functional as F

from cmf_b200 import _lib, kernels as K, ops

pytestmark = pytest.mark.gpu
DEV = "cuda"

SHAPES = [(1, 8, 8), (2, 8, 8), (3, 28, 28), (5, 14, 14), (1, 32, 32), (7, 6, 10), (4, 16, 16), (2, 2, 2),
          (130, 14, 14), (1, 4, 60), (3, 1, 4)]


def _rand(shape, seed, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(DEV)


def _rel(a, ref):
    return ((a.double() - ref).abs().max() / ref.abs().max().clamp_min(1e-30)).item()


@pytest.mark.parametrize("N,H,W", SHAPES)
@pytest.mark.parametrize("bias", [True, False])
def test_forward_and_data_gradient(N, H, W, bias):
    x, w = _rand((N, 64, H, W), 1), _rand((64, 64, 3, 3), 2, 1 / 24)
    b = _rand((64,), 3) if bias else None
    y = K.conv3x3_c64(x, K.conv3x3_c64_prep(w, False), b)
    ref = F.conv2d(x.double(), w.double(), None if b is None else b.double(), padding=1)
    assert _rel(y, ref) < 2e-6
    gy = _rand((N, 64, H, W), 4)
    gx = K.conv3x3_c64(gy, K.conv3x3_c64_prep(w, True), None)
    gref = F.conv_transpose2d(gy.double(), w.double(), padding=1)
    assert _rel(gx, gref) < 2e-6


def _wgrad_ref(x, gy):
    w = torch.zeros(64, 64, 3, 3, device=DEV, dtype=torch.float64)
    return torch.ops.aten.convolution_backward(gy.double(), x.double(), w, None, (1, 1), (1, 1), (1, 1), False, (0, 0), 1,
                                               (False, True, False))[1]


@pytest.mark.parametrize("N,H,W", SHAPES + [(40, 28, 28), (300, 14, 14)])
def test_weight_gradient(N, H, W):
    """pixels-as-K tcgen05 kernel + fixed-order partial reduction vs FP64; several CTAs and tiles
    per CTA in the larger cases (back-to-back launches: the protocol once raced only then)"""
    x, gy = _rand((N, 64, H, W), 20), _rand((N, 64, H, W), 21)
    if not K.conv3x3_c64_wgrad_usable(N, H, W):                # wider than the halo box: rejected, not approximated
        with pytest.raises(_lib.CmfB200Error):
            K.conv3x3_c64_wgrad(x, gy)
        return
    gw = K.conv3x3_c64_wgrad(x, gy)
    assert _rel(gw, _wgrad_ref(x, gy)) < 5e-6
    # two issuer warps add into one TMEM accumulator in arrival order: run-to-run differences are
    # confined to the truncation noise of the accumulate (the partial reduction itself is fixed-order)
    assert _rel(K.conv3x3_c64_wgrad(x, gy), gw.double()) < 2e-6


@pytest.mark.parametrize("H,W", [(28, 28), (14, 14)])
def test_benchmarked_tangent_batch(H, W):
    """N = 3200 = B * d of the benchmarked step (B = 50, d = 64): 148 persistent CTAs each walking
    ~66 (28x28) / ~17 (14x14) tile pairs, weight-gradient chunks of T = 14 / 15 tiles -- forward, data
    gradient (both drain intervals) and weight gradient (both modes) against FP64 on the device."""
    N = 3200
    x, gy, w = _rand((N, 64, H, W), 41), _rand((N, 64, H, W), 42), _rand((64, 64, 3, 3), 43, 1 / 24)
    b = _rand((64,), 44)
    devs = {}
    ref = torch.empty(N, 64, H, W, dtype=torch.float64, device=DEV)
    for i in range(0, N, 400):                                  # FP64 reference in slices (memory)
        ref[i:i + 400] = F.conv2d(x[i:i + 400].double(), w.double(), b.double(), padding=1)
    devs["fwd"] = _rel(K.conv3x3_c64(x, K.conv3x3_c64_prep(w, False), b), ref)
    for i in range(0, N, 400):
        ref[i:i + 400] = F.conv_transpose2d(gy[i:i + 400].double(), w.double(), padding=1)
    wt = K.conv3x3_c64_prep(w, True)
    devs["dgrad drain 1"] = _rel(K.conv3x3_c64(gy, wt, None), ref)
    devs["dgrad drain 2"] = _rel(K.conv3x3_c64(gy, wt, None, drain=2), ref)
    del ref
    gw_ref = torch.zeros(64, 64, 3, 3, dtype=torch.float64, device=DEV)
    for i in range(0, N, 400):
        gw_ref += _wgrad_ref(x[i:i + 400], gy[i:i + 400])
    devs["wgrad"] = _rel(K.conv3x3_c64_wgrad(x, gy, deterministic=False), gw_ref)
    devs["wgrad deterministic"] = _rel(K.conv3x3_c64_wgrad(x, gy, deterministic=True), gw_ref)
    try:
        from helpers import record
        for k, v in devs.items():
            record(f"conv3x3 N=3200 {H}x{W}", k + " vs FP64", v, 2e-5 if k.startswith("wgrad") else 2e-6)
    except ImportError:
        pass
    assert devs["fwd"] < 2e-6 and devs["dgrad drain 1"] < 2e-6 and devs["dgrad drain 2"] < 2e-6, devs
    assert devs["wgrad"] < 2e-5 and devs["wgrad deterministic"] < 2e-5, devs


def test_weight_gradient_same_sign_terms():
    """all-positive terms: the longest truncating accumulation (T = 14 tiles per CTA)"""
    x, gy = _rand((64, 64, 28, 28), 22).abs(), _rand((64, 64, 28, 28), 23).abs()
    gw = K.conv3x3_c64_wgrad(x, gy)
    ref = _wgrad_ref(x, gy)
    assert ((gw.double() - ref) / ref).abs().max().item() < 5e-5


def test_badly_scaled_channels():
    """per-channel magnitudes spread over 1e6: the hi/lo split must stay exact"""
    x, w = _rand((3, 64, 14, 14), 5), _rand((64, 64, 3, 3), 6, 1 / 24)
    scale = torch.logspace(-3, 3, 64, device=DEV).view(1, 64, 1, 1)
    x = x * scale
    y = K.conv3x3_c64(x, K.conv3x3_c64_prep(w, False), None)
    ref = F.conv2d(x.double(), w.double(), padding=1)
    assert _rel(y, ref) < 1e-5


def test_relu_inputs_accumulate_without_bias():
    """all-positive products (post-ReLU inputs, positive weights) are the worst case for the
    truncating tensor-core accumulator: one K = 576 accumulation drifts by ~1e-5; with the FP32
    register drain every K = 96 the systematic part is ~2e-6 (cuDNN FP32: 2.5e-6 max, unbiased)"""
    x, w = _rand((4, 64, 28, 28), 7).abs(), _rand((64, 64, 3, 3), 8, 1 / 24).abs()
    y = K.conv3x3_c64(x, K.conv3x3_c64_prep(w, False), None)
    ref = F.conv2d(x.double(), w.double(), padding=1)
    rel = (y.double() - ref) / ref
    assert rel.abs().max().item() < 1e-6 and abs(rel.mean().item()) < 3e-7


def test_autograd_wrapper_matches_library_conv():
    x, w, b = _rand((6, 64, 14, 14), 9), _rand((64, 64, 3, 3), 10, 1 / 24), _rand((64,), 11)
    g = _rand((6, 64, 14, 14), 12)
    outs = []
    for fn in (lambda a, ww, bb: ops.conv3x3_c64(a, ww, bb),
               lambda a, ww, bb: F.conv2d(a.double(), ww.double(), bb.double(), padding=1)):
        a, ww, bb = (t.clone().requires_grad_(True) for t in (x, w, b))
        y = fn(a, ww, bb)
        y.backward(g.to(y.dtype))
        outs.append((y, a.grad, ww.grad, bb.grad))
    for got, ref in zip(*outs):
        assert _rel(got, ref.double()) < 1e-5


def test_shared_presplit_weights_and_tangent_batch():
    """one prep serves the primal (with bias) and the bias-free tangent call"""
    w, b = _rand((64, 64, 3, 3), 13, 1 / 24), _rand((64,), 14)
    ws = K.conv3x3_c64_prep(w, False)
    h, Hv = _rand((2, 64, 14, 14), 15), _rand((2 * 5, 64, 14, 14), 16)
    y, Y = ops.conv3x3_c64(h, w, b, ws), ops.conv3x3_c64(Hv, w, None, ws)
    assert _rel(y, F.conv2d(h.double(), w.double(), b.double(), padding=1)) < 1e-5
    assert _rel(Y, F.conv2d(Hv.double(), w.double(), None, padding=1)) < 1e-5


@pytest.mark.parametrize("N,H,W,d", [(6, 14, 14, 3), (4, 28, 28, 2), (5, 8, 8, 1), (130, 14, 14, 65)])
@pytest.mark.parametrize("mode", ["mask", "res_second", "mask_res", "res"])
def test_fused_epilogue(N, H, W, d, mode):
    """relu-JVP mask (sign of the primal, one primal image per d tangent images) and skip
    connection folded into the epilogue, against the same composition in FP64"""
    import emu_kernels as E
    x, w = _rand((N, 64, H, W), 30), _rand((64, 64, 3, 3), 31, 1 / 24)
    prim, res = _rand((N // d, 64, H, W), 32), _rand((N, 64, H, W), 33)
    kw = {"mask": dict(mask_src=prim, mask_div=d, mask_first=True),
          "res_second": dict(mask_src=prim, mask_div=d, res=res, second=True),
          "mask_res": dict(mask_src=prim, mask_div=d, mask_first=True, res=res),
          "res": dict(res=res)}[mode]
    got = K.conv3x3_c64(x, K.conv3x3_c64_prep(w, False), None, **kw)
    kw64 = {k: (v.double() if torch.is_tensor(v) else v) for k, v in kw.items()}
    ref = E.conv3x3_c64(x.double(), w.double(), None, **kw64)
    got, ref = (got, ref) if isinstance(got, tuple) else ((got,), (ref,))
    for g_, r_ in zip(got, ref):
        assert _rel(g_, r_) < 2e-6
        assert torch.equal(g_ == 0, r_ == 0) or mode == "res"      # masked entries are exact zeros


@pytest.mark.parametrize("N,H,W,d", [(6, 14, 14, 3), (4, 28, 28, 2), (130, 14, 14, 65), (9, 6, 10, 3)])
def test_packed_relu_mask_is_the_same_mask(N, H, W, d):
    """cmf_relu_mask_pack + cmf_conv3x3_c64_fused_bits: bit-identical outputs to the float-mask epilogue
    (zeros and negative zeros of the primal count as masked in both)"""
    import emu_kernels as E
    x, w = _rand((N, 64, H, W), 50), _rand((64, 64, 3, 3), 51, 1 / 24)
    prim, res = _rand((N // d, 64, H, W), 52), _rand((N, 64, H, W), 53)
    prim[0, :, 0, :2] = 0.0
    prim[0, 5, 1, 1] = -0.0
    bits = K.relu_mask_pack(prim)
    assert bits.dtype == torch.int32 and tuple(bits.shape) == (N // d, H * W, 2)
    assert torch.equal(bits.cpu(), E.relu_mask_pack(prim.cpu()))
    ws = K.conv3x3_c64_prep(w, False)
    for kw in (dict(mask_first=True), dict(res=res, second=True), dict(mask_first=True, res=res)):
        a = K.conv3x3_c64(x, ws, None, mask_src=prim, mask_div=d, **kw)
        b = K.conv3x3_c64(x, ws, None, mask_src=bits, mask_div=d, **kw)
        a, b = (a, b) if isinstance(a, tuple) else ((a,), (b,))
        for u, v in zip(a, b):
            assert torch.equal(u, v)


@pytest.mark.parametrize("drain", [2, 3])
def test_longer_drain_interval_of_the_gradient_launches(drain):
    """drain = 2 (used by the data-gradient launches): K = 64 per FP32 drain, still ~5e-7 of the output scale"""
    x, w = _rand((5, 64, 28, 28), 60), _rand((64, 64, 3, 3), 61, 1 / 24)
    got = K.conv3x3_c64(x, K.conv3x3_c64_prep(w, True), None, drain=drain)
    ref = F.conv_transpose2d(x.double(), w.double(), padding=1)
    assert _rel(got, ref) < 2e-6
    with pytest.raises(_lib.CmfB200Error):
        K.conv3x3_c64(x, K.conv3x3_c64_prep(w, True), None, drain=4)


def test_residual_stack_tangent_matches_autograd():
    """the fused tangent node of a residual-block run vs the per-op autograd path (library convs)"""
    from cmf_b200 import networks as NW
    torch.backends.cudnn.allow_tf32 = False          # the library path must be FP32 to serve as the reference
    torch.manual_seed(0)
    blocks = [NW.ResidualBlock(64).to(DEV) for _ in range(3)]
    B, d = 2, 3
    h0, V0 = _rand((B, 64, 14, 14), 40), _rand((B * d, 64, 14, 14), 41)
    gH, gV = _rand((B, 64, 14, 14), 42), _rand((B * d, 64, 14, 14), 43)
    results = []
    for fused in (True, False):
        NW._TC_CONV = fused
        try:
            for b in blocks:
                b.zero_grad()
            h, V = h0.clone().requires_grad_(True), V0.clone().requires_grad_(True)
            hh, VV = NW._residual_stack_jvp(blocks, h, V, d)
            ((hh * gH).sum() + (VV * gV).sum()).backward()
            results.append([hh.detach(), VV.detach(), h.grad, V.grad] + [p.grad.clone() for b in blocks for p in b.parameters()])
        finally:
            NW._TC_CONV = True
    for a, b in zip(*results):
        assert _rel(a, b.double()) < 2e-5


@pytest.mark.parametrize("N,H,W", [(1, 4, 8), (40, 28, 28), (330, 14, 14), (600, 28, 28)])
def test_run_to_run_reproducibility(N, H, W):
    """Forward / data gradient: fixed accumulation order (FP32 drains per segment), same bits every launch.
    Weight gradient: bitwise with deterministic=True (one in-order MMA stream per accumulator, per-CTA
    partials added chunk by chunk, fixed-order reduction), also when the persistent CTAs walk several
    chunks; the default (two issuers sharing an accumulator) reproduces to a few ulp of the result's
    scale and has the same accuracy against FP64."""
    x, gy, w = _rand((N, 64, H, W), 31), _rand((N, 64, H, W), 32), _rand((64, 64, 3, 3), 33, 1 / 24)
    ws, wt = K.conv3x3_c64_prep(w, False), K.conv3x3_c64_prep(w, True)
    y0, gx0 = K.conv3x3_c64(x, ws), K.conv3x3_c64(gy, wt, drain=2)
    gw_det, gw_fast = K.conv3x3_c64_wgrad(x, gy, deterministic=True), K.conv3x3_c64_wgrad(x, gy, deterministic=False)
    ref = torch.ops.aten.convolution_backward(gy.double(), x.double(), w.double(), None, (1, 1), (1, 1), (1, 1), False,
                                              (0, 0), 1, (False, True, False))[1]
    assert _rel(gw_det, ref) < 2e-5 and _rel(gw_fast, ref) < 2e-5     # measured 2e-6 (N = 40) .. 1.1e-5 (N = 3200)
    for _ in range(5):
        assert torch.equal(K.conv3x3_c64(x, ws), y0)
        assert torch.equal(K.conv3x3_c64(gy, wt, drain=2), gx0)
        assert torch.equal(K.conv3x3_c64_wgrad(x, gy, deterministic=True), gw_det)
        assert _rel(K.conv3x3_c64_wgrad(x, gy, deterministic=False), gw_fast.double()) < 2e-6
    with torch.random.fork_rng():
        torch.use_deterministic_algorithms(True)
        try:
            assert torch.equal(K.conv3x3_c64_wgrad(x, gy), gw_det)      # None follows torch's global switch
        finally:
            torch.use_deterministic_algorithms(False)


def test_unsupported_shapes_are_rejected():
    assert not K.conv3x3_c64_usable(1, 7, 7)        # H*W % 4 != 0: TMA row stride not 16-byte aligned
    assert not K.conv3x3_c64_usable(1, 64, 64)      # halo box wider than 256 pixels
    assert K.conv3x3_c64_wgrad_usable(8, 32, 32) and not K.conv3x3_c64_wgrad_usable(8, 48, 48)
    assert K.conv3x3_c64_usable(3250, 28, 28)
    x = _rand((1, 64, 7, 7), 17)
    ws = K.conv3x3_c64_prep(_rand((64, 64, 3, 3), 18), False)
    with pytest.raises(_lib.CmfB200Error):
        K.conv3x3_c64(x, ws, None)
    with pytest.raises(_lib.CmfB200Error):
        K.conv3x3_c64(x.cpu(), ws, None)


def test_empty_batch():
    x = torch.empty(0, 64, 8, 8, device=DEV)
    y = K.conv3x3_c64(x, K.conv3x3_c64_prep(_rand((64, 64, 3, 3), 19), False), None)
    assert y.shape == x.shape


def test_activation_recompute_gives_identical_gradients():
    """networks.recompute_activations: the residual-stack node keeps only its input and re-runs its
    forward launches in the backward pass -- same launches on the same data, so the same bits (the
    deterministic weight-gradient mode makes that exact), at a fraction of the saved memory"""
    import cmf_b200.networks as NW
    torch.manual_seed(0)
    blocks = [NW.ResidualBlock(64).to(DEV) for _ in range(3)]
    B, d = 3, 5
    h0, V0 = _rand((B, 64, 14, 14), 51), _rand((B * d, 64, 14, 14), 52)
    gH, gV = _rand((B, 64, 14, 14), 53), _rand((B * d, 64, 14, 14), 54)
    results, saved_bytes = [], []
    torch.use_deterministic_algorithms(True)
    try:
        for recompute in (False, True):
            for b in blocks:
                b.zero_grad()
            h, V = h0.clone().requires_grad_(True), V0.clone().requires_grad_(True)
            torch.cuda.synchronize()
            base = torch.cuda.memory_allocated()
            if recompute:
                with NW.recompute_activations():
                    hh, VV = NW._residual_stack_jvp(blocks, h, V, d)
            else:
                hh, VV = NW._residual_stack_jvp(blocks, h, V, d)
            saved_bytes.append(torch.cuda.memory_allocated() - base)
            ((hh * gH).sum() + (VV * gV).sum()).backward()
            results.append([hh.detach(), VV.detach(), h.grad, V.grad] + [p.grad.clone() for b in blocks for p in b.parameters()])
    finally:
        torch.use_deterministic_algorithms(False)
    for a, b in zip(*results):
        assert torch.equal(a, b)
    assert saved_bytes[1] < 0.4 * saved_bytes[0], saved_bytes


# ------------------------------------------------------------------------------------------------
# first convolution of the conditioner (cmf_convin_fwd / _dgrad / _wgrad)
# ------------------------------------------------------------------------------------------------
CONVIN_SHAPES = [  # N, C_in, H, W
    (3, 1, 28, 28), (5, 2, 14, 14), (2, 3, 32, 32), (2, 4, 16, 16), (7, 1, 8, 8), (1, 2, 7, 5), (4, 3, 12, 64), (130, 1, 28, 28), (5, 6, 16, 16), (3, 12, 8, 8), (2, 16, 9, 11), (2, 5, 32, 32),
]


@pytest.mark.parametrize("N,Cin,H,W", CONVIN_SHAPES)
def test_convin_fwd_dgrad_wgrad(N, Cin, H, W):
    """C_in -> 64 3x3 convolution (`networks.py:131-139`) against F.conv2d in fp64: output, data gradient, weight gradient."""
    import torch.nn.functional as F
    import cmf_b200.kernels as K
    g = torch.Generator().manual_seed(N + Cin + H + W)
    x = torch.randn(N, Cin, H, W, generator=g, dtype=torch.float64)
    w = torch.randn(64, Cin, 3, 3, generator=g, dtype=torch.float64) / (3 * Cin ** 0.5)
    gy = torch.randn(N, 64, H, W, generator=g, dtype=torch.float64)
    xr, wr = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
    yr = F.conv2d(xr, wr, padding=1)
    yr.backward(gy)
    xd, wd, gyd = x.float().cuda(), w.float().cuda(), gy.float().cuda()
    assert K.convin_usable(xd, wd)
    rel = lambda a, b: float((a.double().cpu() - b).abs().max() / b.abs().max())
    assert rel(K.convin_fwd(xd, wd), yr.detach()) < 2e-6
    assert rel(K.convin_dgrad(gyd, wd), xr.grad) < 3e-6
    g_w = K.convin_wgrad(xd, gyd)
    assert rel(g_w, wr.grad) < 1e-5
    assert torch.equal(g_w, K.convin_wgrad(xd, gyd))              # fixed reduction order


def test_convin_at_the_step_size_and_in_the_conditioner():
    """N = 3250 images of 28x28 (the step's joint batch) against cuDNN FP32, and ScaledTanhResNet with / without the own
    first convolution through autograd."""
    import torch.nn.functional as F
    import cmf_b200.kernels as K
    from cmf_b200 import networks
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    torch.manual_seed(0)
    x = torch.randn(3250, 1, 28, 28, device="cuda")
    w = torch.randn(64, 1, 3, 3, device="cuda") / 3
    gy = torch.randn(3250, 64, 28, 28, device="cuda")
    rel = lambda a, b: float((a.double() - b.double()).abs().max() / b.double().abs().max())
    assert rel(K.convin_fwd(x, w), F.conv2d(x, w, padding=1)) < 5e-6
    assert rel(K.convin_dgrad(gy, w), torch.nn.grad.conv2d_input(x.shape, w, gy, padding=1)) < 1e-5
    ref_w = torch.nn.grad.conv2d_weight(x.double(), w.shape, gy.double(), padding=1)
    assert rel(K.convin_wgrad(x, gy), ref_w) < 2e-5
    net = networks.ScaledTanhResNet(2, [64, 64], 4).cuda()
    xs = torch.randn(3, 2, 14, 14, device="cuda", requires_grad=True)
    V = torch.randn(3, 5, 2, 14, 14, device="cuda", requires_grad=True)
    res = []
    for own in (True, False):
        networks._CONVIN_OWN = own
        try:
            for t in (xs, V, *net.parameters()):
                t.grad = None
            y, Y = net.jvp_batched(xs, V)
            gen = torch.Generator(device="cuda").manual_seed(1)
            ((y * torch.randn(y.shape, device="cuda", generator=gen)).sum() + (Y * torch.randn(Y.shape, device="cuda", generator=gen)).sum()).backward()
            res.append((y.detach(), Y.detach(), xs.grad.clone(), V.grad.clone(), [q.grad.clone() for q in net.parameters()], net(xs.detach())))
        finally:
            networks._CONVIN_OWN = True
    for a, b in zip(res[0][:4] + (res[0][5],), res[1][:4] + (res[1][5],)):
        assert rel(a, b) < 1e-4
    for a, b in zip(res[0][4], res[1][4]):
        assert float((a - b).abs().max()) <= 1e-4 * float(b.abs().max()) + 1e-6
