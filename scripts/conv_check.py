"""GPU check + timing of the tcgen05 conv3x3 (64->64) against cuDNN FP32 / FP64."""
import sys, time
import torch
import torch.nn.functional as F
from cmf_b200 import kernels as K

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
dev = "cuda"


def check(N, H, W, bias=True, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = torch.randn(N, 64, H, W, generator=g).to(dev)
    w = (torch.randn(64, 64, 3, 3, generator=g) / 24).to(dev)
    b = torch.randn(64, generator=g).to(dev) if bias else None
    y = K.conv3x3_c64(x, K.conv3x3_c64_prep(w, False), b)
    ref = F.conv2d(x.double(), w.double(), None if b is None else b.double(), padding=1)
    y32 = F.conv2d(x, w, b, padding=1)
    err = ((y.double() - ref).abs().max() / ref.abs().max()).item()
    err32 = ((y32.double() - ref).abs().max() / ref.abs().max()).item()
    # data gradient
    gy = torch.randn(N, 64, H, W, generator=g).to(dev)
    gx = K.conv3x3_c64(gy, K.conv3x3_c64_prep(w, True), None)
    gref = F.conv_transpose2d(gy.double(), w.double(), padding=1)
    gerr = ((gx.double() - gref).abs().max() / gref.abs().max()).item()
    print(f"N={N} {H}x{W} bias={bias}: fwd err {err:.2e} (cudnn fp32 {err32:.2e})  dgrad err {gerr:.2e}", flush=True)
    return max(err, gerr)


def bench(N, H, W, iters=10):
    x = torch.randn(N, 64, H, W, device=dev)
    w = torch.randn(64, 64, 3, 3, device=dev) / 24
    ws = K.conv3x3_c64_prep(w, False)
    flops = 2.0 * N * 64 * 64 * 9 * H * W
    res = {}
    d = 65 if N % 65 == 0 else 1
    prim = torch.randn(N // d, 64, H, W, device=dev)
    bits = K.relu_mask_pack(prim)
    skip = torch.randn(N, 64, H, W, device=dev)
    for name, fn in (("tc", lambda: K.conv3x3_c64(x, ws, None)),
                     ("tc+mask", lambda: K.conv3x3_c64(x, ws, None, mask_src=bits, mask_div=d, mask_first=True)),
                     ("tc+mask+res", lambda: K.conv3x3_c64(x, ws, None, mask_src=bits, mask_div=d, mask_first=True, res=skip)),
                     ("tc+mask+res+y2", lambda: K.conv3x3_c64(x, ws, None, mask_src=bits, mask_div=d, res=skip, second=True)),
                     ("tc drain2", lambda: K.conv3x3_c64(x, ws, None, drain=2)),
                     ("tc drain3", lambda: K.conv3x3_c64(x, ws, None, drain=3)),
                     ("cudnn", lambda: F.conv2d(x, w, None, padding=1))):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        res[name] = ms
        print(f"  {name:15s} N={N} {H}x{W}: {ms:.3f} ms  {flops / ms / 1e9:.1f} TFLOP/s", flush=True)
    return res


if __name__ == "__main__":
    worst = 0.0
    for shape in [(1, 8, 8), (2, 8, 8), (3, 28, 28), (5, 14, 14), (1, 32, 32), (7, 6, 10), (4, 16, 16)]:
        worst = max(worst, check(*shape))
    worst = max(worst, check(3, 28, 28, bias=False, seed=1))
    print("worst", worst)
    if "--bench" in sys.argv:
        torch.backends.cudnn.benchmark = True
        bench(3250, 28, 28)
        bench(3250, 14, 14)
        bench(50, 28, 28, iters=50)
    sys.exit(0 if worst < 2e-6 else 1)
