"""GPU check + timing of the tcgen05 conv3x3 weight gradient against FP64 / cuDNN FP32."""
import sys
import torch
from cmf_b200 import kernels as K

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
dev = "cuda"


def ref_wgrad(x, gy, dtype):
    w = torch.zeros(64, 64, 3, 3, device=dev, dtype=dtype)
    return torch.ops.aten.convolution_backward(gy.to(dtype), x.to(dtype), w, None, (1, 1), (1, 1), (1, 1), False, (0, 0), 1,
                                               (False, True, False))[1]


def check(N, H, W, seed=0):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(N, 64, H, W, generator=g).to(dev)
    gy = torch.randn(N, 64, H, W, generator=g).to(dev)
    gw = K.conv3x3_c64_wgrad(x, gy)
    torch.cuda.synchronize()
    r64 = ref_wgrad(x, gy, torch.float64)
    r32 = ref_wgrad(x, gy, torch.float32)
    err = ((gw.double() - r64).abs().max() / r64.abs().max()).item()
    e32 = ((r32.double() - r64).abs().max() / r64.abs().max()).item()
    print(f"N={N} {H}x{W}: wgrad err {err:.2e} (cudnn fp32 {e32:.2e})", flush=True)
    return err


def bench(N, H, W, iters=10):
    x = torch.randn(N, 64, H, W, device=dev)
    gy = torch.randn(N, 64, H, W, device=dev)
    flops = 2.0 * N * 64 * 64 * 9 * H * W
    for name, fn in (("tc", lambda: K.conv3x3_c64_wgrad(x, gy)), ("cudnn", lambda: ref_wgrad(x, gy, torch.float32))):
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
        print(f"  {name:6s} N={N} {H}x{W}: {ms:.3f} ms  {flops / ms / 1e9:.1f} TFLOP/s", flush=True)


if __name__ == "__main__":
    worst = 0.0
    for shape in [(1, 8, 8), (2, 8, 8), (3, 28, 28), (5, 14, 14), (1, 32, 32), (7, 6, 10), (4, 16, 16), (40, 28, 28)]:
        worst = max(worst, check(*shape))
    print("worst", worst)
    if "--bench" in sys.argv:
        torch.backends.cudnn.benchmark = True
        bench(3250, 28, 28)
        bench(3250, 14, 14)
        bench(50, 28, 28, iters=50)
    sys.exit(0 if worst < 1e-5 else 1)
