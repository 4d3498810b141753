import sys, torch
from cmf_b200 import kernels as K
N, H, W, iters = (int(a) for a in sys.argv[1:5])
x = torch.randn(N, 64, H, W, device="cuda"); gy = torch.randn(N, 64, H, W, device="cuda")
for _ in range(3):
    K.conv3x3_c64_wgrad(x, gy)
torch.cuda.synchronize()
print("warm ok", flush=True)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(iters):
    K.conv3x3_c64_wgrad(x, gy)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / iters
print(f"N={N} {H}x{W}: {ms:.3f} ms  {2.0*N*64*64*9*H*W/ms/1e9:.1f} TFLOP/s", flush=True)
