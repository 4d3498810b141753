#!/usr/bin/env python
"""First-conv kernels against cuDNN FP32 at the step's shapes (N = 3250 joint batch)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from cmf_b200 import kernels as K
torch.backends.cudnn.allow_tf32 = False
torch.backends.cudnn.benchmark = True
def timed(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
for N, Cin, H in ((3250, 1, 28), (3250, 2, 14), (3120, 3, 32)):
    x = torch.randn(N, Cin, H, H, device="cuda"); w = torch.randn(64, Cin, 3, 3, device="cuda") / 3
    gy = torch.randn(N, 64, H, H, device="cuda")
    mb = N * 64 * H * H * 4 / 1e6
    print(f"N={N} Cin={Cin} {H}x{H} ({mb:.0f} MB of 64-channel tensor)")
    print(f"   fwd   own {timed(lambda: K.convin_fwd(x, w)):8.1f} us   cudnn {timed(lambda: F.conv2d(x, w, padding=1)):8.1f} us")
    print(f"   dgrad own {timed(lambda: K.convin_dgrad(gy, w)):8.1f} us   cudnn {timed(lambda: torch.nn.grad.conv2d_input(x.shape, w, gy, padding=1)):8.1f} us")
    print(f"   wgrad own {timed(lambda: K.convin_wgrad(x, gy)):8.1f} us   cudnn {timed(lambda: torch.nn.grad.conv2d_weight(x, w.shape, gy, padding=1)):8.1f} us")
