"""standalone launches of the conditioner-tail kernels at the step's 28x28 checkerboard shape (B = 50, d = 64; for ncu)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cmf_b200 import kernels as K
torch.manual_seed(0)
B, d, C, H, O = 50, 64, 64, 28, 2
HW = H * H
h = torch.randn(B, C, HW, device="cuda"); Hv = torch.randn(B * d, C, HW, device="cuda")
keep = ((torch.arange(HW) // H + torch.arange(HW) % H) % 2 == 0)
inv = torch.full((HW,), -1, dtype=torch.int32); inv[keep] = torch.arange(int(keep.sum()), dtype=torch.int32)
inv = inv.cuda(); P = int(keep.sum())
w2, b2, sw, sb = torch.randn(O, C, device="cuda") / 8, torch.randn(O, device="cuda"), torch.ones(O, device="cuda"), torch.zeros(O, device="cuda")
gy, gY = torch.randn(B, O, P, device="cuda"), torch.randn(B, d, O, P, device="cuda")
def timed(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
print("fwd us", timed(lambda: K.resnet_tail_fwd(h, Hv, inv, P, w2, b2, sw, sb)))
print("bwd us", timed(lambda: K.resnet_tail_bwd(h, Hv, inv, P, w2, b2, sw, sb, gy, gY)))
