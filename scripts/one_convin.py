"""standalone launches of the first-convolution kernels at the step's joint batch (N = 3250, C_in = 1, 28x28; for ncu)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cmf_b200 import kernels as K
torch.manual_seed(0)
x = torch.randn(3250, 1, 28, 28, device="cuda"); w = torch.randn(64, 1, 3, 3, device="cuda") / 3
gy = torch.randn(3250, 64, 28, 28, device="cuda")
for _ in range(3):
    K.convin_fwd(x, w); K.convin_dgrad(gy, w); K.convin_wgrad(x, gy)
torch.cuda.synchronize()
print("ok")
