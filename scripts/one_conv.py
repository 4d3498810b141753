"""standalone launches of the two tcgen05 conv kernels at the tangent-batch shape of the MNIST-shape step (for ncu)"""
import torch
from cmf_b200 import kernels as K
N, H, W = 3250, 28, 28
x = torch.randn(N, 64, H, W, device="cuda"); gy = torch.randn(N, 64, H, W, device="cuda")
w = torch.randn(64, 64, 3, 3, device="cuda") / 24
ws = K.conv3x3_c64_prep(w, False)
for _ in range(3):
    y = K.conv3x3_c64(x, ws, None)
    gw = K.conv3x3_c64_wgrad(x, gy)
torch.cuda.synchronize()
print("ok")
