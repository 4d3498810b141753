"""standalone launches of the MLP conditioner kernels at the power-shape coupler (B = 5000, d = 6; for ncu)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cmf_b200 import networks
torch.manual_seed(0)
B, d, nin, nout = 5000, 6, 3, 6
net = networks.TangentMLP(nin, [128] * 4, nout, "tanh").cuda()
x = torch.randn(B, nin, device="cuda", requires_grad=True)
V = torch.randn(B, d, nin, device="cuda", requires_grad=True)
for _ in range(3):
    y, Y = net.jvp_batched(x, V)
    (y.sum() + Y.sum()).backward()
torch.cuda.synchronize()
print("ok")
