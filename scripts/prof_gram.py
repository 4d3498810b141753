import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cmf_b200 import kernels as K
B, d, n = 512, 64, 392
V = torch.randn(B, d, n, device="cuda"); gG = torch.randn(B, d, d, device="cuda")
for _ in range(3):
    G = K.gram_fwd(V, path="tc"); g = K.gram_bwd(V, gG, path="tc")
torch.cuda.synchronize()
print("ok")
