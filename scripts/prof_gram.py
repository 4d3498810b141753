"""One shape for ncu captures of the tcgen05 Gram kernels (B, d, n from the environment)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cmf_b200 import kernels as K
B, d, n = int(os.environ.get("B", 512)), int(os.environ.get("D", 64)), int(os.environ.get("N", 392))
V = torch.randn(B, d, n, device="cuda"); gG = torch.randn(B, d, d, device="cuda")
for _ in range(3):
    G = K.gram_fwd(V, path="tc"); g = K.gram_bwd(V, gG, path="tc")
torch.cuda.synchronize()
print("ok")
