"""One shape for ncu captures of the Cholesky kernels (B, d from the environment)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cmf_b200 import kernels as K
B, d = int(os.environ.get("B", 512)), int(os.environ.get("D", 64))
torch.manual_seed(0)
A = torch.randn(B, d, 2 * d, device="cuda")
G = A @ A.transpose(1, 2) + 0.5 * torch.eye(d, device="cuda")
for _ in range(3):
    L, logdet, l1o, l1d, info = K.chol_logdet_fwd(G)
    gG = K.chol_logdet_bwd(L, G, torch.ones_like(logdet), torch.ones_like(l1o), "offdiag")
torch.cuda.synchronize()
print("ok")
