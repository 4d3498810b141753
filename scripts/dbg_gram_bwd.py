import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cmf_b200 import kernels as K
torch.set_printoptions(linewidth=250, precision=3, sci_mode=False)
B, d, n = 1, 32, 128
V = (torch.arange(d).view(1, d, 1) * 1000 + torch.arange(n).view(1, 1, n)).float().cuda()
gG = (0.5 * torch.eye(d)).view(1, d, d).cuda().contiguous()
out = K.gram_bwd(V, gG, path="tc")
ref = K.gram_bwd(V, gG, path="fma")
print("ref[0,:4,:8]\n", ref[0, :4, :8]); print("out[0,:4,:8]\n", out[0, :4, :8])
print("out rows 0..3, cols 32..40\n", out[0, :4, 32:40])
print("max abs out", out.abs().max().item(), "nonzero frac", (out != 0).float().mean().item())
print("err", (out - ref).abs().max().item())
# S = single off-diagonal entry
gG = torch.zeros(1, d, d).cuda(); gG[0, 3, 5] = 1.0
out = K.gram_bwd(V, gG, path="tc"); ref = K.gram_bwd(V, gG, path="fma")
print("offdiag: nonzero rows out", out.abs().sum(2).nonzero().flatten().tolist(), "ref", ref.abs().sum(2).nonzero().flatten().tolist())
print(out[0, 3, :6], ref[0, 3, :6]); print(out[0, 5, :6], ref[0, 5, :6])
