"""standalone launches of the coupling kernels at the scope-A shape (B = 512, d = 64, M = 392; for ncu)"""
import torch
from cmf_b200 import kernels as K
B, d, M = 512, 64, 392
g = torch.Generator(device="cuda").manual_seed(0)
r = lambda *s: torch.rand(*s, device="cuda", generator=g) * 2 - 1
z, V, st, ST, gx, gV = r(B, M), r(B, d, M), r(B, 2 * M) * 0.5, r(B, d, 2 * M), r(B, M), r(B, d, M)
for _ in range(3):
    K.coupling_jvp_fwd(z, V, st, ST, None)
    K.coupling_jvp_bwd(z, V, st, ST, None, None, gx, gV)
torch.cuda.synchronize()
print("ok")
