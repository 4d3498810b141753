import torch, sys
from cmf_b200 import kernels as K
N, H, W = (int(a) for a in sys.argv[1:4])
x = torch.randn(N, 64, H, W, device="cuda")
w = torch.randn(64, 64, 3, 3, device="cuda") / 24
ws = K.conv3x3_c64_prep(w, False)
torch.cuda.synchronize()
print("prep ok", flush=True)
y = K.conv3x3_c64(x, ws, None)
torch.cuda.synchronize()
print("conv ok", float(y.abs().max()), flush=True)
