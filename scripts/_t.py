import torch, torch.nn.functional as F
from cmf_b200 import kernels as K
torch.backends.cudnn.allow_tf32 = False
g = torch.Generator().manual_seed(7)
x = torch.randn(4, 64, 28, 28, generator=g).abs().cuda()
g = torch.Generator().manual_seed(8)
w = (torch.randn(64, 64, 3, 3, generator=g) / 24).abs().cuda()
y = K.conv3x3_c64(x, K.conv3x3_c64_prep(w, False), None)
ref = F.conv2d(x.double(), w.double(), padding=1)
y32 = F.conv2d(x, w, padding=1)
print("tc   max rel", ((y.double() - ref) / ref).abs().max().item(), "mean signed", ((y.double() - ref) / ref).mean().item())
print("cudnn max rel", ((y32.double() - ref) / ref).abs().max().item(), "mean signed", ((y32.double() - ref) / ref).mean().item())
