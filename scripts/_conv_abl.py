import torch, sys
from cmf_b200 import kernels as K
N, H, W = 3250, 28, 28
x = torch.randn(N, 64, H, W, device="cuda")
w = torch.randn(64, 64, 3, 3, device="cuda") / 24
ws = K.conv3x3_c64_prep(w, False)
prim = torch.randn(N // 65, 64, H, W, device="cuda"); res = torch.randn_like(x)
for mode, kw in (("plain", {}), ("mask", dict(mask_src=prim, mask_div=65, mask_first=True)),
                 ("res+second", dict(mask_src=prim, mask_div=65, res=res, second=True)),
                 ("mask+res", dict(mask_src=prim, mask_div=65, mask_first=True, res=res))):
    for _ in range(3):
        K.conv3x3_c64(x, ws, None, **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        K.conv3x3_c64(x, ws, None, **kw)
    e1.record()
    torch.cuda.synchronize()
    print(f"{mode:12s} {e0.elapsed_time(e1) / 10:.3f} ms", flush=True)
