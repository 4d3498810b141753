#!/usr/bin/env python
"""Timing of the tcgen05 convolution launches of the benchmarked step in isolation (N = 3200 tangent images)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from cmf_b200 import kernels as K
torch.manual_seed(0)
out = {}
for H in (28, 14):
    N = 3200
    x = torch.randn(N, 64, H, H, device="cuda")
    w = torch.randn(64, 64, 3, 3, device="cuda") / 24
    ws = K.conv3x3_c64_prep(w, False)
    prim = torch.randn(N // 64, 64, H, H, device="cuda")
    bits = K.relu_mask_pack(prim)
    res = torch.randn_like(x)
    y, y2 = torch.empty_like(x), torch.empty_like(x)
    def timed(fn, n=6):
        fn(); fn(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(n):
            fn()
        b.record(); torch.cuda.synchronize()
        return a.elapsed_time(b) / n
    fl = 2.0 * N * 64 * 64 * 9 * H * H
    for name, fn in [("plain", lambda: K.conv3x3_c64(x, ws, None, out=y)),
                     ("mask", lambda: K.conv3x3_c64(x, ws, None, mask_src=bits, mask_div=64, mask_first=True, out=y)),
                     ("mask+res+second", lambda: K.conv3x3_c64(x, ws, None, mask_src=bits, mask_div=64, res=res, second=True, out=y, out2=y2)),
                     ("mask drain2", lambda: K.conv3x3_c64(x, ws, None, mask_src=bits, mask_div=64, mask_first=True, drain=2, out=y)),
                     ("mask+res drain2", lambda: K.conv3x3_c64(x, ws, None, mask_src=bits, mask_div=64, mask_first=True, res=res, drain=2, out=y))]:
        ms = timed(fn)
        out[f"{H}x{H} {name}"] = (round(ms, 4), round(fl / ms / 1e9, 1))
    gy = torch.randn_like(x)
    ms = timed(lambda: K.conv3x3_c64_wgrad(x, gy))
    out[f"{H}x{H} wgrad"] = (round(ms, 4), round(fl / ms / 1e9, 1))
    del x, res, y, y2, gy
print(json.dumps({k: v for k, v in out.items() if "wgrad" not in k}))
