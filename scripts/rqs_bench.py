#!/usr/bin/env python
"""Spline coupling kernel timings (B = 512, d = 64, M = 392, K = 8) against the byte model of DESIGN 3b."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from cmf_b200 import kernels as K, splines
B, d, M, Kb, tb = 512, 64, 392, 8, 3.0
g = torch.Generator(device="cuda").manual_seed(0)
z = (torch.rand(B, M, device="cuda", generator=g) * 2 - 1) * tb * 1.1
V = torch.randn(B, d, M, device="cuda", generator=g)
p = torch.randn(B, M, 3 * Kb - 1, device="cuda", generator=g)
P = torch.randn(B, d, M, 3 * Kb - 1, device="cuda", generator=g)
cw, ch, dv, CW, CH, DV = splines.knot_tables(p, P, Kb, tb, hidden_features=32)
del p, P
gx, gV = torch.randn_like(z), torch.randn_like(V)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def timed(fn, n=10):
    fn(); torch.cuda.synchronize()
    tot = 0.0
    for _ in range(n):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        tot += a.elapsed_time(b)
    return tot / n
f_ms = timed(lambda: K.rqs_coupling_jvp_fwd(z, V, cw, ch, dv, CW, CH, DV, tb))
b_ms = timed(lambda: K.rqs_coupling_jvp_bwd(z, V, cw, ch, dv, CW, CH, DV, tb, gx, gV))
tab = 3 * (Kb + 1)
f_bytes = 4 * B * M * (1 + d) * (2 + tab)                 # read z/V + tables, write x/V'
b_bytes = 4 * B * M * (1 + d) * (3 + 2 * tab)             # read z/V/g + tables, write g_z/g_V + dense table adjoints
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
print(json.dumps({"shape": [B, d, M, Kb], "fwd_ms": f_ms, "bwd_ms": b_ms, "fwd_gbs": f_bytes / f_ms / 1e6, "bwd_gbs": b_bytes / b_ms / 1e6,
                  "fwd_frac_hbm": f_bytes / f_ms / 1e6 / peak, "bwd_frac_hbm": b_bytes / b_ms / 1e6 / peak}))
