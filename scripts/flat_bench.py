#!/usr/bin/env python
"""Whole-stack kernel (cmf_flat_stack_fwd/bwd) timings on the 2-D config models vs the per-layer program."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
torch.backends.cuda.matmul.allow_tf32 = False
from cmf_b200 import kernels as K
from cmf_b200.config import get_config, get_schema
from cmf_b200.factory import get_density
import cmf_b200.head as H

def timed(fn, n=20):
    fn(); fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3

for ds, x_shape, d in (("sphere", (3,), 3), ("sinusoid-1-6", (6,), 6)):
    cfg = {**get_config(ds), "g_ij_loss": True, "latent_dimension": d, "log_jacobian_method": "cholesky"}
    torch.manual_seed(0)
    density = get_density(get_schema(cfg), x_shape).cuda().train()
    head = density.module if hasattr(density, "module") else density
    plan = head.program.flat_plan
    out = {"config": ds, "P": plan.P}
    for B in (1000, 8000):
        z = torch.randn(B, d, device="cuda")
        params = plan.params().detach()
        gG, gx = torch.randn(B, d, d, device="cuda"), torch.randn(B, x_shape[0], device="cuda")
        out[f"B={B} fwd_us"] = round(timed(lambda: K.flat_stack_fwd(plan.desc, z, params)), 1)
        out[f"B={B} bwd_us"] = round(timed(lambda: K.flat_stack_bwd(plan.desc, z, params, gG, gx)), 1)
    print(json.dumps(out))
