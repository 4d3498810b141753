#!/usr/bin/env python
"""Kernel-time table (kineto) of one eager step of the tabular shapes (power D=6 d=6 B=5000, hepmass D=21 d=10 B=750)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from torch.profiler import profile, ProfilerActivity
from cmf_b200.config import get_config, get_schema
from cmf_b200.factory import get_density
from cmf_b200.non_square_helpers import get_non_square_train_metrics
torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.allow_tf32 = False
dev = torch.device("cuda")
for ds, D, B, d in (("power", 6, 5000, 6), ("hepmass", 21, 750, 10)):
    cfg = {**get_config(ds), "g_ij_loss": True, "latent_dimension": d, "log_jacobian_method": "cholesky",
           "likelihood_warmup": False}
    torch.manual_seed(0)
    density = get_density(get_schema(cfg), (D,)).to(dev).train()
    tm = get_non_square_train_metrics(cfg)[0]
    x = torch.randn(B, D, device=dev)
    def step():
        density.zero_grad(set_to_none=False)
        tm(density.module, x, 0)["loss"].backward()
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        step()
        torch.cuda.synchronize()
    print("==", ds, "B", B, "d", d)
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=28, max_name_column_width=90))
