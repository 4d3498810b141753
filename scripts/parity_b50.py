#!/usr/bin/env python
"""Diagnostic: per-parameter gradient deviation of the B = 50 MNIST-shape step (8 selected samples)
from the fp64 oracle on the device, for the product and for the fp32 oracle.  Prints the worst
parameters by relative L2 error."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch
torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.allow_tf32 = False
from conftest import load_golden
from cases import CASES, make_input
from detfill import fill_state_dict
from helpers import build_product, strip_wrappers
import cmf_oracle as O
import test_parity_gpu as T

g = load_golden("mnist_gij_d64")
case = dict(CASES["mnist_gij_d64"], batch=50)
density = build_product(g, "cpu")
filled = fill_state_dict(density.state_dict())
density = density.cuda().train()
x = make_input(case)
idx = torch.tensor([0, 7, 13, 24, 25, 31, 42, 49])
top = strip_wrappers(density)
elbo = top.elbo(x.cuda(), **O.elbo_kwargs(g["config"], case["epoch"]))["elbo"]
(-elbo[idx.cuda()].mean()).backward()
_, p64, _, _ = T._oracle_on_device(g, filled, x[idx], case["epoch"], chunk=4)
_, p32, _, _ = T._oracle_on_device(g, filled, x[idx], case["epoch"], chunk=8, dtype=torch.float32)
rows = []
for k, got in T._named_grads(density).items():
    ref = p64[k].grad
    if ref is None:
        continue
    n = float(ref.norm())
    e_prod = float((got.double() - ref).norm()) / n
    e_32 = float((p32[k].grad.double() - ref).norm()) / n
    rows.append((e_prod, e_32, k, tuple(ref.shape), n))
rows.sort(reverse=True)
print("env CMF_CONV_SEG =", os.environ.get("CMF_CONV_SEG"))
print("worst by product rel-L2 error:   product   fp32-oracle   param")
for e_prod, e_32, k, shp, n in rows[:12]:
    print(f"   {e_prod:9.2e} {e_32:9.2e}  {k} {shp} |g|={n:.3e}")
rows.sort(key=lambda r: -r[1])
print("worst by fp32-oracle rel-L2 error:")
for e_prod, e_32, k, shp, n in rows[:6]:
    print(f"   {e_prod:9.2e} {e_32:9.2e}  {k} {shp} |g|={n:.3e}")
import statistics
print("median product", statistics.median(r[0] for r in rows), "median fp32 oracle", statistics.median(r[1] for r in rows))
tot64 = torch.cat([p64[r[2]].grad.flatten() for r in rows]); 
totp = torch.cat([T._named_grads(density)[r[2]].double().flatten() for r in rows]); tot32 = torch.cat([p32[r[2]].grad.double().flatten() for r in rows])
print("whole-gradient rel-L2: product", float((totp - tot64).norm() / tot64.norm()), "fp32 oracle", float((tot32 - tot64).norm() / tot64.norm()))
