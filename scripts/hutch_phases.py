#!/usr/bin/env python
"""Where the MNIST-shape hutch_with_cg step spends its time (synchronised wall-clock per phase)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cmf_b200.config import get_config, get_schema
from cmf_b200.factory import get_density
from cmf_b200.non_square_helpers import get_non_square_train_metrics
import cmf_b200.head as H
import cmf_b200.cg as CG
torch.backends.cuda.matmul.allow_tf32 = False; torch.backends.cudnn.allow_tf32 = False; torch.backends.cudnn.benchmark = True
cfg = {**get_config("mnist"), "latent_dimension": 64, "g_ij_loss": False, "log_jacobian_method": "hutch_with_cg", "hutchinson_samples": 1,
       "likelihood_warmup": False}
torch.manual_seed(0)
density = get_density(get_schema(cfg), (1, 28, 28)).cuda().train()
tm = get_non_square_train_metrics(cfg)[0]
x = (torch.randint(0, 256, (50, 1, 28, 28)).float() + torch.rand(50, 1, 28, 28)).cuda()
inner = density.module.density
T = {}
def timed(name, fn):
    def w(*a, **k):
        if torch.cuda.is_current_stream_capturing():
            return fn(*a, **k)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        r = fn(*a, **k)
        torch.cuda.synchronize(); T[name] = T.get(name, 0.0) + time.perf_counter() - t0
        return r
    return w
orig_cg = H.linear_cg_batched
H.linear_cg_batched = timed("cg solve", orig_cg)
head = inner
while type(head).__name__ != "FusedNonSquareHeadDensity":
    head = head.prior
orig_jtj = head._jac_transpose_jac_vec
def jtj(lat, vec, create_graph):
    return timed("final J^T J product (create_graph)" if create_graph else "eager products (inside cg / capture)", orig_jtj)(lat, vec, create_graph)
head._jac_transpose_jac_vec = jtj
for it in range(4):
    T.clear()
    density.zero_grad(set_to_none=False)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    loss = tm(inner, x, 0)["loss"]
    torch.cuda.synchronize(); t1 = time.perf_counter()
    loss.backward()
    torch.cuda.synchronize(); t2 = time.perf_counter()
    print(f"iter {it}: forward {1e3 * (t1 - t0):.1f} ms (", ", ".join(f"{k} {1e3 * v:.1f}" for k, v in T.items()), f") backward {1e3 * (t2 - t1):.1f} ms; cg replays so far {head.cg_graph_replays}")
