#!/usr/bin/env python
"""MNIST-shape hutch_with_cg step as ONE CUDA graph (explicit-G CG path, sync-free CG under capture) vs eager."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cmf_b200.config import get_config, get_schema
from cmf_b200.factory import get_density
from cmf_b200.graphed import GraphedTrainStep
from cmf_b200.non_square_helpers import get_non_square_train_metrics
from cmf_b200.parallel import GradientBucket
torch.backends.cuda.matmul.allow_tf32 = False; torch.backends.cudnn.allow_tf32 = False; torch.backends.cudnn.benchmark = True
cfg = {**get_config("mnist"), "latent_dimension": 64, "g_ij_loss": False, "log_jacobian_method": "hutch_with_cg", "hutchinson_samples": 1,
       "likelihood_warmup": False}
torch.manual_seed(0)
density = get_density(get_schema(cfg), (1, 28, 28)).cuda().train()
tm = get_non_square_train_metrics(cfg)[0]
x = (torch.randint(0, 256, (50, 1, 28, 28)).float() + torch.rand(50, 1, 28, 28)).cuda()
inner = density.module.density
bucket = GradientBucket(density.parameters())
def timed(fn, n=5):
    for _ in range(2): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
def eager():
    bucket.zero(); tm(inner, x, 0)["loss"].backward()
print("eager ms", timed(eager))
g = GraphedTrainStep(inner, tm, bucket, warmup_iters=1)
losses = [float(g(x, 0)) for _ in range(4)]
print("graph losses (fresh probes per replay):", losses, "fallbacks", g.eager_fallbacks)
print("graph ms", timed(lambda: g(x, 0)))
print("randn after capture ok:", float(torch.randn(3, device="cuda").sum()))
