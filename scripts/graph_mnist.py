#!/usr/bin/env python
"""MNIST-shape step eager vs one CUDA graph (cmf_b200.graphed.GraphedTrainStep)."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.allow_tf32 = False
torch.backends.cudnn.benchmark = True
from cmf_b200.config import get_config, get_schema
from cmf_b200.factory import get_density
from cmf_b200.graphed import GraphedTrainStep
from cmf_b200.non_square_helpers import get_non_square_train_metrics
from cmf_b200.parallel import GradientBucket
cfg = {**get_config("mnist"), "g_ij_loss": True, "latent_dimension": 64, "log_jacobian_method": "cholesky", "likelihood_warmup": False}
torch.manual_seed(0)
density = get_density(get_schema(cfg), (1, 28, 28)).cuda().train()
bucket = GradientBucket(density.parameters())
tm = get_non_square_train_metrics(cfg)[0]
inner = density.module.density
x = (torch.randint(0, 256, (50, 1, 28, 28)).float() + torch.rand(50, 1, 28, 28)).cuda()
def eager():
    bucket.zero(); tm(inner, x, 0)["loss"].backward()
def timed(fn, n):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
for _ in range(2): eager()
e = timed(eager, 4)
ref = bucket.flat.clone()
step = GraphedTrainStep(inner, tm, bucket, warmup_iters=1)
step(x, 0)
err = float((bucket.flat - ref).abs().max() / ref.abs().max())
g = timed(lambda: step(x, 0), 4)
print(json.dumps({"eager_ms": e, "graph_ms": g, "grad_rel_diff": err, "fallbacks": step.eager_fallbacks, "peak_gib": torch.cuda.max_memory_allocated() / 2**30}))
