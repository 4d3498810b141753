#!/usr/bin/env python
"""Throughput of the flat BASELINE.json configs (sphere, sinusoid-1-6, power-shape, hepmass-shape)
on one GPU: eager step vs CUDA-graph replay of the same step.  One JSON line per config."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from cmf_b200.config import get_config, get_schema
from cmf_b200.factory import get_density
from cmf_b200.non_square_helpers import get_non_square_train_metrics
from cmf_b200.parallel import GradientBucket
from cmf_b200.graphed import GraphedTrainStep

CONFIGS = [("sphere", (3,), 1000, dict(g_ij_loss=True, latent_dimension=3)),
           ("sinusoid-1-6", (6,), 1000, dict(g_ij_loss=True, latent_dimension=6)),
           ("power", (6,), 5000, dict(g_ij_loss=True, latent_dimension=6)),
           ("hepmass", (21,), 750, dict(g_ij_loss=True, latent_dimension=10))]


def timed(fn, iters):
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); a.record()
    for _ in range(iters):
        fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / iters, (time.perf_counter() - t0) * 1e3 / iters


def main():
    torch.backends.cuda.matmul.allow_tf32 = False
    dev = "cuda"
    for ds, x_shape, B, ov in CONFIGS:
        cfg = {**get_config(ds), **ov, "log_jacobian_method": "cholesky", "likelihood_warmup": False}
        torch.manual_seed(0)
        density = get_density(get_schema(cfg), x_shape).to(dev).train()
        bucket = GradientBucket(density.parameters())
        tm = get_non_square_train_metrics(cfg)[0]
        x = torch.randn(B, *x_shape, device=dev)
        if ds == "sphere":
            x = x / x.norm(dim=1, keepdim=True)

        def eager():
            bucket.zero(); tm(density, x, 0)["loss"].backward()
        for _ in range(3):
            eager()
        ms_e, wall_e = timed(eager, 10)
        g_ref = bucket.flat.clone()
        step = GraphedTrainStep(density, tm, bucket)
        loss = step(x, 0)
        err = float((bucket.flat - g_ref).abs().max() / g_ref.abs().max())
        ms_g, wall_g = timed(lambda: step(x, 0), 50)
        print(json.dumps(dict(config=ds, D=x_shape[0], d=cfg["latent_dimension"], batch=B,
                              eager_ms=round(ms_e, 3), graph_ms=round(ms_g, 3),
                              eager_samples_per_s=round(B / ms_e * 1e3), graph_samples_per_s=round(B / ms_g * 1e3),
                              speedup=round(ms_e / ms_g, 2), grad_rel_diff=err, fallbacks=step.eager_fallbacks)), flush=True)


if __name__ == "__main__":
    main()
