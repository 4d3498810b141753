#!/usr/bin/env python
"""Timing of the one-launch MLP conditioner at the tabular step's shapes against the F.linear stack."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from cmf_b200 import networks
torch.backends.cuda.matmul.allow_tf32 = False
dev = "cuda"
for B, d, nin, nout in ((5000, 6, 3, 6), (750, 10, 11, 20), (5000, 0, 3, 6)):
    net = networks.TangentMLP(nin, [128] * 4, nout, "tanh").to(dev)
    x = torch.randn(B, nin, device=dev, requires_grad=True)
    V = torch.randn(B, d, nin, device=dev, requires_grad=True) if d else None
    gy = torch.randn(B, nout, device=dev)
    gY = torch.randn(B, d, nout, device=dev) if d else None
    for fused in (True, False):
        networks._MLP_FUSED = fused
        def fwd():
            if d:
                return net.jvp_batched(x, V)
            return net(x), None
        def both():
            y, Y = fwd()
            l = (y * gy).sum() if not d else (y * gy).sum() + (Y * gY).sum()
            l.backward()
        def timed(fn, n=20):
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(n):
                fn()
            b.record(); torch.cuda.synchronize()
            return a.elapsed_time(b) / n * 1e3
        with torch.no_grad():
            t_inf = timed(fwd)
        t_all = timed(both)
        flops = 2.0 * B * (1 + d) * (nin * 128 + 3 * 128 * 128 + 128 * nout)
        print(f"B={B} d={d} fused={fused}: forward (no grad) {t_inf:8.1f} us ({flops / t_inf / 1e6:6.1f} TFLOP/s)   fwd+bwd {t_all:8.1f} us ({3 * flops / t_all / 1e6:6.1f} TFLOP/s)")
networks._MLP_FUSED = True

# per-kernel device times (kineto), fused path, power shape
from torch.profiler import profile, ProfilerActivity
for B, d, nin, nout in ((5000, 6, 3, 6), (750, 10, 11, 20)):
    net = networks.TangentMLP(nin, [128] * 4, nout, "tanh").to(dev)
    x = torch.randn(B, nin, device=dev, requires_grad=True)
    V = torch.randn(B, d, nin, device=dev, requires_grad=True)
    def both():
        y, Y = net.jvp_batched(x, V)
        (y.sum() + Y.sum()).backward()
    for _ in range(3):
        both()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(5):
            both()
        torch.cuda.synchronize()
    print("== B", B, "d", d)
    for e in sorted(prof.key_averages(), key=lambda e: -e.device_time_total)[:6]:
        print(f"   {e.device_time_total / e.count:9.1f} us x {e.count:3d}  {e.key[:90]}")
