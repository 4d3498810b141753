#!/usr/bin/env python
"""Kernel micro-benchmarks through the C ABI (CUDA events, L2 flushed between iterations).

    python scripts/kbench.py gram|coupling|chol|all
Prints one JSON line per case: time, algorithmic GB/s and TFLOP/s, fraction of the measured peaks."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from cmf_b200 import kernels as K
K_ = K

DEV = "cuda"
PEAKS = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}
_flush = None


def timeit(fn, iters=10, warmup=3):
    global _flush
    if _flush is None:
        _flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
    for _ in range(warmup):
        fn()
    ts = []
    for _ in range(iters):
        _flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2] * 1e-3


def report(name, t, nbytes, flops=0.0, **kw):
    gbs, tf = nbytes / t / 1e9, flops / t / 1e12
    print(json.dumps(dict(case=name, us=round(t * 1e6, 2), gbs=round(gbs, 1), frac_hbm=round(gbs / PEAKS["hbm_gbs"], 3),
                          tflops=round(tf, 2), **kw)), flush=True)


def bench_gram():
    for (B, d, n) in [(512, 64, 392), (50, 64, 392), (256, 128, 1536), (64, 256, 1536), (32, 512, 1536), (64, 512, 3072)]:
        V = torch.randn(B, d, n, device=DEV)
        gG = torch.randn(B, d, d, device=DEV)
        ref = torch.bmm(V.double(), V.double().transpose(1, 2))
        for path in ("fma", "tc"):
            G = K.gram_fwd(V, path=path)
            err = float(((G.double() - ref).abs().amax() / ref.abs().amax()))
            t = timeit(lambda: K.gram_fwd(V, path=path))
            report(f"gram_fwd[{path}] B={B} d={d} n={n}", t, 4.0 * B * d * n + 4.0 * B * d * d, 2.0 * B * d * d * n, rel_err=err)
        refb = torch.bmm(gG.double() + gG.double().transpose(1, 2), V.double())
        for path in ("fma", "tc"):
            err = float((K.gram_bwd(V, gG, path=path).double() - refb).abs().amax() / refb.abs().amax())
            t = timeit(lambda: K.gram_bwd(V, gG, path=path))
            report(f"gram_bwd[{path}] B={B} d={d} n={n}", t, 8.0 * B * d * n, 2.0 * B * d * d * n, rel_err=err)
        tb = timeit(lambda: torch.bmm(V, V.transpose(1, 2)))
        report(f"torch.bmm fp32 B={B} d={d} n={n}", tb, 4.0 * B * d * n, 2.0 * B * d * d * n)


def bench_coupling():
    for (B, d, M, Q) in [(512, 64, 392, 392), (512, 64, 392, 784), (512, 64, 196, 392), (50, 64, 392, 392), (64, 512, 1536, 1536)]:
        z, V = torch.randn(B, M, device=DEV), torch.randn(B, d, M, device=DEV)
        st, ST = torch.randn(B, 2 * Q, device=DEV) * 0.3, torch.randn(B, d, 2 * Q, device=DEV)
        gx, gV = torch.randn(B, M, device=DEV), torch.randn(B, d, M, device=DEV)
        map_ = inv = None
        if Q != M:
            q = torch.randperm(Q)[:M].sort().values
            inv = torch.full((Q,), -1, dtype=torch.int32); inv[q] = torch.arange(M, dtype=torch.int32)
            map_, inv = q.int().to(DEV), inv.to(DEV)
        t = timeit(lambda: K.coupling_jvp_fwd(z, V, st, ST, map_))
        report(f"coupling_fwd B={B} d={d} M={M} Q={Q}", t, 16.0 * M * (1 + d) * B)
        t = timeit(lambda: K.coupling_jvp_bwd(z, V, st, ST, map_, inv, gx, gV))
        report(f"coupling_bwd B={B} d={d} M={M} Q={Q}", t, 24.0 * M * (1 + d) * B)


def bench_chol():
    for (B, d) in [(512, 64), (50, 64), (256, 128), (64, 256), (64, 512), (8, 512), (5000, 6), (1000, 3)]:
        A = torch.randn(B, d, 2 * d, device=DEV)
        G = torch.bmm(A, A.transpose(1, 2)) + 0.1 * torch.eye(d, device=DEV)
        t = timeit(lambda: K.chol_logdet_fwd(G))
        report(f"chol_fwd B={B} d={d}", t, 8.0 * d * d * B, d ** 3 / 3.0 * B)
        L, logdet, l1o, l1d, info = K.chol_logdet_fwd(G)
        t = timeit(lambda: K.chol_logdet_bwd(L, G, logdet, l1o, "offdiag"))
        report(f"chol_bwd B={B} d={d}", t, 12.0 * d * d * B, d ** 3 * B)
        t = timeit(lambda: torch.linalg.cholesky(G))
        report(f"torch.linalg.cholesky B={B} d={d}", t, 8.0 * d * d * B, d ** 3 / 3.0 * B)


def bench_rqs():
    from cmf_b200 import splines
    for (B, d, M, K, tb) in [(512, 64, 392, 8, 3.0), (64, 16, 392, 8, 3.0), (5000, 6, 3, 8, 3.0)]:
        z = (torch.rand(B, M, device=DEV) * 2 - 1) * tb * 1.1
        p = torch.randn(B, M, 3 * K - 1, device=DEV)
        V = torch.randn(B, d, M, device=DEV)
        P = torch.randn(B, d, M, 3 * K - 1, device=DEV)
        cw, ch, dv, CW, CH, DV = [t.contiguous() for t in splines.knot_tables(p, P, K, tb, hidden_features=32)]
        gx, gV = torch.randn(B, M, device=DEV), torch.randn(B, d, M, device=DEV)
        tab = 3 * (K + 1) * 4.0
        t = timeit(lambda: K_.rqs_coupling_jvp_fwd(z, V, cw, ch, dv, CW, CH, DV, tb))
        report(f"rqs_fwd B={B} d={d} M={M} K={K}", t, B * M * ((4 + tab + 8) + d * (4 + tab + 4)))
        t = timeit(lambda: K_.rqs_coupling_jvp_bwd(z, V, cw, ch, dv, CW, CH, DV, tb, gx, gV))
        report(f"rqs_bwd B={B} d={d} M={M} K={K}", t, B * M * ((8 + tab + 4 + tab) + d * (8 + tab + 4 + tab)))


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    torch.backends.cuda.matmul.allow_tf32 = False
    if what in ("gram", "all"): bench_gram()
    if what in ("coupling", "all"): bench_coupling()
    if what in ("chol", "all"): bench_chol()
    if what in ("rqs", "all"): bench_rqs()
