"""Cholesky log-det forward / adjoint timings (CUDA events, us per launch) and the adjoint's error against fp64.
CMF_CHOL_BWD_BLOCKED=0 selects the substitution form of the adjoint (read once per process)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cmf_b200 import kernels as K

torch.manual_seed(0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
tag = "substitution" if os.environ.get("CMF_CHOL_BWD_BLOCKED", "1")[:1] == "0" else "blocked"
for B, d in ((512, 32), (512, 64), (50, 64), (2048, 64), (512, 128), (64, 128)):
    A = torch.randn(B, d, 2 * d, device="cuda")
    G = A @ A.transpose(1, 2) / (2 * d) + 1e-3 * torch.eye(d, device="cuda")
    gl, g1 = torch.randn(B, device="cuda"), torch.randn(B, device="cuda")
    L, logdet, l1o, l1d, info = K.chol_logdet_fwd(G)
    gG = K.chol_logdet_bwd(L, G, gl, g1, "offdiag")
    ref = gl.double().view(B, 1, 1) * torch.linalg.inv(G.double())
    off = 1.0 - torch.eye(d, device="cuda", dtype=torch.float64)
    ref = ref + g1.double().view(B, 1, 1) * torch.sign(G.double()) * off
    err = float(((gG.double() - ref).flatten(1).norm(dim=1) / ref.flatten(1).norm(dim=1)).max())
    # 10 launches per CUDA-graph replay (the Python wrapper costs as much as the kernel at these sizes)
    def graphed(fn):
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            for _ in range(3): fn()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(10): fn()
        ts = []
        for _ in range(9):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); g.replay(); e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 100.0)
        ts.sort()
        return ts
    tf = graphed(lambda: K.chol_logdet_fwd(G))
    tb = graphed(lambda: K.chol_logdet_bwd(L, G, gl, g1, "offdiag"))
    print(f"{tag:12s} B={B:5d} d={d:4d}  fwd {tf[len(tf)//2]:7.1f} us  bwd {tb[len(tb)//2]:7.1f} us  bwd rel-F err vs fp64 {err:.2e}")
