"""Where the tcgen05 Gram kernels spend their time: timings over batch sizes (waves of 148 CTAs) and column counts
(stages per CTA).  Each figure is a CUDA-graph replay of 12 launches rotating over three operand sets (no launch finds
its input in L2, no host launch latency inside the events), divided by 12."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cmf_b200 import kernels as K

NL = 12


def timed(fn, sets, reps=10):
    for s in sets:
        fn(*s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(NL):
            fn(*sets[i % len(sets)])
    g.replay(); torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in ev:
        a.record(); g.replay(); b.record()
    torch.cuda.synchronize()
    t = sorted(a.elapsed_time(b) for a, b in ev)
    return 1e3 * t[len(t) // 2] / NL


d = int(os.environ.get("D", 64))
S = 128 // d if d <= 128 else 1
print(f"d = {d}")
if __name__ == "__main__":
    import hashlib
    torch.manual_seed(0)
    Vh = torch.randn(37, d, 784, device="cuda") * torch.rand(37, d, 1, device="cuda") * 3
    print("G hash", hashlib.sha1(K.gram_fwd(Vh, path="tc").cpu().numpy().tobytes()).hexdigest()[:16])
which = os.environ.get("WHICH", "fwd,bwd") if __name__ == "__main__" else ""
if "fwd" in which:
    for B, n in [(148 * S, 392), (148 * S, 784), (148 * S, 1568), (148 * S, 3136), (512, 784), (2 * 148 * S, 784), (512, 392), (2048, 784)]:
        sets = [(torch.randn(B, d, n, device="cuda"), torch.empty(B, d, d, device="cuda")) for _ in range(3)]
        us = timed(lambda V, G: K.gram_fwd(V, path="tc"), sets)
        mb = B * d * n * 4 / 1e6
        print(f"fwd B={B:5d} n={n:5d}: {us:8.1f} us  {mb / us / 1e3:7.3f} TB/s read   ctas={-(-B // S)}")
if "bwd" in which:
    for B, n in [(512, 392), (512, 384), (512, 512), (512, 128), (148, 128), (148, 392), (512, 784), (2048, 392)]:
        sets = [(torch.randn(B, d, n, device="cuda"), torch.randn(B, d, d, device="cuda")) for _ in range(3)]
        us = timed(lambda V, g: K.gram_bwd(V, g, path=os.environ.get("BWD_PATH", "tc")), sets)
        mb = (2 * B * d * n + B * d * d) * 4 / 1e6
        print(f"bwd B={B:5d} n={n:5d}: {us:8.1f} us  {mb / us / 1e3:7.3f} TB/s r+w    ctas={B * -(-n // 128)}")
