import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch
from cmf_b200 import kernels as K
from kbench import timeit
for (B, d, n) in [(512, 64, 392), (64, 512, 3072)]:
    V = torch.randn(B, d, n, device="cuda")
    t = timeit(lambda: K.gram_fwd(V, path="tc"))
    print(os.environ.get("CMF_DBG", "0"), (B, d, n), round(t * 1e6, 1), "us")
