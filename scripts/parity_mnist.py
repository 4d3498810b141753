"""grad-norm deviations of the MNIST-shape golden case vs the fp64 oracle probes and vs the fp32 reference"""
import sys, os
sys.path.insert(0, "tests"); sys.path.insert(0, "oracle")
import torch
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
from conftest import load_golden
from helpers import build_product, run_loss
from cases import grad_probe
g = load_golden(sys.argv[1] if len(sys.argv) > 1 else "mnist_gij_d64")
density = build_product(g, "cuda")
loss, grads = run_loss(g, density, "cuda")
print("CMF_CONV", os.environ.get("CMF_CONV", "tc"), "loss", float(loss), "golden", float(g["loss"]), "rel", abs(float(loss) - float(g["loss"])) / abs(float(g["loss"])))
w32 = w64 = 0.0
for k, probe in g["grads"].items():
    mine = grad_probe(k, grads[k])
    e32 = abs(mine["norm"] - probe["norm"]) / max(probe["norm"], 1e-30)
    e64 = abs(mine["norm"] - g["grads_fp64"][k]["norm"]) / max(g["grads_fp64"][k]["norm"], 1e-30) if "grads_fp64" in g else float("nan")
    w32, w64 = max(w32, e32), max(w64, e64)
print("worst grad-norm rel dev vs fp32 reference %.3e   vs fp64 oracle %.3e" % (w32, w64))
