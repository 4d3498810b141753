"""Augment tests/golden/*.pt with an fp64 evaluation of the same algorithm by the CPU oracle
(``loss_fp64``, ``grads_fp64``).  TEST INFRASTRUCTURE ONLY.

Why: for ill-conditioned metrics (MNIST-shape G has cond ~ 5e3..3e4) the gradient of log det G
carries O(cond * eps_fp32) rounding noise in ANY fp32 implementation, the reference's included.
The fp64 value lets the parity test measure the reference's own fp32 error and require the CUDA
path to be within 1e-4 relative OR no worse than a small multiple of that error.

    python oracle/add_fp64_probes.py [case ...]
"""
import os
import sys
import time

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)

import cmf_oracle as O
from cases import CASES, make_input, grad_probe
from detfill import fill_state_dict


def main(argv):
    gdir = os.path.join(os.path.dirname(HERE), "tests", "golden")
    for name in argv or list(CASES):
        path = os.path.join(gdir, name + ".pt")
        g = torch.load(path, weights_only=False)
        t0 = time.time()
        params = fill_state_dict(O.init_params(g["schema"], g["case"]["x_shape"]))
        params = {k: (v.double().requires_grad_(True) if v.is_floating_point() else v) for k, v in params.items()}
        flow = O.OracleFlow(g["schema"], params, g["case"]["x_shape"])
        kw = {"hutchinson_eps": g["hutch_eps"].double()} if "hutch_eps" in g else {}
        loss = O.train_loss(flow, g["config"], make_input(CASES[name]).double(), g["case"]["epoch"], **kw)
        loss.backward()
        g["loss_fp64"] = loss.detach().clone()
        g["grads_fp64"] = {k: grad_probe(k, v.grad) for k, v in params.items()
                           if v.is_floating_point() and v.grad is not None}
        torch.save(g, path)
        worst = max(abs(g["grads"][k]["norm"] - p["norm"]) / max(p["norm"], 1e-30)
                    for k, p in g["grads_fp64"].items() if k in g["grads"])
        print(f"{name}: loss fp64 {float(loss):.6f} vs reference fp32 {float(g['loss']):.6f}; "
              f"worst reference grad-norm error vs fp64 {worst:.2e}  ({time.time() - t0:.1f}s)", flush=True)


if __name__ == "__main__":
    main(sys.argv[1:])
