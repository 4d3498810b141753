"""Golden vectors for the spline oracle from the UNMODIFIED nsf submodule of the reference
(build container only).  TEST INFRASTRUCTURE ONLY.

    PYTHONDONTWRITEBYTECODE=1 python oracle/gen_rqs_golden.py
"""
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.dont_write_bytecode = True
REF = os.environ.get("CMF_REFERENCE", "/root/reference")
sys.path.insert(0, os.path.join(REF, "gitmodules", "nsf"))
from nde.transforms.splines.rational_quadratic import unconstrained_rational_quadratic_spline  # noqa: E402
import utils as nsf_utils  # noqa: E402
from detfill import det_tensor  # noqa: E402


def main():
    out = {}
    for tag, (N, K, tb, scale) in {"k8": (257, 8, 3.0, 2.0), "k10": (64, 10, 1.0, 1.5), "k31": (33, 31, 4.0, 1.0),
                                   "k1": (16, 1, 2.0, 1.0)}.items():
        x = det_tensor("rqs/x/" + tag, (N,), 1.0, salt=1) * (tb * 1.3)
        x[0], x[1], x[2] = tb, -tb, 0.0                       # interval ends are inside
        p = det_tensor("rqs/p/" + tag, (N, 3 * K - 1), 1.0, salt=2) * scale
        uw, uh, ud = p[..., :K].clone(), p[..., K:2 * K].clone(), p[..., 2 * K:].clone()
        res = {}
        for inverse in (False, True):
            y, lad = unconstrained_rational_quadratic_spline(x.clone(), uw.clone(), uh.clone(), ud.clone(),
                                                              inverse=inverse, tails="linear", tail_bound=tb)
            res["inv" if inverse else "fwd"] = dict(y=y, logabsdet=lad)
        # the searchsorted semantics on a fixed knot vector (utils/torchutils.py:133-138)
        kn = torch.sort(det_tensor("rqs/knots/" + tag, (N, K + 1), 1.0, salt=3), dim=-1).values
        res["search"] = dict(knots=kn.clone(), idx=nsf_utils.searchsorted(kn.clone(), x / (tb * 1.3)))
        out[tag] = dict(N=N, K=K, tail_bound=tb, x=x, params=p, **res)
    path = os.path.join(os.path.dirname(HERE), "tests", "golden", "rqs_nsf.pt")
    torch.save(out, path)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
