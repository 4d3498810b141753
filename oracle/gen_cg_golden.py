"""Generate ``tests/golden/cg_gpytorch.pt``: outputs of the reference's OWN conjugate-gradient
dependency -- ``gpytorch.utils.linear_cg`` vendored at /root/reference/gitmodules/gpytorch (fc2053b),
called exactly as `cmf/models/components/densities/non_square.py:241-247` calls it (no preconditioner,
``max_tridiag_iter = max_iter``) -- on small seeded SPD systems in fp64.  Build-container only (the
GPU box has no /root/reference).  TEST INFRASTRUCTURE ONLY.

    PYTHONDONTWRITEBYTECODE=1 python oracle/gen_cg_golden.py
"""
import os
import sys
import warnings

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
REF = os.environ.get("CMF_REFERENCE", "/root/reference")


def systems():
    """(name, A (B,n,n) SPD fp64, rhs (B,n,m), max_iter, tolerance): a zero right-hand side column,
    early exit (tolerance 1 after >= 10 iterations), a tight tolerance, an iteration cap.  Moderate
    condition numbers on purpose: CG amplifies rounding differences ~30x per iteration once it has
    converged (measured: two bit-different but algebraically identical implementations drift from
    1e-16 to 1e-4 over 16 iterations at cond 100), which would hide control-flow differences."""
    from detfill import det_tensor
    out = []
    for name, B, n, m, cond, max_iter, tol in [("well_n6", 3, 6, 1, 10.0, 6, 1.0), ("zero_rhs_n10_m2", 4, 10, 2, 30.0, 10, 1.0),
                                               ("n24_early_exit", 2, 24, 3, 50.0, 24, 1.0), ("n12_tight", 2, 12, 1, 20.0, 12, 1e-6),
                                               ("n12_capped", 3, 12, 2, 1e2, 5, 1.0)]:
        Q, _ = torch.linalg.qr(det_tensor("cg/q/" + name, (B, n, n), 1.0, dtype=torch.float64))
        ev = torch.logspace(0, torch.log10(torch.tensor(cond)).item(), n, dtype=torch.float64)
        A = Q @ torch.diag_embed(ev.expand(B, n)) @ Q.transpose(1, 2)
        rhs = det_tensor("cg/rhs/" + name, (B, n, m), 1.5, dtype=torch.float64)
        if name == "zero_rhs_n10_m2":
            rhs[1, :, 0] = 0.0
        out.append((name, A, rhs, max_iter, tol))
    return out


def main():
    warnings.filterwarnings("ignore")
    sys.dont_write_bytecode = True
    sys.path.insert(0, os.path.join(REF, "gitmodules", "gpytorch"))
    from gpytorch.utils import linear_cg
    sys.path.pop(0)
    res = {}
    for name, A, rhs, max_iter, tol in systems():
        sol = linear_cg(lambda t, A=A: A @ t, rhs, max_iter=max_iter, max_tridiag_iter=max_iter, tolerance=tol)
        res[name] = dict(A=A, rhs=rhs, max_iter=max_iter, tolerance=tol, solution=sol.clone())
        exact = torch.linalg.solve(A, rhs)
        print(f"{name}: |cg - exact| / |exact| = {float((sol - exact).abs().max() / exact.abs().max()):.2e}")
    path = os.path.join(os.path.dirname(HERE), "tests", "golden", "cg_gpytorch.pt")
    torch.save(res, path)
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
