"""Batched conjugate gradients for the ``hutch_with_cg`` estimator (`non_square.py:241-247`).

Host-driven loop over device tensors, reproducing the control flow of the call the reference
makes -- ``gpytorch.utils.linear_cg(closure, rhs, max_iter, max_tridiag_iter, tolerance)`` with
no preconditioner and ``n_tridiag = 0`` (`gitmodules/gpytorch/gpytorch/utils/linear_cg.py:89-330`
at fc2053b): rhs normalised per column, "safe division" for alpha/beta, converged columns frozen,
at most ``min(max_iter, rows)`` iterations, early exit once ``k >= min(10, max_iter - 1)`` and the
mean residual norm is below ``tolerance``.  Each iteration costs one J^T J product = one m-column
run of kernel (a) plus its adjoint.
"""
from __future__ import annotations

import torch

Tensor = torch.Tensor


def linear_cg_batched(matmul, rhs: Tensor, max_iter: int, tolerance: float,
                      eps: float = 1e-10, stop_updating_after: float = 1e-10) -> Tensor:
    """Solve A X = rhs column-wise; rhs (B, rows, m), ``matmul`` maps (B, rows, m) -> same."""
    norm = rhs.norm(2, dim=-2, keepdim=True)
    rhs_is_zero = norm.lt(eps)
    norm = norm.masked_fill(rhs_is_zero, 1)
    rhs = rhs / norm
    sol = torch.zeros_like(rhs)
    resid = rhs - matmul(sol)
    if not torch.equal(resid, resid):
        raise RuntimeError("NaNs encountered when trying to perform matrix-vector multiplication")
    frozen = resid.norm(2, dim=-2, keepdim=True).lt(stop_updating_after)
    if bool(frozen.all()):
        return sol * norm
    direction = resid.clone()
    rr = (resid * resid).sum(-2, keepdim=True)
    for k in range(min(max_iter, rhs.size(-2))):
        a_dir = matmul(direction)
        denom = (direction * a_dir).sum(-2, keepdim=True)
        tiny = denom.lt(eps)
        alpha = (rr / denom.masked_fill(tiny, 1)).masked_fill(tiny, 0).masked_fill(frozen, 0)
        resid = resid - alpha * a_dir
        sol = sol + alpha * direction
        rr_prev, rr = rr, (resid * resid).sum(-2, keepdim=True)
        tiny = rr_prev.lt(eps)
        beta = (rr / rr_prev.masked_fill(tiny, 1)).masked_fill(tiny, 0)
        direction = direction * beta + resid
        res_norm = resid.norm(2, dim=-2, keepdim=True).masked_fill(rhs_is_zero, 0)
        frozen = res_norm.lt(stop_updating_after)
        if k >= min(10, max_iter - 1) and bool(res_norm.mean() < tolerance):
            break
    return sol * norm
