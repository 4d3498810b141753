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
    """Solve A X = rhs column-wise; rhs (B, rows, m), ``matmul`` maps (B, rows, m) -> same.

    While a CUDA graph is being captured the host cannot read the residual, so the data-dependent exits
    (`linear_cg.py:286-300`) become a device-side ``done`` flag: every one of the min(max_iter, rows) iterations is
    issued and the solution stops changing from the iteration at which the eager loop would have left -- the same
    result, no synchronisation (tests/test_cg_golden.py compares the two forms bit for bit)."""
    sync_free = rhs.is_cuda and torch.cuda.is_current_stream_capturing()
    return _cg(matmul, rhs, max_iter, tolerance, eps, stop_updating_after, sync_free)


def _cg(matmul, rhs: Tensor, max_iter: int, tolerance: float, eps: float, stop_updating_after: float, sync_free: bool) -> Tensor:
    norm = rhs.norm(2, dim=-2, keepdim=True)
    rhs_is_zero = norm.lt(eps)
    norm = norm.masked_fill(rhs_is_zero, 1)
    rhs = rhs / norm
    sol = torch.zeros_like(rhs)
    resid = rhs - matmul(sol)
    if not sync_free and not torch.equal(resid, resid):
        raise RuntimeError("NaNs encountered when trying to perform matrix-vector multiplication")
    frozen = resid.norm(2, dim=-2, keepdim=True).lt(stop_updating_after)
    done = frozen.all() if sync_free else None            # 0-dim bool on the device: the loop has been left
    if not sync_free and bool(frozen.all()):
        return sol * norm
    direction = resid.clone()
    rr = (resid * resid).sum(-2, keepdim=True)
    for k in range(min(max_iter, rhs.size(-2))):
        a_dir = matmul(direction)
        denom = (direction * a_dir).sum(-2, keepdim=True)
        tiny = denom.lt(eps)
        alpha = (rr / denom.masked_fill(tiny, 1)).masked_fill(tiny, 0).masked_fill(frozen, 0)
        new_sol = sol + alpha * direction
        sol = torch.where(done, sol, new_sol) if sync_free else new_sol
        resid = resid - alpha * a_dir
        rr_prev, rr = rr, (resid * resid).sum(-2, keepdim=True)
        tiny = rr_prev.lt(eps)
        beta = (rr / rr_prev.masked_fill(tiny, 1)).masked_fill(tiny, 0)
        direction = direction * beta + resid
        res_norm = resid.norm(2, dim=-2, keepdim=True).masked_fill(rhs_is_zero, 0)
        frozen = res_norm.lt(stop_updating_after)
        if k >= min(10, max_iter - 1):
            if sync_free:
                done = done | (res_norm.mean() < tolerance)
            elif bool(res_norm.mean() < tolerance):
                break
    return sol * norm
