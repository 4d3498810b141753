"""Batch-sharded data parallelism: one process per GPU, the model replicated once, each rank's
shard pushed through the same kernels, and ONE NCCL all-reduce of a flat fp32 gradient bucket after
backward (sum, then / world).  Replaces the single-process ``nn.DataParallel`` of the reference
(`wrapper.py:52-68`, `factory.py:76-81`), which re-broadcasts every parameter and gathers outputs
through GPU 0 on every step.

Sharding is exact: every quantity on the path is per-sample and the only cross-sample op is the
final ``.mean()`` of the loss (`non_square_helpers.py:120`); with equal shards the mean of the
rank means is the global mean.  No data-path collective exists.
"""
from __future__ import annotations

from typing import Iterable, List, Optional

import torch
import torch.distributed as dist
import torch.nn as nn

from .densities import Density


class DataParallelDensity(Density):
    """Keeps the reference's wrapper position and its ``module.`` state-dict prefix; the
    data-parallel work itself is :func:`allreduce_gradients` called by the training step."""

    def __init__(self, module: Density):
        super().__init__()
        self.module = module

    def forward(self, mode, *args, **kwargs):
        return self.module(mode, *args, **kwargs)


def _world() -> int:
    return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


def shard_batch(x: torch.Tensor, rank: Optional[int] = None, world: Optional[int] = None) -> torch.Tensor:
    """Equal contiguous shard of the global batch for this rank (global batch % world must be 0)."""
    world = _world() if world is None else world
    if world == 1:
        return x
    rank = dist.get_rank() if rank is None else rank
    if x.shape[0] % world:
        raise ValueError(f"global batch {x.shape[0]} is not divisible by world size {world}")
    n = x.shape[0] // world
    return x[rank * n:(rank + 1) * n]


class GradientBucket:
    """Flat fp32 buffer aliasing every parameter's ``.grad`` so the reduction is one collective
    (23.9 MB for the 5.98 M-parameter MNIST-shape model)."""

    def __init__(self, params: Iterable[nn.Parameter]):
        self.params: List[nn.Parameter] = [p for p in params if p.requires_grad]
        if not self.params:
            raise ValueError("no trainable parameters")
        dev, dtype = self.params[0].device, self.params[0].dtype
        total = sum(p.numel() for p in self.params)
        self.flat = torch.zeros(total, dtype=dtype, device=dev)
        off = 0
        for p in self.params:
            n = p.numel()
            p.grad = self.flat[off:off + n].view_as(p)
            off += n

    def zero(self):
        self.flat.zero_()

    def sync_aliases(self) -> int:
        """Make every ``p.grad`` a view of ``flat`` again and return how many had to be repaired.

        ``torch.optim.Optimizer.zero_grad()`` defaults to ``set_to_none=True`` (torch >= 2.0; the
        reference's step `cmf/trainer.py:211` calls it), after which backward allocates fresh
        ``.grad`` tensors that no longer alias the bucket.  A fresh gradient is copied into its
        slice, a missing one (parameter that received no gradient) zeroes it, and the view is put
        back -- so a reduction never ships stale data.  Steps that clear gradients with
        :meth:`zero` or ``zero_grad(set_to_none=False)`` never pay this copy."""
        repaired, off = 0, 0
        for p in self.params:
            n = p.numel()
            view = self.flat[off:off + n].view_as(p)
            off += n
            if p.grad is None:
                view.zero_()
            elif p.grad.data_ptr() == view.data_ptr():
                continue
            else:
                view.copy_(p.grad)
            p.grad = view
            repaired += 1
        return repaired

    def gather_fresh(self) -> None:
        """For a backward that ran with every ``p.grad`` cleared to ``None`` (autograd then keeps the incoming gradient
        tensor instead of launching one ``grad += new`` per parameter: 685 launches in the MNIST-shape step): copy the
        fresh gradients into their slices with one multi-tensor copy, zero the slices of parameters that received
        none, and put the views back.  Used inside the CUDA-graph capture of the training step."""
        src, dst, off = [], [], 0
        for p in self.params:
            n = p.numel()
            view = self.flat[off:off + n].view_as(p)
            off += n
            if p.grad is None:
                view.zero_()
            elif p.grad.data_ptr() != view.data_ptr():
                src.append(p.grad)
                dst.append(view)
            p.grad = view
        if src:
            torch._foreach_copy_(dst, src)

    def sync_aliases_after_replay(self) -> None:
        """After a CUDA-graph replay the gradients ARE in ``flat`` (the captured kernels wrote there):
        only the views are put back, nothing is copied or zeroed."""
        off = 0
        for p in self.params:
            n = p.numel()
            p.grad = self.flat[off:off + n].view_as(p)
            off += n

    def allreduce(self, async_op: bool = False):
        """sum over ranks, then divide by the world size (mean of per-rank mean-losses)."""
        self.sync_aliases()
        world = _world()
        if world == 1:
            return None
        if dist.get_backend() == "nccl":
            # the division by the world size is folded into the reduction (NCCL's AVG): one kernel, no extra pass
            return dist.all_reduce(self.flat, op=dist.ReduceOp.AVG, async_op=async_op) if async_op else \
                (dist.all_reduce(self.flat, op=dist.ReduceOp.AVG), None)[1]
        work = dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, async_op=async_op)   # gloo (CPU tests): no AVG
        if async_op:
            return work
        self.flat.div_(world)
        return None


def allreduce_gradients(bucket: GradientBucket, loss: Optional[torch.Tensor] = None):
    """Reduce gradients (and, for logging, the scalar loss) across ranks."""
    bucket.allreduce()
    if loss is not None and _world() > 1:
        loss = loss.detach().clone()
        dist.all_reduce(loss, op=dist.ReduceOp.SUM)
        loss /= _world()
    return loss
