"""CUDA-graph capture of the training step ``loss = train_metrics(density, x, epoch)["loss"];
loss.backward()`` (`cmf/trainer.py:213-216`).

The flat configs (2-D and tabular: D <= 21, d <= 10, MLP conditioners) issue a few thousand
microsecond-sized kernels per step and are bound by launch latency, not by the GPU; replaying the
step as one graph removes that bound without changing a single kernel (SURVEY 8f N1-ii asks for a
whole-stack kernel; the graph is the step that does not need new device code).  The image configs
gain nothing (GPU-bound) but are supported the same way.

Contract
  * one graph per distinct ``(epoch-derived loss weights, batch shape)``: the weights are Python
    floats baked into the captured kernels' arguments (`non_square_helpers.py:99-123`);
  * parameters' ``.grad`` must be persistent tensors (``GradientBucket`` provides that), and no
    eager autograd graph over these parameters may be alive at capture time (its AccumulateGrad
    nodes are bound to the stream it was built on and would be reused inside the capture);
  * the Cholesky ``info`` word is checked after the replay; a non-positive-definite batch re-runs
    the step eagerly so the reference's ``+eps*I`` retry loop and its WARNING line still happen.
"""
from __future__ import annotations

import os
from typing import Callable, Dict, Tuple

import torch

from . import ops
from .parallel import GradientBucket


# CMF_GRAPH_STEAL_GRADS=0: accumulate into the zeroed bucket inside the graph (one add kernel per parameter) instead
_STEAL_GRADS = os.environ.get("CMF_GRAPH_STEAL_GRADS", "1") != "0"


class GraphedTrainStep:
    def __init__(self, density: torch.nn.Module, train_metrics: Callable, bucket: GradientBucket,
                 weights_key: Callable[[int], Tuple] = None, warmup_iters: int = 3):
        self.density, self.train_metrics, self.bucket = density, train_metrics, bucket
        self.weights_key = weights_key or (lambda epoch: (epoch,))
        self.warmup_iters = warmup_iters
        self._graphs: Dict[Tuple, tuple] = {}
        self.eager_fallbacks = 0

    def _eager(self, x, epoch):
        self.bucket.zero()
        loss = self.train_metrics(self.density, x, epoch)["loss"]
        loss.backward()
        return loss.detach()

    def _capture(self, x, epoch):
        static_x = x.clone()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):                 # warm-up off the capture stream (allocator, cuDNN/cuBLAS plans)
            for _ in range(self.warmup_iters):
                self._eager(static_x, epoch)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with ops.defer_pd_check() as pending:
            with torch.cuda.graph(graph):
                if _STEAL_GRADS:
                    for p in self.bucket.params:      # autograd keeps the incoming tensors: no per-parameter add kernels
                        p.grad = None
                else:
                    self.bucket.flat.zero_()
                loss = self.train_metrics(self.density, static_x, epoch)["loss"]
                loss.backward()
                if _STEAL_GRADS:
                    self.bucket.gather_fresh()        # one multi-tensor copy into bucket.flat
                static_loss = loss.detach()
            infos = list(pending)
        info_all = torch.stack([i.view(()) for i in infos]) if infos else None
        return graph, static_x, static_loss, infos, info_all

    def __call__(self, x: torch.Tensor, epoch: int) -> torch.Tensor:
        """Runs one step; gradients land in ``bucket.flat`` (aliased by every ``p.grad``).
        Returns the detached loss (a persistent device scalar: clone it to keep it)."""
        key = (self.weights_key(epoch), tuple(x.shape), x.dtype)
        entry = self._graphs.get(key)
        if entry is None:
            entry = self._graphs[key] = self._capture(x, epoch)
        graph, static_x, static_loss, infos, _ = entry
        static_x.copy_(x)
        graph.replay()
        for p in self.bucket.params:                  # the replay wrote bucket.flat: an optimizer's
            if p.grad is None:                        # zero_grad(set_to_none=True) must not hide it
                self.bucket.sync_aliases_after_replay()
                break
        if infos and int(torch.stack([i.view(()) for i in infos]).max().item()):
            self.eager_fallbacks += 1                 # not PD: eager path replays the retry loop
            return self._eager(x, epoch)
        return static_loss
