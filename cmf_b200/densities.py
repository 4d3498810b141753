"""Density-tree nodes around the fused head: same module/buffer names, mode dispatch and result
dicts as the reference (`cmf/models/components/densities/*.py`), so state dicts and callers are
interchangeable.  Only what ``--model non-square`` reaches is here.
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn

Tensor = torch.Tensor


class Density(nn.Module):
    """Mode dispatch of `density.py:7-74`."""

    def forward(self, mode, *args, **kwargs):
        if mode == "elbo":
            return self._elbo(*args, **kwargs)
        if mode == "sample":
            return self._sample(*args)
        if mode == "fixed-sample":
            return self._fixed_sample(*args)
        if mode == "ood":
            return self._ood(*args)
        if mode == "extract-latent":
            return self._extract_latent(*args, **kwargs)
        assert False, f"Invalid mode {mode}"

    def elbo(self, x, **kwargs): return self("elbo", x, **kwargs)
    def sample(self, num_samples): return self("sample", num_samples)
    def fixed_sample(self, noise=None): return self("fixed-sample", noise)
    def ood(self, x): return self("ood", x)
    def extract_latent(self, x, **kwargs): return self("extract-latent", x, **kwargs)

    def _elbo(self, x, **kwargs): raise NotImplementedError
    def _sample(self, num_samples): raise NotImplementedError
    def _fixed_sample(self, noise): raise NotImplementedError
    def _ood(self, x): raise NotImplementedError
    def _extract_latent(self, x, **kwargs): raise NotImplementedError


class BijectionDensity(Density):
    """`exact.py:8-47`."""

    def __init__(self, prior: Density, bijection: nn.Module):
        super().__init__()
        self.bijection = bijection
        self.prior = prior

    def _elbo(self, x, **kwargs):
        res = self.bijection.x_to_z(x)
        prior_dict = self.prior.elbo(res["z"], **kwargs)
        return {"elbo": prior_dict["elbo"] + res["log-jac"], "bijection-info": res, "prior-dict": prior_dict}

    def _sample(self, num_samples):
        return self.bijection.z_to_x(self.prior.sample(num_samples))["x"]

    def _fixed_sample(self, noise):
        return self.bijection.z_to_x(self.prior.fixed_sample(noise=noise))["x"]

    def _extract_latent(self, x, **kwargs):
        return self.prior.extract_latent(self.bijection.x_to_z(x)["z"], **kwargs)

    def _ood(self, x):
        return self.prior.ood(self.bijection.x_to_z(x)["z"])


class DiagonalGaussianDensity(Density):
    """`gaussian.py:44-87`: buffers ``mean``, ``stddev``, ``_fixed_samples``."""

    def __init__(self, mean: Tensor, stddev: Tensor, num_fixed_samples: int = 0):
        super().__init__()
        assert mean.shape == stddev.shape
        self.register_buffer("mean", mean)
        self.register_buffer("stddev", stddev)
        if num_fixed_samples > 0:
            self.register_buffer("_fixed_samples", self.sample(num_fixed_samples))

    def _elbo(self, z):
        if getattr(self, "_standard", False):       # N(0, I): the quadratic form on the fused kernel
            from . import ops
            dim = z[0].numel()
            return {"elbo": -0.5 * dim * math.log(2 * math.pi) - 0.5 * ops.sumsq_diff(z), "z": z}
        w = z.flatten(1)
        mu = self.mean.expand_as(z).flatten(1)
        var = self.stddev.expand_as(z).flatten(1) ** 2
        dim = w.shape[1]
        log_prob = (-0.5 * dim * math.log(2 * math.pi)
                    - 0.5 * torch.sum(torch.log(var), dim=1, keepdim=True)
                    - 0.5 * torch.sum((w - mu) ** 2 / var, dim=1, keepdim=True))
        return {"elbo": log_prob, "z": z}

    def _sample(self, num_samples):
        shape = (num_samples, *self.mean.shape)
        return self.stddev.expand(shape) * torch.randn(shape, dtype=self.mean.dtype,
                                                        device=self.mean.device) + self.mean.expand(shape)

    def _fixed_sample(self, noise):
        return noise if noise is not None else self._fixed_samples

    def _extract_latent(self, x, **kwargs):
        return x


def standard_gaussian(shape) -> DiagonalGaussianDensity:
    """`factory.py:200-205`."""
    density = DiagonalGaussianDensity(torch.zeros(tuple(shape)), torch.ones(tuple(shape)), num_fixed_samples=64)
    density._standard = True     # buffers stay (state-dict parity); their values are known: 0 and 1
    return density


class SplitDensity(Density):
    """`split.py:6-52`; in non-square mode the second half is dropped on encode and zero-padded
    on decode / JVP."""

    def __init__(self, density_1: Density, density_2: Density, dim: int, non_square: bool = False):
        super().__init__()
        self.density_1, self.density_2 = density_1, density_2
        self.dim, self.non_square = dim, non_square

    def _elbo(self, x):
        x1, x2 = torch.chunk(x, 2, dim=self.dim)
        d1, d2 = self.density_1.elbo(x1), self.density_2.elbo(x2)
        return {"elbo": d1["elbo"] + d2["elbo"], "prior-dict": d1, "prior-dict-2": d2}

    def pad_inputs(self, x1):
        return {"x": torch.cat((x1, torch.zeros_like(x1)), dim=self.dim)}

    def _join(self, x1, x2_fn):
        if self.non_square:
            return self.pad_inputs(x1)["x"]
        return torch.cat((x1, x2_fn()), dim=self.dim)

    def _sample(self, num_samples):
        return self._join(self.density_1.sample(num_samples), lambda: self.density_2.sample(num_samples))

    def _fixed_sample(self, noise):
        return self._join(self.density_1.fixed_sample(noise=noise), lambda: self.density_2.fixed_sample(noise=noise))


class WrapperDensity(Density):
    """`wrapper.py:7-24`."""

    def __init__(self, density: Density):
        super().__init__()
        self.density = density

    def _elbo(self, x, **kwargs): return self.density.elbo(x, **kwargs)
    def _sample(self, num_samples): return self.density.sample(num_samples)
    def _fixed_sample(self, noise): return self.density.fixed_sample(noise=noise)
    def _ood(self, x): return self.density.ood(x)
    def _extract_latent(self, x, **kwargs): return self.density.extract_latent(x, **kwargs)


class DequantizationDensity(WrapperDensity):
    """`wrapper.py:27-30`: uniform noise added IN PLACE to the caller's batch, like the reference."""

    def _elbo(self, x, **kwargs):
        chain = self._logit_chain()
        if chain is None or x.requires_grad or not x.is_contiguous():
            return super()._elbo(x.add_(torch.rand_like(x)), **kwargs)
        # dequantise + scalar-mult + scalar-add + logit (`wrapper.py:27-30`, `math.py:41-105`,
        # `config/schemas.py:30-50`) in one pass over the batch; the three BijectionDensity nodes keep
        # their place in the tree (state dict, sampling, ood) and their result dicts are rebuilt here
        from . import kernels
        mult_node, add_node, logit_node = chain
        mult, add = mult_node.bijection.value, add_node.bijection.value
        y, lj = kernels.preproc_logit_fwd(x, torch.rand_like(x), mult, add)
        inner = logit_node.prior.elbo(y, **kwargs)
        if "elbo" not in inner:
            return inner                                  # ood: {"likelihood", "reconstruction-error"}
        lj_mult = x[0].numel() * math.log(abs(mult))
        d_logit = {"elbo": inner["elbo"] + (lj - lj_mult), "prior-dict": inner}
        d_add = {"elbo": d_logit["elbo"], "prior-dict": d_logit}
        return {"elbo": d_add["elbo"] + lj_mult, "prior-dict": d_add}

    def _logit_chain(self):
        """the three elementwise nodes `schemas.py:30-50` puts below the dequantiser, or None"""
        from .bijections import LogitBijection, ScalarAdditionBijection, ScalarMultiplicationBijection
        nodes, node = [], self.density
        for cls in (ScalarMultiplicationBijection, ScalarAdditionBijection, LogitBijection):
            if not (isinstance(node, BijectionDensity) and isinstance(node.bijection, cls)):
                return None
            nodes.append(node)
            node = node.prior
        return nodes


class NonSquareTailDensity(Density):
    """`non_square.py:365-421`: permute + keep the first d coordinates on encode; buffers ``mask``,
    ``permutation``, ``inverse_permutation`` keep the reference's names."""

    def __init__(self, prior: Density, x_shape, latent_dimension: int, detach_before_prior: bool):
        super().__init__()
        self.prior = prior
        self.detach_before_prior = detach_before_prior
        self.x_shape = tuple(x_shape)
        self.latent_dimension = latent_dimension
        self.flattened_dims = int(math.prod(x_shape))
        self.register_buffer("mask", torch.arange(self.flattened_dims) < latent_dimension)
        self.register_buffer("permutation", torch.randperm(self.flattened_dims))
        self.register_buffer("inverse_permutation", torch.argsort(self.permutation))

    def _elbo(self, x):
        # == x[:, permutation][:, mask] of `non_square.py:383-384` (mask = first latent_dimension
        # entries); one index_select, no boolean-mask compaction and hence no host sync, so the step
        # can be captured into a CUDA graph
        low_dim_x = x.flatten(1).index_select(1, self.permutation[: self.latent_dimension])
        prior_dict = self.prior.elbo(low_dim_x.detach() if self.detach_before_prior else low_dim_x)
        return {"elbo": prior_dict["elbo"], "low-dim-x": low_dim_x, "prior-dict": prior_dict}

    def low_dim_to_masked(self, low_dim_x):
        padded = torch.zeros(low_dim_x.shape[0], self.flattened_dims, dtype=low_dim_x.dtype,
                             device=low_dim_x.device)
        padded[:, : self.latent_dimension] = low_dim_x
        return {"x": padded[:, self.inverse_permutation].view(low_dim_x.shape[0], *self.x_shape)}

    def _sample(self, num_samples): return self.low_dim_to_masked(self.prior.sample(num_samples))["x"]
    def _fixed_sample(self, noise): return self.low_dim_to_masked(self.prior.fixed_sample(noise))["x"]
    def _extract_latent(self, x, **kwargs): return self.prior.extract_latent(x, **kwargs)
