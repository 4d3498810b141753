"""Deterministic, RNG-free tensor fills shared by the golden generator, the oracle tests and
the GPU parity tests (TEST INFRASTRUCTURE ONLY -- never imported by ``cmf_b200``).

Weights are a pure function of (state-dict key, shape, salt): a splitmix64 hash of the
element index, mapped to [-1, 1) in float64 and scaled.  This lets the real reference (in the
build container), the oracle and the CUDA path all be loaded with bit-identical "trained-like"
weights without committing multi-megabyte state dicts.
"""
from __future__ import annotations

import zlib
from typing import Dict

import numpy as np
import torch

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def _splitmix64(x: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        x = (x + np.uint64(0x9E3779B97F4A7C15)) & _M64
        z = x
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
        return z ^ (z >> np.uint64(31))


def unit_uniform(name: str, n: int, salt: int = 0) -> np.ndarray:
    """n float64 values in [-1, 1), a pure function of (name, salt, index)."""
    seed = np.uint64(zlib.crc32(name.encode()) + (salt << 32))
    idx = np.arange(n, dtype=np.uint64)
    with np.errstate(over="ignore"):
        h = _splitmix64(idx * np.uint64(0x2545F4914F6CDD1D) + seed)
    return (h >> np.uint64(11)).astype(np.float64) * (2.0 / (1 << 53)) - 1.0


def det_tensor(name: str, shape, scale: float = 1.0, salt: int = 0, dtype=torch.float32) -> torch.Tensor:
    n = int(np.prod(shape)) if len(shape) else 1
    return torch.from_numpy(unit_uniform(name, n, salt) * scale).reshape(tuple(shape)).to(dtype)


def det_permutation(name: str, n: int, salt: int = 0) -> torch.Tensor:
    return torch.from_numpy(np.argsort(unit_uniform(name, n, salt), kind="stable")).long()


def fill_state_dict(state: Dict[str, torch.Tensor], salt: int = 0, gain: float = 1.0) -> Dict[str, torch.Tensor]:
    """Return a new state dict with the same keys/shapes and deterministic values.

    * conv / linear ``weight``: uniform(+-gain/sqrt(fan_in))   (PyTorch-default-like scale)
    * ``.bias`` of conv/linear, affine ``shift``/``log_scale``: uniform(+-0.1)
    * scaled-tanh ``weights`` (shape (C,1,1)): 1 + uniform(+-0.2); its ``bias``: uniform(+-0.1)
    * tail ``permutation``: hash-ordered permutation, ``inverse_permutation`` = argsort
    * every other buffer (masks, Gaussian mean/stddev, fixed samples) is kept as is.
    Keys are matched after stripping a leading ``module.`` so wrapped / unwrapped dicts agree.
    """
    out = {}
    for key, val in state.items():
        name = key[len("module."):] if key.startswith("module.") else key
        leaf = name.rsplit(".", 1)[-1]
        if leaf == "permutation":
            out[key] = det_permutation(name, val.numel(), salt)
        elif leaf == "inverse_permutation":
            perm = det_permutation(name[: -len("inverse_permutation")] + "permutation", val.numel(), salt)
            out[key] = torch.argsort(perm)
        elif not val.is_floating_point() or leaf in ("mask", "mean", "stddev", "_fixed_samples"):
            out[key] = val.clone()
        elif leaf == "weight":
            fan_in = int(np.prod(val.shape[1:]))
            out[key] = det_tensor(name, val.shape, gain / np.sqrt(fan_in), salt, val.dtype)
        elif leaf == "weights":
            out[key] = 1.0 + det_tensor(name, val.shape, 0.2, salt, val.dtype)
        else:  # bias, shift, log_scale
            out[key] = det_tensor(name, val.shape, 0.1, salt, val.dtype)
    return out
