"""``FusedNonSquareHeadDensity`` -- drop-in for the reference's ``NonSquareHeadDensity``
(`cmf/models/components/densities/non_square.py:22-340`): same constructor, same ``_elbo``
signature / result dicts / error behaviour, same place in the density tree
(`cmf/models/factory.py:121-144`).  What changes is how ``J``, ``G = J^T J``, ``log det G`` and
the metric penalty are produced: one compiled layer program on hand-written sm_100a kernels with
all d tangent columns in flight, instead of d Python-level JVP passes.

There is no CPU fallback: the ops raise if the CUDA library is missing or a tensor is not on a GPU.
"""
from __future__ import annotations

import os
from typing import Optional

import numpy as np
import torch

from . import ops
from .cg import linear_cg_batched
from .densities import Density
from .networks import library_convs
from .program import LayerProgram

Tensor = torch.Tensor

# CMF_FLAT_STACK=0 keeps the small flat configs on the per-layer program (A/B measurements)
_FLAT_STACK = os.environ.get("CMF_FLAT_STACK", "1") != "0"
# The CG loop of hutch_with_cg needs ~12 products J^T J v (gpytorch's exit rule with the shipped cg_tolerance); each
# is a full network pass on B m images.  For small latents ONE d-column pass under no_grad builds G = J^T J outright
# (the tangent-batched program this package is about) and the products become 64 x 64 mat-vecs.  auto: d <= 128.
_CG_EXPLICIT = os.environ.get("CMF_CG_EXPLICIT", "auto")
# CMF_CG_GRAPH=0: the J^T J product inside the CG loop of hutch_with_cg is launched eagerly instead of replayed
_CG_GRAPH = os.environ.get("CMF_CG_GRAPH", "1") != "0"


class _GraphSlot:
    """holder of a captured CUDA graph that copy.deepcopy / pickling of the module leave behind (empty)"""

    def __init__(self):
        self.entry = None

    def __deepcopy__(self, memo):
        return _GraphSlot()

    def __reduce__(self):
        return (_GraphSlot, ())


class FusedNonSquareHeadDensity(Density):
    _VALID_LOG_JACOBIAN_METHODS = ["cholesky", "hutch_with_cg"]
    # BASELINE.json / README spell the second method "hutchinson"; accepted as an alias
    _ALIASES = {"hutchinson": "hutch_with_cg"}

    def __init__(self, prior, regularization_param, log_jacobian_method, x_shape,
                 hutchinson_distribution, num_hutchinson_samples=1, max_cg_iterations=None,
                 cg_tolerance=1):
        super().__init__()
        self.prior = prior
        self.regularization_param = regularization_param
        self.x_shape = tuple(x_shape)
        self.hutchinson_distribution = hutchinson_distribution
        self.num_hutchinson_samples = num_hutchinson_samples
        self.max_cg_iterations = max_cg_iterations
        self.cg_tolerance = cg_tolerance
        log_jacobian_method = self._ALIASES.get(log_jacobian_method, log_jacobian_method)
        if log_jacobian_method not in self._VALID_LOG_JACOBIAN_METHODS:
            raise ValueError(f"{log_jacobian_method} not a valid Jacobian calculation method")
        self.log_jacobian_method = log_jacobian_method
        self._program: Optional[LayerProgram] = None
        self.last_cholesky_attempts = 1
        self._cg_slot = _GraphSlot()   # entry: {key: (graph, static latent, static vector, static result)}
        self.cg_graph_replays = 0

    # ------------------------------------------------------------------ program cache
    def _load_from_state_dict(self, *args, **kwargs):
        self._program = None      # the tail permutation (a buffer) is baked into the program
        return super()._load_from_state_dict(*args, **kwargs)

    @property
    def program(self) -> LayerProgram:
        if self._program is None:
            self._program = LayerProgram(self.prior)
        return self._program

    # ------------------------------------------------------------------ entry points
    def _ood(self, x):
        return self("elbo", x, ood=True)

    def _sample(self, num_samples):
        return self.prior.sample(num_samples)

    def _fixed_sample(self, noise):
        return self.prior.fixed_sample(noise=noise)

    def _extract_latent(self, x, **kwargs):
        prior_dict = self.prior.elbo(x)
        low_dim_latent, _ = self._low_dim(prior_dict)
        if kwargs.get("earliest_latent"):
            return self.program.tail.extract_latent(low_dim_latent)
        return low_dim_latent

    @staticmethod
    def _low_dim(prior_dict):
        """`non_square.py:157-176`: descend the encode-pass result to the tail's entry.  Only the
        d-dimensional prior's elbo is kept; log-jacobians of the layers above the tail are not."""
        while "low-dim-x" not in prior_dict:
            prior_dict = prior_dict["prior-dict"]
        return prior_dict["low-dim-x"], prior_dict["elbo"]

    def _elbo(self, x, add_reconstruction=True, add_diagonal_metric_reg=False,
              add_offdiagonal_metric_reg=False, likelihood_wt=1., metric_wt=1.,
              visualization=False, ood=False, test_metric=False):
        prior_dict = self.prior.elbo(x)
        low_dim_latent, low_dim_elbo = self._low_dim(prior_dict)

        if ood:
            assert self.log_jacobian_method == "cholesky"

        if not np.isclose(likelihood_wt, 0.):
            penalty = "diag" if add_diagonal_metric_reg else ("offdiag" if add_offdiagonal_metric_reg else "none")
            log_det_jacobian, reconstructed_x, jtj, metric_l1_loss = self._log_det_recon_penalty(
                low_dim_latent, penalty)
            likelihood_term = low_dim_elbo - log_det_jacobian / 2.
            if penalty == "none":
                metric_l1_loss = 0
        else:
            # likelihood warm-up: reconstruction only, no tangents at all (non_square.py:105-109)
            metric_l1_loss = 0
            likelihood_term = 0
            reconstructed_x = self.flow_forward(low_dim_latent)

        if add_reconstruction:
            assert not visualization
            reconstruction_loss = ops.sumsq_diff(reconstructed_x, x)          # non_square.py:111-114
        else:
            reconstruction_loss = 0

        if ood:
            assert add_reconstruction and not np.isclose(likelihood_wt, 0.)
            return {"likelihood": likelihood_term, "reconstruction-error": reconstruction_loss}

        return {
            "elbo": likelihood_wt * likelihood_term - self.regularization_param * reconstruction_loss
                    - metric_wt * metric_l1_loss,
            "prior-dict": prior_dict,
        }

    # ------------------------------------------------------------------ the hot path
    def _dense(self, blocks, B: int, like: Tensor) -> Tensor:
        """Assemble the primal output (B, *x_shape) from compact blocks (primal only: B*D floats)."""
        out = torch.zeros(B, self.program.D, dtype=like.dtype, device=like.device)
        for pos, xb, _ in blocks:
            out = out.index_copy(1, pos, xb)
        return out.view(B, *self.x_shape)

    def flow_forward(self, z: Tensor) -> Tensor:
        """`non_square.py:313-320`, primal only (kernel (a) with d = 0)."""
        blocks, _ = self.program.run(z, None)
        return self._dense(blocks, z.shape[0], z)

    def jvp_forward(self, z: Tensor, v: Tensor):
        """`non_square.py:322-329` for one tangent v (B,d): returns (x, J v) dense."""
        blocks, _ = self.program.run(z, v.unsqueeze(1).contiguous())
        B = z.shape[0]
        jv = torch.zeros(B, self.program.D, dtype=z.dtype, device=z.device)
        for pos, _, Vb in blocks:
            jv = jv.index_copy(1, pos, Vb[:, 0])
        return self._dense(blocks, B, z), jv.view(B, *self.x_shape)

    def _get_full_jac_transpose_jac(self, latent: Tensor, create_graph: bool = True):
        """`non_square.py:298-311`: (G (B,d,d), reconstruction).  One program run with V0 = I."""
        B, d = latent.shape
        plan = self.program.flat_plan
        if plan is not None and _FLAT_STACK:
            # small flat flow: tail pad, all coupling layers with their MLP conditioners and J^T J in one launch
            gram, x = ops.flat_stack_gram(latent, plan.params(), plan.desc)
            return gram, x.view(B, *self.x_shape)
        eye = torch.eye(d, dtype=latent.dtype, device=latent.device).expand(B, d, d).contiguous()
        blocks, _ = self.program.run(latent, eye)
        gram = ops.gram([Vb for _, _, Vb in blocks])
        return gram, self._dense(blocks, B, latent)

    def _log_det_recon_penalty(self, latent: Tensor, penalty: str):
        if not self.training or self.log_jacobian_method == "cholesky":
            return self._exact_log_det_jac_and_reconstruction(latent, penalty)
        log_det, recon, jtj_eps = self._approx_log_det_jac_and_reconstruction(latent)
        # The reference takes the penalty over J^T J eps here, not over G (non_square.py:98-99,258;
        # SURVEY 3.2 note) -- reproduced as is.
        if penalty == "diag":
            l1 = torch.diagonal(jtj_eps, dim1=-2, dim2=-1).abs().sum(1, keepdim=True)
        elif penalty == "offdiag":
            off = ~torch.eye(jtj_eps.shape[1], device=jtj_eps.device, dtype=torch.bool)
            l1 = jtj_eps.masked_select(off).view(jtj_eps.shape[0], -1).abs().sum(1, keepdim=True)
        else:
            l1 = 0
        return log_det, recon, jtj_eps, l1

    def _get_log_det_jac_and_reconstruction(self, latent, visualization=False):
        """Reference-shaped return (log det, reconstruction, G): `non_square.py:131-138,296`."""
        return self._log_det_recon_penalty(latent, "none")[:3]

    def _exact_log_det_jac_and_reconstruction(self, latent: Tensor, penalty: str = "none"):
        """`non_square.py:262-296` with the penalty of `:87-100` fused into kernel (c)."""
        eps, eps_factor = 1e-6, 10
        gram, recon = self._get_full_jac_transpose_jac(latent, self.training)
        attempts = 1
        eye = None
        while True:
            try:
                log_det, l1 = ops.chol_logdet_l1(gram, penalty)
                break
            except ops.NotPositiveDefiniteError:
                # HACK kept from the reference: add eps*I to J^T J for the whole batch and retry
                if eye is None:
                    eye = torch.eye(gram.shape[1], dtype=gram.dtype, device=gram.device).expand_as(gram)
                gram = gram + eps * eye
                attempts += 1
                eps *= eps_factor
        if attempts > 1:
            print(f"WARNING: Numerical non-invertibility in JtJ observed - {attempts} attempts needed to fix")
        self.last_cholesky_attempts = attempts
        return log_det, recon, gram, l1

    # ------------------------------------------------------------------ hutch_with_cg
    def _draw_hutchinson(self, shape, device, dtype):
        if self.hutchinson_distribution == "normal":
            return torch.randn(*shape, device=device, dtype=dtype)
        if self.hutchinson_distribution == "rademacher":
            return torch.bernoulli(0.5 * torch.ones(*shape, device=device, dtype=dtype)).mul_(2.).subtract_(1.)
        raise ValueError(f"Unknown hutchinson distribution {self.hutchinson_distribution}")

    def _jac_transpose_jac_vec(self, latent: Tensor, vec: Tensor, create_graph: bool):
        """`non_square.py:190-201`: (J^T J vec, reconstruction) for one tangent per row."""
        def jtj(lat):
            blocks, _ = self.program.run(lat, vec.unsqueeze(1).contiguous())
            with torch.enable_grad():
                lat_p = lat if lat.requires_grad else lat.detach().requires_grad_(True)
                if create_graph:
                    # this VJP is differentiated again by loss.backward(): the primal coupling / gather
                    # ops record their adjoints (ops._CouplingPrimalVJP); the conditioner convolutions
                    # take the library path, which autograd can differentiate twice
                    with library_convs():
                        prim, _ = self.program.run(lat_p, None)
                else:
                    prim, _ = self.program.run(lat_p, None)
                outs = [xb for _, xb, _ in prim]
                cots = [Vb[:, 0] for _, _, Vb in blocks]
                (res,) = torch.autograd.grad(outs, lat_p, cots, create_graph=create_graph)
            return res, self._dense(blocks, lat.shape[0], lat)
        if not create_graph:
            with torch.no_grad():
                return jtj(latent.detach())
        return jtj(latent)

    def _cg_product(self, latent: Tensor):
        """v -> J^T J v at ``latent`` for the CG loop (`non_square.py:241-247`: detached, no graph).  Every CG
        iteration issues the same ~1000 small launches (conditioners on B m images): on a GPU they are captured
        once into a CUDA graph (static latent / vector / result buffers; the parameters are read in place, so
        the graph stays valid across optimizer steps) and replayed per iteration -- bit-identical kernels."""
        eager = lambda v: self._jac_transpose_jac_vec(latent, v, create_graph=False)[0]
        if not (_CG_GRAPH and latent.is_cuda) or torch.cuda.is_current_stream_capturing():
            return eager
        key = (tuple(latent.shape), latent.dtype, latent.device, self.training,
               tuple(p.data_ptr() for p in self.parameters()))
        cache = self._cg_slot.entry = self._cg_slot.entry or {}        # a few shapes / modes (train and validation batches)
        if key not in cache:
            lat = latent.detach().clone()
            vec = torch.zeros_like(lat)
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):                 # warm-up off the capture stream (allocator, library plans)
                for _ in range(2):
                    self._jac_transpose_jac_vec(lat, vec, create_graph=False)
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            try:
                with torch.cuda.graph(graph):
                    out = self._jac_transpose_jac_vec(lat, vec, create_graph=False)[0]
            except Exception as e:
                # a failed capture can leave torch's CUDA RNG / allocator in capture state: fail loudly, do not limp on
                raise RuntimeError("cmf_b200: CUDA-graph capture of the hutch_with_cg J^T J product failed "
                                   f"({type(e).__name__}: {e}); set CMF_CG_GRAPH=0 to launch it eagerly") from e
            if len(cache) >= 4:
                cache.pop(next(iter(cache)))
            cache[key] = (graph, lat, vec, out)
        graph, lat, vec, out = cache[key]
        lat.copy_(latent.detach())

        def replay(v):
            vec.copy_(v)
            graph.replay()
            self.cg_graph_replays += 1
            return out.clone()
        return replay

    def _approx_log_det_jac_and_reconstruction(self, latent: Tensor, hutchinson_samples: Optional[Tensor] = None):
        """`non_square.py:203-258`.  ``hutchinson_samples`` (B,d,m) may be injected for parity tests."""
        B, d = latent.shape
        m = self.num_hutchinson_samples
        if hutchinson_samples is None:
            hutchinson_samples = self._draw_hutchinson((B, d, m), latent.device, latent.dtype)
        repeated = latent.repeat_interleave(m, dim=0)

        def to_vec(t): return t.transpose(1, 2).reshape(B * m, d)
        def to_ten(v): return v.reshape(B, m, d).transpose(1, 2)

        explicit = _CG_EXPLICIT == "1" or (_CG_EXPLICIT == "auto" and d <= 128 and latent.is_cuda)
        if explicit:
            # same operator, formed once: G v instead of J^T (J v) per iteration (rounding-level differences only;
            # the solution is detached either way, `non_square.py:247`)
            with torch.no_grad():
                G = self._get_full_jac_transpose_jac(latent.detach(), create_graph=False)[0]

            def closure(t):
                return torch.bmm(G, t)
        else:
            product = self._cg_product(repeated)

            def closure(t):
                return to_ten(product(to_vec(t)))

        inv = linear_cg_batched(closure, hutchinson_samples, max_iter=self.max_cg_iterations,
                                tolerance=self.cg_tolerance).detach()
        jtj_vec, recon_rep = self._jac_transpose_jac_vec(repeated, to_vec(hutchinson_samples),
                                                         create_graph=self.training)
        jtj_eps = to_ten(jtj_vec)
        approx = torch.mean(torch.sum(inv * jtj_eps, dim=1, keepdim=True), dim=2)
        return approx, recon_rep[::m], jtj_eps


# the reference's name for the same role
NonSquareHeadDensity = FusedNonSquareHeadDensity
