"""Generate ``tests/golden/*.pt`` by running the UNMODIFIED reference (k-flouris/cmf) from
/root/reference.  Build-container only: the GPU box has no /root/reference, which is why the
outputs are committed as fixtures.  TEST INFRASTRUCTURE ONLY.

    PYTHONDONTWRITEBYTECODE=1 python oracle/gen_golden.py [case ...]

Import recipe: SURVEY.md Appendix A (vendored gpytorch first, then pyro stubs needed only by
`cmf/models/components/bijections/sos.py:6-7`).  No reference code is copied: the reference
modules are imported, loaded with the deterministic weights of ``oracle/detfill.py`` and driven
through their own public entry points (`config.get_config/get_schema`, `cmf.models.get_density`,
`cmf.non_square_helpers.get_non_square_train_metrics`).
"""
from __future__ import annotations

import json
import os
import sys
import time
import types
import warnings

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
REF = os.environ.get("CMF_REFERENCE", "/root/reference")


def import_reference():
    warnings.filterwarnings("ignore")
    sys.setrecursionlimit(3000)
    sys.dont_write_bytecode = True
    sys.path.insert(0, os.path.join(REF, "gitmodules", "gpytorch"))
    from gpytorch.utils import linear_cg  # noqa: F401
    sys.path.pop(0)
    for name in ["pyro", "pyro.nn", "pyro.distributions", "pyro.distributions.transforms",
                 "pyro.distributions.transforms.polynomial"]:
        mod = types.ModuleType(name)
        mod.__path__ = []
        sys.modules[name] = mod
    sys.modules["pyro.distributions.transforms.polynomial"].Polynomial = object
    sys.modules["pyro.nn"].AutoRegressiveNN = object
    sys.path.insert(0, REF)
    from config import get_config, get_schema, expand_grid
    from cmf.models import get_density
    from cmf.non_square_helpers import get_non_square_train_metrics
    _give_spline_bijection_a_jvp()
    return get_config, get_schema, expand_grid, get_density, get_non_square_train_metrics


def _give_spline_bijection_a_jvp():
    """The reference's ``RationalQuadraticSplineBijection`` (`bijections/nsf.py:19-42`) defines
    ``_x_to_z`` / ``_z_to_x`` only, so ``Bijection.forward("jvp")`` (`bijection.py:26-35`) cannot run and an
    ``nsf-c`` layer cannot sit on the non-square JVP stack.  For the ``*_nsfc_*`` fixtures
    the class is given the one definition a JVP can have -- the autograd JVP of its own ``flow.inverse``
    (nsf's unmodified transform) -- at generation time, in this process only; nothing under
    /root/reference is touched and every other line of the reference runs as it is."""
    import torch.autograd.functional as AF
    from cmf.models.components.bijections.nsf import RationalQuadraticSplineBijection

    def _jvp(self, z, v):
        x, jv = AF.jvp(lambda t: self.flow.inverse(t)[0], z, v, create_graph=torch.is_grad_enabled())
        return {"x": x, "jvp": jv}
    RationalQuadraticSplineBijection._jvp = _jvp


def jsonable(obj):
    return json.loads(json.dumps(obj, default=str))


class inject_hutchinson:
    """For ``hutch=True`` cases: while active, ``torch.randn`` (the reference's only draw inside
    ``_approx_log_det_jac_and_reconstruction``, non_square.py:206) returns the deterministic probes
    of ``cases.hutch_eps`` and records them.  The reference itself is not modified."""

    def __init__(self, name, case):
        self.name, self.on, self.drawn = name, bool(case.get("hutch")), []

    def __enter__(self):
        if self.on:
            from cases import hutch_eps
            self._orig = torch.randn

            def fake(*shape, **kw):
                eps = hutch_eps(self.name, shape)
                self.drawn.append(eps)
                return eps
            torch.randn = fake
        return self.drawn

    def __exit__(self, *exc):
        if self.on:
            torch.randn = self._orig
        return False


def run_case(name, case, api):
    from cases import make_input, grad_probe
    from detfill import fill_state_dict
    get_config, get_schema, expand_grid, get_density, get_train_metrics = api

    ds = case["dataset"]
    cfg = expand_grid({"model": "non-square", "dataset": ds,
                       **get_config(ds, "non-square", False), **case["overrides"]})[0]
    schema = get_schema(cfg)
    torch.manual_seed(0)
    x = make_input(case)
    density = get_density(schema, x)
    # checkerboard masks are expand()ed buffers (acl.py:73) that load_state_dict cannot write into
    filled = {k: v for k, v in fill_state_dict(density.state_dict()).items() if not k.endswith(".mask")}
    missing, unexpected = density.load_state_dict(filled, strict=False)
    assert not unexpected and all(k.endswith(".mask") for k in missing)
    density.train()

    # strip DataParallel (CPU: a plain passthrough) and the dequantiser (in-place rand, wrapper.py:27-30)
    top = density.module if hasattr(density, "module") else density
    if type(top).__name__ == "DequantizationDensity":
        top = top.density

    train_metrics = get_train_metrics(cfg)[0]
    t0 = time.time()
    with inject_hutchinson(name, case) as drawn:
        loss = train_metrics(top, x.clone(), case["epoch"])["loss"]
    loss.backward()
    dt = time.time() - t0

    grads = {}
    for pname, p in density.named_parameters():
        if p.grad is not None:
            key = pname[len("module."):] if pname.startswith("module.") else pname
            grads[key] = grad_probe(key, p.grad)

    # intermediates, through the head's own methods (no graph needed)
    head, xin = top, x.clone()
    while type(head).__name__ != "NonSquareHeadDensity":
        xin = head.bijection.x_to_z(xin)["z"]
        head = head.prior
    out = dict(name=name, case=jsonable({k: v for k, v in case.items()}), config=jsonable(cfg),
               schema=jsonable(schema), loss=loss.detach().clone(), grads=grads, ref_seconds=dt,
               x_pre=xin.detach().clone() if xin.numel() <= 4096 else None)
    with torch.no_grad():
        prior_dict = head.prior.elbo(xin)
        z, low_dim_elbo, _ = head._traverse_backward(xin, prior_dict)
        out.update(z=z.clone(), low_dim_elbo=low_dim_elbo.clone())
        logdet, recon, gram = head._exact_log_det_jac_and_reconstruction(z)
        out.update(logdet=logdet.clone(), gram=gram.clone(),
                   recon_loss=((recon - xin).flatten(1) ** 2).sum(-1, keepdim=True))
        off = ~torch.eye(gram.shape[1], dtype=torch.bool)
        out["l1_offdiag"] = gram.masked_select(off).view(gram.shape[0], -1).abs().sum(1, keepdim=True)
        out["l1_diag"] = torch.diagonal(gram, dim1=-2, dim2=-1).abs().sum(1, keepdim=True)
    if case.get("hutch"):
        assert len(drawn) == 1, "the reference draws its probes exactly once per elbo call"
        out["hutch_eps"] = drawn[0].clone()
        # the estimator's own outputs (non_square.py:203-258), no graph; same injected probes
        with torch.no_grad(), inject_hutchinson(name, case):
            est, recon, jtj_eps = head._approx_log_det_jac_and_reconstruction(z)
        out.update(hutch_estimate=est.clone(), hutch_jtj_eps=jtj_eps.clone())
    return out


def main(argv):
    from cases import CASES
    api = import_reference()
    out_dir = os.path.join(os.path.dirname(HERE), "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    names = argv or [n for n, c in CASES.items()]
    for name in names:
        t0 = time.time()
        res = run_case(name, CASES[name], api)
        path = os.path.join(out_dir, name + ".pt")
        torch.save(res, path)
        print(f"{name}: loss={float(res['loss']):.6f}  {os.path.getsize(path) / 1024:.1f} KiB  "
              f"{time.time() - t0:.1f}s", flush=True)


if __name__ == "__main__":
    main(sys.argv[1:])
