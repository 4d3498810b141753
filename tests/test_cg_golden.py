"""The two restatements of the reference's conjugate-gradient dependency -- the product's host loop
``cmf_b200.cg.linear_cg_batched`` and the oracle's ``batched_cg`` -- against outputs of
``gpytorch.utils.linear_cg`` itself (vendored in the reference, called as `non_square.py:241-247`
calls it), stored by ``oracle/gen_cg_golden.py``.  fp64, so the comparison is of control flow
(normalisation, safe division, frozen columns, early exit), not of rounding."""
import pytest
import torch

from conftest import load_golden
import cmf_oracle as O
from cmf_b200.cg import linear_cg_batched

CASES = ["well_n6", "zero_rhs_n10_m2", "n24_early_exit", "n12_tight", "n12_capped"]


@pytest.mark.parametrize("name", CASES)
@pytest.mark.parametrize("impl", ["product", "oracle"])
def test_cg_matches_gpytorch(name, impl):
    c = load_golden("cg_gpytorch")[name]
    fn = linear_cg_batched if impl == "product" else O.batched_cg
    sol = fn(lambda t: c["A"] @ t, c["rhs"], max_iter=c["max_iter"], tolerance=c["tolerance"])
    err = float((sol - c["solution"]).abs().max() / c["solution"].abs().max())
    assert err < 1e-8, err


@pytest.mark.parametrize("name", CASES)
def test_cg_sync_free_form_equals_the_host_loop(name):
    """Under CUDA-graph capture the early exits become a device-side flag (cmf_b200.cg._cg, sync_free=True):
    bit-identical solutions, on the gpytorch cases (early exit, frozen columns, zero right-hand side, capped)."""
    from cmf_b200.cg import _cg
    c = load_golden("cg_gpytorch")[name]
    mm = lambda t: c["A"] @ t
    a = _cg(mm, c["rhs"], c["max_iter"], c["tolerance"], 1e-10, 1e-10, False)
    b = _cg(mm, c["rhs"], c["max_iter"], c["tolerance"], 1e-10, 1e-10, True)
    assert torch.equal(a, b)
    a32 = _cg(lambda t: c["A"].float() @ t, c["rhs"].float(), c["max_iter"], c["tolerance"], 1e-10, 1e-10, False)
    b32 = _cg(lambda t: c["A"].float() @ t, c["rhs"].float(), c["max_iter"], c["tolerance"], 1e-10, 1e-10, True)
    assert torch.equal(a32, b32)
