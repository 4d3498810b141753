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
