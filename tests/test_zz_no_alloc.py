"""``cmf_coupling_jvp_bwd_static`` -- the allocation-free variant of the coupling adjoint (scratch in a
file-scope ``static __device__`` array; no counterpart in the reference)
-- against the existing ``cmf_coupling_jvp_bwd`` entry point, through the C ABI.

Collected last on purpose (``zz``).  Passes alone in a fresh
process and after the other tests (it relies on no state but the library's own).
"""
import pytest
import torch

from cmf_b200 import _lib

DEV = "cuda"


def _rand(*shape, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed + sum(shape))
    return ((torch.rand(*shape, generator=g) * 2 - 1) * scale).float()


def _inputs(B, d, M, Q, seed):
    t = dict(z=_rand(B, M, seed=seed + 1), st=_rand(B, 2 * Q, seed=seed + 2, scale=0.7),
             V=_rand(B, d, M, seed=seed + 3), ST=_rand(B, d, 2 * Q, seed=seed + 4),
             gx=_rand(B, M, seed=seed + 5), gV=_rand(B, d, M, seed=seed + 6))
    map_ = inv = None
    if M != Q:
        g = torch.Generator().manual_seed(seed)
        q = torch.randperm(Q, generator=g)[:M].sort().values
        inv = torch.full((Q,), -1, dtype=torch.int32)
        inv[q] = torch.arange(M, dtype=torch.int32)
        map_, inv = q.int().to(DEV), inv.to(DEV)
    return {k: v.to(DEV) for k, v in t.items()}, map_, inv


def _outputs(B, d, M, Q, fill):
    mk = lambda *s: torch.full(s, fill, dtype=torch.float32, device=DEV)
    return mk(B, M), mk(B, d, M), mk(B, 2 * Q), mk(B, d, 2 * Q)


def _ptr(t):
    return None if t is None else t.data_ptr()


def _run_pair(lib, B, d, M, Q, seed):
    """one call of each entry point on the same inputs; returns (existing outputs, static outputs)"""
    t, map_, inv = _inputs(B, d, M, Q, seed)
    stream = torch.cuda.current_stream().cuda_stream
    nbytes = lib.cmf_coupling_jvp_bwd_workspace(B, d, M, Q)
    assert 0 < nbytes <= lib.cmf_coupling_jvp_bwd_static_max_bytes(), "shape must exercise the scratch array"
    ws = torch.empty(nbytes // 4, dtype=torch.float32, device=DEV)
    ref = _outputs(B, d, M, Q, float("nan"))
    _lib.check(lib.cmf_coupling_jvp_bwd(B, d, M, Q, _ptr(t["z"]), _ptr(t["V"]), _ptr(t["st"]), _ptr(t["ST"]),
                                        _ptr(map_), _ptr(inv), _ptr(t["gx"]), _ptr(t["gV"]),
                                        *[_ptr(o) for o in ref], _ptr(ws), nbytes, stream), "cmf_coupling_jvp_bwd")
    got = _outputs(B, d, M, Q, float("nan"))      # outputs are the caller's; the call itself allocates nothing
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    _lib.check(lib.cmf_coupling_jvp_bwd_static(B, d, M, Q, _ptr(t["z"]), _ptr(t["V"]), _ptr(t["st"]), _ptr(t["ST"]),
                                               _ptr(map_), _ptr(inv), _ptr(t["gx"]), _ptr(t["gV"]),
                                               *[_ptr(o) for o in got], stream), "cmf_coupling_jvp_bwd_static")
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    return ref, got


def _check(ref, got):
    for name, a, b in zip(("g_z", "g_Vin", "g_st", "g_ST"), got, ref):
        assert not torch.isnan(a).any(), name
        err = float((a.double() - b.double()).abs().max())
        scale = float(b.double().abs().max())
        # the tolerance of tests/test_kernels_gpu.py::test_coupling_fwd_bwd; the same kernels in the same
        # order give the same bits, which is what is observed
        assert err <= 2e-5 * scale + 2e-6, f"{name}: err {err:.3e} scale {scale:.3e}"


@pytest.mark.gpu
def test_static_scratch_matches_workspace_entry_point_two_calls_in_a_row():
    lib = _lib.load()
    # two calls in a row on different shapes (the second, smaller and mapped, finds the first call's
    # partials in the shared array and must not be affected by them), then the first shape again
    # (the 128-bit path has since learnt to reduce its tangent chunks inside the CTA and leaves the scratch alone at this
    # shape; the mapped call below is the one that writes and reads it)
    first = _run_pair(lib, 50, 64, 392, 392, seed=10)       # the MNIST-shape layer, 128-bit path
    second = _run_pair(lib, 7, 64, 196, 392, seed=20)       # checkerboard map, 32-bit path
    third = _run_pair(lib, 50, 64, 392, 392, seed=30)
    for ref, got in (first, second, third):
        _check(ref, got)


@pytest.mark.gpu
def test_static_scratch_without_tangent_split_needs_no_scratch():
    lib = _lib.load()
    B, d, M, Q = 4000, 2, 64, 64                           # one tangent chunk: workspace query says 0 bytes
    assert lib.cmf_coupling_jvp_bwd_workspace(B, d, M, Q) == 0
    t, _, _ = _inputs(B, d, M, Q, 40)
    stream = torch.cuda.current_stream().cuda_stream
    ref, got = _outputs(B, d, M, Q, float("nan")), _outputs(B, d, M, Q, float("nan"))
    _lib.check(lib.cmf_coupling_jvp_bwd(B, d, M, Q, _ptr(t["z"]), _ptr(t["V"]), _ptr(t["st"]), _ptr(t["ST"]), None, None,
                                        _ptr(t["gx"]), _ptr(t["gV"]), *[_ptr(o) for o in ref], None, 0, stream), "bwd")
    _lib.check(lib.cmf_coupling_jvp_bwd_static(B, d, M, Q, _ptr(t["z"]), _ptr(t["V"]), _ptr(t["st"]), _ptr(t["ST"]), None,
                                               None, _ptr(t["gx"]), _ptr(t["gV"]), *[_ptr(o) for o in got], stream), "static")
    torch.cuda.synchronize()
    _check(ref, got)


def test_static_scratch_rejects_sizes_beyond_the_maximum():
    """argument validation precedes every CUDA call: runs on the build box too"""
    lib = _lib.load()
    assert lib.cmf_coupling_jvp_bwd_static_max_bytes() == 4 << 20
    B, d, M, Q = 512, 64, 392, 392
    assert lib.cmf_coupling_jvp_bwd_workspace(B, d, M, Q) > lib.cmf_coupling_jvp_bwd_static_max_bytes()
    rc = lib.cmf_coupling_jvp_bwd_static(B, d, M, Q, *([None] * 13))
    assert rc < 0
    msg = lib.cmf_last_error_string().decode()
    assert "static array holds" in msg and "needs" in msg, msg
    with pytest.raises(_lib.CmfB200Error):
        _lib.check(rc, "cmf_coupling_jvp_bwd_static")
