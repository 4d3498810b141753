"""The C-ABI library loads and exports every symbol include/cmf_b200.h declares (no compute)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib_path():
    from cmf_b200 import build
    return build.build()          # no-op when the in-tree .so is current


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "cmf_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(cmf_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported(lib_path):
    lib = ctypes.CDLL(lib_path)
    names = declared_symbols()
    assert len(names) >= 10
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/cmf_b200.h but not exported"


def test_binding_table_matches_header(lib_path):
    from cmf_b200 import _lib
    assert sorted(_lib.SIGNATURES) == declared_symbols()
    lib = _lib.load()
    assert lib.cmf_version() == _lib.ABI_VERSION


def test_bad_arguments_return_negative_without_gpu(lib_path):
    """Argument validation happens before any CUDA call, so it is testable on the build box."""
    from cmf_b200 import _lib
    lib = _lib.load()
    rc = lib.cmf_coupling_jvp_fwd(4, 2, 8, 4, None, None, None, None, None, None, None, None)   # Q < M
    assert rc < 0 and b"bad sizes" in lib.cmf_last_error_string()
    rc = lib.cmf_chol_logdet_fwd(4, 1000, None, None, None, None, None, None, None)
    assert rc < 0
    with pytest.raises(_lib.CmfB200Error):
        _lib.check(rc, "cmf_chol_logdet_fwd")


def test_sass_is_sm100a(lib_path):
    import shutil, subprocess
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    out = subprocess.run([cuobjdump, "-lelf", lib_path], capture_output=True, text=True).stdout
    assert "sm_100a" in out
