"""CPU: libmplb200.so builds for sm_100a, loads, and exports every symbol include/mplb200.h declares.
No compute calls here (no GPU); the product path must fail loudly without a device."""
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as g
    g.build()
    from matchinglib_poselib_b200 import _lib
    return _lib.load()


def _declared():
    names = set()
    for fn in os.listdir(os.path.join(ROOT, "include")):
        if fn.endswith(".h"):
            src = open(os.path.join(ROOT, "include", fn)).read()
            src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
            names |= set(re.findall(r"\b(mpl_[a-z0-9_]+)\s*\(", src))
    return names


def test_every_declared_symbol_is_exported(lib):
    from matchinglib_poselib_b200 import _lib
    declared = _declared()
    assert len(declared) >= 15
    for name in sorted(declared):
        assert hasattr(lib, name), "libmplb200.so does not export %s" % name
    assert declared == set(_lib.SIGNATURES), "ctypes table and header disagree: %s" % (declared ^ set(_lib.SIGNATURES))


def test_sm100a_cubin_present():
    import subprocess
    from matchinglib_poselib_b200 import _lib
    out = subprocess.run(["cuobjdump", "-lelf", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out


def test_fails_loudly_without_gpu(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import matchinglib_poselib_b200 as m
    with pytest.raises(m._lib.MplError):
        m.knn2_hamming(np.zeros((20, 32), np.uint8), np.zeros((20, 32), np.uint8))
    kp = [None] * 20
    err, _ = m.getMatchesArray(kp, kp, np.zeros((20, 32), np.uint8), np.zeros((20, 32), np.uint8), None, "LINEAR")
    assert err == -3                      # device failure maps to "matching failed", never to a CPU result
    err, _ = m.getMatchesArray(kp, kp, np.zeros((20, 32), np.uint8), np.zeros((20, 32), np.uint8), None, "HNSW")
    assert err == -2                      # unsupported matcher: no fallback
    err, _ = m.getMatchesArray(kp[:10], kp, np.zeros((10, 32), np.uint8), np.zeros((20, 32), np.uint8), None, "LINEAR")
    assert err == -4


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "matchinglib_poselib_b200")
    for dp, _, fns in os.walk(pkg):
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dp, fn)).read()
                assert "from oracle" not in txt and "import oracle" not in txt and "_oracle.so" not in txt, fn
