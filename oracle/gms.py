"""ctypes front end of oracle/gms_oracle.c and of the reference's own GMS build (oracle/_ref/libgms_ref.so).
TEST INFRASTRUCTURE ONLY (see gms_oracle.c)."""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libgms_oracle.so")
REF_LIB = os.path.join(HERE, "_ref", "libgms_ref.so")
_lib = None
_ref = None


def build(ref=True):
    src = os.path.join(HERE, "gms_oracle.c")
    if not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", HERE, "libgms_oracle.so"])
    if ref and not os.path.exists(REF_LIB) and os.path.isdir("/root/reference/matchinglib_poselib"):
        subprocess.check_call(["make", "-s", "-C", HERE, "ref_gms"])


def lib():
    global _lib
    if _lib is None:
        build(ref=False)
        _lib = ctypes.CDLL(LIB)
    return _lib


def ref_lib():
    global _ref
    if _ref is None and os.path.exists(REF_LIB):
        _ref = ctypes.CDLL(REF_LIB)
    return _ref


def _call(fn, kp1, size1, kp2, size2, matches_qt, use_scale, use_rotation):
    kp1 = np.ascontiguousarray(kp1, np.float32); kp2 = np.ascontiguousarray(kp2, np.float32)
    m = np.ascontiguousarray(matches_qt, np.int32).reshape(-1, 2)
    mask = np.zeros(max(len(m), 1), np.uint8)
    p = lambda a: ctypes.c_void_p(a.ctypes.data)
    n = fn(p(kp1), len(kp1), int(size1[0]), int(size1[1]), p(kp2), len(kp2), int(size2[0]), int(size2[1]), p(m), len(m),
           int(bool(use_scale)), int(bool(use_rotation)), p(mask))
    return int(n), mask[: len(m)]


def filter_matches(kp1, size1, kp2, size2, matches_qt, use_scale=False, use_rotation=False):
    """filterMatchesGMS (gms.cpp:54-60).  kp: [n, 2] float32 pixel coordinates; size = (width, height); matches_qt [m, 2]
    (queryIdx, trainIdx).  Returns (n_inliers, mask uint8 [m])."""
    return _call(lib().orc_gms_filter, kp1, size1, kp2, size2, matches_qt, use_scale, use_rotation)


def ref_filter_matches(kp1, size1, kp2, size2, matches_qt, use_scale=False, use_rotation=False):
    """The same through the reference's own MatchGMS.cpp (None when oracle/_ref/libgms_ref.so was not built)."""
    R = ref_lib()
    if R is None:
        return None
    return _call(R.ref_gms_filter, kp1, size1, kp2, size2, matches_qt, use_scale, use_rotation)
