"""VFC oracle: the reference's OWN vector-field-consensus filter -- matchinglib/source/vfcMatches.cpp (filterWithVFC) and
source/vfc.cpp (class VFC, SparseVFC: Ma et al., "Robust point matching via vector field consensus", TIP 2014) -- compiled
unmodified into oracle/_ref/libvfc_ref.so (`make -C oracle ref_vfc`) against a float cv::Mat / cv::solve stand-in
(oracle/ref_vfc/stub).  TEST INFRASTRUCTURE ONLY.

Parity of the GPU filter against it is within float tolerance, not bitwise: the EM iteration is float32 with long sequential
sums (the reference) against tree sums (the device) and expf implementations that differ in the last ulp, so a match whose
posterior probability ends within ~1e-4 of the 0.75 decision threshold can flip."""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_LIB = os.path.join(HERE, "_ref", "libvfc_ref.so")
_ref = None


def build():
    if not os.path.exists(REF_LIB) and os.path.isdir("/root/reference/matchinglib_poselib"):
        subprocess.check_call(["make", "-s", "-C", HERE, "ref_vfc"])


def ref_lib():
    global _ref
    if _ref is None and os.path.exists(REF_LIB):
        _ref = ctypes.CDLL(REF_LIB)
    return _ref


def filter_with_vfc(kp1, kp2, matches_qt, seed=1):
    """filterWithVFC(keypL, keypR, matches_in, matches_out) (vfcMatches.cpp:49-86) after srand(seed).  kp1 / kp2: (n, 2) float32 pixel
    coordinates, matches_qt: (m, 2) int32 (queryIdx, trainIdx).  Returns (code, keep uint8 [m]) or None without the reference build."""
    R = ref_lib()
    if R is None:
        return None
    kp1 = np.ascontiguousarray(kp1, np.float32); kp2 = np.ascontiguousarray(kp2, np.float32)
    mt = np.ascontiguousarray(matches_qt, np.int32).reshape(-1, 2)
    keep = np.zeros(max(len(mt), 1), np.uint8); nk = ctypes.c_int(0)
    rc = R.vfc_ref_filter(ctypes.c_void_p(kp1.ctypes.data), len(kp1), ctypes.c_void_p(kp2.ctypes.data), len(kp2), ctypes.c_void_p(mt.ctypes.data), len(mt),
                          ctypes.c_uint(int(seed) & 0xffffffff), ctypes.c_void_p(keep.ctypes.data), ctypes.byref(nk))
    return int(rc), keep[:len(mt)]
