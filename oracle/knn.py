"""ctypes front end of oracle/knn_oracle.c (+ the reference-NMSLIB build in oracle/_ref/).

TEST INFRASTRUCTURE ONLY -- see oracle/knn_oracle.c for the reference file:line each function follows.
"""
import ctypes
import os
import subprocess
import threading

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libknn_oracle.so")
REF_LIB = os.path.join(HERE, "_ref", "libnmslib_ref.so")
DMATCH_DTYPE = np.dtype([("queryIdx", "<i4"), ("trainIdx", "<i4"), ("imgIdx", "<i4"), ("distance", "<f4")])

_lib = None
_ref = None


def build(ref=True):
    """Compile the C restatement; compile the reference's own NMSLIB sources when /root/reference exists."""
    src = os.path.join(HERE, "knn_oracle.c")
    if not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", HERE, "libknn_oracle.so"])
    if ref and not os.path.exists(REF_LIB) and os.path.isdir("/root/reference/matchinglib_poselib"):
        subprocess.check_call(["make", "-s", "-C", HERE, "ref"])


def lib():
    global _lib
    if _lib is None:
        build(ref=False)
        _lib = ctypes.CDLL(LIB)
    return _lib


def ref_lib():
    """The reference's NMSLIB (oracle/_ref/libnmslib_ref.so) or None when it was never built."""
    global _ref
    if _ref is None and os.path.exists(REF_LIB):
        _ref = ctypes.CDLL(REF_LIB)
    return _ref


def _p(a):
    return ctypes.c_void_p(a.ctypes.data)


def _run_striped(fn, nq, threads):
    """Query striping over host threads (ctypes releases the GIL).  Contiguous stripes: results are
    per-query independent, so the partition does not change them (nmslib_matchers.h:112-113 uses q % T)."""
    if threads <= 1 or nq < 2 * threads:
        fn(0, nq)
        return
    bounds = np.linspace(0, nq, threads + 1).astype(int)
    ts = [threading.Thread(target=fn, args=(int(bounds[i]), int(bounds[i + 1]))) for i in range(threads)]
    [t.start() for t in ts]
    [t.join() for t in ts]


def knn_hamming(q, t, k=2, mode=0, threads=1):
    """mode 0 = LINEAR (cvflann HammingLUT, all bytes); 1 = BRUTEFORCENMS (NMSLIB, last packed word dropped)."""
    q = np.ascontiguousarray(q, np.uint8)
    t = np.ascontiguousarray(t, np.uint8)
    nq, nb = q.shape
    nt = t.shape[0]
    idx = np.full((nq, k), -1, np.int32)
    dist = np.full((nq, k), np.iinfo(np.int32).max, np.int32)
    L = lib()

    def work(q0, q1):
        L.orc_knn_hamming(_p(q), nq, ctypes.c_size_t(nb), _p(t), nt, ctypes.c_size_t(nb), nb, mode, k, q0, q1, _p(idx), _p(dist))
    if nq:
        _run_striped(work, nq, threads)
    return idx, dist


def knn_l2(q, t, k=2, mode=0, threads=1):
    """mode 0 = LINEAR (cvflann L2<float>, squared); 1 = BRUTEFORCENMS (NMSLIB l2, sqrt)."""
    q = np.ascontiguousarray(q, np.float32)
    t = np.ascontiguousarray(t, np.float32)
    nq, d = q.shape
    nt = t.shape[0]
    idx = np.full((nq, k), -1, np.int32)
    dist = np.full((nq, k), np.finfo(np.float32).max, np.float32)
    L = lib()

    def work(q0, q1):
        L.orc_knn_l2(_p(q), nq, ctypes.c_size_t(4 * d), _p(t), nt, ctypes.c_size_t(4 * d), d, mode, k, q0, q1, _p(idx), _p(dist))
    if nq:
        _run_striped(work, nq, threads)
    return idx, dist


def ratio_matches(idx, dist, ratio_test=True):
    """matchers.cpp:601-625 / 677-701: DMatch array in ascending queryIdx."""
    nq, k = idx.shape
    out = np.zeros(max(nq, 1), DMATCH_DTYPE)
    idx = np.ascontiguousarray(idx, np.int32)
    if dist.dtype == np.int32:
        n = lib().orc_ratio_matches_i32(_p(idx), _p(np.ascontiguousarray(dist)), nq, k, int(ratio_test), _p(out))
    else:
        n = lib().orc_ratio_matches_f32(_p(idx), _p(np.ascontiguousarray(dist, np.float32)), nq, k, int(ratio_test), _p(out))
    return out[:n].copy()


def get_matches(desc1, desc2, matcher_name="LINEAR", ratio_test=True, n_kp1=None, n_kp2=None, threads=1):
    """getMatches restated for LINEAR / BRUTEFORCENMS (matchers.cpp:115-135, 476-519, 525-714): (err, matches)."""
    n_kp1 = desc1.shape[0] if n_kp1 is None else n_kp1
    n_kp2 = desc2.shape[0] if n_kp2 is None else n_kp2
    empty = np.zeros(0, DMATCH_DTYPE)
    if n_kp1 < 15 or n_kp2 < 15:
        return -4, empty
    if n_kp1 != desc1.shape[0] or n_kp2 != desc2.shape[0]:
        return -1, empty
    if matcher_name not in ("LINEAR", "BRUTEFORCENMS"):
        return -2, empty
    if desc1.dtype not in (np.uint8, np.float32):
        return -1, empty
    mode = 0 if matcher_name == "LINEAR" else 1
    k = (2 if ratio_test else 1) if mode == 0 else 2
    if desc1.dtype == np.uint8:
        idx, dist = knn_hamming(desc1, desc2, k, mode, threads)
    else:
        idx, dist = knn_l2(desc1, desc2, k, mode, threads)
    m = ratio_matches(idx, dist, ratio_test)
    if mode == 0 and len(m) < 2:
        return -3, m
    return 0, m


def ref_nmslib_hamming(q, t, threads=8):
    R = ref_lib()
    q = np.ascontiguousarray(q, np.uint8); t = np.ascontiguousarray(t, np.uint8)
    nq, nb = q.shape
    idx = np.zeros((nq, 2), np.int32); dist = np.zeros((nq, 2), np.int32)
    R.ref_nmslib_knn2_hamming(_p(q), nq, ctypes.c_size_t(nb), _p(t), t.shape[0], ctypes.c_size_t(nb), nb, threads, _p(idx), _p(dist))
    return idx, dist


def ref_nmslib_l2(q, t, threads=8):
    R = ref_lib()
    q = np.ascontiguousarray(q, np.float32); t = np.ascontiguousarray(t, np.float32)
    nq, d = q.shape
    idx = np.zeros((nq, 2), np.int32); dist = np.zeros((nq, 2), np.float32)
    R.ref_nmslib_knn2_l2(_p(q), nq, ctypes.c_size_t(4 * d), _p(t), t.shape[0], ctypes.c_size_t(4 * d), d, threads, _p(idx), _p(dist))
    return idx, dist
