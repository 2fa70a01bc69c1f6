"""Host-side mirror of the reference's poselib API for the hot path.

`estimateEssentialMat` keeps the reference's name, argument meaning and bool return
(matchinglib_poselib/source/poselib/include/poselib/pose_estim.h:204-210, source/pose_estim.cpp:857-890) for
method == "RANSAC" (findEssentialMat, five-point-nister/five-point.cpp:69-148).  cv::Mat -> numpy arrays:
p1, p2 are n x 2 (camera-normalised coordinates); E is returned 3 x 3 float64, mask 1 x n uint8.
All O(models x correspondences) arithmetic runs in libmplb200.so on the GPU.
"""
import ctypes

import numpy as np

from . import _lib


def _pts(p):
    a = np.asarray(p)
    if a.ndim == 3 and a.shape[1] == 1:
        a = a.reshape(a.shape[0], a.shape[2])
    if a.ndim != 2 or a.shape[1] != 2:
        raise AssertionError("CV_Assert(npoints >= 0 && points.checkVector(2) == npoints)")
    return np.ascontiguousarray(a, np.float64)            # findEssentialMat converts to CV_64F (five-point.cpp:89-90)


def ransac5pt(p1, p2, samples, threshold, max_iters=1000, confidence=0.999, refine=False, trace=False):
    """runRANSAC over PRE-DRAWN minimal samples (h x 5 int32).  Returns a dict: ok, E (3x3), mask (n,), count,
    iters and, with trace=True, n_models (h,), models (M x 9), counts (M,), errsums (M,) for every sample."""
    lib = _lib.load()
    p1, p2 = _pts(p1), _pts(p2)
    n = p1.shape[0]
    if p2.shape[0] != n:
        raise AssertionError("CV_Assert(points2.checkVector(2) == npoints)")
    s = None if samples is None else np.ascontiguousarray(samples, np.int32)
    h = 0 if s is None else s.shape[0]
    E = np.zeros(9); mask = np.zeros(n, np.uint8); cnt = ctypes.c_int(0); iters = ctypes.c_int(0)
    hh = h if s is not None else max_iters
    tm = np.zeros((10 * hh, 9)) if trace else None
    tn = np.zeros(hh, np.int32) if trace else None
    tc = np.zeros(10 * hh, np.int32) if trace else None
    te = np.zeros(10 * hh) if trace else None
    _lib.check(lib.mpl_ransac5pt(_lib.ptr(p1), _lib.ptr(p2), n, _lib.ptr(s), h, float(threshold), int(max_iters),
                                 float(confidence), int(bool(refine)), _lib.ptr(E), _lib.ptr(mask), ctypes.byref(cnt),
                                 ctypes.byref(iters), _lib.ptr(tm), _lib.ptr(tn), _lib.ptr(tc), _lib.ptr(te)), "mpl_ransac5pt")
    out = dict(ok=cnt.value > 0, E=E.reshape(3, 3), mask=mask, count=cnt.value, iters=iters.value)
    if trace:
        tot = int(tn.sum())
        out.update(n_models=tn, models=tm[:tot], counts=tc[:tot], errsums=te[:tot])
    return out


def estimateEssentialMat(p1, p2, method="ARRSAC", threshold=1.6, refine=True, want_mask=True):
    """poselib::estimateEssentialMat.  Returns (ok, E, mask).  Only method "RANSAC" is served (the reference's
    default "ARRSAC" and "LMEDS"/"USAC" are other algorithms; there is no CPU fallback)."""
    lib = _lib.load()
    p1, p2 = _pts(p1), _pts(p2)
    n = p1.shape[0]
    if n < 5 or p2.shape[0] != n:
        raise AssertionError("CV_Assert(npoints >= 5 && points2.checkVector(2) == npoints)")
    E = np.zeros((10, 9)); n_E = ctypes.c_int(0)
    mask = np.zeros(n, np.uint8)
    rc = lib.mpl_estimate_essential_api(_lib.ptr(p1), _lib.ptr(p2), n, method.encode(), float(threshold), int(bool(refine)),
                                        _lib.ptr(E), ctypes.byref(n_E), _lib.ptr(mask))
    if rc == -103:
        raise AssertionError(_lib.last_error())
    if rc < 0:
        raise _lib.MplError("estimateEssentialMat failed (%d): %s" % (rc, _lib.last_error()))
    if n == 5:
        return bool(rc), E[: n_E.value].reshape(-1, 3), mask.reshape(1, -1)
    return bool(rc), E[0].reshape(3, 3), (mask.reshape(1, -1) if want_mask else None)


def getPoseTriangPts(E, p1, p2, mask=None, dist=50.0, translatE=False):
    """poselib::getPoseTriangPts (pose_estim.h:192-200, pose_estim.cpp:913-946 -> recoverPose, five-point.cpp:150-338).
    Returns (n_valid, R 3x3, t 3x1, Q n x 3, mask 1 x n uint8).  `mask` (optional, 1 x n or n) is the inlier mask
    of estimateEssentialMat: only its non-zero entries count and the returned mask is its AND with the cheirality /
    distance tests; without it every correspondence counts.  The per-point work runs on the GPU."""
    lib = _lib.load()
    E = np.ascontiguousarray(np.asarray(E, np.float64).reshape(3, 3))
    p1, p2 = _pts(p1), _pts(p2)
    n = p1.shape[0]
    if p2.shape[0] != n:
        raise AssertionError("CV_Assert(npoints >= 0 && points2.checkVector(2) == npoints)")
    given = mask is not None
    m = np.ascontiguousarray(np.asarray(mask, np.uint8).reshape(-1)).copy() if given else np.zeros(n, np.uint8)
    if m.shape[0] != n:
        raise AssertionError("mask must hold one byte per correspondence")
    R = np.zeros((3, 3)); t = np.zeros((3, 1)); Q = np.zeros((max(n, 1), 3))
    rc = lib.mpl_get_pose_triang_pts_api(_lib.ptr(E), _lib.ptr(p1), _lib.ptr(p2), n, _lib.ptr(m), int(given), float(dist),
                                         int(bool(translatE)), _lib.ptr(R), _lib.ptr(t), _lib.ptr(Q))
    if rc == -103:
        raise AssertionError(_lib.last_error())
    if rc < 0:
        raise _lib.MplError("getPoseTriangPts failed (%d): %s" % (rc, _lib.last_error()))
    return rc, R, t, Q[:n], m.reshape(1, -1)


DMATCH = np.dtype([("queryIdx", "<i4"), ("trainIdx", "<i4"), ("imgIdx", "<i4"), ("distance", "<f4")])


def matchAndEstimatePose(desc1, desc2, pts1, pts2, matcher_name="LINEAR", ratioTest=True, threshold=1.6, n_samples=1000,
                         confidence=0.999, refine=True, seed=1):
    """Both halves of the hot path for one image pair with the matches kept on the GPU in between
    (mpl_match_pose_pair; the reference driver's getCorrespondences -> estimateEssentialMat sequence,
    tests/poselib-test/main.cpp:1522-1599).  desc: uint8 [n, nbytes]; pts: [n, 2] camera-normalised keypoint coordinates.
    Returns dict(matches (cv::DMatch records), samples [h, 5] drawn on the device, ok, E, mask (over matches), count, iters)."""
    lib = _lib.load()
    d1 = np.ascontiguousarray(desc1, np.uint8); d2 = np.ascontiguousarray(desc2, np.uint8)
    p1, p2 = _pts(pts1), _pts(pts2)
    if d1.ndim != 2 or d2.ndim != 2 or d1.shape[1] != d2.shape[1] or p1.shape[0] != d1.shape[0] or p2.shape[0] != d2.shape[0]:
        raise AssertionError("descriptor / keypoint shapes do not agree")
    n1 = d1.shape[0]
    matches = np.zeros(max(n1, 1), DMATCH); samples = np.zeros((max(n_samples, 1), 5), np.int32)
    E = np.zeros(9); mask = np.zeros(max(n1, 1), np.uint8)
    nm, ns, cnt, iters = ctypes.c_int(0), ctypes.c_int(0), ctypes.c_int(0), ctypes.c_int(0)
    _lib.check(lib.mpl_match_pose_pair(_lib.ptr(d1), n1, d1.strides[0], _lib.ptr(d2), d2.shape[0], d2.strides[0], d1.shape[1],
                                       _lib.ptr(p1), _lib.ptr(p2), matcher_name.encode(), int(bool(ratioTest)), float(threshold),
                                       int(n_samples), float(confidence), int(bool(refine)), int(seed), _lib.ptr(matches),
                                       ctypes.byref(nm), _lib.ptr(samples), ctypes.byref(ns), _lib.ptr(E), _lib.ptr(mask),
                                       ctypes.byref(cnt), ctypes.byref(iters)), "mpl_match_pose_pair")
    return dict(matches=matches[: nm.value], samples=samples[: ns.value], ok=cnt.value > 0, E=E.reshape(3, 3),
                mask=mask[: nm.value], count=cnt.value, iters=iters.value)


def _smoke():
    """One small RANSAC run on cuda:0, checked against the oracle (called by __graft_entry__.smoke)."""
    import os
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(root, "tests"))
    import datagen_pose
    from oracle import ransac as orc
    sc = datagen_pose.make_scene(400, 0.3, seed=5)
    s = datagen_pose.draw_samples(sc["p1"], sc["p2"], 64, seed=6)
    got = ransac5pt(sc["p1"], sc["p2"], s, sc["th"], refine=True)
    exp = orc.ransac5pt(sc["p1"], sc["p2"], s, sc["th"], refine=True)
    assert got["count"] == exp["count"] and np.array_equal(got["mask"], exp["mask"]), "RANSAC inlier set differs from oracle"
    d = min(np.linalg.norm(got["E"] - exp["E"]), np.linalg.norm(got["E"] + exp["E"]))
    assert d < 1e-6, "E differs from oracle by %g" % d
