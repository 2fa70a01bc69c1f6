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


def lmeds5pt(p1, p2, samples=None, confidence=0.999, max_iters=2000, trace=False):
    """runLMeDS (modelest.cpp:483-564) through mpl_lmeds5pt; samples: PRE-DRAWN h x 5 int32, or None to draw the reference's fixed
    iteration count inside.  Returns dict(ok, E, mask, count, min_median, sigma[, medians, n_models])."""
    lib = _lib.load()
    p1, p2 = _pts(p1), _pts(p2)
    n = p1.shape[0]
    if p2.shape[0] != n:
        raise AssertionError("CV_Assert(m1.size() == m2.size() && m1.cols == mask.cols)")
    s = None if samples is None else np.ascontiguousarray(samples, np.int32)
    h = 0 if s is None else s.shape[0]
    hh = h if s is not None else max_iters
    E = np.zeros(9); mask = np.zeros(n, np.uint8); cnt = ctypes.c_int(0)
    mm, sg = ctypes.c_double(0), ctypes.c_double(0)
    med = np.zeros(10 * max(hh, 1)) if trace else None
    nm = np.zeros(max(hh, 1), np.int32) if trace else None
    _lib.check(lib.mpl_lmeds5pt(_lib.ptr(p1), _lib.ptr(p2), n, _lib.ptr(s), h, float(confidence), int(max_iters), _lib.ptr(E),
                                _lib.ptr(mask), ctypes.byref(cnt), ctypes.byref(mm), ctypes.byref(sg), _lib.ptr(med), _lib.ptr(nm)),
               "mpl_lmeds5pt")
    out = dict(ok=cnt.value >= 5, E=E.reshape(3, 3), mask=mask, count=cnt.value, min_median=mm.value, sigma=sg.value)
    if trace:
        out.update(n_models=nm, medians=med[: int(nm.sum())])
    return out


def estimateEssentialMat(p1, p2, method="ARRSAC", threshold=1.6, refine=True, want_mask=True):
    """poselib::estimateEssentialMat.  Returns (ok, E, mask).  Methods "ARRSAC" (the reference's default; mpl_arrsac5pt),
    "RANSAC" and "LMEDS" (no refinement, like the reference) are served; "USAC" has its own function
    (estimateEssentialOrPoseUSAC), anything else raises (no CPU fallback)."""
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


def arrsac5pt(p1, p2, threshold, refine=True, seed=None):
    """runARRSAC (modelest.cpp:197-341) through mpl_arrsac5pt.  seed = (prosac_state, random_state) restarts the two sampler
    streams (None: they keep running, like the reference's static generators).  Returns dict(ok, E, mask, count,
    n_hypotheses, n_batches)."""
    lib = _lib.load()
    p1, p2 = _pts(p1), _pts(p2)
    n = p1.shape[0]
    if n < 5 or p2.shape[0] != n:
        raise AssertionError("CV_Assert(npoints >= 5 && points2.checkVector(2) == npoints)")
    if seed is not None:
        lib.mpl_arrsac_set_seed(int(seed[0]), int(seed[1]))
    E = np.zeros(9); mask = np.zeros(n, np.uint8)
    cnt, nh, nb = ctypes.c_int(0), ctypes.c_int(0), ctypes.c_int(0)
    _lib.check(lib.mpl_arrsac5pt(_lib.ptr(p1), _lib.ptr(p2), n, float(threshold), int(bool(refine)), _lib.ptr(E), _lib.ptr(mask),
                                 ctypes.byref(cnt), ctypes.byref(nh), ctypes.byref(nb)), "mpl_arrsac5pt")
    return dict(ok=cnt.value > 0, E=E.reshape(3, 3), mask=mask, count=cnt.value, n_hypotheses=nh.value, n_batches=nb.value)


def usac5pt(p1, p2, threshold, sorted_idx=None, seed=1, estimator="NISTER", **overrides):
    """estimateEssentialMatUsac (usac/usac_estimations.cpp:283-735) through mpl_usac5pt: USAC with SPRT verification, LOSAC local
    optimisation (8-point with Torr weights) and uniform (sorted_idx=None) or PROSAC sampling (sorted_idx = correspondence
    indices by increasing match distance).  `overrides` set fields of mpl_usac_params (sprt_delta, sprt_epsilon, prosac_beta,
    sprt_mS, ...).  Returns dict(ok, E, mask, count, hyp_count, model_count, rejected_sample_count, rejected_model_count,
    total_points_verified, num_local_optimizations, sprt_delta, sprt_epsilon, n_batches, n_launches)."""
    lib = _lib.load()
    p1, p2 = _pts(p1), _pts(p2)
    n = p1.shape[0]
    if p2.shape[0] != n:
        raise AssertionError("CV_Assert((p1.cols == 2) && (p2.cols == 2) && (p1.rows == p2.rows))")
    prm = _lib.UsacParams()
    lib.mpl_usac_params_default(ctypes.byref(prm), float(threshold))
    est = {"NISTER": 0, "STEWENIUS": 2}.get(str(estimator).upper())
    if est is None:
        raise _lib.MplError("usac5pt: estimator %r is not built (NISTER / STEWENIUS only; no CPU fallback)" % (estimator,))
    prm.estimator = est
    if est == 2:
        prm.sprt_tM, prm.sprt_mS = 2736.0, 6.0                     # usac_estimations.cpp:392-421 (first call: avgModels = 6)
    prm.seed = int(seed) & 0xffffffff
    for k, v in overrides.items():
        if not hasattr(prm, k):
            raise TypeError("usac5pt: unknown parameter %r" % (k,))
        setattr(prm, k, v)
    idx = None
    if sorted_idx is not None:
        idx = np.ascontiguousarray(np.asarray(sorted_idx, np.uint32).reshape(-1))
        if idx.shape[0] != n:
            raise AssertionError("CV_Assert(sortedMatchIdx.empty() || (p1.rows == (int)sortedMatchIdx.size()))")
    E = np.zeros(9); mask = np.zeros(max(n, 1), np.uint8)
    res = _lib.UsacResult()
    _lib.check(lib.mpl_usac5pt(_lib.ptr(p1), _lib.ptr(p2), n, _lib.ptr(idx) if idx is not None else None, ctypes.byref(prm),
                               _lib.ptr(E), _lib.ptr(mask), ctypes.byref(res)), "mpl_usac5pt")
    out = dict(ok=bool(res.ok), E=E.reshape(3, 3), mask=mask[:n], count=int(res.best_inlier_count))
    for k, _t in _lib.UsacResult._fields_[1:]:
        out.setdefault(k, getattr(res, k))
    return out


# enums of poselib/pose_estim.h:61-91
DEGEN_NO_CHECK, DEGEN_QDEGSAC, DEGEN_USAC_INTERNAL = 0, 1, 2
POSE_NISTER, POSE_EIG_KNEIP, POSE_STEWENIUS = 0, 1, 2
(REF_WEIGHTS, REF_8PT_PSEUDOHUBER, REF_EIG_KNEIP, REF_EIG_KNEIP_WEIGHTS, REF_STEWENIUS, REF_STEWENIUS_WEIGHTS, REF_NISTER,
 REF_NISTER_WEIGHTS) = range(8)
SPRT_DEFAULT_INIT, SPRT_DELTA_AUTOM_INIT, SPRT_EPSILON_AUTOM_INIT = 0, 1, 2


class ConfigUSAC:
    """poselib::ConfigUSAC (pose_estim.h:94-132) with the reference's defaults.  Served: estimator POSE_NISTER / POSE_STEWENIUS,
    refinealg REF_WEIGHTS, degeneracyCheck DEGEN_NO_CHECK (poselib-test's cfgUSAC digits 4-6 = 0, 0|2, 0); every other value makes
    estimateEssentialOrPoseUSAC raise -- there is no CPU fallback.  `matches` is the DMatch record array getMatches returns (or
    None: uniform sampling instead of PROSAC), keypoints1/2 are (n, 2) pixel coordinates (only read for SPRT_DELTA_AUTOM_INIT).
    seed / lo_mode are this library's additions (the reference seeds with time(); see mpl_usac_params in include/mplb200.h)."""

    def __init__(self, **kw):
        self.focalLength = 800.0; self.th_pixels = 0.8
        self.degeneracyCheck = DEGEN_USAC_INTERNAL; self.estimator = POSE_STEWENIUS; self.refinealg = REF_STEWENIUS_WEIGHTS
        self.prevalidateSample = False; self.noAutomaticProsacParamters = False
        self.automaticSprtInit = SPRT_DELTA_AUTOM_INIT | SPRT_EPSILON_AUTOM_INIT
        self.matches = None; self.keypoints1 = None; self.keypoints2 = None
        self.nrMatchesVfcFiltered = 0; self.imgSize = (800, 600); self.degenDecisionTh = 0.85
        self.seed = 1; self.lo_mode = 0
        for k, v in kw.items():
            if not hasattr(self, k):
                raise TypeError("ConfigUSAC has no field %r" % (k,))
            setattr(self, k, v)


def _convex_hull_area(pts):
    """cv::contourArea(cv::convexHull(points)) (pose_helper.cpp:2848-2853): monotone chain + shoelace."""
    p = np.unique(np.asarray(pts, np.float64).reshape(-1, 2), axis=0)
    if len(p) < 3:
        return 0.0
    p = p[np.lexsort((p[:, 1], p[:, 0]))]

    def half(seq):
        h = []
        for q in seq:
            while len(h) >= 2 and (h[-1][0] - h[-2][0]) * (q[1] - h[-2][1]) - (h[-1][1] - h[-2][1]) * (q[0] - h[-2][0]) <= 0:
                h.pop()
            h.append(q)
        return h
    hull = np.array(half(p)[:-1] + half(p[::-1])[:-1])
    x, y = hull[:, 0], hull[:, 1]
    return 0.5 * abs(float(np.dot(x, np.roll(y, -1)) - np.dot(y, np.roll(x, -1))))


def estimateSprtDeltaInit(matches, kp1, kp2, th, imgSize):
    """poselib::estimateSprtDeltaInit (pose_helper.cpp:2830-2867): area around the longest possible epipolar line over the smaller
    convex-hull area of the matched keypoints, at least 0.001."""
    kp1 = np.asarray(kp1, np.float64).reshape(-1, 2); kp2 = np.asarray(kp2, np.float64).reshape(-1, 2)
    area = min(_convex_hull_area(kp1[matches["queryIdx"]]), _convex_hull_area(kp2[matches["trainIdx"]]))
    max_epi = float(np.sqrt(float(imgSize[0]) ** 2 + float(imgSize[1]) ** 2)) * 2 * th
    area = max(area, 6 * max_epi)
    return max(max_epi / area, 0.001)


def estimateSprtEpsilonInit(matches, nrMatchesVfcFiltered):
    """poselib::estimateSprtEpsilonInit (pose_helper.cpp:2881-2889)."""
    return float(min(max(0.8 * float(nrMatchesVfcFiltered) / float(len(matches)), 0.1), 0.4))


def getSortedMatchIdx(matches):
    """poselib::getSortedMatchIdx (pose_helper.cpp:2902-2926): positions of the matches by ascending distance (std::sort is not
    stable; ties are broken by position here)."""
    return np.argsort(np.asarray(matches["distance"]), kind="stable").astype(np.uint32)


def estimateEssentialOrPoseUSAC(p1, p2, th, cfg, verbose=False):
    """poselib::estimateEssentialOrPoseUSAC (pose_estim.h:212-223, pose_estim.cpp:1737-2262) for the configurations ConfigUSAC's
    docstring lists.  Returns (code, E 3x3 or None, isDegenerate, inliers 1 x n uint8, info): code 0 success, -1 unsupported
    estimator / refinement (raised instead: no fallback), -2 "USAC failed!".  The SPRT / PROSAC start values follow the reference's
    first call (:1810-1976); its cross-call running statistics (function statics, :1752-1758, 2226-2259) are not kept -- every call
    starts like the first."""
    if cfg.estimator not in (POSE_NISTER, POSE_STEWENIUS):
        raise _lib.MplError("estimateEssentialOrPoseUSAC: estimator %r is not built (POSE_NISTER / POSE_STEWENIUS only)" % (cfg.estimator,))
    if cfg.refinealg != REF_WEIGHTS:
        raise _lib.MplError("estimateEssentialOrPoseUSAC: refinealg %r is not built (REF_WEIGHTS only)" % (cfg.refinealg,))
    if cfg.degeneracyCheck != DEGEN_NO_CHECK:
        raise _lib.MplError("estimateEssentialOrPoseUSAC: degeneracyCheck %r is not built (DEGEN_NO_CHECK only)" % (cfg.degeneracyCheck,))
    if cfg.automaticSprtInit not in (0, 1, 2, 3):
        print("Method for SPRT initialization not supported!")
        return -1, None, False, None, {}
    sprt_delta, sprt_epsilon, prosac_beta = 0.05, 0.15, 0.09                       # :1748
    if cfg.automaticSprtInit & SPRT_DELTA_AUTOM_INIT:
        sprt_delta = estimateSprtDeltaInit(cfg.matches, cfg.keypoints1, cfg.keypoints2, cfg.th_pixels, cfg.imgSize)
    if cfg.automaticSprtInit & SPRT_EPSILON_AUTOM_INIT:
        sprt_epsilon = estimateSprtEpsilonInit(cfg.matches, cfg.nrMatchesVfcFiltered)
    if not cfg.noAutomaticProsacParamters:
        prosac_beta = sprt_delta                                                   # :1972-1975
    idx = getSortedMatchIdx(cfg.matches) if cfg.matches is not None else None     # :1977-1981
    r = usac5pt(p1, p2, th, sorted_idx=idx, seed=cfg.seed, estimator="NISTER" if cfg.estimator == POSE_NISTER else "STEWENIUS",
                sprt_delta=float(sprt_delta), sprt_epsilon=float(sprt_epsilon), prosac_beta=float(prosac_beta), lo_mode=int(cfg.lo_mode))
    if not r["ok"]:
        print("USAC failed!")
        return -2, None, False, None, r
    if verbose:
        print("Number of hypotheses/models: %d/%d\nMax inliers/total points: %d/%d" % (r["hyp_count"], r["model_count"], r["count"], len(r["mask"])))
    return 0, r["E"], False, r["mask"].reshape(1, -1), r


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


def pixToCamFact(K0, K1):
    """th = th_pix * pixToCamFact(K0, K1): the driver's pixel -> camera-unit threshold (tests/poselib-test/main.cpp:1597-1599)."""
    K0, K1 = np.asarray(K0, np.float64), np.asarray(K1, np.float64)
    return 4.0 / (np.sqrt(2.0) * (K0[0, 0] + K0[1, 1] + K1[0, 0] + K1[1, 1]))


class PairBatch:
    """A batch of image pairs marshalled once for mpl_match_pose_batch (arguments as matchAndEstimatePoseBatch): run() is
    exactly one C call, so a caller that streams batches (or a benchmark) pays the Python marshalling outside it."""

    def __init__(self, pairs, matcher_name="LINEAR", ratioTest=True, threshold=1.6, n_samples=1000, confidence=0.999,
                 refine=True, seeds=None, K0=None, K1=None, dist0=None, dist1=None, want_samples=False):
        self.lib = _lib.load()
        n_pairs = self.n_pairs = len(pairs)
        pix = K0 is not None
        self.io = (_lib.PairIO * max(n_pairs, 1))()
        self.res = (_lib.PairResult * max(n_pairs, 1))()
        self.keep, self.outs = [], []
        cols, is_float = None, None
        for i, (d1, d2, p1, p2) in enumerate(pairs):
            d1, d2 = np.asarray(d1), np.asarray(d2)
            fl = d1.dtype == np.float32
            if is_float is None:
                is_float, cols = fl, d1.shape[1]
            if fl != is_float or d2.dtype != d1.dtype or d1.shape[1] != cols or d2.shape[1] != cols or (not fl and d1.dtype != np.uint8):
                raise AssertionError("all pairs of a batch need descriptors of one type and length")
            if d1.strides[1] != d1.itemsize:
                d1 = np.ascontiguousarray(d1)
            if d2.strides[1] != d2.itemsize:
                d2 = np.ascontiguousarray(d2)
            if pix:
                p1 = np.ascontiguousarray(p1, np.float32); p2 = np.ascontiguousarray(p2, np.float32)
            else:
                p1, p2 = _pts(p1), _pts(p2)
            if p1.shape != (d1.shape[0], 2) or p2.shape != (d2.shape[0], 2):
                raise AssertionError("descriptor / keypoint shapes do not agree")
            n1 = d1.shape[0]
            m = np.zeros(max(n1, 1), DMATCH); k = np.zeros(max(n1, 1), np.uint8)
            sm = np.zeros((max(n_samples, 1), 5), np.int32) if want_samples else None
            self.keep.append((d1, d2, p1, p2))
            self.outs.append((m, k, sm))
            e = self.io[i]
            e.desc1, e.desc2 = d1.ctypes.data, d2.ctypes.data
            e.step1 = d1.strides[0] if n1 > 1 else cols * d1.itemsize
            e.step2 = d2.strides[0] if d2.shape[0] > 1 else cols * d2.itemsize
            e.pts1, e.pts2, e.n1, e.n2 = p1.ctypes.data, p2.ctypes.data, n1, d2.shape[0]
            e.seed = int(seeds[i]) if seeds is not None else 1 + i
            e.matches, e.mask = m.ctypes.data, k.ctypes.data
            e.samples = sm.ctypes.data if sm is not None else None
        prm = self.prm = _lib.PairParams()
        prm.cv_type, prm.cols = (5 if is_float else 0), int(cols or 1)
        prm.matcher_name, prm.ratio_test, prm.pts_mode = matcher_name.encode(), int(bool(ratioTest)), int(pix)
        if pix:
            for dst, src, nn in ((prm.K0, K0, 9), (prm.K1, K1 if K1 is not None else K0, 9), (prm.dist0, dist0, 8), (prm.dist1, dist1, 8)):
                v = np.zeros(nn) if src is None else np.asarray(src, np.float64).reshape(-1)[:nn]
                for j in range(len(v)):
                    dst[j] = float(v[j])
        prm.th, prm.h, prm.conf, prm.refine = float(threshold), int(n_samples), float(confidence), int(bool(refine))

    def run(self):
        _lib.check(self.lib.mpl_match_pose_batch(self.io, self.n_pairs, ctypes.byref(self.prm), self.res), "mpl_match_pose_batch")
        return self

    def results(self):
        out = []
        for i in range(self.n_pairs):
            r, (m, k, sm) = self.res[i], self.outs[i]
            d = dict(matches=m[: r.n_matches], ok=r.n_inl > 0, E=np.array(list(r.E)).reshape(3, 3), mask=k[: r.n_matches], count=r.n_inl,
                     iters=r.iters_run, n_corr=r.n_corr, status=r.status)
            if sm is not None:
                d["samples"] = sm[: r.n_samples]
            out.append(d)
        return out


def matchAndEstimatePoseBatch(pairs, matcher_name="LINEAR", ratioTest=True, threshold=1.6, n_samples=1000, confidence=0.999,
                              refine=True, seeds=None, K0=None, K1=None, dist0=None, dist1=None, want_samples=False):
    """mpl_match_pose_batch: both halves of the hot path for a list of image pairs in ONE call (device resident between
    the scan and the final E; sub-batches pipelined on two streams).  pairs: sequence of (desc1, desc2, pts1, pts2) with
    desc uint8 [n, nbytes] or float32 [n, dim]; pts [n, 2] float64 camera-normalised coordinates, or -- when K0 / K1 are
    given -- float32 pixel coordinates (cv::KeyPoint::pt) that the library normalises (ImgToCamCoordTrans) and undistorts
    (Remove_LensDist, dist0 / dist1: 8 OpenCV coefficients) on the device.  `threshold` is in camera units
    (pixToCamFact).  Returns one dict per pair: matches, ok, E, mask (over matches), count, iters, n_corr, status[, samples]."""
    return PairBatch(pairs, matcher_name, ratioTest, threshold, n_samples, confidence, refine, seeds, K0, K1, dist0, dist1,
                     want_samples).run().results()
