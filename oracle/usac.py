"""USAC oracle: the reference's OWN estimator -- poselib/include/usac/estimators/{USAC.h,EssentialMatEstimator.h}, its utils and
the vendored opengv solvers -- compiled where they lie into oracle/_ref/libusac_ref.so (`make -C oracle ref_usac`;
oracle/ref_usac/driver.cc restates only the configuration estimateEssentialMatUsac fills), plus numpy restatements of the
pieces the GPU path is checked against exactly.  TEST INFRASTRUCTURE ONLY.

Reference (paths relative to matchinglib_poselib/source/poselib):
  estimateEssentialMatUsac                  source/usac/usac_estimations.cpp:283-735
  USAC::solve, designSPRTTest               include/usac/estimators/USAC.h:345-612, 918-942
  evaluateModel (Sampson error)             include/usac/estimators/EssentialMatEstimator.h:1110-1174
  generateRefinedModel (REFINE_WEIGHTS)     include/usac/estimators/EssentialMatEstimator.h:540-597
  findWeights (Torr)                        include/usac/estimators/EssentialMatEstimator.h:2376-2389
  normalizePoints / singulF                 source/usac/utils/FundmatrixFunctions.cpp:7-60, 334-360
Parity of USAC is statistical (randomised, adaptive; the minimal solver's roundings and model order decide near-ties): tests
compare success, inlier counts, hypothesis counts and the distance of E to the ground truth, not bits."""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_LIB = os.path.join(HERE, "_ref", "libusac_ref.so")
_ref = None


def build():
    if not os.path.exists(REF_LIB) and os.path.isdir("/root/reference/matchinglib_poselib"):
        subprocess.check_call(["make", "-s", "-j8", "-C", HERE, "ref_usac"])


def ref_lib():
    global _ref
    if _ref is None and os.path.exists(REF_LIB):
        _ref = ctypes.CDLL(REF_LIB)
        dp = ctypes.c_void_p
        _ref.usac_ref_essential.restype = ctypes.c_int
        _ref.usac_ref_essential.argtypes = [dp, dp, ctypes.c_int] + [ctypes.c_double] * 7 + [ctypes.c_int, dp, ctypes.c_uint, dp, dp, dp, dp]
    return _ref


def usac(p1, p2, th, sorted_idx=None, seed=1, estimator="NISTER", prosac_beta=0.09, sprt_delta=0.05, sprt_epsilon=0.15, mS=0.0,
         focal_length=800.0, th_pixels=0.8):
    """The reference's estimateEssentialMatUsac configuration (REFINE_WEIGHTS, no degeneracy test) on its own estimator.
    Returns dict(ok, E, mask, count, hyp_count, model_count, rejected_sample_count, rejected_model_count,
    total_points_verified, num_local_optimizations, sprt_delta, sprt_epsilon) or None when the reference build is absent."""
    R = ref_lib()
    if R is None:
        return None
    p1 = np.ascontiguousarray(p1, np.float64); p2 = np.ascontiguousarray(p2, np.float64)
    n = len(p1)
    idx = None if sorted_idx is None else np.ascontiguousarray(sorted_idx, np.uint32)
    E = np.zeros(9); mask = np.zeros(max(n, 1), np.uint8); stats = np.zeros(8, np.uint32); sp = np.zeros(2)
    rc = R.usac_ref_essential(p1.ctypes.data, p2.ctypes.data, n, th, focal_length, th_pixels, prosac_beta, sprt_delta, sprt_epsilon, mS,
                              {"NISTER": 0, "STEWENIUS": 2}[estimator], None if idx is None else idx.ctypes.data, int(seed) & 0xffffffff,
                              E.ctypes.data, mask.ctypes.data, stats.ctypes.data, sp.ctypes.data)
    return dict(ok=rc == 0, E=E.reshape(3, 3), mask=mask[:n], count=int(stats[6]), hyp_count=int(stats[0]), model_count=int(stats[1]),
                rejected_sample_count=int(stats[2]), rejected_model_count=int(stats[3]), total_points_verified=int(stats[4]),
                num_local_optimizations=int(stats[5]), sprt_delta=float(sp[0]), sprt_epsilon=float(sp[1]))


def ref_singulF(F):
    """The reference's FTools::singulF on a 3 x 3 matrix (None when the reference build is absent)."""
    R = ref_lib()
    if R is None:
        return None
    out = np.ascontiguousarray(np.asarray(F, np.float64).reshape(9)).copy()
    R.usac_ref_singulF(ctypes.c_void_p(out.ctypes.data))
    return out.reshape(3, 3)


# ---- numpy restatements ---------------------------------------------------------------------------------------------

def sampson(E, p1, p2):
    """evaluateModel's error of every correspondence (EssentialMatEstimator.h:1135-1143): r^2 / (|E^T x2|_xy^2 + |E x1|_xy^2)."""
    m = np.asarray(E, np.float64).reshape(9)
    x1, y1 = np.asarray(p1, np.float64).T; x2, y2 = np.asarray(p2, np.float64).T
    rxc = m[0] * x2 + m[3] * y2 + m[6]; ryc = m[1] * x2 + m[4] * y2 + m[7]; rwc = m[2] * x2 + m[5] * y2 + m[8]
    r = x1 * rxc + y1 * ryc + rwc
    rx = m[0] * x1 + m[1] * y1 + m[2]; ry = m[3] * x1 + m[4] * y1 + m[5]
    return r * r / (rxc * rxc + ryc * ryc + rx * rx + ry * ry)


def normalize_points(p1, p2):
    """FTools::normalizePoints: (p1n, p2n, T1, T2) with mean distance sqrt(2) from the centroid."""
    out = []
    for p in (np.asarray(p1, np.float64), np.asarray(p2, np.float64)):
        m = p.sum(0) / len(p)
        s = np.sqrt(2.0) / (np.sqrt(((p - m) ** 2).sum(1)).sum() / len(p))
        T = np.array([[s, 0, -s * m[0]], [0, s, -s * m[1]], [0, 0, 1.0]])
        out.append((p * s + T[:2, 2], T))
    return out[0][0], out[1][0], out[0][1], out[1][1]


def torr_weights(Fn, p1n, p2n):
    """findWeights, REFINE_WEIGHTS: 1 / sqrt(|F^T x2|_xy^2 + |F x1|_xy^2) with the normalised model and points."""
    m = np.asarray(Fn, np.float64).reshape(9)
    x1, y1 = p1n.T; x2, y2 = p2n.T
    rxc = m[0] * x2 + m[3] * y2 + m[6]; ryc = m[1] * x2 + m[4] * y2 + m[7]
    rx = m[0] * x1 + m[1] * y1 + m[2]; ry = m[3] * x1 + m[4] * y1 + m[5]
    return 1.0 / np.sqrt(rxc * rxc + ryc * ryc + rx * rx + ry * ry)


def refined_model(p1n, p2n, T1, T2, weights=None):
    """generateRefinedModel, REFINE_WEIGHTS: smallest eigenvector of the (weighted) data matrix' covariance, singulF (drop the
    smallest singular value), then models_denorm_ = T2^T F T1.  Returns (F normalised, E denormalised)."""
    x1, y1 = p1n.T; x2, y2 = p2n.T
    A = np.stack([x2 * x1, x2 * y1, x2, y2 * x1, y2 * y1, y2, x1, y1, np.ones(len(x1))], 1)
    if weights is not None:
        A = A * np.asarray(weights)[:, None]
    w, V = np.linalg.eigh(A.T @ A)
    F = V[:, 0].reshape(3, 3)
    U, s, Vt = np.linalg.svd(F)
    F = U @ np.diag([s[0], s[1], 0.0]) @ Vt
    return F, T2.T @ F @ T1


def design_sprt(delta, epsilon, tM, mS):
    """designSPRTTest: the decision threshold A."""
    C = (1 - delta) * np.log((1 - delta) / (1 - epsilon)) + delta * np.log(delta / epsilon)
    K = (tM * C) / mS + 1
    An_1 = K
    An = K
    for _ in range(10):
        An = K + np.log(An_1)
        if An - An_1 < 1.5e-8:
            break
        An_1 = An
    return An
