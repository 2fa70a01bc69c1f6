"""ARRSAC oracle: the reference's OWN theia::Arrsac template (oracle/_ref/libarrsac_ref.so, compiled from /root/reference by
`make -C oracle ref_arrsac`; oracle/ref_arrsac/driver.cc restates the OpenCV/Eigen-typed estimator around it) plus a numpy
restatement of robustEssentialRefine.  TEST INFRASTRUCTURE ONLY.

Reference (paths relative to matchinglib_poselib/source/poselib):
  runARRSAC                  source/five-point-nister/modelest.cpp:197-341
  robustEssentialRefine      source/pose_estim.cpp:337-837 (model 0, normalizeCorrs false, makeClosestE true, iters 0)
  SampsonL1                  source/pose_helper.cpp:115-143      costPseudoHuber  source/BA_driver.cpp:2639-2648
  getClosestE                source/pose_helper.cpp:152-176
Parity of ARRSAC is statistical (randomised, near-ties decided by solver roundings): tests compare success, inlier counts and
the distance of E to the ground truth, not bits."""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_LIB = os.path.join(HERE, "_ref", "libarrsac_ref.so")
_ref = None


def build():
    if not os.path.exists(REF_LIB) and os.path.isdir("/root/reference/matchinglib_poselib"):
        subprocess.check_call(["make", "-s", "-C", HERE, "ref_arrsac"])


def ref_lib():
    global _ref
    if _ref is None and os.path.exists(REF_LIB):
        _ref = ctypes.CDLL(REF_LIB)
    return _ref


def arrsac(p1, p2, th):
    """Arrsac::Estimate + findInliers + acceptance rule (modelest.cpp:269-277).  Returns dict(ok, E, mask, count) or None
    when the reference build is absent."""
    R = ref_lib()
    if R is None:
        return None
    p1 = np.ascontiguousarray(p1, np.float64); p2 = np.ascontiguousarray(p2, np.float64)
    n = len(p1)
    E = np.zeros(9); mask = np.zeros(n, np.uint8); cnt = ctypes.c_int(0)
    ok = R.ref_arrsac(ctypes.c_void_p(p1.ctypes.data), ctypes.c_void_p(p2.ctypes.data), n, ctypes.c_double(th),
                      ctypes.c_void_p(E.ctypes.data), ctypes.c_void_p(mask.ctypes.data), ctypes.byref(cnt))
    return dict(ok=bool(ok), E=E.reshape(3, 3), mask=mask, count=cnt.value)


def robust_refine(p1, p2, E_init, th, max_iters=50):
    """robustEssentialRefine(points1, points2, E_init, E_refined, th, 0, true, ...) on the given (inlier) correspondences."""
    p1 = np.asarray(p1, np.float64); p2 = np.asarray(p2, np.float64)
    n = len(p1)
    if n < 50:
        return np.array(E_init, np.float64).reshape(3, 3)
    F3 = np.array(E_init, np.float64).reshape(3, 3).copy()
    X1 = np.concatenate([p1, np.ones((n, 1))], 1); X2 = np.concatenate([p2, np.ones((n, 1))], 1)
    min_diff, min_err = th / 10, th * th / 100 * n
    err_old = 1e12
    A = np.stack([p2[:, 0] * p1[:, 0], p2[:, 0] * p1[:, 1], p2[:, 0], p2[:, 1] * p1[:, 0], p2[:, 1] * p1[:, 1], p2[:, 1], p1[:, 0], p1[:, 1], np.ones(n)], 1)
    for j in range(max_iters):
        xpE = X2 @ F3                                          # rows: x2^T E
        num = (xpE * X1).sum(1)
        Ex1 = X1 @ F3.T
        denom1 = 1.0 / (np.sqrt(Ex1[:, 0] ** 2 + Ex1[:, 1] ** 2 + xpE[:, 0] ** 2 + xpE[:, 1] ** 2) + 1e-8)
        d_abs = np.abs(num * denom1) + 1e-12
        w = np.sqrt(2 * th * th * (np.sqrt(1 + (d_abs / th) ** 2) - 1)) / d_abs
        wn = np.sqrt(((w * denom1) ** 2).sum())
        A1 = A * (denom1 * w / wn)[:, None]
        ev, evec = np.linalg.eig(A1.T @ A1)
        ev = ev.real
        if (np.abs(ev) < np.finfo(float).eps).sum() >= 2:       # pose_estim.cpp:601-611 (stated on the count, the solver's order is not defined)
            return np.array(E_init, np.float64).reshape(3, 3)
        last = evec[:, np.argmin(np.abs(ev))].real
        F2 = last.reshape(3, 3)
        U, s, Vt = np.linalg.svd(F2)
        if not (-1e-3 < s[2] < 1e-3) or s[0] / s[1] > 1.5 or s[0] / s[1] < 0.66:     # getClosestE
            break
        F3 = U @ np.diag([s[0], s[1], 0.0]) @ Vt
        err = float(((A1 @ last) ** 2).sum())
        if j > 1 and (abs(err_old - err) < min_diff or err < min_err):
            break
        err_old = err
    return F3


def estimate(p1, p2, th, refine=True):
    """findEssentialMat(..., ARRSAC, ..., refine, robustEssentialRefine): dict(ok, E, mask, count)."""
    r = arrsac(p1, p2, th)
    if r is None or not r["ok"]:
        return r
    if refine and len(p1) > 5 and r["count"] >= 50:
        inl = r["mask"].astype(bool)
        r["E"] = robust_refine(np.asarray(p1)[inl], np.asarray(p2)[inl], r["E"], th / 50.0)
    return r
