"""ctypes front end of oracle/ransac5pt_oracle.c.  TEST INFRASTRUCTURE ONLY (see that file's header)."""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libransac_oracle.so")
_lib = None


def build():
    src = os.path.join(HERE, "ransac5pt_oracle.c")
    if not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", HERE, "libransac_oracle.so"])


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(LIB)
        _lib.orc_ransac5pt.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_double,
                                       ctypes.c_double, ctypes.c_int, ctypes.c_int] + [ctypes.c_void_p] * 8 + [ctypes.c_int]
        _lib.orc_sampson.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_double,
                                     ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        _lib.orc_lmeds_niters.argtypes = [ctypes.c_double, ctypes.c_int]
        _lib.orc_lmeds5pt.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int] + [ctypes.c_void_p] * 2 + \
                                     [ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double),
                                      ctypes.c_void_p, ctypes.c_void_p]
    return _lib


def _p(a):
    return None if a is None else ctypes.c_void_p(a.ctypes.data)


def solve_poly(coeffs, max_iters=300):
    c = np.ascontiguousarray(coeffs, np.float64)
    n0 = len(c) - 1
    roots = np.zeros((n0, 2))
    lib().orc_solve_poly(_p(c), n0, _p(roots), max_iters)
    return roots


def run5point(q1, q2):
    q1 = np.ascontiguousarray(q1, np.float64); q2 = np.ascontiguousarray(q2, np.float64)
    models = np.zeros((10, 9))
    n = lib().orc_run5point(_p(q1), _p(q2), q1.shape[0], _p(models))
    return models[:n].copy()


def sampson(E, p1, p2, th):
    """(count, errsum, err float32[n], mask uint8[n])"""
    E = np.ascontiguousarray(E, np.float64).reshape(9)
    p1 = np.ascontiguousarray(p1, np.float64); p2 = np.ascontiguousarray(p2, np.float64)
    n = p1.shape[0]
    err = np.zeros(n, np.float32); mask = np.zeros(n, np.uint8); es = ctypes.c_double(0)
    cnt = lib().orc_sampson(_p(E), _p(p1), _p(p2), n, th, _p(err), _p(mask), ctypes.byref(es))
    return cnt, es.value, err, mask


def ransac5pt(p1, p2, samples, th, conf=0.999, max_iters=1000, refine=False, trace=False):
    """runRANSAC over pre-drawn samples.  Returns dict(ok, E, mask, count, iters[, n_models, models, counts, errsums])."""
    p1 = np.ascontiguousarray(p1, np.float64); p2 = np.ascontiguousarray(p2, np.float64)
    samples = np.ascontiguousarray(samples, np.int32)
    n, h = p1.shape[0], samples.shape[0]
    E = np.zeros(9); mask = np.zeros(n, np.uint8); cnt = ctypes.c_int(0); iters = ctypes.c_int(0)
    tm = np.zeros((10 * h, 9)) if trace else None
    tc = np.zeros(10 * h, np.int32) if trace else None
    te = np.zeros(10 * h) if trace else None
    nm = np.zeros(h, np.int32)
    ok = lib().orc_ransac5pt(_p(p1), _p(p2), n, _p(samples), h, th, conf, max_iters, int(refine), _p(E), _p(mask),
                             ctypes.byref(cnt), ctypes.byref(iters), _p(tm), _p(tc), _p(te), _p(nm), int(trace))
    out = dict(ok=bool(ok), E=E.reshape(3, 3), mask=mask, count=cnt.value, iters=iters.value, n_models=nm)
    if trace:
        tot = int(nm.sum())
        out.update(models=tm[:tot], counts=tc[:tot], errsums=te[:tot])
    return out


def lmeds_niters(conf=0.999, max_iters=2000):
    """Iteration count of runLMeDS (modelest.cpp:518-519): fixed, from an assumed outlier ratio of 0.45."""
    return int(lib().orc_lmeds_niters(ctypes.c_double(conf), int(max_iters)))


def lmeds5pt(p1, p2, samples):
    """runLMeDS over pre-drawn samples (modelest.cpp:483-564).  Returns dict(ok, E, mask, count, min_median, sigma, medians, n_models)."""
    p1 = np.ascontiguousarray(p1, np.float64); p2 = np.ascontiguousarray(p2, np.float64)
    samples = np.ascontiguousarray(samples, np.int32)
    n, h = p1.shape[0], samples.shape[0]
    E = np.zeros(9); mask = np.zeros(n, np.uint8); cnt = ctypes.c_int(0)
    mm = ctypes.c_double(0); sg = ctypes.c_double(0)
    med = np.zeros(10 * max(h, 1)); nm = np.zeros(max(h, 1), np.int32)
    ok = lib().orc_lmeds5pt(_p(p1), _p(p2), n, _p(samples), h, _p(E), _p(mask), ctypes.byref(cnt), ctypes.byref(mm), ctypes.byref(sg),
                            _p(med), _p(nm))
    return dict(ok=bool(ok), E=E.reshape(3, 3), mask=mask, count=cnt.value, min_median=mm.value, sigma=sg.value,
                medians=med[:int(nm[:h].sum())], n_models=nm[:h])


def match_models(a, b):
    """For each model in a (k x 9) the distance to the nearest model of b (sign-free Frobenius)."""
    if len(a) == 0:
        return np.zeros(0)
    if len(b) == 0:
        return np.full(len(a), np.inf)
    d = np.minimum(np.linalg.norm(a[:, None, :] - b[None], axis=2), np.linalg.norm(a[:, None, :] + b[None], axis=2))
    return d.min(1)
