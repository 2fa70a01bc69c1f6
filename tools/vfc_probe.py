"""GPU VFC filter vs the reference's own vfc.cpp on synthetic scenes: kept-set differences, iterations, wall clock per call."""
import sys, time, ctypes
import numpy as np
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import datagen_vfc
from oracle import vfc as ovfc
from matchinglib_poselib_b200 import matchinglib as M, _lib
lib = _lib.load(); _lib.check(lib.mpl_init(0), "mpl_init")
for n, outl, seed in [(3000, 0.4, 1), (500, 0.2, 2), (10000, 0.6, 3), (2000, 0.85, 6), (6000, 0.5, 8), (20000, 0.5, 9)]:
    k1, k2, mt, inl = datagen_vfc.scene(n, outl, seed=seed)
    dm = np.zeros(n, M.DMATCH_DTYPE); dm["queryIdx"] = mt[:, 0]; dm["trainIdx"] = mt[:, 1]; dm["distance"] = 1
    for s in (1, 7, 1234):
        lib.mpl_vfc_set_seed(s)
        keep = np.zeros(n, np.uint8); nk, code, it = ctypes.c_int(0), ctypes.c_int(0), ctypes.c_int(0)
        t0 = time.perf_counter()
        lib.mpl_vfc_filter(_lib.ptr(k1), n, 8, _lib.ptr(k2), n, 8, _lib.ptr(dm), n, _lib.ptr(keep), ctypes.byref(nk), ctypes.byref(code), ctypes.byref(it))
        tg = time.perf_counter() - t0
        t0 = time.perf_counter(); r = ovfc.filter_with_vfc(k1, k2, mt, seed=s); tr = time.perf_counter() - t0
        d = keep.astype(bool) != r[1].astype(bool)
        print("n=%d outl=%.2f seed=%d: GPU kept %d (code %d, %d iterations, %.2f ms) reference kept %d (code %d, %.1f ms) differ %d, of them true inliers %d" %
              (n, outl, s, nk.value, code.value, it.value, tg * 1e3, r[1].sum(), r[0], tr * 1e3, d.sum(), (d & inl).sum()))
