"""Per-class device time of the pair-batch pipeline at the C4 shape (library CUDA-event timers): python tools/c4_prof.py [pairs]"""
import ctypes, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import bench
import matchinglib_poselib_b200 as m
from matchinglib_poselib_b200 import poselib
lib = m._lib.load(); m._lib.check(lib.mpl_init(0), "init")
n_pairs = int(sys.argv[1]) if len(sys.argv) > 1 else 512
pool = [bench.make_c4_pair(1004 + i) for i in range(4)]
pairs = [tuple(pool[i % 4][0]) for i in range(n_pairs)]
th = pool[0][1]
PROF = dict(hamming=0, l2_exact=1, l2_tc=2, l2_rerank=3, merge=4, solve=5, score=6, l2_prep=7)
for path in (1, 3, 2):
    lib.mpl_set_hamming_path(path)
    for h in (0, 1000):
        poselib.matchAndEstimatePoseBatch(pairs[:64], "LINEAR", True, th, h, 0.999, True, seeds=list(range(64)))
        t0 = time.perf_counter()
        poselib.matchAndEstimatePoseBatch(pairs, "LINEAR", True, th, h, 0.999, True, seeds=list(range(n_pairs)))
        wall = time.perf_counter() - t0
        lib.mpl_prof_enable(1)
        for c in PROF.values():
            ms, n = ctypes.c_double(0), ctypes.c_longlong(0); lib.mpl_prof_read(c, ctypes.byref(ms), ctypes.byref(n))
        poselib.matchAndEstimatePoseBatch(pairs, "LINEAR", True, th, h, 0.999, True, seeds=list(range(n_pairs)))
        res = {}
        for k, c in PROF.items():
            ms, n = ctypes.c_double(0), ctypes.c_longlong(0); lib.mpl_prof_read(c, ctypes.byref(ms), ctypes.byref(n))
            if n.value: res[k] = (round(ms.value * 1e3 / n_pairs, 2), n.value)
        lib.mpl_prof_enable(0)
        print("path", path, "h", h, "wall us/pair %.1f" % (wall * 1e6 / n_pairs), "device us/pair by class (launches):", res)
lib.mpl_set_hamming_path(0)
