"""Per-role clock64 trace of l2_tc_rank_kernel at config C2 (build with MPL_EXTRA_NVCC_FLAGS=-DMPL_TC_TRACE first):
    MPL_EXTRA_NVCC_FLAGS=-DMPL_TC_TRACE python matchinglib_poselib_b200/build.py --force && python tools/tc_trace.py [signed]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import datagen
import matchinglib_poselib_b200 as m
m._lib.check(m._lib.load().mpl_init(0), "init")
q, t = (datagen.signed_float if "signed" in sys.argv else datagen.sift_like)(20000, 20000, 128, seed=1002)
for _ in range(2):
    m.knn2_l2(q, t, 2, False)
