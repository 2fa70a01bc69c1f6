"""Latency of ONE C4-unit pair through mpl_match_pose_batch (batch of one), by stage:  python tools/pair_latency.py"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import bench
from matchinglib_poselib_b200 import poselib, _lib
lib = _lib.load(); _lib.check(lib.mpl_init(0), "init")
(q, t, p1, p2), th = bench.make_c4_pair(1004)
def med(h, refine, early=0):
    lib.mpl_set_solver_early_exit(early)
    b = poselib.PairBatch([(q, t, p1, p2)], "LINEAR", True, th, h, 0.999, refine, seeds=[5])
    ts = []
    for i in range(12):
        torch.cuda.synchronize(); t0 = time.perf_counter(); b.run(); ts.append(time.perf_counter() - t0)
    lib.mpl_set_solver_early_exit(0)
    return round(float(np.median(ts[3:])) * 1e3, 3)
print(dict(match_only=med(0, False), ransac64=med(64, False), ransac1000=med(1000, False), ransac1000_refit=med(1000, True),
           ransac1000_refit_early=med(1000, True, 1), ransac64_refit_early=med(64, True, 1)))
