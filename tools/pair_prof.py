"""Where does one C4-unit pair spend its time?  Device time per kernel class vs wall (diagnostic, GPU box)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import bench, datagen_pose
from matchinglib_poselib_b200 import poselib
g = bench.Gpu(0)
n, nb = 5000, 61
sc = datagen_pose.make_scene(n, 0.0, seed=1004)
rng = np.random.default_rng(1004)
t = rng.integers(0, 256, (n, nb), dtype=np.uint8)
q = rng.integers(0, 256, (n, nb), dtype=np.uint8)
good = rng.random(n) < 0.5
flips = np.packbits(rng.random((n, nb * 8)) < 0.08, axis=1)
q[good] = t[good] ^ flips[good]
def one(h=1000, refine=True):
    return poselib.matchAndEstimatePose(q, t, sc["p1"], sc["p2"], "LINEAR", True, sc["th"], h, 0.999, refine, seed=5)
for h, refine in ((1000, True), (1000, False), (100, False), (0, False)):
    for _ in range(3): one(h, refine)
    ts = []
    for _ in range(10):
        t0 = time.perf_counter(); one(h, refine); ts.append(time.perf_counter() - t0)
    g.lib.mpl_prof_enable(1)
    for c in bench.PROF.values(): g.prof_read(c)
    one(h, refine)
    g.lib.mpl_prof_enable(0)
    prof = {k: g.prof_read(c) for k, c in bench.PROF.items()}
    print("h=%d refine=%s wall %.3f ms; device: %s" % (h, refine, np.median(ts) * 1e3,
          ", ".join("%s %.3f ms x%d" % (k, v[0], v[1]) for k, v in prof.items() if v[1])))
