"""One USAC scene, a few calls (for `ncu --metrics gpu__time_duration.sum`): python tools/usac_one.py n outlier_fraction [prosac]"""
import sys
import numpy as np
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import datagen_pose
from oracle import usac as ousac
from matchinglib_poselib_b200 import poselib, _lib
_lib.check(_lib.load().mpl_init(0), "mpl_init")
n, outl = int(sys.argv[1]), float(sys.argv[2])
sc = datagen_pose.make_scene(n, outl, seed=1)
idx = None
if len(sys.argv) > 3:
    out = ousac.sampson(sc["E"], sc["p1"], sc["p2"]) >= sc["th"] ** 2
    idx = np.argsort(out + 0.6 * np.random.default_rng(1).random(n), kind="stable").astype(np.uint32)
for seed in range(1, 4):
    g = poselib.usac5pt(sc["p1"], sc["p2"], sc["th"], sorted_idx=idx, seed=seed)
    print(g["hyp_count"], g["count"], g["n_batches"], g["num_local_optimizations"], g["n_launches"])
