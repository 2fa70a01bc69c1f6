"""Debug: GPU LMedS medians vs medians recomputed on the host from the GPU's own models."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, datagen_pose
import matchinglib_poselib_b200 as m
from oracle import ransac as orc
m._lib.check(m._lib.load().mpl_init(0), "init")
for n in (50, 1001, 2000):
    sc = datagen_pose.make_scene(n, 0.3, seed=5)
    s = datagen_pose.draw_samples_fast(n, 134, 55)
    got = m.lmeds5pt(sc["p1"], sc["p2"], s, trace=True)
    tr = m.ransac5pt(sc["p1"], sc["p2"], s, sc["th"], max_iters=134, trace=True)
    exp = orc.lmeds5pt(sc["p1"], sc["p2"], s)
    M = tr["models"]
    print(n, len(M), len(got["medians"]), len(exp["medians"]))
    host = []
    for E in M:
        _, _, err, _ = orc.sampson(E, sc["p1"], sc["p2"], 1.0)
        e = np.sort(err.view(np.int32)).view(np.float32)
        host.append(float(e[n // 2]) if n % 2 else float(np.float32(e[n // 2 - 1] + e[n // 2])) * 0.5)
    host = np.array(host)
    print(" gpu == host-from-gpu-models:", (got["medians"] == host).mean(), " gpu == oracle:", (got["medians"] == exp["medians"]).mean(),
          " host-from-gpu-models == oracle:", (host == exp["medians"]).mean())
    bad = np.nonzero(got["medians"] != host)[0][:5]
    for b in bad: print("  model", b, got["medians"][b], host[b], exp["medians"][b])
