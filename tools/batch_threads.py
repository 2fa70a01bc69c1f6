"""C4 unit throughput vs number of host threads (diagnostic, GPU box)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from concurrent.futures import ThreadPoolExecutor
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
th = sc["th"]
p1, p2 = sc["p1"], sc["p2"]
if os.environ.get("PIN"):
    keep = [torch.from_numpy(np.ascontiguousarray(a)).pin_memory() for a in (q, t, p1, p2)]
    q, t, p1, p2 = [k.numpy() for k in keep]
def one(i):
    return poselib.matchAndEstimatePose(q, t, p1, p2, "LINEAR", True, th, 1000, 0.999, True, seed=5 + i)["count"]
for nt in (1, 16, 32):
    with ThreadPoolExecutor(nt) as ex:
        list(ex.map(one, range(2 * nt)))
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        P = 16 * nt if nt < 16 else 256
        list(ex.map(one, range(P)))
        dt = time.perf_counter() - t0
    print("threads %3d: %.1f pairs/s" % (nt, P / dt), flush=True)
