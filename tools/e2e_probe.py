"""Where does the end-to-end time of getMatches(LINEAR, C2) go?  (diagnostic, run on the GPU box)"""
import ctypes, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import bench
g = bench.Gpu(0)
q, t = bench.make_c2(0)
(tq, hq), (tt, ht) = bench.pinned(q, torch), bench.pinned(t, torch)
d = torch.empty_like(tq, device="cuda")
for n in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter(); d.copy_(tq, non_blocking=True); torch.cuda.synchronize(); dt = time.perf_counter() - t0
print("H2D 10.24 MB pinned: %.3f ms  (%.1f GB/s)" % (dt * 1e3, tq.numel() * 4 / dt / 1e9))
pq = torch.from_numpy(q)          # pageable
for n in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter(); d.copy_(pq); torch.cuda.synchronize(); dt = time.perf_counter() - t0
print("H2D 10.24 MB pageable: %.3f ms" % (dt * 1e3))
m, lib = g.m, g.lib
idx = np.empty((20000, 2), np.int32); dist = np.empty((20000, 2), np.float32)
P = m._lib.ptr
def raw():
    m._lib.check(lib.mpl_knn2_l2f32(P(hq), 20000, 512, P(ht), 20000, 512, 128, 0, 2, P(idx), P(dist)), "knn")
raw()
ts = []
for _ in range(10):
    torch.cuda.synchronize(); t0 = time.perf_counter(); raw(); ts.append(time.perf_counter() - t0)
print("mpl_knn2_l2f32 (host ptrs, pinned in, pageable out): median %.3f ms  min %.3f" % (np.median(ts) * 1e3, min(ts) * 1e3))
kp = range(20000)
ts = []
for _ in range(10):
    torch.cuda.synchronize(); t0 = time.perf_counter(); m.getMatchesArray(kp, kp, hq, ht, None, "LINEAR", False, True); ts.append(time.perf_counter() - t0)
print("getMatchesArray: median %.3f ms  min %.3f" % (np.median(ts) * 1e3, min(ts) * 1e3))
