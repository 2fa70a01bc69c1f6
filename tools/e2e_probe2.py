"""e2e breakdown, C2, medians over 30 reps (diagnostic, GPU box)."""
import ctypes, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import bench
g = bench.Gpu(0)
q, t = bench.make_c2(0)
(tq, hq), (tt, ht) = bench.pinned(q, torch), bench.pinned(t, torch)
dq = torch.empty_like(tq, device="cuda"); dt_ = torch.empty_like(tt, device="cuda")
def med(f, n=30):
    for _ in range(40): f()
    ts = []
    for _ in range(n):
        torch.cuda.synchronize(); t0 = time.perf_counter(); f(); ts.append(time.perf_counter() - t0)
    return np.median(ts) * 1e3, min(ts) * 1e3
def c1():
    dq.copy_(tq, non_blocking=True); torch.cuda.synchronize()
def c2():
    dq.copy_(tq, non_blocking=True); dt_.copy_(tt, non_blocking=True); torch.cuda.synchronize()
print("one 10.24 MB copy + sync: median %.3f min %.3f ms" % med(c1))
print("two copies + sync:        median %.3f min %.3f ms" % med(c2))
m, lib = g.m, g.lib
idx = np.empty((20000, 2), np.int32); dist = np.empty((20000, 2), np.float32)
(ti, hidx), (td, hdist) = bench.pinned(idx, torch), bench.pinned(dist, torch)
P = m._lib.ptr
def raw():
    m._lib.check(lib.mpl_knn2_l2f32(P(hq), 20000, 512, P(ht), 20000, 512, 128, 0, 2, P(idx), P(dist)), "knn")
def rawp():
    m._lib.check(lib.mpl_knn2_l2f32(P(hq), 20000, 512, P(ht), 20000, 512, 128, 0, 2, P(hidx), P(hdist)), "knn")
print("mpl_knn2_l2f32 pageable out: median %.3f min %.3f ms" % med(raw))
print("mpl_knn2_l2f32 pinned out:   median %.3f min %.3f ms" % med(rawp))
kp = range(20000)
def gm():
    m.getMatchesArray(kp, kp, hq, ht, None, "LINEAR", False, True)
print("getMatchesArray:             median %.3f min %.3f ms" % med(gm))
def gmf():
    g.flush_l2(); torch.cuda.synchronize()
    t0 = time.perf_counter(); gm(); return time.perf_counter() - t0
ts = [gmf() for _ in range(20)]
print("getMatchesArray after L2 flush: median %.3f min %.3f ms" % (np.median(ts) * 1e3, min(ts) * 1e3))
