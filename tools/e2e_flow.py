"""Does running bench_l2 first slow the end-to-end call?  (diagnostic, GPU box)"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import bench
g = bench.Gpu(0)
m = g.m
def h2d(tag, q, t):
    (tq, hq), (tt, ht) = bench.pinned(q, torch), bench.pinned(t, torch)
    dq = torch.empty_like(tq, device="cuda"); dt = torch.empty_like(tt, device="cuda")
    ts = []
    for _ in range(30):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        dq.copy_(tq, non_blocking=True); dt.copy_(tt, non_blocking=True); torch.cuda.synchronize()
        ts.append(time.perf_counter() - t0)
    print("%s: H2D 20 MB median %.3f ms (last 10: %.3f)" % (tag, np.median(ts) * 1e3, np.median(ts[-10:]) * 1e3), flush=True)
mode = sys.argv[1] if len(sys.argv) > 1 else "full"
if mode == "full":
    r = bench.bench_l2(g, 20, 3)
    q, t = r["q"], r["t"]
else:
    q, t = bench.make_c2(0)
h2d("after " + mode, q, t)
for i in range(3):
    s, n, bi, bo = bench.e2e_l2(g, q, t, 20)
    print("e2e %.3f ms" % (s * 1e3), flush=True)
h2d("end", q, t)
