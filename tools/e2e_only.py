"""getMatches(LINEAR) at C2 with pinned host buffers: time series of the call (diagnostic, GPU box)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import bench
g = bench.Gpu(0)
m = g.m
q, t = bench.make_c2(0)
kp = range(20000)
for trial in range(2):
    (_, hq), (_, ht) = bench.pinned(q, torch), bench.pinned(t, torch)
    for mode in ("flush",):
        ts = []
        for i in range(40):
            if mode == "flush":
                g.flush_l2()
            torch.cuda.synchronize()
            if mode == "sleep1ms":
                time.sleep(0.001)
            t0 = time.perf_counter()
            m.getMatchesArray(kp, kp, hq, ht, None, "LINEAR", False, True)
            ts.append((time.perf_counter() - t0) * 1e3)
        print(trial, mode, "chunks", os.environ.get("MPL_L2_UPLOAD_CHUNKS"), "median %.3f min %.3f" % (np.median(ts[5:]), min(ts)))
