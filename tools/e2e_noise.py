"""Which component is slow when the end-to-end call is slow?  Interleaved time series (diagnostic, GPU box)."""
import ctypes, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import bench
g = bench.Gpu(0)
m = g.m
q, t = bench.make_c2(0)
(tq, hq), (tt, ht) = bench.pinned(q, torch), bench.pinned(t, torch)
dq = torch.empty_like(tq, device="cuda"); dt = torch.empty_like(tt, device="cuda")
kp = range(20000)
rows = []
def ev():
    return torch.cuda.Event(enable_timing=True)
for it in range(60):
    a, b, c = ev(), ev(), ev()
    torch.cuda.synchronize()
    a.record(); dq.copy_(tq, non_blocking=True); b.record(); dt.copy_(tt, non_blocking=True); c.record(); torch.cuda.synchronize()
    h1, h2 = a.elapsed_time(b), b.elapsed_time(c)
    t0 = time.perf_counter(); m.getMatchesArray(kp, kp, hq, ht, None, "LINEAR", False, True); e2e = (time.perf_counter() - t0) * 1e3
    t0 = time.perf_counter(); x = hq.sum(); cpu = (time.perf_counter() - t0) * 1e3          # host memory read of 10 MB
    rows.append((h1, h2, e2e, cpu))
    time.sleep(0.03)
r = np.array(rows)
print("columns: h2d_q h2d_t e2e host_read_10MB (ms)")
for x in r[::3]:
    print(" ".join("%.3f" % v for v in x))
print("median", " ".join("%.3f" % v for v in np.median(r, axis=0)), " max", " ".join("%.3f" % v for v in r.max(axis=0)))
