"""Timeline of the chunked upload + scan (diagnostic, GPU box): event timestamps on the copy and compute streams."""
import ctypes, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import bench
g = bench.Gpu(0)
lib, m = g.lib, g.m
q, t = bench.make_c2(0)
(tq, hq), (tt, ht) = bench.pinned(q, torch), bench.pinned(t, torch)
nq, nt, dim = 20000, 20000, 128
dq = torch.empty_like(tq, device="cuda"); dt = torch.empty_like(tt, device="cuda")
scr_bytes = int(lib.mpl_knn_l2_scratch_bytes(nq, nt, dim))
scr = torch.empty(scr_bytes, dtype=torch.uint8, device="cuda")
cs = torch.cuda.Stream(); ms = torch.cuda.current_stream()
def run(nch, trace=False):
    parts = torch.empty((nch, nq, 2), dtype=torch.int64, device="cuda")
    keys = torch.empty((nq, 2), dtype=torch.int64, device="cuda")
    E = lambda: torch.cuda.Event(enable_timing=True)
    e0 = E(); e0.record(ms); cs.wait_event(e0)
    marks = []
    with torch.cuda.stream(cs):
        dq.copy_(tq, non_blocking=True)
        e = E(); e.record(cs); marks.append(("q up", e))
    per = ((nt + nch - 1) // nch + 127) // 128 * 128
    c = 0
    for r0 in range(0, nt, per):
        rows = min(per, nt - r0)
        with torch.cuda.stream(cs):
            dt[r0:r0 + rows].copy_(tt[r0:r0 + rows], non_blocking=True)
            e = E(); e.record(cs); marks.append(("t%d up" % c, e))
        ms.wait_event(e)
        m._lib.check(lib.mpl_knn2_l2f32_dev(dq.data_ptr(), nq, dim, dt[r0:].data_ptr(), rows, dim, dim, 0, r0, scr.data_ptr(), scr_bytes,
                                            parts[c].data_ptr(), ctypes.c_void_p(ms.cuda_stream)), "dev")
        e = E(); e.record(ms); marks.append(("scan%d" % c, e))
        c += 1
    m._lib.check(lib.mpl_merge_top2_dev(parts.data_ptr(), c, nq, 2, 1, None, None, keys.data_ptr(), ctypes.c_void_p(ms.cuda_stream)), "merge")
    e = E(); e.record(ms); marks.append(("merge", e))
    torch.cuda.synchronize()
    if trace:
        print("chunks=%d: " % nch + "  ".join("%s %.3f" % (n, e0.elapsed_time(e)) for n, e in marks))
    return e0.elapsed_time(marks[-1][1])
for nch in (1, 2, 3, 4):
    for _ in range(30): run(nch)
    ts = [run(nch) for _ in range(20)]
    run(nch, True)
    print("chunks=%d median %.3f ms" % (nch, np.median(ts)))

print("---- query chunks (train rows first, one piece)")
def runq(nch, trace=False):
    keys = torch.empty((nq, 2), dtype=torch.int64, device="cuda")
    E = lambda: torch.cuda.Event(enable_timing=True)
    e0 = E(); e0.record(ms); cs.wait_event(e0)
    marks = []
    with torch.cuda.stream(cs):
        dt.copy_(tt, non_blocking=True)
        e = E(); e.record(cs); marks.append(("t up", e))
    per = ((nq + nch - 1) // nch + 127) // 128 * 128
    c = 0
    for r0 in range(0, nq, per):
        rows = min(per, nq - r0)
        with torch.cuda.stream(cs):
            dq[r0:r0 + rows].copy_(tq[r0:r0 + rows], non_blocking=True)
            e = E(); e.record(cs); marks.append(("q%d up" % c, e))
        ms.wait_event(e)
        m._lib.check(lib.mpl_knn2_l2f32_dev(dq[r0:].data_ptr(), rows, dim, dt.data_ptr(), nt, dim, dim, 0, 0, scr.data_ptr(), scr_bytes,
                                            keys[r0:].data_ptr(), ctypes.c_void_p(ms.cuda_stream)), "dev")
        e = E(); e.record(ms); marks.append(("scan%d" % c, e))
        c += 1
    torch.cuda.synchronize()
    if trace:
        print("qchunks=%d: " % nch + "  ".join("%s %.3f" % (n, e0.elapsed_time(e)) for n, e in marks))
    return e0.elapsed_time(marks[-1][1])
for nch in (1, 2, 3, 4):
    for _ in range(30): runq(nch)
    ts = [runq(nch) for _ in range(20)]
    runq(nch, True)
    print("qchunks=%d median %.3f ms" % (nch, np.median(ts)))
