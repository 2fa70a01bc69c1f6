"""Why are two back-to-back pinned H2D copies slow?  (diagnostic, GPU box)"""
import os, sys, time, subprocess
import numpy as np
import torch
print("cpus allowed:", len(os.sched_getaffinity(0)), "of", os.cpu_count())
print(subprocess.run("lscpu | grep -i -E 'numa|socket|model name'; nvidia-smi topo -m | head -8", shell=True, capture_output=True, text=True).stdout)
dev = torch.device("cuda:0")
n = 20000 * 128
a = torch.randn(n).pin_memory(); b = torch.randn(n).pin_memory()
da = torch.empty(n, device=dev); db = torch.empty(n, device=dev)
def ev_time(f, reps=20):
    f(); f()
    ts = []
    for _ in range(reps):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter(); e0.record(); f(); e1.record(); torch.cuda.synchronize(); w = time.perf_counter() - t0
        ts.append((e0.elapsed_time(e1), w * 1e3))
    ts = np.array(ts); return np.median(ts[:, 0]), np.median(ts[:, 1])
for rnd in range(3):
    print('--- round', rnd)
    print("a->da            : ev %.3f wall %.3f" % ev_time(lambda: da.copy_(a, non_blocking=True)))
    print("b->db            : ev %.3f wall %.3f" % ev_time(lambda: db.copy_(b, non_blocking=True)))
    print("a->da, a->db     : ev %.3f wall %.3f" % ev_time(lambda: (da.copy_(a, non_blocking=True), db.copy_(a, non_blocking=True))))
    print("a->da, b->db     : ev %.3f wall %.3f" % ev_time(lambda: (da.copy_(a, non_blocking=True), db.copy_(b, non_blocking=True))))
    ab = torch.randn(2 * n).pin_memory(); dab = torch.empty(2 * n, device=dev)
    print("ab->dab (20 MB)  : ev %.3f wall %.3f" % ev_time(lambda: dab.copy_(ab, non_blocking=True)))
    def split4():
        c = n // 2
        for i in range(4):
            dab[i * c:(i + 1) * c].copy_(ab[i * c:(i + 1) * c], non_blocking=True)
    print("ab->dab 4 chunks : ev %.3f wall %.3f" % ev_time(split4))
    # with a kernel touching the destination in between (as in the real step)
    def with_kernel():
        da.copy_(a, non_blocking=True); db.copy_(b, non_blocking=True); (da[:1024] + db[:1024]).sum()
    print("a,b + tiny kernel: ev %.3f wall %.3f" % ev_time(with_kernel))
    def read_all():
        da.copy_(a, non_blocking=True); db.copy_(b, non_blocking=True); torch.dot(da, db)
    print("a,b + dot kernel : ev %.3f wall %.3f" % ev_time(read_all))
