"""Pinned-memory flavours vs H2D rate (diagnostic, GPU box): torch pin_memory vs mmap + MADV_HUGEPAGE + cudaHostRegister."""
import ctypes, mmap, os, sys, time
import numpy as np
import torch
print("THP:", open("/sys/kernel/mm/transparent_hugepage/enabled").read().strip(), "| defrag:", open("/sys/kernel/mm/transparent_hugepage/defrag").read().strip())
torch.cuda.init()
n = 20000 * 128 * 4
dev = torch.empty(n, dtype=torch.uint8, device="cuda")
junk = [np.random.default_rng(i).standard_normal(3_000_000) for i in range(40)]   # fragment the heap a little (1 GB)
junk = junk[::2]
def rate(tag, t):
    ts = []
    for _ in range(40):
        torch.cuda.synchronize(); t0 = time.perf_counter(); dev.copy_(t, non_blocking=True); torch.cuda.synchronize(); ts.append(time.perf_counter() - t0)
    print("%-40s %.3f ms  %.1f GB/s" % (tag, np.median(ts[10:]) * 1e3, n / np.median(ts[10:]) / 1e9), flush=True)
for trial in range(3):
    a = torch.empty(n, dtype=torch.uint8).pin_memory(); a.fill_(1)
    rate("torch pin_memory #%d" % trial, a)
libc = ctypes.CDLL("libc.so.6", use_errno=True)
rt = torch.cuda.cudart()
for trial in range(3):
    size = ((n + (2 << 20) - 1) >> 21 << 21) + (2 << 20)
    mm = mmap.mmap(-1, size, flags=mmap.MAP_PRIVATE | mmap.MAP_ANONYMOUS)
    buf = (ctypes.c_char * size).from_buffer(mm)
    base = ctypes.addressof(buf)
    al = (base + (2 << 20) - 1) >> 21 << 21
    r = libc.madvise(ctypes.c_void_p(al), ctypes.c_size_t(size - (al - base)), 14)
    arr = np.frombuffer(mm, dtype=np.uint8, count=n, offset=al - base)
    arr[:] = 1
    err = rt.cudaHostRegister(al, n, 0)
    t = torch.from_numpy(arr)
    rate("mmap+MADV_HUGEPAGE+register #%d (madvise %d, reg %s)" % (trial, r, err), t)
