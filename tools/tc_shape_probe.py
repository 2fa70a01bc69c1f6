"""Ranking-kernel time at the C2 geometry (20000 x 20000: 79 query blocks x 157 train tiles) for the Hamming forms, to separate the
effect of the problem shape from the MMA kind: python tools/tc_shape_probe.py"""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import matchinglib_poselib_b200 as m
lib = m._lib.load(); m._lib.check(lib.mpl_init(0), "init")
dev = torch.device("cuda:0")
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
for nq, nt, nb in ((20000, 20000, 32), (20000, 20000, 16), (20000, 20000, 64), (20000, 200000, 32)):
    g = torch.Generator(device=dev); g.manual_seed(1)
    dq = torch.randint(0, 256, (nq, nb), dtype=torch.uint8, device=dev, generator=g)
    dt = torch.randint(0, 256, (nt, nb), dtype=torch.uint8, device=dev, generator=g)
    w16 = nb // 16
    sb = int(lib.mpl_knn_hamming_scratch_bytes(nq, nt, w16)); scr = torch.empty(sb + 256, dtype=torch.uint8, device=dev)
    k = torch.empty((nq, 2), dtype=torch.int64, device=dev)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    for path in (3, 2):
        lib.mpl_set_hamming_path(path)
        for i in range(3):
            lib.mpl_knn2_hamming_dev(dq.data_ptr(), nq, dt.data_ptr(), nt, w16, 0, scr.data_ptr(), sb, k.data_ptr(), st)
        torch.cuda.synchronize()
        lib.mpl_prof_enable(1)
        for c in (2, 3, 7):
            ms, n = ctypes.c_double(0), ctypes.c_longlong(0); lib.mpl_prof_read(c, ctypes.byref(ms), ctypes.byref(n))
        for i in range(5):
            flush.fill_(i)
            lib.mpl_knn2_hamming_dev(dq.data_ptr(), nq, dt.data_ptr(), nt, w16, 0, scr.data_ptr(), sb, k.data_ptr(), st)
        torch.cuda.synchronize()
        ms, n = ctypes.c_double(0), ctypes.c_longlong(0); lib.mpl_prof_read(2, ctypes.byref(ms), ctypes.byref(n))
        lib.mpl_prof_enable(0)
        per = ms.value / 5
        tiles = ((nq + 255) // 256) * ((nt + 127) // 128)
        print("%d x %d x %d B  %s: rank %.4f ms, %.0f clk/tile at 1.965 GHz on 148 SMs, %.0f TFLOP/s" %
              (nq, nt, nb, "e4m3" if path == 2 else "bf16", per, per * 1e-3 * 1.965e9 / (tiles / 148.0), 2.0 * nb * 8 * nq * nt / (per * 1e-3) / 1e12))
    lib.mpl_set_hamming_path(0)
