"""Hamming k-NN: popc scan (path 1) vs tensor cores (path 2) through mpl_knn2_hamming_dev, device-resident inputs, CUDA events.
   python tools/hamming_tc_time.py [nq nt nbytes]..."""
import ctypes, os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import matchinglib_poselib_b200 as m
lib = m._lib.load(); m._lib.check(lib.mpl_init(0), "init")
dev = torch.device("cuda:0")
shapes = [(8192, 8192, 32), (8192, 8192, 64), (5000, 5000, 64), (10000, 1250000, 32), (10000, 12500000, 32)]
if len(sys.argv) > 3:
    a = list(map(int, sys.argv[1:])); shapes = [tuple(a[i:i + 3]) for i in range(0, len(a), 3)]
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
out = []
for nq, nt, nb in shapes:
    g = torch.Generator(device=dev); g.manual_seed(nq + nt)
    dq = torch.randint(0, 256, (nq, nb), dtype=torch.uint8, device=dev, generator=g)
    dt = torch.randint(0, 256, (nt, nb), dtype=torch.uint8, device=dev, generator=g)
    w16 = nb // 16
    sb = int(lib.mpl_knn_hamming_scratch_bytes(nq, nt, w16)); scr = torch.empty(sb + 256, dtype=torch.uint8, device=dev)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    res = {}; keys = {}
    for path in (1, 3, 2):
        lib.mpl_set_hamming_path(path)
        k = torch.empty((nq, 2), dtype=torch.int64, device=dev)
        ts = []
        reps = 3 if nq * nt > 1e10 else 12
        for i in range(reps):
            flush.fill_(i)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            m._lib.check(lib.mpl_knn2_hamming_dev(dq.data_ptr(), nq, dt.data_ptr(), nt, w16, 0, scr.data_ptr(), sb, k.data_ptr(), st), "dev")
            e1.record(); torch.cuda.synchronize()
            if i >= 2 or reps == 3 and i >= 1: ts.append(e0.elapsed_time(e1))
        res[path] = float(np.median(ts)); keys[path] = k
    lib.mpl_set_hamming_path(0)
    same = bool(torch.equal(keys[1], keys[2])) and bool(torch.equal(keys[1], keys[3]))
    r = dict(nq=nq, nt=nt, nbytes=nb, popc_ms=res[1], tc_bf16_ms=res[3], tc_e4m3_ms=res[2], speedup_e4m3=res[1] / res[2], keys_identical=same,
             popc_pairs_per_s=nq * nt / res[1] * 1e3, tc_e4m3_pairs_per_s=nq * nt / res[2] * 1e3, scratch_mb=sb / 2**20)
    print(json.dumps(r)); out.append(r)
    del dq, dt, scr
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "hamming_tc_time.json"), "w"), indent=1)
