"""Device time of the ranking kernel at C2 (library timers), for A/B builds:  python tools/tc_time.py [signed] [dim=200] [path=1]"""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch, datagen
import matchinglib_poselib_b200 as m
lib = m._lib.load(); m._lib.check(lib.mpl_init(0), "init")
D = int(([a[4:] for a in sys.argv if a.startswith("dim=")] or [128])[0])
lib.mpl_set_l2_path(int(([a[5:] for a in sys.argv if a.startswith("path=")] or [0])[0]))
q, t = (datagen.signed_float if "signed" in sys.argv else datagen.sift_like)(20000, 20000, D, seed=1002)
dev = torch.device("cuda:0")
dq, dt = torch.from_numpy(q).to(dev), torch.from_numpy(t).to(dev)
sb = int(lib.mpl_knn_l2_scratch_bytes(20000, 20000, D)); scr = torch.empty(sb, dtype=torch.uint8, device=dev)
keys = torch.empty((20000, 2), dtype=torch.int64, device=dev)
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
lib.mpl_prof_enable(1)
for i in range(12):
    lib.mpl_knn2_l2f32_dev(dq.data_ptr(), 20000, D, dt.data_ptr(), 20000, D, D, 0, 0, scr.data_ptr(), sb, keys.data_ptr(), st)
    torch.cuda.synchronize()
    if i == 1:
        for c in range(8):
            ms, n = ctypes.c_double(0), ctypes.c_longlong(0); lib.mpl_prof_read(c, ctypes.byref(ms), ctypes.byref(n))
res = {}
for name, c in dict(l2_exact=1, l2_tc=2, rerank=3, merge=4, prep=7).items():
    ms, n = ctypes.c_double(0), ctypes.c_longlong(0); lib.mpl_prof_read(c, ctypes.byref(ms), ctypes.byref(n))
    res[name] = round(ms.value / max(n.value, 1), 5)
print(os.environ.get("TAG", ""), "dim", D, res)
