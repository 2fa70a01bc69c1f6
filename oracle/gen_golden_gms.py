"""Golden vectors for the GMS filter: inputs + the masks the REFERENCE's own MatchGMS.cpp returns (oracle/_ref/libgms_ref.so,
compiled from /root/reference by `make -C oracle ref_gms`).  Run in the build container:  python oracle/gen_golden_gms.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import datagen_gms
from oracle import gms

out = {}
for name, kw in datagen_gms.CASES.items():
    k1, k2, m, sz = datagen_gms.scene(**kw["scene"])
    r = gms.ref_filter_matches(k1, sz, k2, sz, m, kw["use_scale"], kw["use_rotation"])
    assert r is not None, "oracle/_ref/libgms_ref.so missing: make -C oracle ref_gms"
    out[name + "_n"] = np.int32(r[0]); out[name + "_mask"] = r[1]
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "gms_golden.npz"), **out)
print({k: int(v) for k, v in out.items() if k.endswith("_n")})
