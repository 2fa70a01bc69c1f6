"""Generate tests/golden/knn_golden.npz from the REAL reference arithmetic, run in this container:

  * LINEAR     : python cv2.flann_Index(train, {'algorithm': 0}[, FLANN_DIST_HAMMING]).knnSearch -- the same
                 cvflann linear index the reference calls (matchers.cpp:567-588, 634-664), OpenCV 4.13 module.
  * BRUTEFORCENMS : oracle/_ref/libnmslib_ref.so = the reference's own vendored NMSLIB sources compiled from
                 /root/reference (oracle/Makefile target `ref`) behind the restated wrapper.

The inputs are stored with the outputs so the fixtures do not depend on numpy's RNG stream.
Run:  python oracle/gen_golden_knn.py      (needs cv2 and /root/reference; the GPU box needs neither)
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import cv2  # noqa: E402
import datagen  # noqa: E402
from oracle import knn as orc  # noqa: E402

FLANN_DIST_HAMMING = 9


def main():
    orc.build(ref=True)
    out = {}
    cases = [("orb32", 32, 257, 700), ("akaze61", 61, 130, 300), ("brisk64", 64, 100, 260), ("odd7", 7, 40, 90)]
    for name, nb, nq, nt in cases:
        q, t = datagen.binary_descriptors(nq, nt, nb, seed=100 + nb)
        fi = cv2.flann_Index(t, {"algorithm": 0}, FLANN_DIST_HAMMING)
        ci, cd = fi.knnSearch(q, 2)
        ri, rd = orc.ref_nmslib_hamming(q, t, threads=8)
        out.update({f"{name}_q": q, f"{name}_t": t, f"{name}_lin_idx": ci.astype(np.int32), f"{name}_lin_dist": cd.astype(np.int32),
                    f"{name}_nms_idx": ri, f"{name}_nms_dist": rd})
    fcases = [("sift128", 128, 150, 500, "sift"), ("surf64", 64, 120, 400, "signed"), ("odd61", 61, 50, 200, "signed")]
    for name, d, nq, nt, kind in fcases:
        q, t = (datagen.sift_like(nq, nt, d, seed=200 + d) if kind == "sift" else datagen.signed_float(nq, nt, d, seed=200 + d))
        fi = cv2.flann_Index(t, {"algorithm": 0})
        ci, cd = fi.knnSearch(q, 2)
        ri, rd = orc.ref_nmslib_l2(q, t, threads=8)
        out.update({f"{name}_q": q, f"{name}_t": t, f"{name}_lin_idx": ci.astype(np.int32), f"{name}_lin_dist": cd.astype(np.float32),
                    f"{name}_nms_idx": ri, f"{name}_nms_dist": rd})
    path = os.path.join(ROOT, "tests", "golden", "knn_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes; cv2", cv2.__version__)


if __name__ == "__main__":
    main()
