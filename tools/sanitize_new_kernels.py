"""Small invocations of the kernels added late in round 2, small shapes, odd sizes (compute-sanitizer is closed on the GPU pool, so this is a plain run that must exit 0):
tensor-core Hamming (single, chunked, batch; e4m3 and bf16), wide float rows, LMedS, VFC, USAC."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import datagen, datagen_pose, datagen_vfc
import matchinglib_poselib_b200 as m
from matchinglib_poselib_b200 import poselib, matchinglib
lib = m._lib.load(); m._lib.check(lib.mpl_init(0), "init")
what = sys.argv[1:] or ["hamming", "batch", "wide", "lmeds", "vfc", "usac"]
if "hamming" in what:
    for path in (2, 3):
        lib.mpl_set_hamming_path(path)
        for nq, nt, nb in ((300, 700, 32), (130, 400, 64), (70, 300, 16), (257, 129, 61)):
            q, t = datagen.binary_descriptors(nq, nt, nb, seed=nq)
            m.knn2_hamming(q, t, 2, 0)
    lib.mpl_set_hamming_path(0)
    print("hamming ok")
if "batch" in what:
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from test_pair_batch_gpu import make_pair
    pairs = []
    for i, (n1, n2) in enumerate(((300, 400), (130, 129), (5, 300), (0, 50), (400, 2))):
        q, t, p1, p2, th = make_pair(n1, n2, 61, seed=i)
        pairs.append((q, t, p1, p2))
    for path in (2, 3):
        lib.mpl_set_hamming_path(path)
        poselib.matchAndEstimatePoseBatch(pairs, "LINEAR", True, th, 32, 0.999, True, seeds=list(range(len(pairs))))
    lib.mpl_set_hamming_path(0)
    print("batch ok")
if "wide" in what:
    lib.mpl_set_l2_path(2)
    for kind in (datagen.signed_float, datagen.sift_like):
        q, t = kind(300, 500, 200, seed=3)
        m.knn2_l2(q, t, 2, False)
    lib.mpl_set_l2_path(0)
    print("wide ok")
if "lmeds" in what:
    sc = datagen_pose.make_scene(501, 0.3, seed=1)
    poselib.lmeds5pt(sc["p1"], sc["p2"], datagen_pose.draw_samples_fast(501, 20, 2))
    print("lmeds ok")
if "vfc" in what:
    k1, k2, mt, inl = datagen_vfc.scene(300, 0.4, seed=1)
    dm = np.zeros(len(mt), matchinglib.DMATCH_DTYPE); dm["queryIdx"] = mt[:, 0]; dm["trainIdx"] = mt[:, 1]; dm["distance"] = 1
    matchinglib.filterWithVFC(k1, k2, dm, seed=1)
    print("vfc ok")
if "usac" in what:
    sc = datagen_pose.make_scene(400, 0.3, seed=2)
    poselib.usac5pt(sc["p1"], sc["p2"], sc["th"], seed=1)
    print("usac ok")
