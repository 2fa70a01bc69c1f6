"""Golden vectors for getPoseTriangPts: the reference's recoverPose steps (five-point.cpp:150-338) executed with
OpenCV's OWN primitives -- cv2.SVDecomp for decomposeEssentialMat, cv2.triangulatePoints for the DLT -- on seeded
scenes.  Run here (cv2 python build 4.13; the reference pins OpenCV 4.2.0, same algorithms) and commit the output:
    python oracle/gen_golden_pose.py  ->  tests/golden/pose_golden.npz
"""
import os
import sys

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import datagen_pose
from oracle import pose as orc


def cv_svd(E):
    w, U, Vt = cv2.SVDecomp(E)
    return w.ravel(), U, Vt


def cv_triang(P0, P1, x1, x2):
    return cv2.triangulatePoints(P0, P1, np.ascontiguousarray(x1), np.ascontiguousarray(x2))


def main():
    out = {}
    cases = [("a", 600, 0.3, 11, 50.0, True), ("b", 257, 0.0, 12, 130.0, False), ("c", 1000, 0.5, 13, 8.0, True)]
    for name, n, outl, seed, dist, with_mask in cases:
        sc = datagen_pose.make_scene(n, outl, seed=seed)
        rng = np.random.default_rng(seed)
        mask = (rng.random(n) < 0.8).astype(np.uint8) if with_mask else None
        good, R, t, Q, m = orc.recover_pose(sc["E"], sc["p1"], sc["p2"], mask=mask if with_mask else np.ones(n, np.uint8),
                                            dist=dist, triang=cv_triang, svd=cv_svd)
        out.update({name + "_E": sc["E"], name + "_p1": sc["p1"], name + "_p2": sc["p2"], name + "_dist": dist,
                    name + "_mask_in": mask if with_mask else np.zeros(0, np.uint8), name + "_good": good, name + "_R": R,
                    name + "_t": t, name + "_Q": Q, name + "_mask": m, name + "_Rtrue": sc["R"], name + "_ttrue": sc["t"]})
    # translational essential matrix (R = I)
    sc = datagen_pose.make_scene(300, 0.2, seed=21)
    t = sc["t"]; E = np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]])
    z = np.random.default_rng(3).uniform(2, 20, 300)
    X = np.concatenate([sc["p1"] * z[:, None], z[:, None]], 1) + t
    p2 = X[:, :2] / X[:, 2:3]
    good, R, tt, Q, m = orc.recover_pose(E, sc["p1"], p2, mask=np.ones(300, np.uint8), dist=50.0, translatE=True, triang=cv_triang)
    out.update(dict(t_E=E, t_p1=sc["p1"], t_p2=p2, t_good=good, t_R=R, t_t=tt, t_Q=Q, t_mask=m))
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "pose_golden.npz"), **out)
    print({k: (v if np.ndim(v) == 0 else np.shape(v)) for k, v in out.items() if k.endswith("good")})


if __name__ == "__main__":
    main()
