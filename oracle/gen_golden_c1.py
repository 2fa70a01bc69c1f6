"""Fixture for BASELINE config C1: ORB (8000 features, cv2) on the reference's bundled stereo pair
tests/poselib-test/imgs/stereo/{left,right}_0000000000.png, plus the calibration of cameras 00/01
(calib_cam_to_cam.txt: K, D, R_01, T_01).  The images cannot travel to the GPU box, the descriptors can:
    python oracle/gen_golden_c1.py  ->  tests/golden/c1_stereo_orb.npz
Keypoints are stored in pixels and as camera-normalised, undistorted coordinates (cv2.undistortPoints; the reference
driver does Remove_LensDist + ImgToCamCoordTrans, tests/poselib-test/main.cpp:1522-1599)."""
import os

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STEREO = "/root/reference/matchinglib_poselib/source/tests/poselib-test/imgs/stereo"


def calib():
    d = {}
    for line in open(os.path.join(STEREO, "calib_cam_to_cam.txt")):
        k, v = line.split(":", 1)
        if k in ("K_00", "D_00", "K_01", "D_01", "R_01", "T_01"):
            d[k] = np.array([float(x) for x in v.split()])
    return d


def main():
    c = calib()
    K0, K1 = c["K_00"].reshape(3, 3), c["K_01"].reshape(3, 3)
    orb = cv2.ORB_create(nfeatures=8000)
    out = {}
    for side, K, D in (("l", K0, c["D_00"]), ("r", K1, c["D_01"])):
        img = cv2.imread(os.path.join(STEREO, ("left" if side == "l" else "right") + "_0000000000.png"), cv2.IMREAD_GRAYSCALE)
        kp, desc = orb.detectAndCompute(img, None)
        pts = np.array([k.pt for k in kp], np.float32)
        norm = cv2.undistortPoints(pts.reshape(-1, 1, 2).astype(np.float64), K, D).reshape(-1, 2)
        out[side + "_desc"] = desc
        out[side + "_pts"] = pts
        out[side + "_norm"] = norm
    out.update(K0=K0, K1=K1, R01=c["R_01"].reshape(3, 3), T01=c["T_01"])
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "c1_stereo_orb.npz"), **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
