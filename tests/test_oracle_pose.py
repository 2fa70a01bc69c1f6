"""CPU: the numpy restatement of getPoseTriangPts (oracle/pose.py) against the golden vectors produced with cv2's own
SVDecomp / triangulatePoints (oracle/gen_golden_pose.py -> tests/golden/pose_golden.npz)."""
import os

import numpy as np
import pytest

from oracle import pose as orc

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "pose_golden.npz"))


def check_against_golden(name, good, R, t, Q, mask):
    assert good == int(G[name + "_good"])
    assert np.array_equal(np.asarray(mask).reshape(-1) != 0, G[name + "_mask"].reshape(-1) != 0)
    assert np.allclose(R, G[name + "_R"], atol=1e-9) and np.allclose(np.asarray(t).reshape(3), G[name + "_t"].reshape(3), atol=1e-9)
    gq = G[name + "_Q"]
    fin = np.isfinite(gq).all(1) & (np.abs(gq).max(1) < 1e6)      # points near infinity are ill-conditioned by nature
    assert np.allclose(Q[fin], gq[fin], rtol=1e-7, atol=1e-9)


@pytest.mark.parametrize("name", ["a", "b", "c"])
def test_recover_pose_matches_cv2_golden(name):
    mi = G[name + "_mask_in"]
    mask = mi if mi.size else np.ones(G[name + "_p1"].shape[0], np.uint8)
    good, R, t, Q, m = orc.recover_pose(G[name + "_E"], G[name + "_p1"], G[name + "_p2"], mask=mask, dist=float(G[name + "_dist"]))
    check_against_golden(name, good, R, t, Q, m)
    # the recovered pose is the scene's: rotation within 0.5 deg of the ground truth, t parallel to it
    Rt = G[name + "_Rtrue"]
    ang = np.degrees(np.arccos(np.clip((np.trace(R @ Rt.T) - 1) / 2, -1, 1)))
    assert ang < 0.5 and float(t.reshape(3) @ G[name + "_ttrue"]) > 0.99


def test_translational_essential():
    good, R, t, Q, m = orc.recover_pose(G["t_E"], G["t_p1"], G["t_p2"], mask=np.ones(300, np.uint8), translatE=True)
    check_against_golden("t", good, R, t, Q, m)
    assert np.array_equal(R, np.eye(3))
