"""GPU parity of getPoseTriangPts (csrc/recover_pose.cu, through the C ABI) against the golden vectors made with cv2's
primitives and against the numpy oracle on fresh scenes; plus the estimateEssentialMat -> getPoseTriangPts chain."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(__file__))
import datagen_pose
from oracle import pose as orc
from test_oracle_pose import G, check_against_golden

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pl():
    import __graft_entry__ as g
    g.build()
    from matchinglib_poselib_b200 import poselib, _lib
    _lib.check(_lib.load().mpl_init(0), "mpl_init")
    return poselib


@pytest.mark.parametrize("name", ["a", "b", "c"])
def test_golden(pl, name):
    mi = G[name + "_mask_in"]
    good, R, t, Q, m = pl.getPoseTriangPts(G[name + "_E"], G[name + "_p1"], G[name + "_p2"], mask=mi if mi.size else None,
                                           dist=float(G[name + "_dist"]))
    check_against_golden(name, good, R, t, Q, m)


def test_golden_translational(pl):
    good, R, t, Q, m = pl.getPoseTriangPts(G["t_E"], G["t_p1"], G["t_p2"], translatE=True)
    check_against_golden("t", good, R, t, Q, m)


@pytest.mark.parametrize("n,seed", [(5, 1), (33, 2), (10000, 3)])
def test_against_oracle(pl, n, seed):
    sc = datagen_pose.make_scene(n, 0.3, seed=seed)
    mask = (np.random.default_rng(seed).random(n) < 0.7).astype(np.uint8)
    good, R, t, Q, m = pl.getPoseTriangPts(sc["E"], sc["p1"], sc["p2"], mask=mask, dist=130.0)
    eg, eR, et, eQ, em = orc.recover_pose(sc["E"], sc["p1"], sc["p2"], mask=mask, dist=130.0)
    assert good == eg and np.array_equal(m.reshape(-1) != 0, em != 0)
    assert np.allclose(R, eR, atol=1e-9) and np.allclose(t, et, atol=1e-9)
    fin = np.isfinite(eQ).all(1) & (np.abs(eQ).max(1) < 1e6)
    assert np.allclose(Q[fin], eQ[fin], rtol=1e-7, atol=1e-9)
    assert m.reshape(-1)[mask == 0].max(initial=0) == 0        # AND with the caller's mask


def test_chain_ransac_then_pose(pl):
    """estimateEssentialMat("RANSAC") -> getPoseTriangPts recovers the scene's R, t (the driver's sequence,
    tests/poselib-test/main.cpp:1597-1660)."""
    sc = datagen_pose.make_scene(2000, 0.3, seed=9)
    ok, E, mask = pl.estimateEssentialMat(sc["p1"], sc["p2"], "RANSAC", sc["th"], True)
    assert ok
    good, R, t, Q, m = pl.getPoseTriangPts(E, sc["p1"], sc["p2"], mask=mask, dist=130.0)
    ang = np.degrees(np.arccos(np.clip((np.trace(R @ sc["R"].T) - 1) / 2, -1, 1)))
    assert good > 1000 and ang < 0.5 and float(t.reshape(3) @ sc["t"]) > 0.99


def test_missing_outputs_and_empty(pl):
    from matchinglib_poselib_b200 import _lib
    lib = _lib.load()
    E = np.eye(3)
    assert lib.mpl_get_pose_triang_pts_api(_lib.ptr(E), None, None, 0, None, 0, 50.0, 0, None, None, None) == -1
