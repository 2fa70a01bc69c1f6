"""GPU: poselib::estimateEssentialMat with its DEFAULT method "ARRSAC" (mpl_arrsac5pt: device hypothesis generation +
verification, the reference's schedule on the host, robustEssentialRefine on the device).  Parity with the reference is
STATISTICAL for this randomised, adaptive algorithm (csrc/arrsac.cu says why); the bars here: both sides accept / reject the
same scenes, inlier counts within 10 % of each other and of the ground truth's support, E within 5e-2 of the ground truth,
the returned mask is exactly the inlier set of the returned (pre-refinement) model under the reference's test, and the refined
E equals the oracle's refinement of the same model to 1e-6."""
import numpy as np
import pytest

import datagen_pose
from oracle import arrsac as oarrsac, ransac as oransac

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def P():
    import __graft_entry__ as g
    g.build()
    from matchinglib_poselib_b200 import poselib, _lib
    _lib.check(_lib.load().mpl_init(0), "mpl_init")
    return poselib


def _ed(a, b):
    a = a / np.linalg.norm(a); b = b / np.linalg.norm(b)
    return min(np.linalg.norm(a - b), np.linalg.norm(a + b))


@pytest.mark.parametrize("n,outl,seed", [(1000, 0.3, 1), (3000, 0.5, 2), (300, 0.2, 3), (10000, 0.3, 4), (150, 0.6, 5), (2500, 0.05, 6), (60, 0.1, 7)])
def test_arrsac_vs_oracle_statistics(P, n, outl, seed):
    sc = datagen_pose.make_scene(n, outl, seed=seed)
    truth = oransac.sampson(sc["E"], sc["p1"], sc["p2"], sc["th"])[0]
    g = P.arrsac5pt(sc["p1"], sc["p2"], sc["th"], refine=False, seed=(0, 0))
    assert g["ok"] and g["count"] >= 0.8 * truth, (g["count"], truth)
    # the mask is the reference's findInliers of the returned model (modelest.cpp:273)
    cnt, _, _, omask = oransac.sampson(g["E"], sc["p1"], sc["p2"], sc["th"])
    assert cnt == g["count"] and np.array_equal(omask, g["mask"])
    assert _ed(g["E"], sc["E"]) < 5e-2
    o = oarrsac.arrsac(sc["p1"], sc["p2"], sc["th"])
    if o is not None:
        assert o["ok"] and abs(o["count"] - g["count"]) <= 0.1 * truth, (o["count"], g["count"], truth)
    print("n=%d outliers %.0f%%: GPU %d inliers (%d hypotheses, %d device batches), reference schedule %s, ground-truth support %d" %
          (n, 100 * outl, g["count"], g["n_hypotheses"], g["n_batches"], o["count"] if o else "n/a", truth))


def test_arrsac_refinement_equals_oracle_refinement(P):
    """robustEssentialRefine on the device vs its numpy restatement, from the same starting model and inlier set."""
    for seed in range(5):
        sc = datagen_pose.make_scene(1500 + 300 * seed, 0.3, seed=40 + seed)
        a = P.arrsac5pt(sc["p1"], sc["p2"], sc["th"], refine=False, seed=(seed + 1, seed + 7))
        b = P.arrsac5pt(sc["p1"], sc["p2"], sc["th"], refine=True, seed=(seed + 1, seed + 7))
        assert a["ok"] and b["ok"] and np.array_equal(a["mask"], b["mask"]) and a["count"] == b["count"]     # same schedule, same streams
        inl = a["mask"].astype(bool)
        exp = oarrsac.robust_refine(sc["p1"][inl], sc["p2"][inl], a["E"], sc["th"] / 50.0)
        assert _ed(b["E"], exp) < 1e-6, _ed(b["E"], exp)
        assert _ed(b["E"], sc["E"]) < 5e-2


def test_arrsac_is_the_default_method(P):
    sc = datagen_pose.make_scene(800, 0.3, seed=77)
    ok, E, mask = P.estimateEssentialMat(sc["p1"], sc["p2"], threshold=sc["th"])          # method = "ARRSAC", refine = True
    assert ok and E.shape == (3, 3) and mask.shape == (1, 800) and mask.sum() > 400 and _ed(E, sc["E"]) < 5e-2
    from matchinglib_poselib_b200 import _lib
    with pytest.raises(_lib.MplError):
        P.estimateEssentialMat(sc["p1"], sc["p2"], "MLESAC", sc["th"])                       # unknown method strings are refused


def test_arrsac_rejects_noise_and_small_sets(P):
    rng = np.random.default_rng(0)
    p1 = rng.uniform(-0.8, 0.8, (400, 2)); p2 = rng.uniform(-0.8, 0.8, (400, 2))
    r = P.arrsac5pt(p1, p2, 1e-3, refine=True, seed=(0, 0))
    assert not r["ok"] and r["count"] == 0
    sc = datagen_pose.make_scene(40, 0.0, seed=3)                                         # fewer points than one block (arrsac.h:365-378)
    r = P.arrsac5pt(sc["p1"], sc["p2"], sc["th"], refine=True, seed=(0, 0))
    assert r["ok"] and r["count"] >= 30
    r5 = P.arrsac5pt(sc["p1"][:5], sc["p2"][:5], sc["th"], refine=False)                  # count == modelPoints (modelest.cpp:222-265)
    assert r5["ok"] and r5["count"] == 5


def test_arrsac_deterministic_for_a_seed(P):
    sc = datagen_pose.make_scene(2000, 0.4, seed=11)
    a = P.arrsac5pt(sc["p1"], sc["p2"], sc["th"], seed=(5, 9))
    b = P.arrsac5pt(sc["p1"], sc["p2"], sc["th"], seed=(5, 9))
    assert np.array_equal(a["E"], b["E"]) and np.array_equal(a["mask"], b["mask"])
