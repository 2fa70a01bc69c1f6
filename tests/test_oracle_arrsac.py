"""CPU: the ARRSAC oracle -- the reference's own theia::Arrsac template compiled from /root/reference (oracle/_ref/
libarrsac_ref.so) around the restated estimator, plus the numpy restatement of robustEssentialRefine -- recovers the ground
truth of seeded scenes.  (ARRSAC parity is statistical; this pins that the oracle itself behaves like the algorithm.)"""
import numpy as np
import pytest

import datagen_pose
from oracle import arrsac, ransac


def _ed(a, b):
    a = a / np.linalg.norm(a); b = b / np.linalg.norm(b)
    return min(np.linalg.norm(a - b), np.linalg.norm(a + b))


@pytest.fixture(scope="module", autouse=True)
def _built():
    arrsac.build()
    if arrsac.ref_lib() is None:
        pytest.skip("oracle/_ref/libarrsac_ref.so not built (no /root/reference here)")


@pytest.mark.parametrize("n,outl,seed", [(1000, 0.3, 1), (3000, 0.5, 2), (300, 0.2, 3), (10000, 0.3, 4), (150, 0.6, 5)])
def test_oracle_arrsac_recovers_scene(n, outl, seed):
    sc = datagen_pose.make_scene(n, outl, seed=seed)
    r = arrsac.estimate(sc["p1"], sc["p2"], sc["th"], True)
    truth = ransac.sampson(sc["E"], sc["p1"], sc["p2"], sc["th"])[0]
    assert r["ok"] and r["count"] >= 0.8 * truth and _ed(r["E"], sc["E"]) < 5e-2
    assert r["count"] == int(r["mask"].sum())


def test_oracle_arrsac_rejects_noise():
    rng = np.random.default_rng(0)
    p1 = rng.uniform(-0.8, 0.8, (400, 2)); p2 = rng.uniform(-0.8, 0.8, (400, 2))
    r = arrsac.estimate(p1, p2, 1e-3, True)
    assert not r["ok"]


def test_robust_refine_improves_a_perturbed_model():
    sc = datagen_pose.make_scene(2000, 0.0, seed=9)
    E0 = sc["E"] + 2e-3 * np.random.default_rng(1).standard_normal((3, 3))
    E1 = arrsac.robust_refine(sc["p1"], sc["p2"], E0, sc["th"] / 50.0)
    assert _ed(E1, sc["E"]) < 0.5 * _ed(E0, sc["E"])
