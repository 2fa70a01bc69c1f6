"""GPU parity of the GMS filter kernel (mpl_gms_filter) against the oracle and the reference-made golden masks: bit-exact
inlier masks and counts (integer / grid arithmetic)."""
import os

import numpy as np
import pytest

import datagen_gms
from oracle import gms

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "gms_golden.npz"))


@pytest.fixture(scope="module")
def m():
    import __graft_entry__ as g
    g.build()
    import matchinglib_poselib_b200 as mod
    mod._lib.check(mod._lib.load().mpl_init(0), "mpl_init")
    return mod


@pytest.mark.parametrize("name", sorted(datagen_gms.CASES))
def test_gms_golden(m, name):
    kw = datagen_gms.CASES[name]
    k1, k2, mt, sz = datagen_gms.scene(**kw["scene"])
    n, mask = m.filterMatchesGMS(k1, sz, k2, sz, datagen_gms.dmatch(mt), kw["use_scale"], kw["use_rotation"])
    assert n == int(G[name + "_n"]) and np.array_equal(mask.astype(np.uint8), G[name + "_mask"])


def test_gms_vs_oracle_sweep(m):
    for seed in range(40, 60):
        k1, k2, mt, sz = datagen_gms.scene(n=200 + 577 * (seed - 40), seed=seed, frac=0.2 + 0.04 * (seed % 10), rot=0.05 * (seed % 5), sc=1.0 + 0.07 * (seed % 6),
                                            border=bool(seed % 2))
        for us, ur in ((0, 0), (1, 0), (0, 1)) if seed % 4 == 0 else ((0, 0),):
            n, mask = m.filterMatchesGMS(k1, sz, k2, sz, datagen_gms.dmatch(mt), us, ur)
            on, omask = gms.filter_matches(k1, sz, k2, sz, mt, us, ur)
            assert n == on and np.array_equal(mask.astype(np.uint8), omask), (seed, us, ur)


def test_gms_edge_cases(m):
    k1, k2, mt, sz = datagen_gms.scene(n=100, seed=1, frac=0.0)
    n, mask = m.filterMatchesGMS(k1, sz, k2, sz, datagen_gms.dmatch(mt))
    assert n == 0 and not mask.any()
    n, mask = m.filterMatchesGMS(k1, sz, k2, sz, datagen_gms.dmatch(mt[:0]))
    assert n == 0 and len(mask) == 0
    bad = datagen_gms.dmatch(mt); bad["trainIdx"][3] = 10 ** 6
    with pytest.raises(m._lib.MplError):
        m.filterMatchesGMS(k1, sz, k2, sz, bad)


def test_get_correspondences_gms_refine(m):
    """getCorrespondences(GMSrefine=True) (correspondences.cpp:378-396): the match list is replaced by its GMS inliers."""
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(0)
    img1 = (rng.random((256, 384)) * 255).astype(np.uint8)
    img1 = cv2.GaussianBlur(img1, (0, 0), 2.0)
    img2 = np.roll(img1, (3, 5), (0, 1))
    e0, m0, kp1, kp2 = m.getCorrespondences(img1, img2, "ORB", "ORB", "LINEAR", limitNrfeatures=1500)
    e1, m1, _, _ = m.getCorrespondences(img1, img2, "ORB", "ORB", "LINEAR", limitNrfeatures=1500, GMSrefine=True)
    assert e0 == 0 and e1 == 0 and 2 <= len(m1) <= len(m0)
    pts1 = np.array([k.pt for k in kp1], np.float32); pts2 = np.array([k.pt for k in kp2], np.float32)
    on, omask = gms.filter_matches(pts1, (384, 256), pts2, (384, 256), np.stack([m0["queryIdx"], m0["trainIdx"]], 1))
    assert np.array_equal(m1, m0[omask.astype(bool)] if on >= 2 else m0)
