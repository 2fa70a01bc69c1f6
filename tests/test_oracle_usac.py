"""CPU: the USAC checker (oracle/usac.py) -- the reference's own estimator compiled into oracle/_ref/libusac_ref.so recovers
synthetic scenes, the numpy restatements agree with it, and the host logic of csrc/usac.cu that needs no GPU (glibc rand()
restated) reproduces libc bit for bit."""
import ctypes

import numpy as np
import pytest

import datagen_pose
from oracle import usac as ousac


def _ed(a, b):
    a = a / np.linalg.norm(a); b = b / np.linalg.norm(b)
    return min(np.linalg.norm(a - b), np.linalg.norm(a + b))


def _built():
    ousac.build()
    if ousac.ref_lib() is None:
        pytest.skip("oracle/_ref/libusac_ref.so not built (needs /root/reference)")


@pytest.mark.parametrize("n,outl,seed,est", [(1000, 0.3, 1, "NISTER"), (2000, 0.5, 2, "STEWENIUS"), (300, 0.2, 3, "NISTER")])
def test_reference_usac_recovers_scene(n, outl, seed, est):
    _built()
    sc = datagen_pose.make_scene(n, outl, seed=seed)
    truth = int((ousac.sampson(sc["E"], sc["p1"], sc["p2"]) < sc["th"] ** 2).sum())
    for idx in (None, np.arange(n, dtype=np.uint32)):
        o = ousac.usac(sc["p1"], sc["p2"], sc["th"], sorted_idx=idx, seed=7, estimator=est)
        assert o["ok"] and o["count"] >= 0.75 * truth and _ed(o["E"], sc["E"]) < 5e-2     # USAC stops at 99 % confidence: no final polish
        # the reference's inlier flags are evaluateModel's error of its final model under the threshold: pins oracle.usac.sampson
        assert np.array_equal(ousac.sampson(o["E"], sc["p1"], sc["p2"]) < sc["th"] ** 2, o["mask"].astype(bool))
        assert o["count"] == int(o["mask"].sum())


def test_reference_usac_too_few_points():
    _built()
    sc = datagen_pose.make_scene(50, 0.1, seed=4)
    assert not ousac.usac(sc["p1"][:4], sc["p2"][:4], sc["th"])["ok"]
    assert not ousac.usac(sc["p1"][:12], sc["p2"][:12], sc["th"], sorted_idx=np.arange(12))["ok"]


def test_refined_model_restatement_fits_the_inliers():
    sc = datagen_pose.make_scene(1500, 0.3, seed=5, noise_px=0.2)
    inl = ousac.sampson(sc["E"], sc["p1"], sc["p2"]) < sc["th"] ** 2
    p1n, p2n, T1, T2 = ousac.normalize_points(sc["p1"], sc["p2"])
    assert abs(np.sqrt((p1n ** 2).sum(1)).mean() - np.sqrt(2)) < 1e-9 and abs(p1n.mean()) < 1e-9
    F, E = ousac.refined_model(p1n[inl], p2n[inl], T1, T2)
    assert _ed(E, sc["E"]) < 2e-2 and abs(np.linalg.det(F)) < 1e-12
    w = ousac.torr_weights(F, p1n[inl], p2n[inl])
    F2, E2 = ousac.refined_model(p1n[inl], p2n[inl], T1, T2, w)
    assert _ed(E2, sc["E"]) < 2e-2
    assert (ousac.sampson(E2, sc["p1"], sc["p2"]) < sc["th"] ** 2).sum() >= 0.95 * inl.sum()
    A = ousac.design_sprt(0.05, 0.15, 2314.0, 8.5)
    assert 10 < A < 100


def test_reference_singulF_defect_is_pinned():
    """Why mpl_usac5pt has lo_mode: the reference's rank-2 projection is right for a generic matrix and wrong for an essential-like
    one (two nearly equal singular values), so every model its local optimisation refits is mangled."""
    _built()
    def proj(F):
        U, s, Vt = np.linalg.svd(F)
        return U @ np.diag([s[0], s[1], 0.0]) @ Vt
    G = np.array([[1, 2, 3], [4, 5, 6], [7, 8, 10.0]])
    assert np.linalg.norm(ousac.ref_singulF(G) - proj(G)) < 1e-9
    F = np.array([0.0307373, 0.245357, -0.170661, -0.252817, 0.0102013, 0.646064, 0.0969459, -0.647193, -0.00737131]).reshape(3, 3)
    assert np.linalg.svd(F)[1][2] < 2e-5                                                  # already rank 2 to 1e-5: the projection should change nothing
    assert np.linalg.norm(proj(F) - F) < 2e-5 and np.linalg.norm(ousac.ref_singulF(F) - F) > 0.2
    # and on a scene: an 8-point fit of ALL ground-truth inliers, pushed through the reference's singulF, loses most of them
    sc = datagen_pose.make_scene(800, 0.05, seed=31)
    inl = ousac.sampson(sc["E"], sc["p1"], sc["p2"]) < sc["th"] ** 2
    p1n, p2n, T1, T2 = ousac.normalize_points(sc["p1"], sc["p2"])
    x1, y1 = p1n[inl].T; x2, y2 = p2n[inl].T
    A = np.stack([x2 * x1, x2 * y1, x2, y2 * x1, y2 * y1, y2, x1, y1, np.ones(len(x1))], 1)
    f = np.linalg.eigh(A.T @ A)[1][:, 0].reshape(3, 3)
    good = (ousac.sampson(T2.T @ proj(f) @ T1, sc["p1"], sc["p2"]) < sc["th"] ** 2).sum()
    bad = (ousac.sampson(T2.T @ ousac.ref_singulF(f) @ T1, sc["p1"], sc["p2"]) < sc["th"] ** 2).sum()
    assert good >= 0.98 * inl.sum() and bad < 0.2 * inl.sum(), (good, bad, inl.sum())


def test_rand_stream_is_glibc():
    import __graft_entry__ as g
    g.build()
    from matchinglib_poselib_b200 import _lib
    lib = _lib.load()
    libc = ctypes.CDLL(None)
    for seed in (0, 1, 7, 12345, 2 ** 31 - 1, 2 ** 32 - 5):
        out = np.zeros(1500, np.int32)
        assert lib.mpl_usac_rand_stream(seed, len(out), _lib.ptr(out)) == 0
        libc.srand(ctypes.c_uint(seed))
        assert np.array_equal(out, np.array([libc.rand() for _ in range(len(out))], np.int32))
