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
