"""GPU parity: CUDA RANSAC 5-point path through the C ABI against the oracle and the committed golden vectors
(generated from the reference's own coefficient table + OpenCV primitives).  Identical pre-drawn samples on both
sides.  Bars: inlier counts of a GIVEN model exact; solver models equal to the reference's per sample (nearest-E,
sign-free) to rounding for well-conditioned samples; final count / mask exact, E within 1e-6 Frobenius."""
import os

import numpy as np
import pytest

import datagen_pose
from oracle import ransac as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def m():
    import __graft_entry__ as g
    g.build()
    import matchinglib_poselib_b200 as mod
    mod._lib.check(mod._lib.load().mpl_init(0), "mpl_init")
    return mod


@pytest.fixture(scope="module")
def gold(golden_dir):
    return np.load(os.path.join(golden_dir, "ransac_golden.npz"))


def _edist(a, b):
    return min(np.linalg.norm(a - b), np.linalg.norm(a + b))


@pytest.mark.parametrize("name", ["small", "mid", "clean"])
def test_trace_vs_golden(m, gold, name):
    p1, p2, th, samples = gold[name + "_p1"], gold[name + "_p2"], float(gold[name + "_th"]), gold[name + "_samples"]
    res = m.ransac5pt(p1, p2, samples, th, max_iters=len(samples), trace=True)
    g_nm, g_models, g_counts, g_es = gold[name + "_n_models"], gold[name + "_models"], gold[name + "_counts"], gold[name + "_errsums"]
    assert np.array_equal(res["n_models"], g_nm), "real-root sets differ from cv::solvePoly's"
    o = 0
    dists = []
    for nm in g_nm:
        a, b = res["models"][o:o + nm], g_models[o:o + nm]
        d = np.minimum(np.linalg.norm(a[:, None] - b[None], axis=2), np.linalg.norm(a[:, None] + b[None], axis=2))
        j = d.argmin(1)
        assert sorted(j) == list(range(nm))
        dmin = d.min(1) if nm else np.zeros(0)
        dists.extend(dmin)
        tight = dmin < 1e-10
        assert np.array_equal(res["counts"][o:o + nm][tight], g_counts[o:o + nm][j][tight])
        np.testing.assert_allclose(res["errsums"][o:o + nm][tight], g_es[o:o + nm][j][tight], rtol=1e-6)
        o += nm
    dists = np.array(dists)
    assert np.median(dists) < 1e-12 and np.quantile(dists, 0.9) < 1e-9 and np.quantile(dists, 0.95) < 1e-7 and dists.max() < 1e-3


@pytest.mark.parametrize("name", ["small", "mid", "clean"])
@pytest.mark.parametrize("refine", [0, 1])
def test_end_to_end_vs_golden(m, gold, name, refine):
    p1, p2, th, samples = gold[name + "_p1"], gold[name + "_p2"], float(gold[name + "_th"]), gold[name + "_samples"]
    res = m.ransac5pt(p1, p2, samples, th, refine=bool(refine))
    tag = f"{name}_ref{refine}"
    assert res["ok"] and res["count"] == int(gold[tag + "_count"]) and res["iters"] == int(gold[tag + "_iters"])
    assert np.array_equal(res["mask"], gold[tag + "_mask"])
    assert _edist(res["E"], gold[tag + "_E"].reshape(3, 3)) < 1e-6


def test_scoring_exact_for_given_models(m):
    """Feed the ORACLE's models to the device scorer: counts must be bit-exact, error sums equal to 1e-12."""
    import ctypes
    import torch
    lib = m._lib.load()
    sc = datagen_pose.make_scene(5000, 0.3, seed=31)
    s = datagen_pose.draw_samples_fast(5000, 200, seed=32)
    ref = orc.ransac5pt(sc["p1"], sc["p2"], s, sc["th"], max_iters=200, trace=True)
    models = np.ascontiguousarray(ref["models"])
    M = len(models)
    dev = torch.device("cuda:0")
    d_p1 = torch.from_numpy(sc["p1"]).to(dev); d_p2 = torch.from_numpy(sc["p2"]).to(dev)
    d_corr = torch.empty(5000 * 4, dtype=torch.float64, device=dev)
    d_models = torch.from_numpy(models).to(dev)
    d_counts = torch.zeros(M, dtype=torch.int32, device=dev); d_es = torch.zeros(M, dtype=torch.float64, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    m._lib.check(lib.mpl_ransac_pack_corr_dev(d_p1.data_ptr(), d_p2.data_ptr(), 5000, d_corr.data_ptr(), st), "pack")
    m._lib.check(lib.mpl_ransac_score_dev(d_models.data_ptr(), M, d_corr.data_ptr(), 5000, ctypes.c_double(sc["th"]), d_counts.data_ptr(), d_es.data_ptr(), st), "score")
    torch.cuda.synchronize()
    assert np.array_equal(d_counts.cpu().numpy(), ref["counts"])
    np.testing.assert_allclose(d_es.cpu().numpy(), ref["errsums"], rtol=1e-12)


@pytest.mark.parametrize("n,h,outl,seed", [(6, 10, 0.0, 1), (50, 30, 0.2, 2), (1000, 300, 0.5, 3), (10000, 500, 0.3, 4)])
def test_vs_oracle(m, n, h, outl, seed):
    sc = datagen_pose.make_scene(n, outl, seed=seed)
    s = datagen_pose.draw_samples(sc["p1"], sc["p2"], h, seed=seed + 50) if n <= 1000 else datagen_pose.draw_samples_fast(n, h, seed + 50)
    for refine in (False, True):
        got = m.ransac5pt(sc["p1"], sc["p2"], s, sc["th"], refine=refine)
        exp = orc.ransac5pt(sc["p1"], sc["p2"], s, sc["th"], refine=refine)
        assert got["ok"] == exp["ok"] and got["count"] == exp["count"] and got["iters"] == exp["iters"], (n, h, refine)
        assert np.array_equal(got["mask"], exp["mask"])
        if exp["ok"]:
            assert _edist(got["E"], exp["E"]) < 1e-6
    # recovered pose close to ground truth on the clean-ish scenes
    if outl <= 0.3 and n >= 1000:
        assert _edist(got["E"], sc["E"]) < 5e-2


def test_estimate_essential_api(m):
    sc = datagen_pose.make_scene(800, 0.3, seed=77)
    m._lib.load().mpl_ransac_set_seed(1234)
    ok, E, mask = m.estimateEssentialMat(sc["p1"], sc["p2"], "RANSAC", sc["th"], True)
    assert ok and E.shape == (3, 3) and mask.shape == (1, 800)
    assert abs(np.linalg.norm(E) - 1) < 1e-12 and mask.sum() > 400
    assert _edist(E, sc["E"]) < 5e-2
    # the mask is exactly the inlier set of the returned E under the reference's test
    cnt, _, _, omask = orc.sampson(E, sc["p1"], sc["p2"], sc["th"])
    assert np.array_equal(omask, mask[0])
    # n == 5: solver only, stacked models, mask all ones (five-point.cpp:108-114)
    ok, Es, mask5 = m.estimateEssentialMat(sc["p1"][:5], sc["p2"][:5], "RANSAC", sc["th"], False)
    ref = orc.run5point(sc["p1"][:5], sc["p2"][:5])
    assert ok and Es.shape == (3 * len(ref), 3) and mask5.all()
    assert orc.match_models(Es.reshape(-1, 9), ref).max() < 1e-8
    with pytest.raises(m._lib.MplError):
        m.estimateEssentialMat(sc["p1"], sc["p2"], "USAC", sc["th"])
    with pytest.raises(AssertionError):
        m.estimateEssentialMat(sc["p1"][:4], sc["p2"][:4], "RANSAC", sc["th"])
