"""GPU: estimateEssentialMatUsac / poselib::estimateEssentialOrPoseUSAC (mpl_usac5pt: device minimal solves + model tests +
inlier bit matrix + one-launch local optimisation, the reference's USAC schedule on the host).  The checker is the reference's OWN
estimator compiled into oracle/_ref/libusac_ref.so (USAC.h + EssentialMatEstimator.h + opengv, oracle/ref_usac), seeded like the
GPU path (both draw from glibc's rand()).

Parity is statistical for this randomised, adaptive algorithm -- the 5-point solvers differ in rounding and in the order of a
sample's models, and the evaluation-pool cursor drifts apart after a local optimisation -- but with the same seed most runs are
IDENTICAL in (support, hypotheses) -- often also in (models, rejected models, local optimisations) -- which pins sampler, pool,
SPRT, PROSAC and stopping rules.  The reference's local optimisation works in some scenes and is a no-op in others (a defect of its 3 x 3 SVD
port, include/mplb200.h lo_mode, pinned in tests/test_oracle_usac.py), so every scene is run in both lo_modes and must match the
reference in one of them.  The bars: identical runs in at least half of the seeds, medians of support and hypotheses within
10 % / a factor 1.5, and -- exact -- the returned mask is the inlier set of the returned model under evaluateModel's own error.
lo_mode 0 (the default: local optimisation as stated) must also reach the ground truth's support."""
import numpy as np
import pytest

import datagen_pose
from oracle import usac as ousac

pytestmark = pytest.mark.gpu

KEYS = ("count", "hyp_count", "model_count", "rejected_sample_count", "rejected_model_count", "num_local_optimizations")


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


def _order(sc, seed, mix=0.6):
    """A PROSAC ordering like match distances give: inliers of the ground truth mostly first."""
    rng = np.random.default_rng(seed)
    out = ousac.sampson(sc["E"], sc["p1"], sc["p2"]) >= sc["th"] ** 2
    return np.argsort(out + mix * rng.random(len(out)), kind="stable").astype(np.uint32)


SCENES = [(1000, 0.3, 1, False), (1000, 0.3, 1, True), (3000, 0.5, 2, False), (3000, 0.5, 2, True), (300, 0.2, 3, False), (300, 0.2, 3, True),
          (10000, 0.3, 4, True), (150, 0.6, 5, False), (2500, 0.05, 6, True), (60, 0.1, 7, True), (5000, 0.7, 8, True), (800, 0.05, 31, False),
          (2000, 0.55, 21, False), (2000, 0.55, 21, True)]


@pytest.mark.parametrize("n,outl,seed,prosac", SCENES)
def test_usac_vs_reference_per_seed(P, n, outl, seed, prosac):
    sc = datagen_pose.make_scene(n, outl, seed=seed)
    truth = int((ousac.sampson(sc["E"], sc["p1"], sc["p2"]) < sc["th"] ** 2).sum())
    idx = _order(sc, seed) if prosac else None
    gc, gh, same, full = {0: [], 1: []}, {0: [], 1: []}, {0: 0, 1: 0}, {0: 0, 1: 0}
    oc, oh = [], []
    for s in range(1, 11):
        o = ousac.usac(sc["p1"], sc["p2"], sc["th"], sorted_idx=idx, seed=s)
        if o is not None:
            assert o["ok"]
            oc.append(o["count"]); oh.append(o["hyp_count"])
        for mode in (0, 1):
            g = P.usac5pt(sc["p1"], sc["p2"], sc["th"], sorted_idx=idx, seed=s, lo_mode=mode)
            assert g["ok"] and g["count"] >= 0.5 * truth, (g["count"], truth)
            # exact: mask == {evaluateModel error of the returned model < th^2}, count == its size (storeSolution, USAC.h:1208-1228)
            omask = ousac.sampson(g["E"], sc["p1"], sc["p2"]) < sc["th"] ** 2
            assert np.array_equal(omask, g["mask"].astype(bool)) and int(omask.sum()) == g["count"]
            assert _ed(g["E"], sc["E"]) < 0.15
            gc[mode].append(g["count"]); gh[mode].append(g["hyp_count"])
            if o is not None:
                same[mode] += int(o["count"] == g["count"] and o["hyp_count"] == g["hyp_count"])
                full[mode] += int(all(o[k] == g[k] for k in KEYS))
    m = 0 if same[0] >= same[1] else 1
    print("n=%d outliers %.0f%% %s, ground truth %d: reference inliers %s hypotheses %s | lo_mode 0 inliers %s hypotheses %s, same (support, hypotheses) "
          "in %d/10 runs, all six statistics in %d/10 | lo_mode 1 inliers %s hypotheses %s, same in %d/10, all six in %d/10 -> the reference behaves like lo_mode %d here" %
          (n, 100 * outl, "PROSAC" if prosac else "uniform", truth, oc, oh, gc[0], gh[0], same[0], full[0], gc[1], gh[1], same[1], full[1], m))
    if oc:
        assert same[m] >= 5, same
        assert abs(np.median(gc[m]) - np.median(oc)) <= 0.1 * truth, (gc[m], oc, truth)
        assert 1 / 1.5 <= (np.median(gh[m]) + 1) / (np.median(oh) + 1) <= 1.5, (gh[m], oh)


@pytest.mark.parametrize("n,outl,seed,prosac", SCENES[:8] + SCENES[9:])
def test_usac_local_optimisation_as_stated_reaches_the_ground_truth(P, n, outl, seed, prosac):
    """lo_mode 0: the device local optimisation (8-point refits with Torr weights, USAC.h:948-1060) lifts the support to the ground
    truth model's and never below what the reference reaches; it needs no more hypotheses than lo_mode 1."""
    sc = datagen_pose.make_scene(n, outl, seed=seed)
    truth = int((ousac.sampson(sc["E"], sc["p1"], sc["p2"]) < sc["th"] ** 2).sum())
    idx = _order(sc, seed) if prosac else None
    c0, c1, h0, h1 = [], [], [], []
    for s in range(1, 7):
        g = P.usac5pt(sc["p1"], sc["p2"], sc["th"], sorted_idx=idx, seed=s, lo_mode=0)
        r = P.usac5pt(sc["p1"], sc["p2"], sc["th"], sorted_idx=idx, seed=s, lo_mode=1)
        omask = ousac.sampson(g["E"], sc["p1"], sc["p2"]) < sc["th"] ** 2
        assert g["ok"] and np.array_equal(omask, g["mask"].astype(bool)) and int(omask.sum()) == g["count"]
        c0.append(g["count"]); c1.append(r["count"]); h0.append(g["hyp_count"]); h1.append(r["hyp_count"])
        assert _ed(g["E"], sc["E"]) < 5e-2
    print("n=%d outliers %.0f%% %s, ground truth %d: lo_mode 0 inliers %s hypotheses %s | lo_mode 1 inliers %s hypotheses %s" %
          (n, 100 * outl, "PROSAC" if prosac else "uniform", truth, c0, h0, c1, h1))
    assert np.median(c0) >= 0.97 * truth and np.median(c0) >= np.median(c1)
    assert np.median(h0) <= np.median(h1)


def test_usac_too_few_points_and_noise(P):
    sc = datagen_pose.make_scene(100, 0.2, seed=41)
    assert not P.usac5pt(sc["p1"][:4], sc["p2"][:4], sc["th"])["ok"]                        # USAC.h:351-356
    assert not P.usac5pt(sc["p1"][:15], sc["p2"][:15], sc["th"], sorted_idx=np.arange(15))["ok"]       # PROSAC: n < minStopLen
    assert P.usac5pt(sc["p1"][:15], sc["p2"][:15], sc["th"])["ok"]
    rng = np.random.default_rng(3)
    p1 = rng.uniform(-0.5, 0.5, (400, 2)); p2 = rng.uniform(-0.5, 0.5, (400, 2))
    g = P.usac5pt(p1, p2, sc["th"], seed=5, max_hypotheses=3000)
    assert g["ok"] and g["count"] < 40 and g["hyp_count"] == 3000                          # no structure: the budget is spent
    o = ousac.usac(p1, p2, sc["th"], seed=5)
    if o is not None:
        assert o["count"] < 40


def test_usac_deterministic_and_parameter_checks(P):
    from matchinglib_poselib_b200 import _lib
    sc = datagen_pose.make_scene(1200, 0.4, seed=51)
    for mode in (1, 0):
        a = P.usac5pt(sc["p1"], sc["p2"], sc["th"], seed=9, lo_mode=mode)
        b = P.usac5pt(sc["p1"], sc["p2"], sc["th"], seed=9, lo_mode=mode)
        assert a["hyp_count"] == b["hyp_count"] and a["count"] == b["count"] and np.array_equal(a["mask"], b["mask"]) and np.array_equal(a["E"], b["E"])
    s = P.usac5pt(sc["p1"], sc["p2"], sc["th"], seed=9, estimator="STEWENIUS", lo_mode=0)
    assert s["ok"] and abs(s["count"] - a["count"]) <= 0.05 * a["count"]
    with pytest.raises(_lib.MplError):
        P.usac5pt(sc["p1"], sc["p2"], sc["th"], refine=1)                                  # REFINE_8PT_PSEUDOHUBER: not built, no fallback
    with pytest.raises(_lib.MplError):
        P.usac5pt(sc["p1"], sc["p2"], sc["th"], degeneracy=2)
    with pytest.raises(_lib.MplError):
        P.usac5pt(sc["p1"], sc["p2"], sc["th"], estimator="EIG_KNEIP")


def test_estimateEssentialOrPoseUSAC_mirror(P):
    """The host mirror of poselib::estimateEssentialOrPoseUSAC: SPRT / PROSAC start values from matches and keypoints
    (pose_estim.cpp:1810-1981), PROSAC order from the match distances, the served configurations only."""
    from matchinglib_poselib_b200 import _lib, matchinglib
    sc = datagen_pose.make_scene(2000, 0.4, seed=61)
    n = 2000
    inl = ousac.sampson(sc["E"], sc["p1"], sc["p2"]) < sc["th"] ** 2
    rng = np.random.default_rng(61)
    m = np.zeros(n, matchinglib.DMATCH_DTYPE)
    m["queryIdx"] = np.arange(n); m["trainIdx"] = np.arange(n); m["imgIdx"] = -1
    m["distance"] = np.where(inl, rng.uniform(10, 60, n), rng.uniform(30, 90, n)).astype(np.float32)     # inliers tend to match better
    kp1 = sc["p1"] * 800.0 + np.array([696.0, 256.0]); kp2 = sc["p2"] * 800.0 + np.array([696.0, 256.0])
    cfg = P.ConfigUSAC(estimator=P.POSE_NISTER, refinealg=P.REF_WEIGHTS, degeneracyCheck=P.DEGEN_NO_CHECK, matches=m, keypoints1=kp1,
                       keypoints2=kp2, th_pixels=1.6, imgSize=(1392, 512), nrMatchesVfcFiltered=int(0.55 * n), seed=4)
    d = P.estimateSprtDeltaInit(m, kp1, kp2, 1.6, (1392, 512))
    assert 0.001 <= d <= 1.0 / 6 + 1e-12 and P.estimateSprtEpsilonInit(m, cfg.nrMatchesVfcFiltered) == 0.4
    order = P.getSortedMatchIdx(m)
    assert np.all(np.diff(m["distance"][order]) >= 0)
    code, E, degenerate, inliers, info = P.estimateEssentialOrPoseUSAC(sc["p1"], sc["p2"], sc["th"], cfg)
    assert code == 0 and not degenerate and inliers.shape == (1, n) and inliers.sum() == info["count"] >= 0.95 * inl.sum()
    assert _ed(E, sc["E"]) < 5e-2 and abs(info["sprt_delta"]) > 0
    o = ousac.usac(sc["p1"], sc["p2"], sc["th"], sorted_idx=order, seed=4, prosac_beta=d, sprt_delta=d, sprt_epsilon=0.4)
    if o is not None:
        g1 = P.usac5pt(sc["p1"], sc["p2"], sc["th"], sorted_idx=order, seed=4, prosac_beta=d, sprt_delta=d, sprt_epsilon=0.4, lo_mode=1)
        assert (o["count"], o["hyp_count"]) in ((info["count"], info["hyp_count"]), (g1["count"], g1["hyp_count"])) or \
            abs(o["count"] - info["count"]) <= 0.15 * inl.sum()
    cfg.matches = None; cfg.automaticSprtInit = P.SPRT_DEFAULT_INIT                        # no matches: uniform sampling, default start values
    code, E, _, inliers, info = P.estimateEssentialOrPoseUSAC(sc["p1"], sc["p2"], sc["th"], cfg)
    assert code == 0 and inliers.sum() >= 0.95 * inl.sum()
    for bad in (dict(estimator=P.POSE_EIG_KNEIP), dict(refinealg=P.REF_8PT_PSEUDOHUBER), dict(degeneracyCheck=P.DEGEN_USAC_INTERNAL)):
        c = P.ConfigUSAC(estimator=P.POSE_NISTER, refinealg=P.REF_WEIGHTS, degeneracyCheck=P.DEGEN_NO_CHECK, automaticSprtInit=0)
        for k, v in bad.items():
            setattr(c, k, v)
        with pytest.raises(_lib.MplError):
            P.estimateEssentialOrPoseUSAC(sc["p1"], sc["p2"], sc["th"], c)
    assert P.estimateEssentialOrPoseUSAC(sc["p1"][:4], sc["p2"][:4], sc["th"], cfg)[0] == -2      # "USAC failed!"
