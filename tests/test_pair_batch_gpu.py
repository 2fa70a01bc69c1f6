"""GPU parity of the pair-batch pipeline (mpl_match_pose_batch: device-resident match -> correspondences -> RANSAC with the
selection replayed and the refit done on the device) against (a) the oracle chain on the same device-drawn samples and
(b) the library's own host-replay path (mpl_ransac5pt), which states the same selection rule independently.
Bars: match lists bit-exact; inlier count, iteration count and mask exact; E within 1e-6 Frobenius (sign-free) for the
minimal-sample result, refit E within the conditioning floor measured in test_oracle_ransac.py."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(__file__))
import datagen
import datagen_pose
from oracle import knn as oknn, pose as opose, ransac as oransac

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def P():
    import __graft_entry__ as g
    g.build()
    from matchinglib_poselib_b200 import poselib, _lib
    _lib.check(_lib.load().mpl_init(0), "mpl_init")
    return poselib


def _edist(a, b):
    return min(np.linalg.norm(a - b), np.linalg.norm(a + b))


def make_pair(n1, n2, nb, seed, outliers=0.0, dtype="u8"):
    """Descriptors with planted neighbours (query i <-> train i for the first min(n1, n2) / 2 rows) over a synthetic scene."""
    n = max(n1, n2)
    sc = datagen_pose.make_scene(n, outliers, seed=seed)
    rng = np.random.default_rng(seed)
    if dtype == "u8":
        t = rng.integers(0, 256, (n2, nb), dtype=np.uint8)
        q = rng.integers(0, 256, (n1, nb), dtype=np.uint8)
        k = min(n1, n2)
        good = np.zeros(n1, bool); good[:k] = rng.random(k) < 0.5
        flips = np.packbits(rng.random((n1, nb * 8)) < 0.08, axis=1)[:, :nb]
        idx = np.nonzero(good)[0]
        q[idx] = t[idx] ^ flips[idx]
    else:
        t = np.abs(rng.standard_normal((n2, nb))); t = np.minimum(np.round(t / np.linalg.norm(t, axis=1, keepdims=True) * 512), 255).astype(np.float32)
        q = np.abs(rng.standard_normal((n1, nb))); q = np.minimum(np.round(q / np.linalg.norm(q, axis=1, keepdims=True) * 512), 255).astype(np.float32)
        k = min(n1, n2)
        idx = np.nonzero(rng.random(k) < 0.5)[0]
        q[idx] = np.clip(t[idx] + np.round(rng.standard_normal((len(idx), nb)) * 2), 0, 255).astype(np.float32)
    return q, t, sc["p1"][:n1].copy(), sc["p2"][:n2].copy(), sc["th"]


def oracle_matches(q, t, matcher, ratio=True):
    """k-NN + ratio test of the oracle without getMatches' keypoint-count gate (the pipeline has none)."""
    mode = 0 if matcher == "LINEAR" else 1
    if len(q) == 0 or len(t) == 0:
        return np.zeros(0, oknn.DMATCH_DTYPE)
    idx, dist = (oknn.knn_hamming if q.dtype == np.uint8 else oknn.knn_l2)(q, t, 2, mode, threads=8)
    return oknn.ratio_matches(idx, dist, ratio)


def check_pair_against_oracle(r, q, t, p1, p2, th, matcher, h, refine, tol_E):
    omt = oracle_matches(q, t, matcher)
    assert np.array_equal(r["matches"], omt)
    c1 = np.ascontiguousarray(p1[omt["queryIdx"]]); c2 = np.ascontiguousarray(p2[omt["trainIdx"]])
    if len(omt) < 5:
        assert r["count"] == 0
        return None
    s = r["samples"]
    assert s.shape == (h, 5) and s.min() >= 0 and s.max() < len(omt) and all(len(set(row)) == 5 for row in s[:300].tolist())
    exp = oransac.ransac5pt(c1, c2, s, th, max_iters=h, refine=refine)
    assert r["count"] == exp["count"] and r["iters"] == exp["iters"], (r["count"], exp["count"], r["iters"], exp["iters"])
    assert np.array_equal(r["mask"], exp["mask"])
    if exp["ok"]:
        d = _edist(r["E"], exp["E"])
        assert d < tol_E, d
        return d
    return None


@pytest.mark.parametrize("matcher", ["LINEAR", "BRUTEFORCENMS"])
def test_batch_equals_oracle_and_host_replay(P, matcher):
    """12 pairs of different shapes in ONE call (two pipeline lanes are exercised through the forced sub-batch split)."""
    shapes = [(900, 1100, 61), (1100, 900, 61), (700, 700, 61), (40, 2000, 61), (2000, 40, 61), (16, 16, 61), (3, 500, 61),
              (1500, 1500, 61), (1000, 1200, 61), (1200, 1000, 61), (800, 801, 61), (640, 1024, 61), (512, 512, 61), (1300, 700, 61),
              (300, 300, 61), (1024, 2048, 61)]
    pairs, ths = [], []
    for i, (n1, n2, nb) in enumerate(shapes):
        q, t, p1, p2, th = make_pair(n1, n2, nb, seed=100 + i, outliers=0.2 if i % 3 == 0 else 0.0)
        pairs.append((q, t, p1, p2)); ths.append(th)
    h = 200
    res = P.matchAndEstimatePoseBatch(pairs, matcher, True, ths[0], h, 0.999, True, seeds=[50 + i for i in range(len(pairs))], want_samples=True)
    assert len(res) == len(pairs)
    for i, (r, (q, t, p1, p2)) in enumerate(zip(res, pairs)):
        check_pair_against_oracle(r, q, t, p1, p2, ths[0], matcher, h, True, 1e-6)
        # the single-pair entry point is the same pipeline as a batch of one: identical results
        one = P.matchAndEstimatePose(q, t, p1, p2, matcher, True, ths[0], h, 0.999, True, seed=50 + i)
        assert np.array_equal(one["matches"], r["matches"]) and np.array_equal(one["mask"], r["mask"]) and one["count"] == r["count"] and one["iters"] == r["iters"]
        assert np.array_equal(one["E"], r["E"])
        # host replay (mpl_ransac5pt) over the same correspondences and samples: the independent statement of modelest.cpp:394-415
        if r["n_corr"] >= 5:
            c1 = np.ascontiguousarray(p1[r["matches"]["queryIdx"]]); c2 = np.ascontiguousarray(p2[r["matches"]["trainIdx"]])
            hr = P.ransac5pt(c1, c2, r["samples"], ths[0], max_iters=h, refine=True)
            assert hr["count"] == r["count"] and hr["iters"] == r["iters"] and np.array_equal(hr["mask"], r["mask"])
            if hr["ok"]:
                assert np.array_equal(hr["E"], r["E"])               # same kernels from the selection on: bit-identical


@pytest.mark.parametrize("nb", [16, 32, 61])
def test_batch_hamming_paths_identical(P, nb):
    """The batch's Hamming scan on the tensor cores (one ranking launch per group of pairs; bits as bf16) and the popc batch kernel
    give the same matches, hence the same everything; pairs of different sizes, strides that are no multiple of the tiles."""
    from matchinglib_poselib_b200 import _lib
    shapes = [(1500, 1700), (1700, 1500), (130, 1650), (1650, 129), (5, 900), (1111, 1333), (257, 255), (1700, 1700), (0, 100), (640, 2)]
    pairs, th = [], None
    for i, (n1, n2) in enumerate(shapes):
        q, t, p1, p2, th = make_pair(n1, n2, nb, seed=300 + i)
        if n2 > 300:
            t[n2 - 1] = t[7]; t[n2 // 2] = t[7]                             # duplicates: ties must resolve to the lowest train index
        pairs.append((q, t, p1, p2))
    lib = _lib.load()
    out = {}
    for path in (1, 2, 3):
        lib.mpl_set_hamming_path(path)
        try:
            out[path] = P.matchAndEstimatePoseBatch(pairs * 3, "LINEAR", True, th, 64, 0.999, True, seeds=list(range(len(pairs) * 3)))
        finally:
            lib.mpl_set_hamming_path(0)
    for tc in (2, 3):
        for a, b in zip(out[1], out[tc]):
            assert np.array_equal(a["matches"], b["matches"]) and np.array_equal(a["mask"], b["mask"]) and a["count"] == b["count"]
            assert np.array_equal(a["E"], b["E"])
    for r, (q, t, p1, p2) in zip(out[2][:4], pairs[:4]):
        assert np.array_equal(r["matches"], oracle_matches(q, t, "LINEAR"))


def test_pinned_host_buffers_upload_through_the_gather_kernel(P):
    """MPL_PAIR_ZEROCOPY=1 (experiment, off by default): buffers in mpl_host_alloc memory (matchinglib.pinned_array) are read by one gather
    kernel per sub-batch instead of four copies per pair; plain numpy buffers take the copies.  Same results; odd sizes exercise the
    16-byte body / byte tail of the gather."""
    import os
    from matchinglib_poselib_b200 import matchinglib
    shapes = [(700, 900, 61), (901, 333, 61), (5, 1200, 61), (0, 40, 61), (1201, 1201, 61), (64, 64, 61)]
    plain, pinned, th = [], [], None
    for i, (n1, n2, nb) in enumerate(shapes):
        q, t, p1, p2, th = make_pair(n1, n2, nb, seed=700 + i)
        plain.append((q, t, p1, p2))
        pin = []
        for a in (q, t, p1, p2):
            v = matchinglib.pinned_array(a.shape, a.dtype) if a.size else a
            if a.size:
                v[...] = a
            pin.append(v)
        pinned.append(tuple(pin))
    seeds = list(range(len(shapes)))
    ra = P.matchAndEstimatePoseBatch(plain, "LINEAR", True, th, 96, 0.999, True, seeds=seeds)
    os.environ["MPL_PAIR_ZEROCOPY"] = "1"
    try:
        rb = P.matchAndEstimatePoseBatch(pinned, "LINEAR", True, th, 96, 0.999, True, seeds=seeds)
    finally:
        del os.environ["MPL_PAIR_ZEROCOPY"]
    for a, b in zip(ra, rb):
        assert np.array_equal(a["matches"], b["matches"]) and np.array_equal(a["mask"], b["mask"]) and a["count"] == b["count"] and np.array_equal(a["E"], b["E"])
    assert np.array_equal(rb[0]["matches"], oracle_matches(plain[0][0], plain[0][1], "LINEAR"))


def test_c4_unit_vs_oracle(P):
    """The unit of BASELINE config C4 at its own size: AKAZE-like 61-byte descriptors 5000 x 5000 + RANSAC 1000 samples + refit."""
    q, t, p1, p2, th = make_pair(5000, 5000, 61, seed=1004)
    r = P.matchAndEstimatePoseBatch([(q, t, p1, p2)], "LINEAR", True, th, 1000, 0.999, True, seeds=[5], want_samples=True)[0]
    d = check_pair_against_oracle(r, q, t, p1, p2, th, "LINEAR", 1000, True, 1e-6)
    assert r["count"] > 2000 and r["iters"] < 100
    print("C4 unit: matches %d inliers %d iterations %d  |E_gpu - E_oracle| = %.3g" % (len(r["matches"]), r["count"], r["iters"], d))


def test_hard_scene_runs_all_chunks(P):
    """70 % outliers: the adaptive stop never shrinks far, every solve / score chunk of the 700-sample budget is used."""
    q, t, p1, p2, th = make_pair(1500, 1500, 32, seed=77, outliers=0.7)
    for refine in (False, True):
        r = P.matchAndEstimatePoseBatch([(q, t, p1, p2)] * 3, "LINEAR", True, th, 700, 0.999, refine, seeds=[1, 2, 3], want_samples=True)
        for x in r:
            check_pair_against_oracle(x, q, t, p1, p2, th, "LINEAR", 700, refine, 1e-6)
        assert max(x["iters"] for x in r) > 300


def test_pixel_keypoints_and_undistortion(P):
    """pts_mode 1: float pixel keypoints, ImgToCamCoordTrans + Remove_LensDist on the device (pose_helper.cpp:1100-1109, 1169-1279)
    against the numpy restatement, then RANSAC on exactly those correspondences."""
    K0 = np.array([[721.5, 0, 609.6], [0, 721.5, 172.9], [0, 0, 1.0]]); K1 = np.array([[718.3, 0, 607.2], [0, 719.0, 185.2], [0, 0, 1.0]])
    for dist0, dist1 in ((np.zeros(8), np.zeros(8)), (np.array([-0.37, 0.2, 1e-3, -2e-3, -0.07, 0, 0, 0]), np.array([-0.36, 0.18, -1e-3, 1e-3, -0.05, 0, 0, 0]))):
        q, t, p1, p2, th = make_pair(1400, 1400, 32, seed=11)
        # pixel coordinates: distort the ideal points with the forward model so that undistortion has something to remove
        def to_pix(p, K, d):
            x, y = p[:, 0], p[:, 1]
            r2 = x * x + y * y
            rad = (1 + ((d[4] * r2 + d[1]) * r2 + d[0]) * r2) / (1 + ((d[7] * r2 + d[6]) * r2 + d[5]) * r2)
            xd = x * rad + d[2] * 2 * x * y + d[3] * (r2 + 2 * x * x); yd = y * rad + d[2] * (r2 + 2 * y * y) + d[3] * 2 * x * y
            return np.stack([xd * K[0, 0] + K[0, 2], yd * K[1, 1] + K[1, 2]], 1).astype(np.float32)
        px1, px2 = to_pix(p1, K0, dist0), to_pix(p2, K1, dist1)
        omt = oracle_matches(q, t, "LINEAR")
        px1[omt["queryIdx"][3]] = (1e5, 1e5)                                # a wild keypoint of a matched query: its undistortion fails the proof test
        thc = 1.6 * P.pixToCamFact(K0, K1)
        r = P.matchAndEstimatePoseBatch([(q, t, px1, px2)], "LINEAR", True, thc, 300, 0.999, True, seeds=[3], K0=K0, K1=K1, dist0=dist0, dist1=dist1,
                                        want_samples=True)[0]
        omt = oracle_matches(q, t, "LINEAR")
        assert np.array_equal(r["matches"], omt)
        c1 = opose.img_to_cam(px1[omt["queryIdx"]], K0); c2 = opose.img_to_cam(px2[omt["trainIdx"]], K1)
        ok, u1, u2, keep = opose.remove_lens_dist(c1, c2, dist0, dist1)
        assert ok and r["status"] == 0 and r["n_corr"] == int(keep.sum())
        if dist0.any():
            assert not keep.all()
        exp = oransac.ransac5pt(u1.astype(np.float64), u2.astype(np.float64), r["samples"], thc, max_iters=300, refine=True)
        assert r["count"] == exp["count"] and r["iters"] == exp["iters"]
        full = np.zeros(len(omt), np.uint8); full[np.nonzero(keep)[0]] = exp["mask"]
        assert np.array_equal(r["mask"], full)
        assert _edist(r["E"], exp["E"]) < 1e-6


def test_undistortion_failure_status(P):
    q, t, p1, p2, th = make_pair(300, 300, 32, seed=5)
    K = np.array([[700.0, 0, 600], [0, 700, 200], [0, 0, 1]])
    px = (np.random.default_rng(0).uniform(0, 1, (300, 2)) * 1e6).astype(np.float32)   # nothing survives the proof test
    r = P.matchAndEstimatePoseBatch([(q, t, px, px)], "LINEAR", True, 1e-3, 50, 0.999, True, K0=K, K1=K, dist0=np.array([5.0, 3, 0, 0, 1, 0, 0, 0]),
                                    dist1=np.array([5.0, 3, 0, 0, 1, 0, 0, 0]))[0]
    assert r["status"] == 1 and r["count"] == 0 and r["n_corr"] == 0 and len(r["matches"]) > 50 and not r["mask"].any()


@pytest.mark.parametrize("matcher", ["LINEAR", "BRUTEFORCENMS"])
def test_float_descriptor_pairs(P, matcher):
    """CV_32F descriptors through the same pipeline (tensor-core L2 path per pair: 2100 x 2100 >= 2^22 pairs)."""
    pairs, th = [], None
    for i, (n1, n2) in enumerate([(2100, 2100), (300, 500), (2200, 2050)]):
        q, t, p1, p2, th = make_pair(n1, n2, 128, seed=300 + i, dtype="f32")
        pairs.append((q, t, p1, p2))
    res = P.matchAndEstimatePoseBatch(pairs, matcher, True, th, 150, 0.999, True, seeds=[7, 8, 9], want_samples=True)
    for r, (q, t, p1, p2) in zip(res, pairs):
        omt = oracle_matches(q, t, matcher)
        assert np.array_equal(r["matches"]["queryIdx"], omt["queryIdx"]) and np.array_equal(r["matches"]["trainIdx"], omt["trainIdx"])
        if matcher == "LINEAR":
            assert np.array_equal(r["matches"]["distance"].view(np.int32), omt["distance"].view(np.int32))
        else:
            np.testing.assert_allclose(r["matches"]["distance"], omt["distance"], rtol=1e-5)
        c1 = np.ascontiguousarray(p1[omt["queryIdx"]]); c2 = np.ascontiguousarray(p2[omt["trainIdx"]])
        exp = oransac.ransac5pt(c1, c2, r["samples"], th, max_iters=150, refine=True)
        assert r["count"] == exp["count"] and r["iters"] == exp["iters"] and np.array_equal(r["mask"], exp["mask"])


def test_large_batch_two_lanes(P):
    """600 pairs (3 sub-batches over 2 lanes): every pair's result equals the same pair run alone."""
    base = [make_pair(600, 600, 61, seed=500 + i) for i in range(6)]
    pairs = [base[i % 6][:4] for i in range(600)]
    th = base[0][4]
    res = P.matchAndEstimatePoseBatch(pairs, "LINEAR", True, th, 128, 0.999, True, seeds=[i % 12 for i in range(600)])
    ref = {}
    for i in range(12):
        ref[(i % 6, i % 12)] = P.matchAndEstimatePose(*base[i % 6][:4], "LINEAR", True, th, 128, 0.999, True, seed=i % 12)
    for i, r in enumerate(res):
        e = ref[(i % 6, i % 12)]
        assert np.array_equal(r["matches"], e["matches"]) and np.array_equal(r["mask"], e["mask"]) and r["count"] == e["count"] and np.array_equal(r["E"], e["E"])


def test_empty_and_tiny(P):
    q, t, p1, p2, th = make_pair(64, 64, 32, seed=1)
    assert P.matchAndEstimatePoseBatch([], "LINEAR") == []
    r = P.matchAndEstimatePoseBatch([(q[:0], t, p1[:0], p2), (q, t[:0], p1, p2[:0]), (q[:4], t, p1[:4], p2)], "LINEAR", False, th, 20)
    assert len(r[0]["matches"]) == 0 and len(r[1]["matches"]) == 0 and len(r[2]["matches"]) == 4 and all(x["count"] == 0 for x in r)
    r = P.matchAndEstimatePoseBatch([(q, t, p1, p2)], "LINEAR", True, th, 0)                 # h = 0: matching only
    assert r[0]["count"] == 0 and len(r[0]["matches"]) > 0
    from matchinglib_poselib_b200 import _lib
    with pytest.raises(_lib.MplError):
        P.matchAndEstimatePoseBatch([(q, t, p1, p2)], "HNSW", True, th, 20)
