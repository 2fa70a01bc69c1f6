"""CPU: the k-NN oracle (oracle/knn_oracle.c) against golden vectors produced by the real reference
arithmetic (cv2's cvflann linear index; the reference's own NMSLIB build) -- oracle/gen_golden_knn.py."""
import os

import numpy as np
import pytest

from oracle import knn as orc

BIN = ["orb32", "akaze61", "brisk64", "odd7"]
FLT = ["sift128", "surf64", "odd61"]


@pytest.fixture(scope="module")
def gold(golden_dir):
    return np.load(os.path.join(golden_dir, "knn_golden.npz"))


@pytest.mark.parametrize("name", BIN)
def test_hamming_linear_bit_exact(gold, name):
    idx, dist = orc.knn_hamming(gold[name + "_q"], gold[name + "_t"], 2, mode=0)
    assert np.array_equal(idx, gold[name + "_lin_idx"])
    assert np.array_equal(dist, gold[name + "_lin_dist"])


@pytest.mark.parametrize("name", BIN)
def test_hamming_nmslib_quirk(gold, name):
    """Distances equal the reference NMSLIB build exactly (last packed word dropped); the
    best index is compared where d0 < d1 (the heap's order inside a tie depends on object addresses)."""
    idx, dist = orc.knn_hamming(gold[name + "_q"], gold[name + "_t"], 2, mode=1)
    assert np.array_equal(dist, gold[name + "_nms_dist"])
    strict = dist[:, 0] < dist[:, 1]
    assert np.array_equal(idx[strict, 0], gold[name + "_nms_idx"][strict, 0])
    lin_idx, lin_dist = orc.knn_hamming(gold[name + "_q"], gold[name + "_t"], 2, mode=0)
    assert (lin_dist != dist).any(), "fixture must exercise the dropped tail word"


@pytest.mark.parametrize("name", FLT)
def test_l2_linear_bit_exact(gold, name):
    idx, dist = orc.knn_l2(gold[name + "_q"], gold[name + "_t"], 2, mode=0)
    assert np.array_equal(idx, gold[name + "_lin_idx"])
    assert np.array_equal(dist.view(np.int32), gold[name + "_lin_dist"].view(np.int32))


@pytest.mark.parametrize("name", FLT)
def test_l2_nmslib_tolerance(gold, name):
    """The reference builds NMSLIB with -Ofast: parity is by tolerance (1e-5 relative), indices equal
    except near-ties."""
    idx, dist = orc.knn_l2(gold[name + "_q"], gold[name + "_t"], 2, mode=1)
    g_idx, g_dist = gold[name + "_nms_idx"], gold[name + "_nms_dist"]
    np.testing.assert_allclose(dist, g_dist, rtol=1e-5, atol=1e-6)
    diff = idx != g_idx
    if diff.any():
        rel = np.abs(dist - g_dist)[diff] / np.maximum(g_dist[diff], 1e-12)
        assert rel.max() < 1e-5


def test_ratio_and_codes():
    rng = np.random.default_rng(5)
    d1 = rng.integers(0, 256, (40, 32), dtype=np.uint8)
    d2 = rng.integers(0, 256, (50, 32), dtype=np.uint8)
    assert orc.get_matches(d1[:10], d2)[0] == -4
    assert orc.get_matches(d1, d2, n_kp1=39)[0] == -1
    assert orc.get_matches(d1, d2, "HNSW")[0] == -2
    assert orc.get_matches(d1, d2, "LINEAR")[0] == -3          # random rows never pass the ratio test
    d1[:20] = d2[:20]
    err, m = orc.get_matches(d1, d2, "LINEAR")
    assert err == 0 and len(m) >= 20 and np.all(np.diff(m["queryIdx"]) > 0)
    err, m = orc.get_matches(d1, d2, "LINEAR", ratio_test=False)
    assert err == 0 and len(m) == 40


def test_empty_and_single():
    idx, dist = orc.knn_hamming(np.zeros((0, 32), np.uint8), np.zeros((5, 32), np.uint8))
    assert idx.shape == (0, 2)
    idx, dist = orc.knn_hamming(np.zeros((3, 32), np.uint8), np.zeros((1, 32), np.uint8))
    assert np.array_equal(idx, [[0, -1]] * 3) and np.all(dist[:, 1] == np.iinfo(np.int32).max)
