"""GPU parity: CUDA k-NN through the C ABI against the CPU oracle and the committed golden vectors.
Bar: Hamming bit-exact (idx + dist); float LINEAR distances bit-exact, indices equal (ties resolved to the
lowest train index on both sides); BRUTEFORCENMS float within 1e-5 relative."""
import os

import numpy as np
import pytest

import datagen
from oracle import knn as orc

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
    return np.load(os.path.join(golden_dir, "knn_golden.npz"))


@pytest.mark.parametrize("name", ["orb32", "akaze61", "brisk64", "odd7"])
def test_hamming_golden(m, gold, name):
    q, t = gold[name + "_q"], gold[name + "_t"]
    idx, dist = m.knn2_hamming(q, t, 2, 0)
    assert np.array_equal(idx, gold[name + "_lin_idx"]) and np.array_equal(dist, gold[name + "_lin_dist"])
    nb = q.shape[1]
    idx, dist = m.knn2_hamming(q, t, 2, 1 if nb & 1 else 2)
    assert np.array_equal(dist, gold[name + "_nms_dist"])
    strict = dist[:, 0] < dist[:, 1]
    assert np.array_equal(idx[strict, 0], gold[name + "_nms_idx"][strict, 0])


@pytest.mark.parametrize("nq,nt,nb", [(1, 1, 32), (1, 2, 32), (17, 5000, 32), (1500, 3000, 32), (1025, 257, 64),
                                      (300, 4097, 61), (64, 700, 16), (50, 600, 128), (33, 300, 200), (8192, 8192, 32)])
def test_hamming_vs_oracle(m, nq, nt, nb):
    q, t = datagen.binary_descriptors(nq, nt, nb, seed=nq * 7 + nt)
    for k in (1, 2):
        for mode, ign in ((0, 0), (1, 1 if nb & 1 else 2)):
            idx, dist = m.knn2_hamming(q, t, k, ign)
            oi, od = orc.knn_hamming(q, t, k, mode, threads=8)
            assert np.array_equal(dist, od), (nq, nt, nb, k, mode)
            assert np.array_equal(idx, oi), (nq, nt, nb, k, mode)


@pytest.mark.parametrize("nq,nt,nb", [(1, 2, 32), (17, 5000, 32), (1500, 3000, 32), (1025, 257, 64), (300, 4097, 61), (64, 700, 16),
                                      (257, 129, 32), (700, 70000, 32), (200, 250000, 32), (300, 60000, 64), (8192, 8192, 32)])
@pytest.mark.parametrize("path", [2, 3])
def test_hamming_tensor_core_path_vs_oracle(m, nq, nt, nb, path):
    """mpl_set_hamming_path(2 / 3): the bits as e4m3 bytes / bf16 on the float ranking kernel's exact-integer mode (hamming = |q| + |t| -
    2 q.t); the large shapes walk the train set in chunks, 512-bit rows use 128-row query blocks in bf16.  Same keys as the popc kernel."""
    q, t = datagen.binary_descriptors(nq, nt, nb, seed=nq * 7 + nt)
    if nt >= 5000:
        t[nt // 2] = t[3]; t[nt - 1] = t[3]; q[0] = t[3]                # exact duplicates in different tiles / chunks: ties by index
    import ctypes
    lib = m._lib.load()
    lib.mpl_set_hamming_path(path)
    lib.mpl_prof_enable(1)
    ms, cnt = ctypes.c_double(0), ctypes.c_longlong(0)
    lib.mpl_prof_read(2, ctypes.byref(ms), ctypes.byref(cnt))               # reading clears the ranking kernel's timer
    try:
        for k in (1, 2):
            for mode, ign in ((0, 0), (1, 1 if nb & 1 else 2)):
                if mode == 1 and nt > 100000:
                    continue
                idx, dist = m.knn2_hamming(q, t, k, ign)
                oi, od = orc.knn_hamming(q, t, k, mode, threads=8)
                assert np.array_equal(dist, od), (nq, nt, nb, k, mode)
                assert np.array_equal(idx, oi), (nq, nt, nb, k, mode)
        lib.mpl_prof_read(2, ctypes.byref(ms), ctypes.byref(cnt))
        assert cnt.value >= 2, "the tensor-core ranking kernel did not run"
    finally:
        lib.mpl_set_hamming_path(0)
        lib.mpl_prof_enable(0)


def test_hamming_strided_and_empty(m):
    q, t = datagen.binary_descriptors(100, 300, 32, seed=9)
    big_q = np.zeros((100, 48), np.uint8); big_q[:, :32] = q
    big_t = np.zeros((300, 40), np.uint8); big_t[:, :32] = t
    idx, dist = m.knn2_hamming(big_q[:, :32], big_t[:, :32])        # row stride != row bytes (cv::Mat::step)
    oi, od = orc.knn_hamming(q, t)
    assert np.array_equal(idx, oi) and np.array_equal(dist, od)
    idx, dist = m.knn2_hamming(np.zeros((0, 32), np.uint8), t)
    assert idx.shape == (0, 2)
    idx, dist = m.knn2_hamming(q, np.zeros((0, 32), np.uint8))
    assert (idx == -1).all() and (dist == np.iinfo(np.int32).max).all()


@pytest.mark.parametrize("name", ["sift128", "surf64", "odd61"])
def test_l2_golden(m, gold, name):
    q, t = gold[name + "_q"], gold[name + "_t"]
    idx, dist = m.knn2_l2(q, t, 2, False)
    assert np.array_equal(idx, gold[name + "_lin_idx"])
    assert np.array_equal(dist.view(np.int32), gold[name + "_lin_dist"].view(np.int32))
    idx, dist = m.knn2_l2(q, t, 2, True)
    np.testing.assert_allclose(dist, gold[name + "_nms_dist"], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("path", [1, 2])
@pytest.mark.parametrize("nq,nt,d,kind", [(1, 1, 128, "sift"), (70, 1000, 128, "sift"), (1000, 3000, 128, "sift"),
                                           (129, 777, 64, "signed"), (200, 1500, 128, "signed"), (64, 300, 61, "signed"),
                                           (2500, 20000, 128, "sift")])
def test_l2_vs_oracle(m, path, nq, nt, d, kind):
    """path 1 = exact SIMT scan, 2 = tcgen05 ranking + exact re-rank.  Distances bit-exact in cvflann order;
    indices equal (both sides break exact ties toward the lowest train index)."""
    lib = m._lib.load()
    q, t = datagen.sift_like(nq, nt, d, seed=nq + nt) if kind == "sift" else datagen.signed_float(nq, nt, d, seed=nq + nt)
    lib.mpl_set_l2_path(path)
    try:
        try:
            idx, dist = m.knn2_l2(q, t, 2, False)
        except m._lib.MplError as e:
            if path == 2 and "unsupported" in str(e):
                pytest.skip("tensor-core path does not take this shape")
            raise
    finally:
        lib.mpl_set_l2_path(0)
    oi, od = orc.knn_l2(q, t, 2, 0, threads=8)
    assert np.array_equal(dist.view(np.int32), od.view(np.int32)), (path, nq, nt, d)
    assert np.array_equal(idx, oi)


def test_l2_nmslib_mode(m):
    q, t = datagen.signed_float(150, 900, 64, seed=3)
    idx, dist = m.knn2_l2(q, t, 2, True)
    oi, od = orc.knn_l2(q, t, 2, 1)
    assert np.array_equal(dist.view(np.int32), od.view(np.int32)) and np.array_equal(idx, oi)


@pytest.mark.parametrize("matcher", ["LINEAR", "BRUTEFORCENMS"])
@pytest.mark.parametrize("dtype", ["u8", "f32"])
@pytest.mark.parametrize("ratio", [True, False])
def test_get_matches_api(m, matcher, dtype, ratio):
    if dtype == "u8":
        d1, d2 = datagen.binary_descriptors(400, 500, 32, seed=11)
    else:
        d1, d2 = datagen.sift_like(400, 500, 128, seed=12)
    kp1, kp2 = [None] * 400, [None] * 500
    err, got = m.getMatchesArray(kp1, kp2, d1, d2, None, matcher, False, ratio)
    oerr, exp = orc.get_matches(d1, d2, matcher, ratio)
    assert err == oerr == 0
    assert np.array_equal(got["queryIdx"], exp["queryIdx"])
    if dtype == "u8" or matcher == "LINEAR":
        if ratio:
            assert np.array_equal(got["trainIdx"], exp["trainIdx"])
        assert np.array_equal(got["distance"].view(np.int32), exp["distance"].view(np.int32))
    else:
        np.testing.assert_allclose(got["distance"], exp["distance"], rtol=1e-5)
    lst = []
    assert m.getMatches(kp1, kp2, d1, d2, None, lst, matcher, False, ratio) == 0 and len(lst) == len(exp)


def test_get_matches_error_codes(m):
    d1, d2 = datagen.binary_descriptors(40, 50, 32, seed=13, adversarial=False)
    d1 = np.random.default_rng(0).integers(0, 256, (40, 32), dtype=np.uint8)
    kp1, kp2 = [None] * 40, [None] * 50
    assert m.getMatchesArray(kp1[:10], kp2, d1[:10], d2, None, "LINEAR")[0] == -4
    assert m.getMatchesArray(kp1[:39], kp2, d1, d2, None, "LINEAR")[0] == -1
    assert m.getMatchesArray(kp1, kp2, d1, d2, None, "GMBSOF")[0] == -2
    assert m.getMatchesArray(kp1, kp2, d1, d2, None, "LINEAR")[0] == -3
    assert m.getMatchesArray(kp1, kp2, d1.astype(np.float64), d2.astype(np.float64), None, "LINEAR")[0] == -1
    with pytest.raises(AssertionError):
        m.getMatchesArray(kp1, kp2, d1, d2.astype(np.float32), None, "LINEAR")


# ---- tensor-core float path (knn_l2_tc.cu): ranking on tcgen05, exact re-rank, certificate, exact fallback ----
def _tc(m, q, t, sqrt_out=False):
    import ctypes
    lib = m._lib.load()
    lib.mpl_set_l2_path(2)
    try:
        idx, dist = m.knn2_l2(q, t, 2, sqrt_out)
    finally:
        lib.mpl_set_l2_path(0)
    fb, mode = ctypes.c_uint(0), ctypes.c_uint(0)
    m._lib.check(lib.mpl_knn_l2_last_stats_host(ctypes.byref(fb), ctypes.byref(mode)), "stats")
    return idx, dist, fb.value, mode.value


@pytest.mark.parametrize("nq,nt,d,kind", [(5, 2, 128, "sift"), (256, 128, 128, "sift"), (257, 129, 128, "sift"), (300, 40000, 128, "sift"),
                                           (5000, 5000, 128, "sift"), (700, 9000, 64, "sift"), (640, 3000, 120, "signed"),
                                           (1000, 20000, 128, "signed"), (513, 2050, 33, "signed"), (400, 1111, 7, "signed"),
                                           # 128 < dim <= 256 (DAISY is 200-d): query blocks of 128 rows, up to 8 tf32 K slabs
                                           (129, 300, 129, "signed"), (700, 6000, 200, "signed"), (1000, 9000, 256, "signed"),
                                           (5, 2, 200, "sift"), (300, 5000, 192, "sift"), (1100, 7000, 256, "sift"), (257, 1000, 130, "sift")])
def test_l2_tc_path(m, nq, nt, d, kind):
    q, t = datagen.sift_like(nq, nt, d, seed=nq + nt) if kind == "sift" else datagen.signed_float(nq, nt, d, seed=nq + nt)
    idx, dist, fb, mode = _tc(m, q, t)
    assert mode == (0 if kind == "sift" else 1)           # integer-valued SIFT -> bf16-exact operands, else tf32
    oi, od = orc.knn_l2(q, t, 2, 0, threads=8)
    assert np.array_equal(dist.view(np.int32), od.view(np.int32)), (nq, nt, d)
    assert np.array_equal(idx, oi)
    assert fb <= max(4, nq // 20), "certificate failed for %d of %d queries" % (fb, nq)


def test_l2_wide_rows_take_the_tensor_core_path_by_default(m):
    """DAISY-like 200-d rows, large enough for the default path to choose the tensor cores: LINEAR and BRUTEFORCENMS order vs the oracle."""
    q, t = datagen.signed_float(3000, 4000, 200, seed=200)
    for mode in (0, 1):
        import ctypes
        idx, dist = m.knn2_l2(q, t, 2, bool(mode))
        fb, tf = ctypes.c_uint(0), ctypes.c_uint(0)
        m._lib.check(m._lib.load().mpl_knn_l2_last_stats_host(ctypes.byref(fb), ctypes.byref(tf)), "stats")
        assert tf.value == 1 and fb.value <= 150
        oi, od = orc.knn_l2(q, t, 2, mode, threads=8)
        assert np.array_equal(dist.view(np.int32), od.view(np.int32)) and np.array_equal(idx, oi)


def test_l2_tc_sqrt_mode(m):
    q, t = datagen.signed_float(600, 5000, 64, seed=5)
    idx, dist, fb, mode = _tc(m, q, t, True)
    oi, od = orc.knn_l2(q, t, 2, 1, threads=8)
    assert np.array_equal(dist.view(np.int32), od.view(np.int32)) and np.array_equal(idx, oi)


def test_l2_tc_duplicates_integer_mode_needs_no_fallback(m):
    """Ten identical train rows, integer-valued (SIFT) data: scores are exact, the top-2 lists keep the lowest train
    indices on ties and no query needs the exact scan."""
    q, t = datagen.sift_like(300, 3000, 128, seed=77)
    t[100:110] = t[5]
    q[0] = t[5]; q[1] = t[5]; q[1, 3] += 1
    t[2000] = t[5]; t[2999] = t[5]                       # duplicates in other tiles / the other column half too
    idx, dist, fb, mode = _tc(m, q, t)
    oi, od = orc.knn_l2(q, t, 2, 0, threads=8)
    assert np.array_equal(dist.view(np.int32), od.view(np.int32)) and np.array_equal(idx, oi)
    assert idx[0, 0] == 5 and idx[0, 1] == 100 and mode == 0 and fb == 0


def test_l2_tc_duplicates_take_the_fallback(m):
    """Ten identical train rows, non-integer data (tf32 ranking): the 4-deep candidate lists cannot certify -> exact
    scan, lowest indices win."""
    q, t = datagen.signed_float(300, 3000, 128, seed=77)
    t[100:110] = t[5]
    q[0] = t[5]; q[1] = t[5]; q[1, 3] += 0.25
    idx, dist, fb, mode = _tc(m, q, t)
    oi, od = orc.knn_l2(q, t, 2, 0, threads=8)
    assert np.array_equal(dist.view(np.int32), od.view(np.int32)) and np.array_equal(idx, oi)
    assert idx[0, 0] == 5 and idx[0, 1] == 100 and mode == 1 and fb >= 1


def test_l2_tc_large_integers_leave_exact_mode(m):
    """Integer data whose distances exceed 2^24 are not exact in fp32: the top-2 lists cannot be certified and every
    query must come back through the exact scan, still bit-identical to the reference."""
    q, t = datagen.sift_like(64, 600, 128, seed=5)
    q *= 16; t *= 16                                     # still bf16-exact integers, norms ~ 8192
    idx, dist, fb, mode = _tc(m, q, t)
    oi, od = orc.knn_l2(q, t, 2, 0, threads=8)
    assert np.array_equal(dist.view(np.int32), od.view(np.int32)) and np.array_equal(idx, oi)
    assert mode == 0 and fb == 64


def test_l2_tc_equals_exact_scan_at_c2_size(m):
    """BASELINE config C2 (20 000 x 20 000 x 128): tensor-core path == exact fp32 scan, bit for bit."""
    lib = m._lib.load()
    q, t = datagen.sift_like(20000, 20000, 128, seed=1002)
    idx, dist, fb, mode = _tc(m, q, t)
    lib.mpl_set_l2_path(1)
    try:
        ei, ed = m.knn2_l2(q, t, 2, False)
    finally:
        lib.mpl_set_l2_path(0)
    assert np.array_equal(idx, ei) and np.array_equal(dist.view(np.int32), ed.view(np.int32))
    assert mode == 0 and fb < 200


def test_pinned_array_round_trip(m):
    """mpl_host_alloc memory (huge pages + cudaHostRegister) is ordinary writable host memory that the entry points accept;
    results equal those from pageable arrays, and the allocation is released with the array."""
    import gc
    q, t = datagen.sift_like(700, 900, 128, seed=12)
    pq, pt = m.pinned_array(q.shape, q.dtype), m.pinned_array(t.shape, t.dtype)
    pq[...] = q; pt[...] = t
    assert pq.ctypes.data % (2 << 20) == 0
    i1, d1 = m.knn2_l2(pq, pt, 2, False)
    i2, d2 = m.knn2_l2(q, t, 2, False)
    assert np.array_equal(i1, i2) and np.array_equal(d1.view(np.int32), d2.view(np.int32))
    lib = m._lib.load()
    assert lib.mpl_host_free(pq.ctypes.data + 64) != 0            # not an allocation base
    del pq, pt
    gc.collect()


def test_allgather_merge_over_peer_memory_two_ranks_one_device(m):
    """allgather_merge_kernel with two 'ranks' as two streams of one device (peer pointers are then plain device pointers):
    every rank ends with the merge of both shards' keys, step after step (the buffers alternate), equal to
    mpl_merge_top2_dev over the gathered parts and to the one-piece scan."""
    import ctypes
    import torch
    lib = m._lib.load()
    dev = torch.device("cuda:0")
    nq, nt, nb = 3000, 4096, 32
    q, t = datagen.binary_descriptors(nq, nt, nb, seed=21)
    world = 2
    gbytes = int(lib.mpl_peer_gather_bytes(nq, world))
    gather = [torch.zeros(gbytes, dtype=torch.uint8, device=dev) for _ in range(world)]
    sync = [torch.zeros(64, dtype=torch.int32, device=dev) for _ in range(world)]
    pg = (ctypes.c_void_p * world)(*[g.data_ptr() for g in gather])
    ps = (ctypes.c_void_p * world)(*[s.data_ptr() for s in sync])
    streams = [torch.cuda.Stream() for _ in range(world)]
    d_q = torch.from_numpy(q).to(dev)
    half = nt // 2
    shards = [torch.from_numpy(np.ascontiguousarray(t[r * half:(r + 1) * half])).to(dev) for r in range(world)]
    scr_bytes = int(lib.mpl_knn_hamming_scratch_bytes(nq, half, nb // 16))
    for step in range(3):                                           # three steps: both buffer parities, counters keep growing
        if step == 1:                                               # change the data so that a stale buffer would show
            d_q = torch.from_numpy(np.ascontiguousarray(q[::-1])).to(dev)
        keys = [torch.empty((nq, 2), dtype=torch.int64, device=dev) for _ in range(world)]
        scr = [torch.empty(scr_bytes, dtype=torch.uint8, device=dev) for _ in range(world)]
        idx = [torch.empty((nq, 2), dtype=torch.int32, device=dev) for _ in range(world)]
        dist = [torch.empty((nq, 2), dtype=torch.int32, device=dev) for _ in range(world)]
        torch.cuda.synchronize()
        for r in range(world):
            st = ctypes.c_void_p(streams[r].cuda_stream)
            m._lib.check(lib.mpl_knn2_hamming_dev(d_q.data_ptr(), nq, shards[r].data_ptr(), half, nb // 16, r * half, scr[r].data_ptr(), scr_bytes,
                                                  keys[r].data_ptr(), st), "hamming_dev")
            m._lib.check(lib.mpl_allgather_merge_top2_dev(keys[r].data_ptr(), nq, r, world, pg, ps, 2, 0, idx[r].data_ptr(), dist[r].data_ptr(), None, st),
                         "allgather_merge")
        torch.cuda.synchronize()
        qq = q if step == 0 else np.ascontiguousarray(q[::-1])
        ei, ed = m.knn2_hamming(qq, t, 2)
        for r in range(world):
            assert np.array_equal(idx[r].cpu().numpy(), ei) and np.array_equal(dist[r].cpu().numpy(), ed), (step, r)
    assert int(sync[0][0].item()) == 3 and int(sync[1][0].item()) == 3


def test_hamming_database_shard_beyond_2_pow_20_rows(m):
    """Database regime on one GPU: 2.3 M train rows (more than one 2^20-row split, global indices above 2^21).  Planted
    exact copies and their later duplicates must come back as (0, first index), (0, duplicate index); a few queries are
    checked against the oracle over the whole shard."""
    rng = np.random.default_rng(77)
    nt, nq, nb = 2_300_000, 64, 32
    t = rng.integers(0, 256, (nt, nb), dtype=np.uint8)
    pos = np.sort(rng.choice(nt - 1000, nq, replace=False)); pos[-1] = nt - 1000; pos[0] = 0
    dup = np.minimum(pos + 1 + rng.integers(0, 900, nq), nt - 1)
    t[dup] = t[pos]
    q = t[pos].copy()
    idx, dist = m.knn2_hamming(q, t, 2)
    assert np.array_equal(idx[:, 0], pos) and np.array_equal(idx[:, 1], dup) and not dist.any()
    q2 = rng.integers(0, 256, (6, nb), dtype=np.uint8)
    i2, d2 = m.knn2_hamming(q2, t, 2)
    oi, od = orc.knn_hamming(q2, t, 2, 0, threads=8)
    assert np.array_equal(i2, oi) and np.array_equal(d2, od)


@pytest.mark.parametrize("kind,mode", [("sift", 0), ("sift", 1), ("signed", 0), ("signed", 1)])
def test_l2_default_path_vs_oracle_at_c2_size(m, kind, mode):
    """BASELINE config C2 at its own size (20 000 x 20 000 x 128) through the DEFAULT path (auto -> tensor-core ranking +
    exact re-rank) against the ORACLE: sift = bf16-exact integers (kind::f16 ranking), signed = unit-norm floats (kind::tf32
    ranking + top-4 + certificate + exact re-rank); mode 0 = LINEAR (cvflann order, squared), 1 = BRUTEFORCENMS (NMSLIB
    4-lane order, sqrt).  Distances bit-exact fp32, indices equal."""
    import ctypes
    lib = m._lib.load()
    q, t = datagen.sift_like(20000, 20000, 128, seed=1002) if kind == "sift" else datagen.signed_float(20000, 20000, 128, seed=1002)
    lib.mpl_set_l2_path(0)
    idx, dist = m.knn2_l2(q, t, 2, bool(mode))
    fb, tf = ctypes.c_uint(0), ctypes.c_uint(0)
    m._lib.check(lib.mpl_knn_l2_last_stats_host(ctypes.byref(fb), ctypes.byref(tf)), "stats")
    oi, od = orc.knn_l2(q, t, 2, mode, threads=os.cpu_count() or 8)
    assert np.array_equal(dist.view(np.int32), od.view(np.int32)), (kind, mode)
    assert np.array_equal(idx, oi), (kind, mode)
    assert tf.value == (0 if kind == "sift" else 1)
    print("C2 %s mode %d: tf32 ranking %d, uncertified queries rescanned exactly: %d of 20000" % (kind, mode, tf.value, fb.value))
    assert fb.value < 400


def test_c2_gpu_vs_reference_nmslib_near_ties_only(m):
    """north_star's float bar against the reference's OWN code: the compiled vendored NMSLIB (oracle/_ref, built -Ofast, so
    its fp32 sums are not reproducible bit for bit).  Every query whose first neighbour differs must be a near-tie within
    1e-5 relative distance."""
    if orc.ref_lib() is None:
        pytest.skip("oracle/_ref/libnmslib_ref.so was not built (no /root/reference at build time)")
    q, t = datagen.sift_like(4000, 20000, 128, seed=1002)
    idx, dist = m.knn2_l2(q, t, 2, True)
    ri, rd = orc.ref_nmslib_l2(q, t, threads=os.cpu_count() or 8)
    diff = np.nonzero(idx[:, 0] != ri[:, 0])[0]
    gaps = [abs(float(dist[i, 0]) - float(rd[i, 0])) / max(float(rd[i, 0]), 1e-30) for i in diff]
    print("first-neighbour disagreements with the reference's NMSLIB: %d of %d, max relative distance gap %.3g" % (len(diff), len(q), max(gaps) if gaps else 0.0))
    assert all(g < 1e-5 for g in gaps)
    np.testing.assert_allclose(dist[:, 0], rd[:, 0], rtol=1e-5)


def test_reinit_on_second_device(m):
    """mpl_init(1) after work on device 0: the opt-in shared-memory attributes are per device."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("one GPU visible")
    import datagen_pose
    lib = m._lib.load()
    q, t = datagen.sift_like(2100, 2100, 128, seed=2)
    sc = datagen_pose.make_scene(500, 0.3, seed=5)
    s = datagen_pose.draw_samples_fast(500, 64, seed=6)
    try:
        for dev in (0, 1, 0):
            m._lib.check(lib.mpl_init(dev), "mpl_init")
            idx, dist = m.knn2_l2(q, t, 2, False)
            oi, od = orc.knn_l2(q, t, 2, 0, threads=8)
            assert np.array_equal(idx, oi) and np.array_equal(dist.view(np.int32), od.view(np.int32))
            r = m.ransac5pt(sc["p1"], sc["p2"], s, sc["th"], refine=True)
            assert r["ok"]
    finally:
        m._lib.check(lib.mpl_init(0), "mpl_init")
