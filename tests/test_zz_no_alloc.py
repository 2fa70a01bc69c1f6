"""GPU: mpl_knn2_hamming_static_dev, the entry point that allocates nothing (scratch in a file-scope device array,
csrc/knn_hamming_static.cu; no counterpart in the reference), called through the
C-ABI and compared bit for bit with mpl_knn2_hamming_dev on the same device buffers, on two calls in a row (the second with other
data and another shape, so stale partial keys would show), plus the size limit.  Collected last on purpose: `pytest -x` must not
lose the other tests to a failure here.  Passes alone in a fresh process and after the rest of the suite."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def L():
    import __graft_entry__ as g
    g.build()
    from matchinglib_poselib_b200 import _lib
    lib = _lib.load()
    _lib.check(lib.mpl_init(0), "mpl_init")
    return _lib, lib


def _run_both(_lib, lib, torch, q, t, offset):
    dev = torch.device("cuda:0")
    nq, nt, w16 = q.shape[0], t.shape[0], q.shape[1] // 16
    d_q = torch.from_numpy(q).to(dev); d_t = torch.from_numpy(t).to(dev)
    scr_bytes = int(lib.mpl_knn_hamming_scratch_bytes(nq, nt, w16))
    scr = torch.empty(max(scr_bytes, 16), dtype=torch.uint8, device=dev)
    ref = torch.full((nq, 2), -1, dtype=torch.int64, device=dev); out = torch.full((nq, 2), -2, dtype=torch.int64, device=dev)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    _lib.check(lib.mpl_knn2_hamming_dev(d_q.data_ptr(), nq, d_t.data_ptr(), nt, w16, offset, scr.data_ptr(), scr_bytes, ref.data_ptr(), st), "hamming_dev")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    _lib.check(lib.mpl_knn2_hamming_static_dev(d_q.data_ptr(), nq, d_t.data_ptr(), nt, w16, offset, out.data_ptr(), st), "hamming_static_dev")
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    return ref.cpu().numpy(), out.cpu().numpy(), scr_bytes


def test_static_scratch_entry_point_equals_the_scratch_argument_one_twice_in_a_row(L):
    import torch
    _lib, lib = L
    assert lib.mpl_knn_hamming_static_max_bytes() == 4 << 20
    rng = np.random.default_rng(7)
    # call 1: few queries against many train rows -> the train-split path (partial keys in scratch)
    q1 = rng.integers(0, 256, (96, 32), dtype=np.uint8); t1 = rng.integers(0, 256, (150000, 32), dtype=np.uint8)
    t1[1234] = q1[5]; t1[99999] = q1[5]                                      # a tie: the lower train index must win on both paths
    ref1, out1, scr1 = _run_both(_lib, lib, torch, q1, t1, 1000)
    assert scr1 > 96 * 16, "the case must exercise the train-split path"
    assert np.array_equal(ref1, out1)                                       # bit-exact: integer keys (dist << 32 | train index)
    assert (out1[5] >> 32).tolist() == [0, 0] and (out1[5] & 0xffffffff).tolist() == [1000 + 1234, 1000 + 99999]
    # call 2, right after: other data, more queries, another split count, 61-byte rows padded to 64 (w16 = 4)
    q2 = np.zeros((333, 64), np.uint8); t2 = np.zeros((70001, 64), np.uint8)
    q2[:, :61] = rng.integers(0, 256, (333, 61), dtype=np.uint8); t2[:, :61] = rng.integers(0, 256, (70001, 61), dtype=np.uint8)
    ref2, out2, scr2 = _run_both(_lib, lib, torch, q2, t2, 0)
    assert scr2 > 333 * 16 and np.array_equal(ref2, out2)
    # and the oracle agrees with both (independent check of the pair of them)
    from oracle import knn as oknn
    oi, od = oknn.knn_hamming(q2[:16], t2, 2)
    assert np.array_equal(out2[:16] & 0xffffffff, oi) and np.array_equal(out2[:16] >> 32, od)


def test_static_scratch_entry_point_small_and_empty_cases(L):
    import torch
    _lib, lib = L
    rng = np.random.default_rng(8)
    q = rng.integers(0, 256, (5000, 32), dtype=np.uint8); t = rng.integers(0, 256, (4000, 32), dtype=np.uint8)
    ref, out, scr = _run_both(_lib, lib, torch, q, t, 0)                      # the preferred split count needs more than the array holds:
    assert scr > (4 << 20) and np.array_equal(ref, out)                      # fewer splits are used, the keys are the same
    ref, out, scr = _run_both(_lib, lib, torch, np.tile(q, (8, 1)), t[:600], 0)    # many queries, few rows: one split, the scratch is not touched
    assert np.array_equal(ref, out)
    dev = torch.device("cuda:0")
    d_q = torch.from_numpy(q).to(dev); out = torch.zeros((5000, 2), dtype=torch.int64, device=dev)
    _lib.check(lib.mpl_knn2_hamming_static_dev(d_q.data_ptr(), 5000, None, 0, 2, 0, out.data_ptr(), None), "empty train set")
    torch.cuda.synchronize()
    assert bool((out == -1).all())                                           # all-ones keys, like mpl_knn2_hamming_dev
    assert lib.mpl_knn2_hamming_static_dev(None, 0, None, 0, 2, 0, None, None) == 0


def test_static_scratch_entry_point_rejects_what_does_not_fit(L):
    import torch
    _lib, lib = L
    nq, nt = 150000, (1 << 21) + 5                                            # 20-bit row indices force >= 3 splits: 3 x 150000 x 16 B > 4 MiB
    assert int(lib.mpl_knn_hamming_scratch_bytes(nq, nt, 2)) > (4 << 20)
    dev = torch.device("cuda:0")
    d_q = torch.zeros((nq, 32), dtype=torch.uint8, device=dev); d_t = torch.zeros((nt, 32), dtype=torch.uint8, device=dev)
    out = torch.zeros((nq, 2), dtype=torch.int64, device=dev)
    rc = lib.mpl_knn2_hamming_static_dev(d_q.data_ptr(), nq, d_t.data_ptr(), nt, 2, 0, out.data_ptr(), None)
    assert rc == -102 and "MPL_KNN_STATIC_SCRATCH_BYTES" in _lib.last_error()  # MPL_ERR_LIMIT, nothing launched
    torch.cuda.synchronize()
    assert bool((out == 0).all())
