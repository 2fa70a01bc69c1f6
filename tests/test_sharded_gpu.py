"""GPU: the sharded paths on ONE device -- ranks emulated as successive train shards (global indices through
train_offset), their key lists stacked exactly as the NCCL all-gather lays them out, then merge_top2_kernel.
Result must equal the single scan bit for bit.  Also the batch-of-pairs launch (config C4)."""
import ctypes

import numpy as np
import pytest

import datagen
from oracle import knn as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    import __graft_entry__ as g
    g.build()
    import torch
    import matchinglib_poselib_b200 as m
    m._lib.check(m._lib.load().mpl_init(0), "mpl_init")
    return m, torch, m._lib.load()


def _st(torch):
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


@pytest.mark.parametrize("world", [2, 3, 8])
def test_hamming_train_sharded(env, world):
    m, torch, lib = env
    from matchinglib_poselib_b200.sharded import shard_rows
    q, t = datagen.binary_descriptors(777, 30011, 32, seed=world)
    t[20000] = t[7]; q[5] = t[7]                                     # duplicate across shards -> lower global index first
    dev = torch.device("cuda:0")
    d_q = torch.from_numpy(q).to(dev)
    nq = len(q)
    parts = torch.empty((world, nq, 2), dtype=torch.int64, device=dev)
    for r in range(world):
        a, b = shard_rows(len(t), world, r)
        d_t = torch.from_numpy(t[a:b]).to(dev)
        sb = int(lib.mpl_knn_hamming_scratch_bytes(nq, b - a, 2))
        scr = torch.empty(max(sb, 16), dtype=torch.uint8, device=dev)
        m._lib.check(lib.mpl_knn2_hamming_dev(d_q.data_ptr(), nq, d_t.data_ptr(), b - a, 2, a, scr.data_ptr(), sb, parts[r].data_ptr(), _st(torch)), "knn")
    idx = torch.empty((nq, 2), dtype=torch.int32, device=dev); dist = torch.empty((nq, 2), dtype=torch.int32, device=dev)
    m._lib.check(lib.mpl_merge_top2_dev(parts.data_ptr(), world, nq, 2, 0, idx.data_ptr(), dist.data_ptr(), None, _st(torch)), "merge")
    torch.cuda.synchronize()
    oi, od = orc.knn_hamming(q, t, 2, 0, threads=8)
    assert np.array_equal(idx.cpu().numpy(), oi) and np.array_equal(dist.cpu().numpy(), od)
    assert idx[5].tolist() == [7, 20000]


def test_sharded_index_class_single_rank(env):
    m, torch, lib = env
    from matchinglib_poselib_b200.sharded import ShardedHammingIndex
    q, t = datagen.binary_descriptors(200, 5000, 61, seed=9)          # AKAZE-sized rows go through the packer
    ix = ShardedHammingIndex(t, 0, torch.device("cuda:0"))
    idx, dist = ix.match(q)
    oi, od = orc.knn_hamming(q, t, 2, 0, threads=8)
    assert np.array_equal(idx.cpu().numpy(), oi) and np.array_equal(dist.cpu().numpy(), od)


def test_l2_train_sharded_tc(env):
    m, torch, lib = env
    from matchinglib_poselib_b200.sharded import shard_rows
    q, t = datagen.sift_like(600, 24000, 128, seed=4)
    dev = torch.device("cuda:0")
    d_q = torch.from_numpy(q).to(dev)
    nq, world = len(q), 3
    parts = torch.empty((world, nq, 2), dtype=torch.int64, device=dev)
    lib.mpl_set_l2_path(2)
    try:
        for r in range(world):
            a, b = shard_rows(len(t), world, r)
            d_t = torch.from_numpy(t[a:b]).to(dev)
            sb = int(lib.mpl_knn_l2_scratch_bytes(nq, b - a, 128))
            scr = torch.empty(sb, dtype=torch.uint8, device=dev)
            m._lib.check(lib.mpl_knn2_l2f32_dev(d_q.data_ptr(), nq, 128, d_t.data_ptr(), b - a, 128, 128, 0, a, scr.data_ptr(), sb,
                                                parts[r].data_ptr(), _st(torch)), "l2")
    finally:
        lib.mpl_set_l2_path(0)
    idx = torch.empty((nq, 2), dtype=torch.int32, device=dev); dist = torch.empty((nq, 2), dtype=torch.float32, device=dev)
    m._lib.check(lib.mpl_merge_top2_dev(parts.data_ptr(), world, nq, 2, 1, idx.data_ptr(), dist.data_ptr(), None, _st(torch)), "merge")
    torch.cuda.synchronize()
    oi, od = orc.knn_l2(q, t, 2, 0, threads=8)
    assert np.array_equal(idx.cpu().numpy(), oi) and np.array_equal(dist.cpu().numpy().view(np.int32), od.view(np.int32))


def test_pair_batch(env):
    """Config C4 shape in small: independent pairs in one launch (blockIdx.z), AKAZE 61-byte rows packed to 64."""
    m, torch, lib = env
    n_pairs, nq, nt = 5, 300, 450
    dev = torch.device("cuda:0")
    qs, ts = zip(*[datagen.binary_descriptors(nq, nt, 61, seed=100 + p) for p in range(n_pairs)])
    pack = lambda a: np.concatenate([a, np.zeros((a.shape[0], 3), np.uint8)], 1)
    d_q = torch.from_numpy(np.stack([pack(x) for x in qs])).to(dev); d_t = torch.from_numpy(np.stack([pack(x) for x in ts])).to(dev)
    keys = torch.empty((n_pairs, nq, 2), dtype=torch.int64, device=dev)
    m._lib.check(lib.mpl_knn2_hamming_batch_dev(d_q.data_ptr(), d_t.data_ptr(), n_pairs, nq, nt, 4, keys.data_ptr(), _st(torch)), "batch")
    torch.cuda.synchronize()
    k = keys.cpu().numpy().view(np.uint64)
    for p in range(n_pairs):
        oi, od = orc.knn_hamming(qs[p], ts[p], 2, 0)
        assert np.array_equal((k[p] & np.uint64(0xFFFFFFFF)).astype(np.int32), oi) and np.array_equal((k[p] >> np.uint64(32)).astype(np.int32), od)


def _sharded_worker(rank, world, port, tmp):
    import os
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from matchinglib_poselib_b200.sharded import ShardedHammingIndex, ShardedL2Index, shard_rows
    q, t = np.load(os.path.join(tmp, "q.npy")), np.load(os.path.join(tmp, "t.npy"))
    a, b = shard_rows(len(t), world, rank)
    ix = ShardedHammingIndex(np.ascontiguousarray(t[a:b]), a, dev, group=None)
    assert ix.peers.ok, "peer-memory exchange not available between the two GPUs"
    for rep in range(3):                                            # both buffer parities
        idx, d = ix.match(q)
    fq, ft = np.load(os.path.join(tmp, "fq.npy")), np.load(os.path.join(tmp, "ft.npy"))
    a2, b2 = shard_rows(len(ft), world, rank)
    fx = ShardedL2Index(np.ascontiguousarray(ft[a2:b2]), a2, dev)
    fi, fd = fx.match(fq)
    torch.cuda.synchronize()
    np.save(os.path.join(tmp, "out%d.npy" % rank), np.concatenate([idx.cpu().numpy(), d.cpu().numpy()], 1))
    np.save(os.path.join(tmp, "fout%d.npy" % rank), np.concatenate([fi.cpu().numpy(), fd.cpu().numpy().view(np.int32)], 1))
    dist.barrier()
    dist.destroy_process_group()


def test_two_process_peer_exchange_vs_oracle(env, tmp_path):
    """The REAL multi-process path: two ranks on two GPUs, ShardedHammingIndex / ShardedL2Index exchanging their top-2 keys
    through NVLink peer memory inside the merge kernel; every rank's result must equal the oracle's scan of the whole set."""
    m, torch, lib = env
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import socket
    import torch.multiprocessing as mp
    q, t = datagen.binary_descriptors(3000, 40001, 32, seed=8)
    t[30000] = t[9]; q[4] = t[9]
    fq, ft = datagen.sift_like(2200, 6001, 128, seed=9)
    for name, arr in (("q", q), ("t", t), ("fq", fq), ("ft", ft)):
        np.save(tmp_path / (name + ".npy"), arr)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
    mp.spawn(_sharded_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    oi, od = orc.knn_hamming(q, t, 2, 0, threads=8)
    fi, fd = orc.knn_l2(fq, ft, 2, 0, threads=8)
    for r in range(2):
        got = np.load(tmp_path / ("out%d.npy" % r))
        assert np.array_equal(got[:, :2], oi) and np.array_equal(got[:, 2:], od), r
        fgot = np.load(tmp_path / ("fout%d.npy" % r))
        assert np.array_equal(fgot[:, :2], fi) and np.array_equal(fgot[:, 2:], fd.view(np.int32)), r
