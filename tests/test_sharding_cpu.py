"""CPU (gloo, world_size 2): host-side logic of the sharded paths -- shard bounds, global train indices, the
all-gather layout of the top-2 keys and the (dist, index) merge order.  The per-rank scan and the merge are the
CUDA kernels' job on the GPU box; here the oracle plays the scan and numpy the merge, so that what is tested is
matchinglib_poselib_b200/sharded.py (partitioning + exchange)."""
import os
import socket

import numpy as np
import pytest

import datagen

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_rows_cover_and_balance():
    from matchinglib_poselib_b200.sharded import pairs_of_rank, shard_rows
    for n in (0, 1, 7, 100, 100_000_000):
        for world in (1, 2, 3, 8):
            spans = [shard_rows(n, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    for n_pairs in (0, 5, 4096):
        for world in (1, 2, 8):
            got = sorted(p for r in range(world) for p in pairs_of_rank(n_pairs, world, r))
            assert got == list(range(n_pairs))
    with pytest.raises(ValueError):
        shard_rows(10, 2, 2)


def _keys(idx, dist, offset):
    """The library's 64-bit record: dist << 32 | global trainIdx; missing neighbour = all ones."""
    k = (dist.astype(np.uint64) << np.uint64(32)) | (idx.astype(np.int64) + offset).astype(np.uint64)
    k[idx < 0] = np.uint64(0xFFFFFFFFFFFFFFFF)
    return k.view(np.int64)


def _worker(rank, world, port, q_path, t_path, out_path):
    import torch
    import torch.distributed as dist
    from matchinglib_poselib_b200.sharded import gather_keys, shard_rows
    from oracle import knn as orc
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    q, t = np.load(q_path), np.load(t_path)
    a, b = shard_rows(len(t), world, rank)
    idx, d = orc.knn_hamming(q, t[a:b], 2, 0)                       # stands in for mpl_knn2_hamming_dev(train_offset=a)
    parts = gather_keys(torch.from_numpy(_keys(idx, d, a)))          # [world, nq, 2]
    assert parts.shape == (world, len(q), 2)
    flat = np.sort(parts.numpy().view(np.uint64).transpose(1, 0, 2).reshape(len(q), -1), axis=1)[:, :2]   # merge_top2_kernel's rule
    if rank == 0:
        np.save(out_path, flat)
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_top2_equals_single_scan(tmp_path):
    import torch.multiprocessing as mp
    from oracle import knn as orc
    q, t = datagen.binary_descriptors(300, 1001, 32, seed=5)
    t[700] = t[10]; q[7] = t[10]                                    # cross-shard duplicate: the lower GLOBAL index must win
    np.save(tmp_path / "q.npy", q); np.save(tmp_path / "t.npy", t)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
    mp.spawn(_worker, args=(2, port, str(tmp_path / "q.npy"), str(tmp_path / "t.npy"), str(tmp_path / "out.npy")), nprocs=2, join=True)
    merged = np.load(tmp_path / "out.npy")
    idx, d = orc.knn_hamming(q, t, 2, 0)
    exp = _keys(idx, d, 0).view(np.uint64)
    assert np.array_equal(merged, exp)
    assert (merged[7] & np.uint64(0xFFFFFFFF)).tolist() == [10, 700]
