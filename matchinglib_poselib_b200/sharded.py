"""Multi-GPU host side of the hot path: one process per GPU, torch.distributed (NCCL over NVLink) for the plumbing.

The reference has no distributed layer; the two regimes BASELINE.json names are
  * database / loop closure (C5): train rows sharded in contiguous blocks over the ranks, queries replicated, each
    rank scans its shard with GLOBAL train indices (train_offset), ONE all-gather of the per-query top-2 keys
    (16 bytes x nq per rank) and a merge kernel -- (dist, index) key order makes the result identical to one scan;
  * batches of independent image pairs (C4): pair p belongs to rank p mod world, no collective.

Device memory, streams and collectives come from torch; every kernel is libmplb200's.  Keys are the library's
64-bit records (dist << 32 | trainIdx) carried in int64 tensors.
"""
import ctypes

from . import _lib


def shard_rows(n_rows, world, rank):
    """Contiguous block [start, end) of rank `rank`: start = rank * n // world."""
    if not (0 <= rank < world):
        raise ValueError("rank %d outside world %d" % (rank, world))
    return (rank * n_rows) // world, ((rank + 1) * n_rows) // world


def pairs_of_rank(n_pairs, world, rank):
    """Image pairs of a batch owned by `rank` (static round robin; all pairs have the same shape)."""
    if not (0 <= rank < world):
        raise ValueError("rank %d outside world %d" % (rank, world))
    return range(rank, n_pairs, world)


def gather_keys(keys_local, group=None):
    """All-gather of the per-rank top-2 keys: [nq, 2] int64 -> [world, nq, 2] (the merge kernel's part-major layout)."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return keys_local.reshape(1, *keys_local.shape)
    out = torch.empty((world,) + tuple(keys_local.shape), dtype=keys_local.dtype, device=keys_local.device)
    # concatenated (rank-major) output: the layout every backend takes; `out` views it as [world, nq, 2]
    dist.all_gather_into_tensor(out.view(-1, *keys_local.shape[1:]), keys_local.contiguous(), group=group)
    return out


class ShardedHammingIndex:
    """Train shard of binary descriptors resident on this rank's GPU (database regime).

    train_shard: uint8 [rows, nbytes] host array holding rows [row_start, row_start + rows) of the global train set.
    match(query) returns (idx [nq, 2] int32 with GLOBAL train indices, dist [nq, 2] int32), identical on every rank
    and identical to a single-GPU scan of the whole train set."""

    def __init__(self, train_shard, row_start, device, ignore_tail_bytes=0, group=None):
        import torch
        self.torch, self.lib, self.group, self.device = torch, _lib.load(), group, device
        _lib.check(self.lib.mpl_init(device.index if device.index is not None else 0), "mpl_init")
        nb = train_shard.shape[1]
        self.used = nb - ignore_tail_bytes
        self.nbytes = nb
        w = (self.used + 15) // 16
        self.w16 = 1
        while self.w16 < w:
            self.w16 *= 2
        self.row_start = int(row_start)
        self.nt = int(train_shard.shape[0])
        self.d_train = self._pack(train_shard)

    def _pack(self, host_rows):
        torch = self.torch
        raw = torch.from_numpy(host_rows).to(self.device)
        n = raw.shape[0]
        if self.used == self.w16 * 16 and raw.is_contiguous():
            return raw
        out = torch.empty((max(n, 1), self.w16 * 16), dtype=torch.uint8, device=self.device)
        st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        _lib.check(self.lib.mpl_pack_rows_u8_dev(raw.data_ptr(), n, raw.stride(0), self.used, out.data_ptr(), self.w16, st), "pack_rows")
        return out

    def match(self, query):
        torch, lib = self.torch, self.lib
        d_q = self._pack(query)
        nq = int(query.shape[0])
        st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        scr_bytes = int(lib.mpl_knn_hamming_scratch_bytes(nq, self.nt, self.w16))
        scr = torch.empty(max(scr_bytes, 16), dtype=torch.uint8, device=self.device)
        keys = torch.empty((nq, 2), dtype=torch.int64, device=self.device)
        _lib.check(lib.mpl_knn2_hamming_dev(d_q.data_ptr(), nq, self.d_train.data_ptr(), self.nt, self.w16, self.row_start,
                                            scr.data_ptr(), scr_bytes, keys.data_ptr(), st), "knn2_hamming_dev")
        parts = gather_keys(keys, self.group)                       # NVLink: 16 B x nq per rank
        idx = torch.empty((nq, 2), dtype=torch.int32, device=self.device)
        dist = torch.empty((nq, 2), dtype=torch.int32, device=self.device)
        _lib.check(lib.mpl_merge_top2_dev(parts.data_ptr(), parts.shape[0], nq, 2, 0, idx.data_ptr(), dist.data_ptr(), None, st), "merge_top2")
        return idx, dist
