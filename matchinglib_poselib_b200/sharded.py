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


class _PeerExchange:
    """NVLink peer-memory buffers of mpl_allgather_merge_top2_dev for up to max_queries queries: one gather buffer and one
    sync block per rank (CUDA IPC), handles exchanged once through torch.distributed.  ok is False when any rank could not
    allocate or map (no peer access): every rank then uses the NCCL all-gather + merge pair instead."""

    def __init__(self, lib, device, max_queries, group=None):
        import torch
        import torch.distributed as dist
        self.lib, self.ok, self.max_queries = lib, False, int(max_queries)
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        if self.world == 1:
            return
        ptrs, handles, err = [], [], None
        try:
            for nbytes in (int(lib.mpl_peer_gather_bytes(self.max_queries, self.world)), 256):
                p = ctypes.c_void_p(0)
                h = (ctypes.c_ubyte * 64)()
                _lib.check(lib.mpl_peer_alloc(nbytes, ctypes.byref(p), h), "mpl_peer_alloc")
                ptrs.append(p.value); handles.append(bytes(h))
        except Exception as e:                                     # every rank still reaches the collective below
            err, handles = e, None
        everyone = [None] * self.world
        dist.all_gather_object(everyone, handles, group=group)
        good = err is None and all(h is not None for h in everyone)
        self.gather = (ctypes.c_void_p * self.world)(); self.sync = (ctypes.c_void_p * self.world)()
        if good:
            try:
                for which, arr in enumerate((self.gather, self.sync)):
                    for r in range(self.world):
                        if r == self.rank:
                            arr[r] = ptrs[which]
                        else:
                            p = ctypes.c_void_p(0)
                            h = (ctypes.c_ubyte * 64).from_buffer_copy(everyone[r][which])
                            _lib.check(lib.mpl_peer_open(h, ctypes.byref(p)), "mpl_peer_open")
                            arr[r] = p.value
            except Exception:
                good = False
        flag = torch.tensor([1 if good else 0], device=device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)   # every rank uses the same exchange
        self.ok = bool(int(flag.item()))

    def merge(self, keys, nq, is_float, idx, dist_out, stream):
        """Fused all-gather + merge; False when the peers did not show up in time (caller falls back to NCCL)."""
        _lib.check(self.lib.mpl_allgather_merge_top2_dev(keys.data_ptr(), nq, self.rank, self.world, self.gather, self.sync, 2, int(is_float),
                                                         idx.data_ptr(), dist_out.data_ptr(), None, stream), "allgather_merge")
        to = ctypes.c_int(0)
        _lib.check(self.lib.mpl_peer_status(self.sync[self.rank], ctypes.byref(to), stream), "peer_status")
        return to.value == 0


class _ShardedIndex:
    """Common part of the train-sharded indices: exchange of the per-query top-2 keys and the merge."""

    def _setup(self, device, group, max_queries):
        import torch
        self.torch, self.lib, self.group, self.device = torch, _lib.load(), group, device
        _lib.check(self.lib.mpl_init(device.index if device.index is not None else 0), "mpl_init")
        self.peers = _PeerExchange(self.lib, device, max_queries, group) if device.type == "cuda" else None

    def _exchange(self, keys, nq, is_float):
        torch = self.torch
        st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        idx = torch.empty((nq, 2), dtype=torch.int32, device=self.device)
        dist = torch.empty((nq, 2), dtype=torch.float32 if is_float else torch.int32, device=self.device)
        if self.peers is not None and self.peers.ok and nq <= self.peers.max_queries:
            # one kernel: peer stores of the local keys into every rank's buffer, arrival counters, merge (NVLink, no NCCL)
            if self.peers.merge(keys, nq, is_float, idx, dist, st):
                return idx, dist
            self.peers.ok = False                                   # a peer timed out: NCCL from here on
        parts = gather_keys(keys, self.group)                       # NCCL: 16 B x nq per rank
        _lib.check(self.lib.mpl_merge_top2_dev(parts.data_ptr(), parts.shape[0], nq, 2, int(is_float), idx.data_ptr(), dist.data_ptr(), None, st), "merge_top2")
        return idx, dist


class ShardedL2Index(_ShardedIndex):
    """Train shard of float32 descriptors resident on this rank's GPU (database regime, SIFT/SURF-like descriptors).
    match(query) returns (idx [nq, 2] int32 GLOBAL train indices, dist [nq, 2] float32: squared L2 in cvflann order, or
    NMSLIB's sqrt with sqrt_out), identical on every rank and to a single-GPU scan of the whole train set."""

    def __init__(self, train_shard, row_start, device, sqrt_out=False, group=None, max_queries=32768):
        self._setup(device, group, max_queries)
        self.sqrt_out = int(bool(sqrt_out))
        self.row_start, self.nt, self.dim = int(row_start), int(train_shard.shape[0]), int(train_shard.shape[1])
        self.d_train = self.torch.from_numpy(train_shard).to(device).contiguous()

    def match(self, query):
        torch, lib = self.torch, self.lib
        d_q = torch.from_numpy(query).to(self.device).contiguous()
        nq = int(query.shape[0])
        st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        scr_bytes = int(lib.mpl_knn_l2_scratch_bytes(nq, self.nt, self.dim))
        scr = torch.empty(scr_bytes + 256, dtype=torch.uint8, device=self.device)
        base = (scr.data_ptr() + 255) & ~255
        keys = torch.empty((nq, 2), dtype=torch.int64, device=self.device)
        _lib.check(lib.mpl_knn2_l2f32_dev(d_q.data_ptr(), nq, self.dim, self.d_train.data_ptr(), self.nt, self.dim, self.dim, self.sqrt_out,
                                          self.row_start, base, scr_bytes, keys.data_ptr(), st), "knn2_l2f32_dev")
        return self._exchange(keys, nq, True)                       # the keys already carry the final distances (sqrt applied per mode)


class ShardedHammingIndex(_ShardedIndex):

    """Train shard of binary descriptors resident on this rank's GPU (database regime).

    train_shard: uint8 [rows, nbytes] host array holding rows [row_start, row_start + rows) of the global train set.
    match(query) returns (idx [nq, 2] int32 with GLOBAL train indices, dist [nq, 2] int32), identical on every rank
    and identical to a single-GPU scan of the whole train set."""

    def __init__(self, train_shard, row_start, device, ignore_tail_bytes=0, group=None, max_queries=32768):
        self._setup(device, group, max_queries)
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
        return self._exchange(keys, nq, False)
