"""BASELINE config C5: database / loop-closure regime -- 10 000 ORB queries against 100 M train descriptors, the train set
sharded over the ranks (12.5 M rows = 400 MB per GPU at N = 8; at smaller N every rank still holds 12.5 M rows: weak scaling).

    python tools/c5_bench.py                               # one GPU, one 12.5 M-row shard
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/c5_bench.py

Each rank: Hamming top-2 scan of its shard (global train indices), then the all-gather over NVLink peer memory fused with the
merge (mpl_allgather_merge_top2_dev).  Planted copies (one per query, spread over the ranks) must come back as nearest
neighbours with distance 0.  Rank 0 prints one JSON line.  Synthetic data generated on the device."""
import ctypes, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import bench

NQ, ROWS, NB = 10000, 12_500_000, 32


def main():
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    g = bench.Gpu(local)
    lib, m, dev = g.lib, g.m, g.dev
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
        dist.barrier()
    gen = torch.Generator(device=dev); gen.manual_seed(1005 + rank)
    t = torch.randint(0, 256, (ROWS, NB), dtype=torch.uint8, device=dev, generator=gen)
    gq = torch.Generator(device=dev); gq.manual_seed(424242)
    q = torch.randint(0, 256, (NQ, NB), dtype=torch.uint8, device=dev, generator=gq)          # same on every rank
    # plant query i as train row (i * 1237) % ROWS of rank i % world
    mine = torch.arange(rank, NQ, world, device=dev)
    t[(mine * 1237) % ROWS] = q[mine]
    w16 = NB // 16
    scr_bytes = int(lib.mpl_knn_hamming_scratch_bytes(NQ, ROWS, w16))
    scr = torch.empty(scr_bytes, dtype=torch.uint8, device=dev)
    keys = torch.empty((NQ, 2), dtype=torch.int64, device=dev)
    idx = torch.empty((NQ, 2), dtype=torch.int32, device=dev)
    dst = torch.empty((NQ, 2), dtype=torch.int32, device=dev)
    peers = bench.setup_peers(g, dist, rank, world, NQ) if world > 1 else None

    def step():
        st = g.stream()
        m._lib.check(lib.mpl_knn2_hamming_dev(q.data_ptr(), NQ, t.data_ptr(), ROWS, w16, rank * ROWS, scr.data_ptr(), scr_bytes, keys.data_ptr(), st), "hamming")
        if peers is not None:
            m._lib.check(lib.mpl_allgather_merge_top2_dev(keys.data_ptr(), NQ, rank, world, peers[0], peers[1], 2, 0, idx.data_ptr(), dst.data_ptr(), None, st), "allgather_merge")
        else:
            m._lib.check(lib.mpl_merge_top2_dev(keys.data_ptr(), 1, NQ, 2, 0, idx.data_ptr(), dst.data_ptr(), None, st), "merge")

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    steps = 5
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for a, b in ev:
        a.record(); step(); b.record()
    torch.cuda.synchronize()
    ms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
    if dist is not None:
        tm = torch.tensor([ms], device=dev); dist.all_reduce(tm, op=dist.ReduceOp.MAX); ms = float(tm.item())
    # every query's planted copy: global index = owner * ROWS + (i * 1237) % ROWS, distance 0
    i_all = torch.arange(NQ, device=dev)
    want = (i_all % world) * ROWS + (i_all * 1237) % ROWS
    ok = bool((idx[:, 0].long() == want).all().item()) and bool((dst[:, 0] == 0).all().item()) and bool((dst[:, 1] > 0).all().item())
    if not ok and rank == 0:
        bad = (idx[:, 0].long() != want).nonzero().flatten()
        sys.stderr.write("mismatches: %d of %d; first: q=%s got=%s want=%s d0=%s d1=%s nonzero d0: %d\n" % (
            len(bad), NQ, bad[:5].tolist(), idx[bad[:5], 0].tolist(), want[bad[:5]].tolist(), dst[bad[:5], 0].tolist(), dst[bad[:5], 1].tolist(),
            int((dst[:, 0] != 0).sum().item())))
    if rank == 0:
        pairs = float(NQ) * ROWS * world
        pp = bench.pipe_peaks()
        print(json.dumps({"metric": "descriptor pairs/s (Hamming k=2, database regime C5)", "value": pairs / (ms * 1e-3), "unit": "pairs/s", "n_gpus": world,
                          "ms_per_step": ms, "scaling": "weak", "rows_per_gpu": ROWS, "queries": NQ, "bytes_per_descriptor": NB,
                          "planted_neighbours_found": ok, "popc_frac_algorithmic_per_gpu": pairs / world * (NB // 4) / (ms * 1e-3) / 1e9 / pp["popc_gops"],
                          "exchange": "NVLink peer memory inside the merge kernel" if world > 1 else "none"}))
    if not ok:
        raise SystemExit("rank %d: planted neighbours not found" % rank)


if __name__ == "__main__":
    main()
