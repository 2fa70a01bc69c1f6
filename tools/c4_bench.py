"""Config C4 on 1/2/4/8 GPUs: a batch of 4096 image pairs, AKAZE-like 61-byte descriptors 5000 x 5000 (Hamming k=2 +
ratio test, LINEAR) followed by the per-pair RANSAC 5-point (1000 device-drawn samples, refit), each pair through
mpl_match_pose_pair from HOST buffers.  Pair p runs on rank p mod G (independent units, no data-path
collective); inside a rank the pairs are submitted from --threads host threads (per-thread library contexts).

    python tools/c4_bench.py                                                      # 1 GPU
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/c4_bench.py

Prints one JSON line on rank 0 (whole-job pairs/s: all pairs / wall time between barriers, max over ranks).
The descriptor / correspondence data of a pair is drawn from a pool of --pool distinct synthetic pairs per rank
(seed 1004 + pair index mod pool); every pair gets its own RANSAC sample seed."""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
from concurrent.futures import ThreadPoolExecutor


def make_pair(seed, n=5000, nb=61):
    import datagen_pose
    from matchinglib_poselib_b200 import matchinglib
    sc = datagen_pose.make_scene(n, 0.0, seed=seed)
    rng = np.random.default_rng(seed)
    t = rng.integers(0, 256, (n, nb), dtype=np.uint8)
    q = rng.integers(0, 256, (n, nb), dtype=np.uint8)
    good = rng.random(n) < 0.5
    flips = np.packbits(rng.random((n, nb * 8)) < 0.08, axis=1)
    q[good] = t[good] ^ flips[good]
    out = []
    for a in (q, t, sc["p1"], sc["p2"]):                           # pinned host buffers, as a capture pipeline would hold
        v = matchinglib.pinned_array(a.shape, a.dtype)
        v[...] = a
        out.append(v)
    return out, sc["th"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pairs", type=int, default=4096)
    ap.add_argument("--threads", type=int, default=32)
    ap.add_argument("--pool", type=int, default=4)
    ap.add_argument("--out", default=None)
    ap.add_argument("--samples", type=int, default=1000, help="RANSAC samples per pair (0: matching only; diagnostic)")
    ap.add_argument("--refine", type=int, default=1)
    args = ap.parse_args()
    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    import torch
    import bench
    from matchinglib_poselib_b200 import poselib
    g = bench.Gpu(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=g.dev)
    pool = [make_pair(1004 + i) for i in range(args.pool)]
    mine = list(range(rank, args.pairs, world))                    # pair p -> rank p mod G

    def one(p):
        (q, t, p1, p2), th = pool[(p // world) % args.pool]
        r = poselib.matchAndEstimatePose(q, t, p1, p2, "LINEAR", True, th, args.samples, 0.999, bool(args.refine), seed=5 + p)
        return r["count"]

    sync = (lambda: dist.barrier()) if world > 1 else (lambda: None)
    with ThreadPoolExecutor(args.threads) as ex:
        list(ex.map(one, mine[:2 * args.threads]))                 # warm-up: per-thread contexts, first DMA of the pinned pages
        torch.cuda.synchronize()
        sync()
        t0 = time.perf_counter()
        counts = list(ex.map(one, mine))
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    tm = torch.tensor([dt], device=g.dev, dtype=torch.float64)
    mn = torch.tensor([min(counts)], device=g.dev)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        dist.all_reduce(mn, op=dist.ReduceOp.MIN)
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        res = dict(metric="image pairs/s (AKAZE 61 B 5000x5000 Hamming k=2 + ratio, per-pair RANSAC 5pt with refit)", value=args.pairs / float(tm.item()),
                   unit="image pairs/s", n_gpus=world, pairs=args.pairs, seconds=float(tm.item()), host_threads_per_rank=args.threads,
                   host_cores=os.cpu_count(), ransac_samples=args.samples, refit=bool(args.refine), scaling="strong", min_inliers=int(mn.item()), data="synthetic (pool of %d pairs per rank)" % args.pool,
                   api="poselib.matchAndEstimatePose -> mpl_match_pose_pair, pinned host buffers", descriptor_pairs_per_s=args.pairs * 25e6 / float(tm.item()))
        line = json.dumps(res)
        print(line)
        if args.out:
            open(args.out, "w").write(line + "\n")


if __name__ == "__main__":
    main()
