"""mpl_match_pose_batch throughput at several batch sizes (one rank's share of C4 at N = 1 .. 8): python tools/c4_sizes.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench
g = bench.Gpu(0)
import sys as _s
for n in ([int(a) for a in _s.argv[1:]] or [4096, 2048, 1024, 512]):
    r = bench.bench_pair_batch_c4(g, n_pairs=n)
    print(n, "pairs: %.0f pairs/s, %.1f ms" % (r["value"], r["seconds"] * 1e3), r["stages"])
