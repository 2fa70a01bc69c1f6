"""Diagnostic: C3 RANSAC part of the bench alone (solve / score kernel times)."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench
g = bench.Gpu(0)
r = bench.bench_ransac(g, 3, 2)
print(json.dumps({k: r[k] for k in ("value", "ms_per_step", "kernels", "best_inliers")}))
p = bench.bench_pair_pipeline(g, 5)
print(json.dumps({k: p[k] for k in ("ms_per_pair", "matches", "inliers")}))
