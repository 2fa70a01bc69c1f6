"""Diagnostic: time the tensor-core ranking kernel alone (library CUDA-event timers) under MPL_TC_DEBUG settings."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench
g = bench.Gpu(0)
r = bench.bench_l2(g, 10, 3)
k = r["kern"]
print("MPL_TC_DEBUG=%s  step %.4f ms  tc %.4f ms  rerank %.4f  prep %.4f" % (os.environ.get("MPL_TC_DEBUG"), sum(r["ms"]) / len(r["ms"]),
      k["l2_tc"][0] / max(1, k["l2_tc"][1]), k["l2_rerank"][0] / max(1, k["l2_rerank"][1]), k["l2_prep"][0] / max(1, k["l2_prep"][1])))
