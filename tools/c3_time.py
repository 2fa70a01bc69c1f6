"""C3 solve / score kernel times (library timers) for A/B builds: python tools/c3_time.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench
g = bench.Gpu(0)
r = bench.bench_ransac(g, 5, 3)
print(os.environ.get("TAG", ""), "ms/step %.3f" % r["ms_per_step"], "score kernel ms %.3f frac %.3f" % (r["roofline"]["kernel_ms"], r["roofline"]["frac"]))
