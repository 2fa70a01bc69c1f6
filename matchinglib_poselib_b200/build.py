"""Build libmplb200.so (hand-written CUDA for sm_100a) in-tree with nvcc.

The .so is git-ignored but travels to the GPU box with the gpurun snapshot.  No JIT, no torch
extension machinery: one nvcc command, explicit arch flags (compute_100a / sm_100a).
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_DIR = os.path.join(HERE, "lib")
LIB_PATH = os.path.join(LIB_DIR, "libmplb200.so")
SOURCES = ["capi.cu", "knn_hamming.cu", "knn_hamming_static.cu", "knn_l2_exact.cu", "knn_l2_tc.cu", "ransac5pt.cu", "ransac_batch.cu", "pose_host.cu", "recover_pose.cu", "pipeline.cu", "gms.cu", "arrsac.cu", "usac.cu", "vfc.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "--fmad=false",            # fp contraction is opted into per expression (__fma_rn / fma()); parity code relies on it being off
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fno-fast-math", "-shared",
]


def _stale():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "mplb200.h"), __file__]
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force=False, verbose=False):
    """Compile every CUDA source into lib/libmplb200.so.  Returns the path."""
    if not force and not _stale():
        return LIB_PATH
    os.makedirs(LIB_DIR, exist_ok=True)
    srcs = [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    extra = os.environ.get("MPL_EXTRA_NVCC_FLAGS", "").split()          # e.g. -DMPL_TC_TRACE (per-role clock64 traces of the ranking kernel)
    cmd = ["nvcc"] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB_PATH] + srcs
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("nvcc failed building libmplb200.so")
    if verbose:
        sys.stderr.write(r.stdout + r.stderr)
    return LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
