#!/usr/bin/env python
"""bench.py -- throughput of the B200 hot path on BASELINE.json's metric.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Headline workload (config C2 of BASELINE.json): SIFT-like 128-d float descriptors, 20 000 queries x
20 000 train rows, brute-force L2 k = 2 + Lowe ratio test.  A "step" is one pass of the hot path over
that batch.  `value` = descriptor pairs/s with the inputs resident in HBM; `e2e` = the same metric through
the reference-facing call (getMatches -> mpl_get_matches_api) with HOST buffers, copies inside the timed
region.  `parts` carries the other two legs of the metric on their own configs: Hamming k-NN (C1, ORB
8192 x 8192 x 32 B) and RANSAC 5-point scoring (C3, 10 000 correspondences x 100 000 pre-drawn samples).

N > 1 (torchrun, one rank per GPU): the database regime -- every rank holds its own 20 000-row train
shard, the queries are replicated, per-rank top-2 keys are exchanged over NVLink peer memory inside the merge
kernel (weak scaling: pairs grow with N).

At EVERY N the line also carries the two workloads BASELINE.json names for multi-GPU:
  parts.database_c5    10 000 ORB queries x 12.5 M train rows PER RANK (100 M on 8 GPUs), Hamming top-2 per shard +
                       peer-memory all-gather fused with the merge; planted copies + an oracle spot check (weak scaling)
  parts.pair_batch_c4  4096 image pairs (AKAZE-like 61 B, 5000 x 5000, + per-pair RANSAC 1000 samples, refit), pair p on
                       rank p mod N, each rank ONE host thread calling mpl_match_pose_batch from pinned host buffers
                       (strong scaling)

--impl reference: the reference's CPU path on the host cores (oracle/_ref = the reference's vendored
NMSLIB compiled from its own sources; falls back to the oracle port), bounded sample per step.

Only this file, tests/ and __graft_entry__.smoke() may touch oracle/.
"""
import argparse
import ctypes
import json
import os
import sys
import threading
import time

# Batches of image pairs run one host thread per pair in flight, each on its own CUDA stream: give the driver enough
# hardware queues (default 8) that those streams do not alias.  Must be set before CUDA initialises.
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

C2 = dict(nq=20000, nt=20000, dim=128, seed=1002)
C1 = dict(nq=8192, nt=8192, nbytes=32, seed=1001)
C3 = dict(n=10000, h=100000, seed=1003)
L2_FLUSH_BYTES = 512 << 20       # > 126 MB L2: written between timed iterations
# integer-ALU instructions (LOP3 + IADD3 + LEA + VIMNMX) per descriptor pair in the SASS inner loop of hamming_top2_kernel<W16>
# (cuobjdump -sass, bytes per descriptor -> count): the executed-work side of the Hamming roofline
HAMMING_ALU_PER_PAIR = {32: 19.0}        # 15 LOP3 + 3 VIMNMX + 1 IADD3 (5 POPC and 3 IMAD go to other pipes)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        d["_src"] = "measured"
        return d
    return dict(hbm_gbs=6650.0, bf16_tflops=1590.0, bf16_tflops_sustained=1400.0, sm_max_mhz=1965.0, _src="fallback")


def pipe_peaks():
    """Per-pipe issue rates measured on a B200 of this pool with tools/microbench.cu (profiles/r1_microbench.json):
    roofline denominators where MEASURED_PEAKS.json has none (integer POPC, fp64)."""
    for name in ("r2_microbench.json", "r1_microbench.json"):
        p = os.path.join(ROOT, "profiles", name)
        if os.path.exists(p):
            d = json.load(open(p))
            d["_src"] = "measured: tools/microbench.cu -> profiles/%s" % name
            return d
    # CUDA programming guide throughput table x 148 SMs x 1.965 GHz
    return dict(popc_gops=4650.0, dfma_ginst=18600.0, ddiv_rn_gops=1300.0, _src="fallback: programming-guide rates")


def ncu_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed `ncu --set full` capture of this
    bench command (profiles/r1_ncu_traffic.json); None when there is no capture for the kernel."""
    for name in ("r2_ncu_traffic.json", "r1_ncu_traffic.json"):
        p = os.path.join(ROOT, "profiles", name)
        if os.path.exists(p):
            v = json.load(open(p)).get(kernel, {}).get("dram_bytes_per_launch")
            if v is not None:
                return v
    return None


# ------------------------------------------------------------------------------------------------
# clocks during the timed region (NVML; same fields as the profiling guide's nvidia-smi line)
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    def __init__(self, index=0, period=0.05):
        self.index, self.period, self.samples, self.reasons = index, period, [], set()
        self.max_mhz, self._stop, self._t = None, threading.Event(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {nv.nvmlClocksEventReasonHwSlowdown: "hw_slowdown", nv.nvmlClocksEventReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksEventReasonSwThermalSlowdown: "sw_thermal_slowdown", nv.nvmlClocksEventReasonSwPowerCap: "sw_power_cap"} \
            if hasattr(nv, "nvmlClocksEventReasonHwSlowdown") else \
                {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown", nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown", nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(self.period)

    def __enter__(self):
        if self.nv is not None:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._t:
            self._t.join()

    def summary(self):
        if not self.samples:
            return dict(sm_mhz=None, sm_max_mhz=self.max_mhz, reasons=sorted(self.reasons), samples=0)
        return dict(sm_mhz=float(np.median(self.samples)), sm_max_mhz=self.max_mhz, reasons=sorted(self.reasons), samples=len(self.samples))


# ------------------------------------------------------------------------------------------------
# device-side helpers
# ------------------------------------------------------------------------------------------------
class Gpu:
    """torch supplies device memory, streams and events; every kernel is libmplb200's."""

    def __init__(self, local_rank=0):
        import torch
        import matchinglib_poselib_b200 as m
        self.torch, self.m = torch, m
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a B200: no CUDA device visible and there is no CPU fallback")
        torch.cuda.set_device(local_rank)
        self.dev = torch.device("cuda", local_rank)
        self.lib = m._lib.load()
        m._lib.check(self.lib.mpl_init(local_rank), "mpl_init")
        self.flush_buf = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device=self.dev)

    def stream(self):
        return ctypes.c_void_p(self.torch.cuda.current_stream().cuda_stream)

    def flush_l2(self):
        self.flush_buf.zero_()

    def launches(self):
        return int(self.lib.mpl_kernel_launch_count())

    def prof_read(self, cls):
        ms, n = ctypes.c_double(0), ctypes.c_longlong(0)
        self.lib.mpl_prof_read(cls, ctypes.byref(ms), ctypes.byref(n))
        return ms.value, n.value

    def time_steps(self, step, steps, warmup, barrier=None):
        """W warm-up steps, then K timed steps.  Each timed step is bracketed by CUDA events on the launching
        stream; L2 is flushed between steps (outside the events).  Returns per-step ms list."""
        torch = self.torch
        for _ in range(warmup):
            self.flush_l2()
            step()
        torch.cuda.synchronize()
        if barrier:
            barrier()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        for a, b in ev:
            self.flush_l2()
            a.record()
            step()
            b.record()
        torch.cuda.synchronize()
        if barrier:
            barrier()
        return [a.elapsed_time(b) for a, b in ev]


PROF = dict(hamming=0, l2_exact=1, l2_tc=2, l2_rerank=3, merge=4, solve=5, score=6, l2_prep=7)


def pinned(arr, torch):
    """Pinned host copy of `arr`: (torch tensor, numpy view), in the library's huge-page pinned memory (mpl_host_alloc:
    2 MiB pages + cudaHostRegister; the copy then runs at the link rate, 52 GB/s measured with tools/pin_probe.py, where
    torch.pin_memory()'s 4 KiB-page allocations gave 25-40 GB/s depending on how fragmented the process already was)."""
    arr = np.ascontiguousarray(arr)
    from matchinglib_poselib_b200 import matchinglib
    view = matchinglib.pinned_array(arr.shape, arr.dtype)
    view[...] = arr
    return torch.from_numpy(view), view


# ------------------------------------------------------------------------------------------------
# C2: float L2 k-NN (headline)
# ------------------------------------------------------------------------------------------------
def make_c2(rank=0):
    """Queries are identical on every rank (a quarter of them are noisy copies of rows of shard 0, so the ratio
    test passes there); train shard `rank` is its own seeded 20 000 rows."""
    import datagen

    def shard(r):
        return datagen.sift_like(16, C2["nt"], C2["dim"], seed=C2["seed"] + 7919 * (r + 1))[1]
    q, _ = datagen.sift_like(C2["nq"], 16, C2["dim"], seed=C2["seed"])
    t0 = shard(0)
    rng = np.random.default_rng(C2["seed"])
    sel = rng.choice(C2["nq"], C2["nq"] // 4, replace=False)
    src = rng.integers(0, C2["nt"], len(sel))
    noise = rng.standard_normal((len(sel), C2["dim"])) * (0.05 * 512.0 / np.sqrt(C2["dim"]))
    q[sel] = np.clip(np.round(t0[src] + noise), 0, 255).astype(np.float32)
    return q, (t0 if rank == 0 else shard(rank))


def setup_peers(g, dist, rank, world, nq):
    """One gather buffer + one sync block per rank in CUDA IPC memory, mapped into every rank: (peer_gather[], peer_sync[]).
    Every rank reaches the collective whatever happens locally; raises afterwards if any step failed."""
    lib, m = g.lib, g.m
    ptrs, handles, err = [], [], None
    try:
        for nbytes in (int(lib.mpl_peer_gather_bytes(nq, world)), 256):
            p = ctypes.c_void_p(0)
            h = (ctypes.c_ubyte * 64)()
            m._lib.check(lib.mpl_peer_alloc(nbytes, ctypes.byref(p), h), "mpl_peer_alloc")
            ptrs.append(p.value); handles.append(bytes(h))
    except Exception as e:
        err, handles = e, None
    everyone = [None] * world
    dist.all_gather_object(everyone, handles)
    if err is not None or any(h is None for h in everyone):
        raise RuntimeError("peer allocation failed on some rank (%s)" % err)
    out = []
    for which in range(2):
        arr = (ctypes.c_void_p * world)()
        for r in range(world):
            if r == rank:
                arr[r] = ptrs[which]
            else:
                p = ctypes.c_void_p(0)
                h = (ctypes.c_ubyte * 64).from_buffer_copy(everyone[r][which])
                m._lib.check(lib.mpl_peer_open(h, ctypes.byref(p)), "mpl_peer_open")
                arr[r] = p.value
        out.append(arr)
    return out


def bench_l2(g, steps, warmup, world=1, rank=0, dist=None):
    torch, lib, m = g.torch, g.lib, g.m
    q, t = make_c2(rank)
    nq, nt, dim = q.shape[0], t.shape[0], q.shape[1]
    d_q, d_t = torch.from_numpy(q).to(g.dev), torch.from_numpy(t).to(g.dev)
    scr_bytes = int(lib.mpl_knn_l2_scratch_bytes(nq, nt, dim))
    d_scr = torch.empty(scr_bytes, dtype=torch.uint8, device=g.dev)
    d_keys = torch.empty((nq, 2), dtype=torch.int64, device=g.dev)
    d_all = torch.empty((world, nq, 2), dtype=torch.int64, device=g.dev) if world > 1 else d_keys
    d_idx = torch.empty((nq, 2), dtype=torch.int32, device=g.dev)
    d_dist = torch.empty((nq, 2), dtype=torch.float32, device=g.dev)

    # N > 1: the per-rank top-2 keys are exchanged through NVLink peer memory inside the merge kernel itself
    # (mpl_allgather_merge_top2_dev); MPL_BENCH_NCCL=1 selects the NCCL all-gather + merge pair instead.
    peers = None
    if world > 1 and os.environ.get("MPL_BENCH_NCCL") is None:
        try:
            peers = setup_peers(g, dist, rank, world, nq)
        except Exception as e:                                     # no peer access between some pair of GPUs: NCCL path
            sys.stderr.write("bench: rank %d: peer memory setup failed (%s); using the NCCL all-gather\n" % (rank, e))
            peers = None
        ok = torch.tensor([1 if peers is not None else 0], device=g.dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)                  # every rank uses the same exchange
        if int(ok.item()) == 0:
            peers = None

    def step():
        st = g.stream()
        m._lib.check(lib.mpl_knn2_l2f32_dev(d_q.data_ptr(), nq, dim, d_t.data_ptr(), nt, dim, dim, 0, rank * nt,
                                            d_scr.data_ptr(), scr_bytes, d_keys.data_ptr(), st), "knn2_l2f32_dev")
        if peers is not None:
            m._lib.check(lib.mpl_allgather_merge_top2_dev(d_keys.data_ptr(), nq, rank, world, peers[0], peers[1], 2, 1,
                                                          d_idx.data_ptr(), d_dist.data_ptr(), None, st), "allgather_merge")
            return
        if world > 1:
            dist.all_gather_into_tensor(d_all.view(-1, 2), d_keys)   # NCCL over NVLink: 16 B x nq per rank
        m._lib.check(lib.mpl_merge_top2_dev(d_all.data_ptr(), world, nq, 2, 1, d_idx.data_ptr(), d_dist.data_ptr(), None, st), "merge")

    lib.mpl_prof_enable(0)
    barrier = (lambda: dist.barrier()) if world > 1 else None
    # The step (memset + 7 kernels of the library, all enqueued on the caller's stream, + the all-gather at N > 1) is
    # captured once into a CUDA graph and replayed -- same kernels, same arguments, no per-launch host gaps.
    run = step
    graphed = False
    if os.environ.get("MPL_BENCH_NO_GRAPH") is None:                # N > 1: the NCCL all-gather is captured with the kernels
        try:
            step()
            lc = g.launches()
            step()
            per_step = g.launches() - lc                           # replays do not pass through the library's counter
            torch.cuda.synchronize()
            cg = torch.cuda.CUDAGraph()
            with torch.cuda.graph(cg):
                step()
            torch.cuda.synchronize()
            run, graphed = cg.replay, True
        except Exception as e:                                     # capture not possible: time the plain launches
            sys.stderr.write("bench: CUDA graph capture failed (%s); timing plain launches\n" % e)
            run, graphed = step, False
    if world > 1:                                                  # every rank replays, or none does
        flag = torch.tensor([1 if graphed else 0], device=g.dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 0:
            run, graphed = step, False
    l0 = g.launches()
    with ClockSampler(g.dev.index) as cs:
        ms = g.time_steps(run, steps, warmup, barrier)
        launches = g.launches() - l0                              # warm-up + timed steps only
        # the timed region lasts a few ms: keep the same work running until NVML has sampled it.  Every rank must issue
        # the SAME number of steps (they contain a collective), so the padding is a fixed count, not a deadline.
        for _ in range(600 if world > 1 else 0):
            run()
        t_end = time.time() + 1.0
        while world == 1 and len(cs.samples) < 8 and time.time() < t_end:
            run()
        torch.cuda.synchronize()
    if peers is not None:                                       # the peer-memory exchange must give what NCCL + merge gives
        step()
        torch.cuda.synchronize()
        got_i, got_d = d_idx.clone(), d_dist.clone()
        dist.all_gather_into_tensor(d_all.view(-1, 2), d_keys)
        m._lib.check(lib.mpl_merge_top2_dev(d_all.data_ptr(), world, nq, 2, 1, d_idx.data_ptr(), d_dist.data_ptr(), None, g.stream()), "merge")
        torch.cuda.synchronize()
        if not (torch.equal(got_i, d_idx) and torch.equal(got_d.view(torch.int32), d_dist.view(torch.int32))):
            raise RuntimeError("rank %d: peer-memory all-gather + merge differs from NCCL all-gather + merge" % rank)
    e2e_sharded = None
    if world > 1:
        e2e_sharded = e2e_l2_sharded(g, dist, world, q, t, d_q, d_t, d_idx, d_dist, step, max(5, min(steps, 20)))
    launches = launches * steps // (steps + warmup)             # kernels launched inside the K timed steps
    if graphed:
        launches = per_step * steps
    # dominant kernel timed live with the library's own CUDA events (same stream), separate short pass
    lib.mpl_prof_enable(1)
    for _ in range(max(3, min(steps, 10))):
        g.flush_l2()
        step()
    torch.cuda.synchronize()
    lib.mpl_prof_enable(0)
    kern = {k: g.prof_read(v) for k, v in PROF.items()}
    res = dict(ms=ms, launches_per_step=None, kern=kern, clocks=cs.summary(), nq=nq, nt=nt, dim=dim,
               idx=d_idx.cpu().numpy(), dist=d_dist.cpu().numpy(), q=q, t=t)
    res["launches_timed"] = launches
    res["graphed"] = graphed
    res["peer_exchange"] = peers is not None
    res["e2e_sharded"] = e2e_sharded
    return res


def e2e_l2_sharded(g, dist, world, q, t, d_q, d_t, d_idx, d_dist, step, reps):
    """The sharded database call end to end from HOST buffers, on every rank at once: upload of the queries and of this
    rank's train shard from pinned host memory, local scan, top-2 exchange + merge (`step`, the timed step of `value`),
    download of the merged (idx, dist) and the ratio test / DMatch emission (mpl_ratio_matches_f32).  Wall clock per
    repetition between barriers, max over ranks; returns (median seconds, matches, h2d bytes, d2h bytes) per rank."""
    torch, lib, m = g.torch, g.lib, g.m
    (tq, _), (tt, _) = pinned(q, torch), pinned(t, torch)
    nq = q.shape[0]
    (t_idx, h_idx), (t_dist, h_dist) = pinned(np.zeros((nq, 2), np.int32), torch), pinned(np.zeros((nq, 2), np.float32), torch)
    out = np.zeros(nq * 4, np.int32)                               # mpl_dmatch: 16 bytes
    n_out = ctypes.c_int(0)
    st = torch.cuda.current_stream()

    def call():
        d_q.copy_(tq, non_blocking=True)
        d_t.copy_(tt, non_blocking=True)
        step()
        t_idx.copy_(d_idx, non_blocking=True)
        t_dist.copy_(d_dist, non_blocking=True)
        st.synchronize()
        m._lib.check(lib.mpl_ratio_matches_f32(h_idx.ctypes.data, h_dist.ctypes.data, nq, 2, 1, out.ctypes.data, ctypes.byref(n_out)), "ratio")

    for _ in range(5):
        call()
    ts = []
    for _ in range(reps):
        g.flush_l2()
        torch.cuda.synchronize()
        dist.barrier()
        t0 = time.perf_counter()
        call()
        ts.append(time.perf_counter() - t0)
    tm = torch.tensor(ts, device=g.dev, dtype=torch.float64)
    dist.all_reduce(tm, op=dist.ReduceOp.MAX)                      # slowest rank of each repetition
    return float(tm.median().item()), int(n_out.value), int(q.nbytes + t.nbytes), int(nq * 16)


def e2e_l2(g, q, t, reps):
    """getMatches("LINEAR") through the C ABI with pinned HOST buffers: H2D + kernels + D2H + ratio test."""
    torch, m = g.torch, g.m
    (_, hq), (_, ht) = pinned(q, torch), pinned(t, torch)
    kp1, kp2 = range(len(hq)), range(len(ht))
    for _ in range(10):                                                             # warm-up: allocations, first DMA of new pinned pages
        err, mt = m.getMatchesArray(kp1, kp2, hq, ht, None, "LINEAR", False, True)
        assert err == 0, err
    ts = []
    for _ in range(reps):
        g.flush_l2()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        err, mt = m.getMatchesArray(kp1, kp2, hq, ht, None, "LINEAR", False, True)
        ts.append(time.perf_counter() - t0)
    return float(np.median(ts)), len(mt), hq.nbytes + ht.nbytes, len(hq) * 2 * 8


_FP8_PEAK = {}


def measure_fp8_peak(torch, dev):
    """cuBLASLt e4m3 x e4m3 GEMM rate, measured live (MEASURED_PEAKS.json holds bf16 only): 8192^3 through torch._scaled_mm, best of 10
    (burst) and the mean of a back-to-back train (sustained).  Fallback: twice the measured bf16 rates (the nominal fp8 : bf16 ratio)."""
    if "v" in _FP8_PEAK:
        return _FP8_PEAK["v"]
    pk = peaks()
    out = dict(burst=2.0 * pk["bf16_tflops"], sustained=2.0 * pk.get("bf16_tflops_sustained", pk["bf16_tflops"]),
               src="fallback: 2 x the measured cuBLAS bf16 rates (%s)" % pk["_src"])
    try:
        n = 8192
        a = torch.randn(n, n, device=dev).to(torch.float8_e4m3fn); b = torch.randn(n, n, device=dev).to(torch.float8_e4m3fn).t()
        one = torch.ones((), device=dev)
        f = lambda: torch._scaled_mm(a, b, scale_a=one, scale_b=one, out_dtype=torch.bfloat16)
        for _ in range(3):
            f()
        best = 1e9
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); f(); e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(200):
            f()
        e1.record(); torch.cuda.synchronize()
        out = dict(burst=2.0 * n ** 3 / (best * 1e-3) / 1e12, sustained=2.0 * n ** 3 * 200 / (e0.elapsed_time(e1) * 1e-3) / 1e12,
                   src="measured live: torch._scaled_mm e4m3 8192^3 (cuBLASLt), best of 10 (burst) / 200 back to back (sustained), random data")
        del a, b
    except Exception as e:
        sys.stderr.write("bench: fp8 peak measurement failed (%s); using %s\n" % (e, out["src"]))
    _FP8_PEAK["v"] = out
    return out


def hamming_kernel_roofline(g, nb, pairs, long_step, per="per GPU"):
    """Roofline of the Hamming kernel the profiled step(s) just ran, from the library's per-class CUDA-event timers: the tensor-core
    path (bits as e4m3 bytes on l2_tc_rank_kernel: 2 * bits flop per pair against the cuBLASLt e4m3 rate measured live) or the popc scan
    (executed alu-pipe instructions against the measured issue rate).  Returns (roofline dict, kernel ms, path)."""
    tc_ms, tc_n = g.prof_read(PROF["l2_tc"]); pr_ms, pr_n = g.prof_read(PROF["l2_prep"]); rr_ms, rr_n = g.prof_read(PROF["l2_rerank"])
    k_ms, k_n = g.prof_read(PROF["hamming"])
    pk, pp = peaks(), pipe_peaks()
    if tc_n > 0:
        bits = ((nb + 15) // 16) * 128
        flops = 2.0 * bits * pairs
        f8 = measure_fp8_peak(g.torch, g.dev)
        # denominator: the BURST cuBLASLt e4m3 rate on random data, also for the long C5 step -- on 0 / 1 operands the power-limited
        # tensor pipe can run above the sustained rate of a random-data GEMM, so the sustained figure is no ceiling here (shown beside)
        peak = f8["burst"]
        ach = flops / (tc_ms * 1e-3) / 1e12
        return dict(bound="tensor (fp8)", achieved=ach, peak=peak, unit="TFLOP/s " + per, frac=ach / peak,
                    frac_of_sustained=(ach / f8["sustained"]) if long_step else None, peak_sustained=f8["sustained"],
                    kernel="l2_tc_rank_kernel (tcgen05 kind::f8f6f4 on bits unpacked to e4m3 bytes, exact-integer mode)", kernel_ms=tc_ms, rank_launches=int(tc_n),
                    prep_ms=pr_ms, certify_ms=rr_ms, flops_per_step=flops, traffic=ncu_traffic("l2_tc_rank_kernel_hamming"), peak_src=f8["src"],
                    note="hamming = |q| + |t| - 2 q.t: 2 x bits flop per pair; scores are exact integers, keys identical to the popc kernel's"), tc_ms, "tensor cores"
    kern_ms = k_ms / max(k_n, 1)
    alu = HAMMING_ALU_PER_PAIR[nb]
    # SURVEY 8(d) counts ceil(B/4) POPC per pair (the reference's scalar form); the kernel folds 8 XOR words through 3 carry-save
    # adders and EXECUTES 5 POPC + 19 alu-pipe instructions (15 LOP3 + 3 VIMNMX + 1 IADD3) per pair: the binding pipe is the alu
    return dict(bound="int alu issue (executed LOP3 + VIMNMX + IADD3 per pair)", achieved=pairs * alu / (kern_ms * 1e-3) / 1e9, peak=pp.get("shf_funnel_ginst"),
                unit="G alu instr/s " + per, frac=(pairs * alu / (kern_ms * 1e-3) / 1e9 / pp["shf_funnel_ginst"]) if "shf_funnel_ginst" in pp else None,
                alu_instr_per_pair=alu, popc_executed_per_pair=5, popc_frac_executed=pairs * 5 / (kern_ms * 1e-3) / 1e9 / pp["popc_gops"],
                popc_frac_algorithmic_8_per_pair=pairs * (nb // 4) / (kern_ms * 1e-3) / 1e9 / pp["popc_gops"], traffic=ncu_traffic("hamming_top2_kernel"),
                kernel="hamming_top2_kernel<%d>" % ((nb + 15) // 16), kernel_ms=kern_ms, peak_src=pp["_src"]), kern_ms, "popc scan"


# ------------------------------------------------------------------------------------------------
# C1: Hamming k-NN (ORB 8192 x 8192 x 32 B)
# ------------------------------------------------------------------------------------------------
def bench_hamming(g, steps, warmup):
    import datagen
    torch, lib, m = g.torch, g.lib, g.m
    q, t = datagen.binary_descriptors(C1["nq"], C1["nt"], C1["nbytes"], seed=C1["seed"])
    nq, nt, nb = q.shape[0], t.shape[0], q.shape[1]
    w16 = nb // 16
    d_q, d_t = torch.from_numpy(q).to(g.dev), torch.from_numpy(t).to(g.dev)
    scr_bytes = int(lib.mpl_knn_hamming_scratch_bytes(nq, nt, w16))
    d_scr = torch.empty(scr_bytes, dtype=torch.uint8, device=g.dev)
    d_keys = torch.empty((nq, 2), dtype=torch.int64, device=g.dev)
    d_idx = torch.empty((nq, 2), dtype=torch.int32, device=g.dev)
    d_dist = torch.empty((nq, 2), dtype=torch.int32, device=g.dev)

    def step():
        st = g.stream()
        m._lib.check(lib.mpl_knn2_hamming_dev(d_q.data_ptr(), nq, d_t.data_ptr(), nt, w16, 0, d_scr.data_ptr(), scr_bytes,
                                              d_keys.data_ptr(), st), "knn2_hamming_dev")
        m._lib.check(lib.mpl_merge_top2_dev(d_keys.data_ptr(), 1, nq, 2, 0, d_idx.data_ptr(), d_dist.data_ptr(), None, st), "merge")

    pairs = float(nq) * nt

    def measure():
        ms = g.time_steps(step, steps, warmup)
        for cls in ("hamming", "merge", "l2_tc", "l2_prep", "l2_rerank"):
            g.prof_read(PROF[cls])
        lib.mpl_prof_enable(1)
        g.flush_l2(); step()
        torch.cuda.synchronize()
        lib.mpl_prof_enable(0)
        roof, kern_ms, path = hamming_kernel_roofline(g, nb, pairs, long_step=False, per="")
        g.prof_read(PROF["merge"])
        return float(np.mean(ms)), roof, path

    step_ms, roof, path = measure()                              # the library's default path (auto rule: tensor cores at this size)
    lib.mpl_set_hamming_path(1)                                  # the popc scan beside it
    try:
        popc_ms, popc_roof, _ = measure()
    finally:
        lib.mpl_set_hamming_path(0)
    # end to end through getMatches("LINEAR") with pinned host buffers
    (_, hq), (_, ht) = pinned(q, torch), pinned(t, torch)
    kp = range(nq)
    err, mt = m.getMatchesArray(kp, range(nt), hq, ht, None, "LINEAR", False, True)
    assert err == 0, err
    ts = []
    for _ in range(max(5, min(steps, 20))):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        err, mt = m.getMatchesArray(kp, range(nt), hq, ht, None, "LINEAR", False, True)
        ts.append(time.perf_counter() - t0)
    return dict(
        workload="C1: ORB 8192x8192x32B Hamming k=2 + ratio (LINEAR)", value=pairs / (step_ms * 1e-3), unit="pairs/s", ms_per_step=step_ms,
        dtype="bits as e4m3 0/1 on tcgen05 (exact integers)" if path == "tensor cores" else "u32 popc", path=path,
        roofline=roof,
        popc_path=dict(value=pairs / (popc_ms * 1e-3), unit="pairs/s", ms_per_step=popc_ms, roofline=popc_roof,
                       note="mpl_set_hamming_path(1): the carry-save popc scan, same keys bit for bit"),
        cpu_baseline=cpu_hamming_c1(q, t),
        e2e=dict(value=pairs / float(np.median(ts)), unit="pairs/s", h2d_bytes_per_step=int(hq.nbytes + ht.nbytes), d2h_bytes_per_step=int(nq * 16),
                 matches=int(len(mt))))


# ------------------------------------------------------------------------------------------------
# C4: batch of image pairs (AKAZE-like 61-byte descriptors 5000 x 5000 + per-pair RANSAC, 1000 samples, refit) through
# mpl_match_pose_batch: ONE host thread per rank, pinned host buffers in, (matches, E, mask) out, everything in between on
# the device.  Pair p runs on rank p mod N (independent units, no data-path collective): strong scaling.
# ------------------------------------------------------------------------------------------------
C4 = dict(pairs=4096, n=5000, nbytes=61, samples=1000, seed=1004, pool=4)


def make_c4_pair(seed, n=None, nb=None, pin=True):
    import datagen_pose
    from matchinglib_poselib_b200 import matchinglib
    n = n or C4["n"]; nb = nb or C4["nbytes"]
    sc = datagen_pose.make_scene(n, 0.0, seed=seed)
    rng = np.random.default_rng(seed)
    t = rng.integers(0, 256, (n, nb), dtype=np.uint8)
    q = rng.integers(0, 256, (n, nb), dtype=np.uint8)
    good = rng.random(n) < 0.5                                   # half of the queries are noisy copies of "their" train row
    flips = np.packbits(rng.random((n, nb * 8)) < 0.08, axis=1)
    q[good] = t[good] ^ flips[good]
    arrs = [q, t, np.ascontiguousarray(sc["p1"]), np.ascontiguousarray(sc["p2"])]
    if pin:                                                      # pinned host buffers, as a capture pipeline would hold them
        out = []
        for a in arrs:
            v = matchinglib.pinned_array(a.shape, a.dtype)
            v[...] = a
            out.append(v)
        arrs = out
    return arrs, sc["th"]


def bench_pair_batch_c4(g, world=1, rank=0, dist=None, n_pairs=None):
    from matchinglib_poselib_b200 import poselib
    torch = g.torch
    n_pairs = n_pairs or C4["pairs"]
    pool = [make_c4_pair(C4["seed"] + i) for i in range(C4["pool"])]
    th = pool[0][1]
    mine = list(range(rank, n_pairs, world))                    # pair p -> rank p mod N
    pairs = [tuple(pool[(p // world) % C4["pool"]][0]) for p in mine]
    seeds = [5 + p for p in mine]
    h = C4["samples"]
    sync = (lambda: dist.barrier()) if world > 1 else (lambda: None)
    # unit latency: one pair through the same call (a batch of one)
    lat = []
    for i in range(7):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        one = poselib.matchAndEstimatePoseBatch(pairs[:1], "LINEAR", True, th, h, 0.999, True, seeds=seeds[:1], want_samples=(i == 0))
        lat.append(time.perf_counter() - t0)
    unit = one[0]
    first = poselib.matchAndEstimatePoseBatch(pairs[:1], "LINEAR", True, th, h, 0.999, True, seeds=seeds[:1], want_samples=True)[0]
    # the batch is marshalled once (ctypes pair table); the timed region is exactly one mpl_match_pose_batch call per rank
    full = poselib.PairBatch(pairs, "LINEAR", True, th, h, 0.999, True, seeds=seeds)
    b_match = poselib.PairBatch(pairs, "LINEAR", True, th, 0, 0.999, False, seeds=seeds)
    b_norefit = poselib.PairBatch(pairs, "LINEAR", True, th, h, 0.999, False, seeds=seeds)
    poselib.PairBatch(pairs[:min(len(pairs), 1024)], "LINEAR", True, th, h, 0.999, True, seeds=seeds[:1024]).run()   # warm-up: full-size arenas of all three lanes (4 sub-batches of 256), first DMA of the pinned pages
    l0 = g.launches()
    torch.cuda.synchronize()
    sync()
    t0 = time.perf_counter()
    full.run()
    dt = time.perf_counter() - t0
    launches = g.launches() - l0
    res = full.results()
    # matching only (h = 0) and matching + RANSAC without the refit, same batch: where the time goes
    sync(); t0 = time.perf_counter(); b_match.run(); dt_match = time.perf_counter() - t0
    sync(); t0 = time.perf_counter(); b_norefit.run(); dt_norefit = time.perf_counter() - t0
    tm = torch.tensor([dt, dt_match, dt_norefit], device=g.dev, dtype=torch.float64)
    mn = torch.tensor([min(r["count"] for r in res)], device=g.dev)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)                # slowest rank
        dist.all_reduce(mn, op=dist.ReduceOp.MIN)
    dt, dt_match, dt_norefit = (float(x) for x in tm.tolist())
    out = None
    if rank == 0:
        # oracle check of one pair of the batch (same device-drawn samples): matches bit-exact, count / iterations / mask exact
        from oracle import knn as oknn, ransac as oransac
        q, t, p1, p2 = (np.asarray(a) for a in pairs[0])
        idx, dd = oknn.knn_hamming(q, t, 2, 0, threads=os.cpu_count() or 1)
        t_cpu0 = time.perf_counter()
        idx, dd = oknn.knn_hamming(q, t, 2, 0, threads=1)
        omt = oknn.ratio_matches(idx, dd, True)
        c1 = np.ascontiguousarray(p1[omt["queryIdx"]]); c2 = np.ascontiguousarray(p2[omt["trainIdx"]])
        exp = oransac.ransac5pt(c1, c2, first["samples"], th, max_iters=h, refine=True)
        cpu_pair_s = time.perf_counter() - t_cpu0
        ok = bool(np.array_equal(first["matches"], omt) and first["count"] == exp["count"] and first["iters"] == exp["iters"]
                  and np.array_equal(first["mask"], exp["mask"]))
        if not ok:
            raise RuntimeError("C4: pair 0 differs from the oracle chain")
        npair_bytes = int(sum(np.asarray(a).nbytes for a in pairs[0]))
        out = dict(workload="C4: %d image pairs, AKAZE-like 61 B x 5000 x 5000 Hamming k=2 + ratio (LINEAR), per-pair RANSAC 5pt (%d device-drawn samples, "
                            "adaptive stop, refit); pair p -> rank p mod N; one host thread per rank, mpl_match_pose_batch from pinned host buffers" % (n_pairs, h),
                   value=n_pairs / dt, unit="image pairs/s", n_gpus=world, scaling="strong", seconds=dt, pairs=n_pairs,
                   e2e=dict(value=n_pairs / dt, unit="image pairs/s", h2d_bytes_per_step=npair_bytes * n_pairs,
                            d2h_bytes_per_step=int(n_pairs * (C4["n"] * 17 + 136)), note="the batch call IS the end-to-end path: host buffers in, host results out"),
                   descriptor_pairs_per_s=n_pairs * float(C4["n"]) ** 2 / dt,
                   stages=dict(matching_only_pairs_per_s=n_pairs / dt_match, match_ransac_norefit_pairs_per_s=n_pairs / dt_norefit),
                   unit_latency_ms=float(np.median(lat[2:]) * 1e3), unit_matches=int(len(unit["matches"])), unit_inliers=int(unit["count"]),
                   unit_ransac_iterations=int(unit["iters"]), min_inliers=int(mn.item()), gpu_launches=int(launches),
                   host_threads_per_rank=1, host_cores=os.cpu_count(), oracle_check="pair 0: matches bit-exact, count/iterations/mask exact",
                   cpu_baseline=dict(value=1.0 / cpu_pair_s, unit="image pairs/s", cores=1, kind="port",
                                     sample="one pair: oracle cvflann-order Hamming scan 5000x5000x61 B + ratio + RANSAC 5pt (adaptive stop, refit), single thread as the reference runs it"),
                   data="synthetic (pool of %d distinct pairs per rank, every pair its own sample seed)" % C4["pool"])
    return out


# ------------------------------------------------------------------------------------------------
# C5: database / loop-closure regime -- 10 000 ORB queries against 12.5 M train rows PER RANK (100 M on 8 GPUs; weak scaling),
# Hamming top-2 per shard with global indices, then the all-gather over NVLink peer memory fused with the merge.
# ------------------------------------------------------------------------------------------------
C5 = dict(nq=10000, rows=12_500_000, nbytes=32, seed=1005)


def bench_database_c5(g, steps, world=1, rank=0, dist=None, peers_factory=None):
    torch, lib, m = g.torch, g.lib, g.m
    dev = g.dev
    NQ, ROWS, NB = C5["nq"], C5["rows"], C5["nbytes"]
    gen = torch.Generator(device=dev); gen.manual_seed(C5["seed"] + rank)
    t = torch.randint(0, 256, (ROWS, NB), dtype=torch.uint8, device=dev, generator=gen)
    gq = torch.Generator(device=dev); gq.manual_seed(424242)
    q = torch.randint(0, 256, (NQ, NB), dtype=torch.uint8, device=dev, generator=gq)          # same on every rank
    mine = torch.arange(rank, NQ, world, device=dev)                                          # plant query i as row (i * 1237) % ROWS of rank i % world
    t[(mine * 1237) % ROWS] = q[mine]
    w16 = NB // 16
    scr_bytes = int(lib.mpl_knn_hamming_scratch_bytes(NQ, ROWS, w16))
    scr = torch.empty(scr_bytes, dtype=torch.uint8, device=dev)
    keys = torch.empty((NQ, 2), dtype=torch.int64, device=dev)
    idx = torch.empty((NQ, 2), dtype=torch.int32, device=dev)
    dst = torch.empty((NQ, 2), dtype=torch.int32, device=dev)
    peers = setup_peers(g, dist, rank, world, NQ) if world > 1 else None

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
    steps = max(3, min(steps, 5))
    l0 = g.launches()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for a, b in ev:                                              # train shard (400 MB) > L2: nothing to flush between steps
        a.record(); step(); b.record()
    torch.cuda.synchronize()
    launches = g.launches() - l0
    ms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
    for cls in ("hamming", "merge", "l2_tc", "l2_prep", "l2_rerank"):
        g.prof_read(PROF[cls])
    lib.mpl_prof_enable(1); step(); torch.cuda.synchronize(); lib.mpl_prof_enable(0)
    roof, kern_ms, path = hamming_kernel_roofline(g, NB, float(NQ) * ROWS, long_step=True)
    mg_ms, mg_n = g.prof_read(PROF["merge"])
    # the popc scan on the same shard, for comparison (mpl_set_hamming_path(1)); the exchange is the same
    lib.mpl_set_hamming_path(1)
    try:
        step(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); step(); e1.record(); torch.cuda.synchronize()
        popc_ms = float(e0.elapsed_time(e1))
        keys_popc = keys.clone()
    finally:
        lib.mpl_set_hamming_path(0)
    step(); torch.cuda.synchronize()                             # the checked results below come from the default path
    paths_equal = bool(torch.equal(keys, keys_popc))
    del keys_popc
    if peers is not None:
        to = ctypes.c_int(0)
        m._lib.check(lib.mpl_peer_status(peers[1][rank], ctypes.byref(to), g.stream()), "peer_status")
        if to.value:
            raise RuntimeError("rank %d: allgather_merge timed out waiting for its peers" % rank)
    if dist is not None:
        tm = torch.tensor([ms], device=dev); dist.all_reduce(tm, op=dist.ReduceOp.MAX); ms = float(tm.item())
    # (1) every query's planted copy: global index = owner * ROWS + (i * 1237) % ROWS, distance 0, runner-up further away
    i_all = torch.arange(NQ, device=dev)
    want = (i_all % world) * ROWS + (i_all * 1237) % ROWS
    planted = bool((idx[:, 0].long() == want).all().item()) and bool((dst[:, 0] == 0).all().item()) and bool((dst[:, 1] > 0).all().item())
    # (2) oracle spot check of the MERGED result: every rank runs the oracle over its own shard for 48 unplanted probe queries,
    #     the per-rank oracle top-2 are gathered and merged on the host in (dist, index) order and compared with what the
    #     GPUs returned for those probes (a second, short pass with the probes appended to nothing else)
    from oracle import knn as oknn
    n_probe = 48
    probe = torch.randint(0, 256, (n_probe, NB), dtype=torch.uint8, device=dev, generator=torch.Generator(device=dev).manual_seed(99))
    pk = torch.empty((n_probe, 2), dtype=torch.int64, device=dev); pi = torch.empty((n_probe, 2), dtype=torch.int32, device=dev); pd = torch.empty((n_probe, 2), dtype=torch.int32, device=dev)
    pscr_b = int(lib.mpl_knn_hamming_scratch_bytes(n_probe, ROWS, w16)); pscr = torch.empty(max(pscr_b, 16), dtype=torch.uint8, device=dev)
    if path == "tensor cores":
        lib.mpl_set_hamming_path(2)                              # 48 queries would take the popc scan by the auto rule: check the path the step ran
    try:
        m._lib.check(lib.mpl_knn2_hamming_dev(probe.data_ptr(), n_probe, t.data_ptr(), ROWS, w16, rank * ROWS, pscr.data_ptr(), pscr_b, pk.data_ptr(), g.stream()), "hamming probe")
    finally:
        lib.mpl_set_hamming_path(0)
    if world > 1:
        allk = torch.empty((world, n_probe, 2), dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(allk.view(-1, 2), pk)
    else:
        allk = pk.view(1, n_probe, 2)
    m._lib.check(lib.mpl_merge_top2_dev(allk.data_ptr(), world, n_probe, 2, 0, pi.data_ptr(), pd.data_ptr(), None, g.stream()), "merge probe")
    torch.cuda.synchronize()
    t_cpu0 = time.perf_counter()
    oi, od = oknn.knn_hamming(probe.cpu().numpy(), t.cpu().numpy(), 2, 0, threads=max(1, (os.cpu_count() or 1) // world))
    cpu_s = time.perf_counter() - t_cpu0
    okeys = (od.astype(np.int64) << 32) | (oi.astype(np.int64) + rank * ROWS)
    ok_t = torch.from_numpy(okeys).to(dev)
    if world > 1:
        allo = torch.empty((world, n_probe, 2), dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(allo.view(-1, 2), ok_t)
    else:
        allo = ok_t.view(1, n_probe, 2)
    merged = np.sort(allo.permute(1, 0, 2).reshape(n_probe, -1).cpu().numpy(), axis=1)[:, :2]
    gi, gd = pi.cpu().numpy(), pd.cpu().numpy()
    oracle_ok = bool(np.array_equal(merged & 0xFFFFFFFF, gi.astype(np.int64)) and np.array_equal(merged >> 32, gd.astype(np.int64)))
    ok_all = torch.tensor([1 if (planted and oracle_ok and paths_equal) else 0], device=dev)
    if dist is not None:
        dist.all_reduce(ok_all, op=dist.ReduceOp.MIN)
    if int(ok_all.item()) == 0:
        raise RuntimeError("rank %d: C5 check failed (planted %s, oracle %s, popc == tensor-core keys %s)" % (rank, planted, oracle_ok, paths_equal))
    out = None
    if rank == 0:
        pairs = float(NQ) * ROWS * world
        out = dict(workload="C5: %d ORB queries x %d train rows per rank (%d rows on %d GPU%s), Hamming k=2, train sharded, top-2 merged over NVLink" %
                            (NQ, ROWS, ROWS * world, world, "s" if world > 1 else ""),
                   value=pairs / (ms * 1e-3), unit="pairs/s", n_gpus=world, scaling="weak", ms_per_step=ms,
                   dtype="bits as e4m3 0/1 on tcgen05 (exact integers)" if path == "tensor cores" else "u32 popc", path=path, gpu_launches=int(launches),
                   exchange="NVLink peer memory inside the merge kernel (mpl_allgather_merge_top2_dev)" if world > 1 else "none (one shard)",
                   kernels=dict(hamming_kernel_ms=kern_ms, merge_or_exchange_ms=mg_ms / max(mg_n, 1)),
                   roofline=roof,
                   popc_path=dict(ms_per_step=popc_ms, value_rank0=float(NQ) * ROWS / (popc_ms * 1e-3), unit="pairs/s per GPU",
                                  note="mpl_set_hamming_path(1): hamming_top2_kernel<2> on rank 0's shard, same keys bit for bit"),
                   checks=dict(planted_neighbours_found=planted, oracle_probe_queries=n_probe, oracle_merged_equal=oracle_ok, popc_and_tensor_core_keys_identical=paths_equal),
                   cpu_baseline=dict(value=float(n_probe) * ROWS / cpu_s, unit="pairs/s", cores=max(1, (os.cpu_count() or 1) // world), kind="port",
                                     sample="%d probe queries x %d rows (this rank's shard), oracle cvflann-order Hamming scan; the full config is %d x that work (extrapolate linearly)" %
                                            (n_probe, ROWS, (NQ * world) // n_probe)))
    del t
    torch.cuda.empty_cache()
    return out


# ------------------------------------------------------------------------------------------------
# C3: RANSAC 5-point, 10 000 correspondences x 100 000 pre-drawn samples
# ------------------------------------------------------------------------------------------------
def bench_ransac(g, steps, warmup):
    import datagen_pose
    torch, lib, m = g.torch, g.lib, g.m
    n, h = C3["n"], C3["h"]
    sc = datagen_pose.make_scene(n, 0.3, seed=C3["seed"])
    samples = datagen_pose.draw_samples_fast(n, h, seed=C3["seed"] + 1)
    d_p1, d_p2 = torch.from_numpy(sc["p1"]).to(g.dev), torch.from_numpy(sc["p2"]).to(g.dev)
    d_s = torch.from_numpy(samples).to(g.dev)
    d_corr = torch.empty(n * 4, dtype=torch.float64, device=g.dev)
    d_models = torch.empty(h * 90, dtype=torch.float64, device=g.dev)
    d_dense = torch.empty(h * 90, dtype=torch.float64, device=g.dev)
    d_nm = torch.empty(h, dtype=torch.int32, device=g.dev)
    d_off = torch.empty(h + 1, dtype=torch.int32, device=g.dev)
    d_counts = torch.empty(h * 10, dtype=torch.int32, device=g.dev)
    d_es = torch.empty(h * 10, dtype=torch.float64, device=g.dev)
    th = ctypes.c_double(sc["th"])
    st0 = g.stream()
    m._lib.check(lib.mpl_ransac_pack_corr_dev(d_p1.data_ptr(), d_p2.data_ptr(), n, d_corr.data_ptr(), st0), "pack")
    m._lib.check(lib.mpl_ransac_solve_dev(d_corr.data_ptr(), n, d_s.data_ptr(), h, d_models.data_ptr(), d_nm.data_ptr(), d_off.data_ptr(),
                                          d_dense.data_ptr(), st0), "solve")
    torch.cuda.synchronize()
    n_models = int(d_off[h].item())          # fixed for the workload: known before the timed region

    def step():
        st = g.stream()
        m._lib.check(lib.mpl_ransac_pack_corr_dev(d_p1.data_ptr(), d_p2.data_ptr(), n, d_corr.data_ptr(), st), "pack")
        m._lib.check(lib.mpl_ransac_solve_dev(d_corr.data_ptr(), n, d_s.data_ptr(), h, d_models.data_ptr(), d_nm.data_ptr(), d_off.data_ptr(),
                                              d_dense.data_ptr(), st), "solve")
        m._lib.check(lib.mpl_ransac_score_dev(d_dense.data_ptr(), n_models, d_corr.data_ptr(), n, th, d_counts.data_ptr(), d_es.data_ptr(), st), "score")

    ms = g.time_steps(step, steps, warmup)
    lib.mpl_prof_enable(1)
    for _ in range(3):
        g.flush_l2()
        step()
    torch.cuda.synchronize()
    lib.mpl_prof_enable(0)
    sc_ms, sc_n = g.prof_read(PROF["score"])
    so_ms, so_n = g.prof_read(PROF["solve"])
    # end to end: mpl_ransac5pt with host buffers, every sample solved + scored, counts/errsums read back
    p1, p2 = np.ascontiguousarray(sc["p1"]), np.ascontiguousarray(sc["p2"])
    E = np.zeros(9); mask = np.zeros(n, np.uint8); cnt = ctypes.c_int(0); iters = ctypes.c_int(0)
    tc = np.zeros(10 * h, np.int32); te = np.zeros(10 * h)
    P = m._lib.ptr

    def call():
        m._lib.check(lib.mpl_ransac5pt(P(p1), P(p2), n, P(samples), h, sc["th"], h, 0.999, 0, P(E), P(mask), ctypes.byref(cnt), ctypes.byref(iters),
                                       None, None, P(tc), P(te)), "mpl_ransac5pt")
    call()
    ts = []
    for _ in range(max(3, min(steps, 10))):
        t0 = time.perf_counter()
        call()
        ts.append(time.perf_counter() - t0)
    evals = float(n_models) * n
    step_ms = float(np.mean(ms))
    score_ms = sc_ms / max(sc_n, 1)
    # fp64 issue bound of one Sampson evaluation in the reference's operation order (no FMA contraction): 16 DMUL +
    # 14 DADD + one IEEE divide; the divide is costed at the measured DFMA-rate / DDIV-rate ratio
    pp = pipe_peaks()
    dp_slots = 30.0 + pp["dfma_ginst"] / pp["ddiv_rn_gops"]
    evals_peak = pp["dfma_ginst"] * 1e9 / dp_slots
    return dict(
        roofline=dict(bound="fp64 issue", achieved=evals / (score_ms * 1e-3), peak=evals_peak, unit="Sampson evaluations/s",
                      frac=evals / (score_ms * 1e-3) / evals_peak, dp_issue_slots_per_eval=dp_slots, traffic=ncu_traffic("sampson_score_kernel"),
                      kernel="sampson_score_kernel", kernel_ms=score_ms, peak_src=pp["_src"]),
        workload="C3: RANSAC 5pt, 10000 correspondences x 100000 pre-drawn samples (%d real models)" % n_models,
        value=evals / (step_ms * 1e-3), unit="model*correspondence Sampson evaluations/s", ms_per_step=step_ms, dtype="f64",
        hypotheses_per_ms=h / step_ms, models_per_ms=n_models / step_ms,
        north_star_target="1e6 hypotheses/ms over 10k correspondences = 1e13 exact-fp64 Sampson evaluations/s = %.1fx the measured fp64 issue ceiling (%.3g evaluations/s); "
                          "measured here: %.3g hypotheses/ms, %.3g models/ms" % (1e13 / evals_peak, evals_peak, h / step_ms, n_models / step_ms),
        cpu_baseline=cpu_ransac_c3(sc, samples, n_models, h),
        kernels=dict(sampson_score_ms=score_ms, fivept_solve_ms=so_ms / max(so_n, 1),
                     score_evals_per_s=evals / (score_ms * 1e-3), score_fp64_tflops=evals * 33 / (score_ms * 1e-3) / 1e12),
        best_inliers=int(d_counts[:n_models].max().item()),
        e2e=dict(value=evals / float(np.median(ts)), unit="evaluations/s", h2d_bytes_per_step=int(p1.nbytes + p2.nbytes + samples.nbytes),
                 d2h_bytes_per_step=int(n_models * 12 + n), inliers=int(cnt.value)))


# ------------------------------------------------------------------------------------------------
# C2 second float set: SURF/KAZE-like unit-norm signed floats -> kind::tf32 ranking, top-4 lists, certificate, exact re-rank
# ------------------------------------------------------------------------------------------------
def bench_l2_tf32(g, steps, warmup):
    import datagen
    torch, lib, m = g.torch, g.lib, g.m
    q, t = datagen.signed_float(C2["nq"], C2["nt"], C2["dim"], seed=C2["seed"])
    nq, nt, dim = q.shape[0], t.shape[0], q.shape[1]
    d_q, d_t = torch.from_numpy(q).to(g.dev), torch.from_numpy(t).to(g.dev)
    scr_bytes = int(lib.mpl_knn_l2_scratch_bytes(nq, nt, dim))
    d_scr = torch.empty(scr_bytes, dtype=torch.uint8, device=g.dev)
    d_keys = torch.empty((nq, 2), dtype=torch.int64, device=g.dev)
    d_idx = torch.empty((nq, 2), dtype=torch.int32, device=g.dev); d_dist = torch.empty((nq, 2), dtype=torch.float32, device=g.dev)

    def step():
        st = g.stream()
        m._lib.check(lib.mpl_knn2_l2f32_dev(d_q.data_ptr(), nq, dim, d_t.data_ptr(), nt, dim, dim, 0, 0, d_scr.data_ptr(), scr_bytes, d_keys.data_ptr(), st), "knn2_l2f32_dev")
        m._lib.check(lib.mpl_merge_top2_dev(d_keys.data_ptr(), 1, nq, 2, 1, d_idx.data_ptr(), d_dist.data_ptr(), None, st), "merge")

    ms = g.time_steps(step, steps, warmup)
    fb, tf = ctypes.c_uint(0), ctypes.c_uint(0)
    m._lib.check(lib.mpl_knn_l2_last_stats(d_scr.data_ptr(), ctypes.byref(fb), ctypes.byref(tf), g.stream()), "stats")
    lib.mpl_prof_enable(1)
    for _ in range(5):
        g.flush_l2(); step()
    torch.cuda.synchronize()
    lib.mpl_prof_enable(0)
    kern = {k: g.prof_read(v) for k, v in PROF.items()}
    tc_ms = kern["l2_tc"][0] / max(kern["l2_tc"][1], 1)
    ts = []
    (_, hq), (_, ht) = pinned(q, torch), pinned(t, torch)
    for i in range(12):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        err, mt = m.getMatchesArray(range(nq), range(nt), hq, ht, None, "LINEAR", False, True)
        if i >= 4:
            ts.append(time.perf_counter() - t0)
    pk, pp = peaks(), pipe_peaks()
    flops = 2.0 * dim * nq * nt
    step_ms = float(np.mean(ms))
    ach = flops / (tc_ms * 1e-3) / 1e12
    # tf32 tensor peak, measured the way MEASURED_PEAKS.json measures bf16: cuBLAS matmul 8192^3 with TF32 inputs, best of 10
    # (burst).  The kind::tf32 UMMA issue rate of tools/microbench.cu (one SM's tensor pipe alone, no power limit) is quoted beside it.
    peak, peak_src = measure_tf32_peak(torch, g.dev)
    if not peak:
        peak, peak_src = pk["bf16_tflops"] / 2, "half of MEASURED_PEAKS bf16_tflops (burst)"
    # a small oracle spot check of this very run (the full-size comparison is tests/test_knn_gpu.py::test_l2_default_path_vs_oracle_at_c2_size)
    from oracle import knn as oknn
    oi, od = oknn.knn_l2(q[:256], t, 2, 0, threads=os.cpu_count() or 1)
    gi, gd = d_idx[:256].cpu().numpy(), d_dist[:256].cpu().numpy()
    if not (np.array_equal(gi, oi) and np.array_equal(gd.view(np.int32), od.view(np.int32))):
        raise RuntimeError("tf32-mode float k-NN differs from the oracle")
    return dict(workload="C2 shape, SURF/KAZE-like unit-norm signed f32 (non-integer): 20000 x 20000 x 128, L2 k=2 (LINEAR); kind::tf32 ranking + top-4 lists + certificate + exact re-rank",
                value=float(nq) * nt / (step_ms * 1e-3), unit="pairs/s", ms_per_step=step_ms, dtype="f32 (tf32 tensor-core ranking, exact fp32 re-rank)",
                tf32_mode=int(tf.value), uncertified_queries_rescanned=int(fb.value),
                roofline=dict(bound="tensor", achieved=ach, peak=peak, unit="TFLOP/s", frac=ach / peak, kernel="l2_tc_rank_kernel (tcgen05 kind::tf32)", kernel_ms=tc_ms,
                              peak_src=peak_src, tf32_umma_issue_bound_tflops=pp.get("tf32_tflops_issue_bound"), step_share={k: (v[0] / max(v[1], 1)) for k, v in kern.items() if v[1]}),
                e2e=dict(value=float(nq) * nt / float(np.median(ts)), unit="pairs/s", h2d_bytes_per_step=int(hq.nbytes + ht.nbytes), d2h_bytes_per_step=int(nq * 16), matches=int(len(mt))),
                oracle_check="first 256 queries: idx equal, distances bit-exact")


def measure_tf32_peak(torch, dev):
    try:
        old = torch.backends.cuda.matmul.allow_tf32
        torch.backends.cuda.matmul.allow_tf32 = True
        a = torch.randn(8192, 8192, device=dev); b = torch.randn(8192, 8192, device=dev)
        for _ in range(3):
            a @ b
        best = 1e9
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); a @ b; e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        torch.backends.cuda.matmul.allow_tf32 = old
        del a, b
        return 2.0 * 8192 ** 3 / (best * 1e-3) / 1e12, "measured live: torch.matmul 8192^3 with TF32 inputs (cuBLAS), best of 10 (burst)"
    except Exception as e:
        sys.stderr.write("bench: tf32 peak measurement failed (%s)\n" % e)
        return None, None


# ------------------------------------------------------------------------------------------------
# the "next" rows of SURVEY 8(f) that are built: GMS post-filter (N4) and ARRSAC (N2), each next to the REFERENCE's own code
# compiled into oracle/_ref (MatchGMS.cpp; the theia::Arrsac template)
# ------------------------------------------------------------------------------------------------
def _threaded_rate(fn, n_calls, n_threads):
    """calls/s of fn(i) issued from a pool of host threads (the library keeps one context -- stream, scratch -- per host thread, so
    independent calls overlap on the GPU; ctypes releases the GIL inside the call)."""
    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(n_threads) as ex:
        list(ex.map(fn, range(n_threads)))                      # first call of every thread creates its context
        t0 = time.perf_counter()
        list(ex.map(fn, range(n_calls)))
        return n_calls / (time.perf_counter() - t0)


def bench_next_rows(g):
    import datagen_gms, datagen_pose
    from matchinglib_poselib_b200 import matchinglib, poselib
    from oracle import gms as ogms, arrsac as oarrsac, usac as ousac, vfc as ovfc
    import datagen_vfc
    torch = g.torch
    out = {}
    k1, k2, mt, sz = datagen_gms.scene(n=8000, seed=1, border=True)
    dm = datagen_gms.dmatch(mt)
    ts = []
    for i in range(12):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        n_g, mask = matchinglib.filterMatchesGMS(k1, sz, k2, sz, dm)
        if i >= 2:
            ts.append(time.perf_counter() - t0)
    t0 = time.perf_counter(); r = ogms.ref_filter_matches(k1, sz, k2, sz, mt) or ogms.filter_matches(k1, sz, k2, sz, mt); dt_cpu = time.perf_counter() - t0
    if r[0] != n_g or not np.array_equal(r[1], mask.astype(np.uint8)):
        raise RuntimeError("GMS: GPU mask differs from the reference's")
    out["gms_filter"] = dict(workload="GMS grid filter, %d matches over 8000 keypoints per image (filterMatchesGMS, no scale / rotation), host buffers in and out" % len(mt),
                             value=len(mt) / float(np.median(ts)), unit="matches/s", ms_per_call=float(np.median(ts) * 1e3), inliers=int(n_g),
                             check="mask bit-exact vs the reference's own MatchGMS.cpp",
                             cpu_baseline=dict(value=len(mt) / dt_cpu, unit="matches/s", cores=1, kind="reference" if ogms.ref_lib() is not None else "port",
                                               sample="the same call through the reference's MatchGMS.cpp (oracle/_ref/libgms_ref.so), %.2f ms" % (dt_cpu * 1e3)))
    sc = datagen_pose.make_scene(3000, 0.5, seed=2)
    ts, res = [], None
    for i in range(7):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        res = poselib.arrsac5pt(sc["p1"], sc["p2"], sc["th"], refine=True, seed=(0, 0))
        if i >= 2:
            ts.append(time.perf_counter() - t0)
    cb = None
    if oarrsac.ref_lib() is not None:
        t0 = time.perf_counter(); o = oarrsac.estimate(sc["p1"], sc["p2"], sc["th"], True); dt_cpu = time.perf_counter() - t0
        cb = dict(value=1.0 / dt_cpu, unit="calls/s", cores=1, kind="reference", inliers=int(o["count"]),
                  sample="the reference's theia::Arrsac template + restated estimator and refinement on the same scene, %.1f ms" % (dt_cpu * 1e3))
    out["arrsac"] = dict(workload="estimateEssentialMat(method = ARRSAC, refine), 3000 correspondences, 50 % outliers", value=1.0 / float(np.median(ts)), unit="calls/s",
                         ms_per_call=float(np.median(ts) * 1e3), inliers=int(res["count"]), hypotheses=int(res["n_hypotheses"]), device_batches=int(res["n_batches"]),
                         note="latency-bound: every device batch waits for one 5-point solve (~1.2 ms of dependent fp64 work); parity with the reference is statistical",
                         cpu_baseline=cb)
    # LMedS: the third method of estimateEssentialMat (134 fixed samples, per-model median of all errors)
    from oracle import ransac as oransac
    sc = datagen_pose.make_scene(10000, 0.35, seed=4)
    smp = datagen_pose.draw_samples_fast(10000, oransac.lmeds_niters(0.999, 2000), 9)
    ts, res = [], None
    for i in range(7):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        res = poselib.lmeds5pt(sc["p1"], sc["p2"], smp)
        if i >= 2:
            ts.append(time.perf_counter() - t0)
    t0 = time.perf_counter(); o = oransac.lmeds5pt(sc["p1"], sc["p2"], smp); dt_cpu = time.perf_counter() - t0
    out["lmeds"] = dict(workload="estimateEssentialMat(method = LMEDS): 134 samples, 10000 correspondences, 35 % outliers, host buffers in and out",
                        value=1.0 / float(np.median(ts)), unit="calls/s", ms_per_call=float(np.median(ts) * 1e3), inliers=int(res["count"]),
                        check="same sample selected as the oracle: %s; inliers %d vs %d" % (abs(res["min_median"] - o["min_median"]) <= 1e-6 * o["min_median"], res["count"], o["count"]),
                        cpu_baseline=dict(value=1.0 / dt_cpu, unit="calls/s", cores=1, kind="port",
                                          sample="the C restatement of runLMeDS on the same samples, single thread as the reference runs it, %.1f ms" % (dt_cpu * 1e3)))
    # VFC post-filter (second half): getMatches(VFCrefine = true)
    v1, v2, vmt, vinl = datagen_vfc.scene(8000, 0.5, seed=3)
    vdm = np.zeros(len(vmt), matchinglib.DMATCH_DTYPE); vdm["queryIdx"] = vmt[:, 0]; vdm["trainIdx"] = vmt[:, 1]; vdm["distance"] = 1
    ts, kept = [], None
    for i in range(8):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        code, kept = matchinglib.filterWithVFC(v1, v2, vdm, seed=1)
        if i >= 2:
            ts.append(time.perf_counter() - t0)
    cb = None
    if ovfc.ref_lib() is not None:
        t0 = time.perf_counter(); rr = ovfc.filter_with_vfc(v1, v2, vmt, seed=1); dt_cpu = time.perf_counter() - t0
        keep = np.zeros(len(vmt), bool); keep[np.searchsorted(vdm["queryIdx"], kept["queryIdx"])] = True
        if rr[0] != code or not np.array_equal(keep, rr[1].astype(bool)):
            raise RuntimeError("VFC: GPU kept set differs from the reference's")
        cb = dict(value=len(vmt) / dt_cpu, unit="matches/s", cores=1, kind="reference", sample="the same call through the reference's vfc.cpp / vfcMatches.cpp (oracle/_ref/libvfc_ref.so), %.1f ms" % (dt_cpu * 1e3))
    out["vfc_filter"] = dict(workload="filterWithVFC (SparseVFC, 16 control points, 50 EM iterations), %d matches, 50 %% outliers, host buffers in and out" % len(vmt),
                             value=len(vmt) / float(np.median(ts)), unit="matches/s", ms_per_call=float(np.median(ts) * 1e3), kept=int(len(kept)),
                             check="kept set identical to the reference's own vfc.cpp for the same rand() seed",
                             note="bound by the reference's sequential float sums: 172 single-thread add chains per iteration, kept to reproduce its (numerically chaotic) result", cpu_baseline=cb)
    r32 = _threaded_rate(lambda i: matchinglib.filterWithVFC(v1, v2, vdm, seed=1), 128, 32)
    out["vfc_filter"]["calls_per_s_32_host_threads"] = r32
    out["vfc_filter"]["matches_per_s_32_host_threads"] = r32 * len(vmt)
    # USAC: uniform sampling needs a few hundred hypotheses on this scene, PROSAC a handful
    sc = datagen_pose.make_scene(10000, 0.5, seed=2)
    outl = ousac.sampson(sc["E"], sc["p1"], sc["p2"]) >= sc["th"] ** 2
    order = np.argsort(outl + 0.6 * np.random.default_rng(2).random(len(outl)), kind="stable").astype(np.uint32)
    for name, idx in (("usac_uniform", None), ("usac_prosac", order)):
        ts, tc, res, ref = [], [], None, None
        for i in range(9):
            torch.cuda.synchronize(); t0 = time.perf_counter()
            res = poselib.usac5pt(sc["p1"], sc["p2"], sc["th"], sorted_idx=idx, seed=i + 1)
            if i >= 2:
                ts.append(time.perf_counter() - t0)
            if ousac.ref_lib() is not None:
                t0 = time.perf_counter(); ref = ousac.usac(sc["p1"], sc["p2"], sc["th"], sorted_idx=idx, seed=i + 1)
                if i >= 2:
                    tc.append(time.perf_counter() - t0)
        cb = None
        if tc:
            cb = dict(value=1.0 / float(np.median(tc)), unit="calls/s", cores=1, kind="reference", inliers=int(ref["count"]), hypotheses=int(ref["hyp_count"]),
                      sample="the reference's own USAC.h + EssentialMatEstimator.h + opengv (oracle/_ref/libusac_ref.so) on the same scene and seeds, median %.1f ms" % (np.median(tc) * 1e3))
        out[name] = dict(workload="estimateEssentialMatUsac (NISTER, REF_WEIGHTS, no degeneracy test, %s sampling), 10000 correspondences, 50 %% outliers, host buffers in and out" %
                         ("PROSAC" if idx is not None else "uniform"), value=1.0 / float(np.median(ts)), unit="calls/s", ms_per_call=float(np.median(ts) * 1e3),
                         inliers=int(res["count"]), hypotheses=int(res["hyp_count"]), device_batches=int(res["n_batches"]), launches=int(res["n_launches"]),
                         note="latency-bound like ARRSAC: each device batch costs one 5-point solve latency; parity per seed in tests/test_usac_gpu.py", cpu_baseline=cb)
        out[name]["calls_per_s_16_host_threads"] = _threaded_rate(lambda i: poselib.usac5pt(sc["p1"], sc["p2"], sc["th"], sorted_idx=idx, seed=i + 1), 128, 16)
    return out


# ------------------------------------------------------------------------------------------------
# CPU baselines of the parts (BASELINE.md section 3 / SURVEY 8(d)): reported next to the GPU numbers, never the target
# ------------------------------------------------------------------------------------------------
def cpu_hamming_c1(q, t):
    """C1 on the host: (i) cvflann LINEAR through cv2.flann_Index -- the very code the reference calls, single thread as in the
    reference (matchers.cpp:567-588); (ii) the reference's vendored NMSLIB seq_search/bit_hamming (oracle/_ref) with the
    wrapper's default 8 threads (NMSLIB_SEARCH_THREADS, matchers.cpp:60); (iii) the oracle port as a fallback."""
    from oracle import knn as oknn
    out = dict(unit="pairs/s")
    pairs = float(len(q)) * len(t)
    try:
        import cv2
        cv2.setNumThreads(1)
        n1 = min(len(q), 2048)
        t0 = time.perf_counter()
        ix = cv2.flann_Index(t, dict(algorithm=0), 9)                          # FLANN_DIST_HAMMING
        ix.knnSearch(q[:n1], 2, params=dict(checks=64))
        dt = time.perf_counter() - t0
        out.update(value=float(n1) * len(t) / dt, cores=1, kind="reference",
                   sample="cv2.flann_Index linear (cvflann HammingLUT, OpenCV %s), first %d of %d queries x %d train rows, single thread, %.2f s" % (cv2.__version__, n1, len(q), len(t), dt))
    except Exception as e:                                                    # no cv2 on the box: the C port of the same scan
        t0 = time.perf_counter(); oknn.knn_hamming(q, t, 2, 0, threads=1); dt = time.perf_counter() - t0
        out.update(value=pairs / dt, cores=1, kind="port", sample="oracle port of cvflann's linear Hamming scan, full C1, single thread (%s)" % type(e).__name__)
    if oknn.ref_lib() is not None:
        t0 = time.perf_counter(); oknn.ref_nmslib_hamming(q, t, threads=8); dt = time.perf_counter() - t0
        out["nmslib_8_threads"] = dict(value=pairs / dt, unit="pairs/s", cores=8, kind="reference", sample="reference NMSLIB seq_search/bit_hamming, full C1, 8 threads (NMSLIB_SEARCH_THREADS), %.2f s" % dt)
    return out


def cpu_ransac_c3(sc, samples, n_models_gpu, h, n_sub=1000):
    """C3 on the host: the C restatement of run5Point + computeReprojError3 (oracle/ransac5pt_oracle.c), single thread like the
    reference, full trace over every n-th sample; evaluations/s is per-sample work, so the whole config extrapolates linearly."""
    from oracle import ransac as oransac
    sub = np.ascontiguousarray(samples[:: max(1, h // n_sub)][:n_sub])
    t0 = time.perf_counter()
    r = oransac.ransac5pt(sc["p1"], sc["p2"], sub, sc["th"], max_iters=len(sub), trace=True)
    dt = time.perf_counter() - t0
    evals = float(r["n_models"].sum()) * len(sc["p1"])
    return dict(value=evals / dt, unit="model*correspondence Sampson evaluations/s (5-point solves included in the time)", cores=1, kind="port",
                sample="%d of %d samples (stratified), %d models x %d correspondences, %.1f s single thread; whole C3 extrapolated: %.0f s" %
                       (len(sub), h, int(r["n_models"].sum()), len(sc["p1"]), dt, dt * h / len(sub)),
                hypotheses_per_ms=len(sub) / (dt * 1e3))


# ------------------------------------------------------------------------------------------------
# CPU legs: the reference's own CPU path on the host cores (the ONLY place bench.py executes oracle/)
# ------------------------------------------------------------------------------------------------
def cpu_l2_runner():
    """Returns (kind, fn(q, t, threads) -> (idx, dist)).  "reference" = the reference's vendored NMSLIB
    (seq_search + l2: the BRUTEFORCENMS matcher's hot loop) compiled from its own sources into oracle/_ref;
    "port" = the C restatement of cvflann's linear L2 scan, query-striped over threads."""
    from oracle import knn as orc
    if orc.ref_lib() is not None:
        return "reference", lambda q, t, th: orc.ref_nmslib_l2(q, t, threads=th)
    return "port", lambda q, t, th: orc.knn_l2(q, t, 2, 0, threads=th)


def cpu_l2_sample(q, t, budget_s=12.0):
    """Bounded sample of the C2 workload: a prefix of the queries against the full train set, all host threads."""
    kind, fn = cpu_l2_runner()
    cores = os.cpu_count() or 1
    n0 = min(len(q), 64 * cores)
    t0 = time.perf_counter(); fn(q[:n0], t, cores); dt0 = time.perf_counter() - t0
    n1 = int(min(len(q), max(n0, n0 * budget_s / max(dt0, 1e-6))))
    t0 = time.perf_counter(); idx, dist = fn(q[:n1], t, cores); dt = time.perf_counter() - t0
    return dict(value=float(n1) * len(t) / dt, unit="pairs/s", cores=cores, kind=kind,
                sample="first %d of %d queries x %d train rows, %.1f s, %d threads (%s)" %
                       (n1, len(q), len(t), dt, cores, "reference NMSLIB seq_search/l2, q%%T striping" if kind == "reference" else "oracle port of cvflann linear L2")), idx, dist


def run_reference(args):
    """--impl reference: K steps, each a bounded sample of C2 on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    q, t = make_c2(0)
    kind, fn = cpu_l2_runner()
    cores = os.cpu_count() or 1
    # size a step at ~2 s so that steps + warm-up stay within a few minutes
    n0 = min(len(q), 32 * cores)
    t0 = time.perf_counter(); fn(q[:n0], t, cores); dt0 = time.perf_counter() - t0
    per_step_s = min(2.0, 150.0 / max(args.steps + args.warmup, 1))
    n1 = int(min(len(q), max(n0, n0 * per_step_s / max(dt0, 1e-6))))
    for _ in range(args.warmup):
        fn(q[:n1], t, cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        fn(q[:n1], t, cores)
    dt = (time.perf_counter() - t0) / max(args.steps, 1)
    v = float(n1) * len(t) / dt
    sample = "each step: first %d of %d queries x %d train rows on %d host threads" % (n1, len(q), len(t), cores)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "pairs/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": dict(workload=WORKLOAD, sample=sample),
        "cpu_baseline": dict(value=v, unit="pairs/s", cores=cores, kind=kind, sample=sample),
        "e2e": dict(value=v, unit="pairs/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0), "gpu_launches": 0}))


METRIC = "descriptor pairs/s (brute-force L2 k=2 + ratio test)"
WORKLOAD = "C2: SIFT-like 128-d f32, 20000 queries x 20000 train rows per GPU, L2 k=2 + Lowe ratio (LINEAR)"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-parts", action="store_true", help="skip parts: C5 database, C4 pair batch (every N); Hamming C1, RANSAC C3, tf32 float set (N = 1)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        run_reference(args)
        return

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world == 1 and args.gpus > 1:
        raise SystemExit("launch N > 1 with torchrun (python -m torch.distributed.run --nproc-per-node N bench.py --gpus N)")
    import __graft_entry__ as ge
    if rank == 0:
        ge.build()                                   # no-op when the .so files travelled with the snapshot
    g = Gpu(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")     # keep stdout to the one JSON line
        dist.init_process_group("nccl", device_id=g.dev)
        dist.barrier()

    r = bench_l2(g, args.steps, args.warmup, world, rank, dist)
    torch = g.torch
    step_ms = float(np.mean(r["ms"]))
    if world > 1:
        tm = torch.tensor([step_ms], device=g.dev)
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)    # device-timed, max over ranks
        step_ms = float(tm.item())
    pairs = float(r["nq"]) * r["nt"] * world
    pk = peaks()
    out = None
    if rank == 0:
        kern = r["kern"]
        tc_ms, tc_n = kern["l2_tc"]
        ex_ms, ex_n = kern["l2_exact"]
        if tc_n:
            kname, k_ms, peak, peak_name = "l2_tc_rank_kernel (tcgen05)", tc_ms / tc_n, pk["bf16_tflops"], "bf16_tflops (burst, cuBLAS bf16)"
        else:
            kname, k_ms, peak, peak_name = "l2_exact_top2_kernel (fp32 SIMT)", ex_ms / max(ex_n, 1), pk["bf16_tflops"], "bf16_tflops (burst, cuBLAS bf16)"
        flops = 2.0 * r["dim"] * r["nq"] * r["nt"]           # per launch: one rank's shard
        ach = flops / (k_ms * 1e-3) / 1e12
        if world > 1:
            e2e_s, n_matches, bi, bo = r["e2e_sharded"]
            e2e_api = "sharded database call: pinned host buffers -> mpl_knn2_l2f32_dev + mpl_allgather_merge_top2_dev -> host, mpl_ratio_matches_f32; every rank at once, max over ranks"
        else:
            e2e_s, n_matches, bi, bo = e2e_l2(g, r["q"], r["t"], max(5, min(args.steps, 20)))
            e2e_api = "getMatches(LINEAR) -> mpl_get_matches_api, pinned host buffers, 1 GPU"
        out = {
            "metric": METRIC, "value": pairs / (step_ms * 1e-3), "unit": "pairs/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32 (bf16/tf32 tensor-core ranking, exact fp32 re-rank)",
            "data": "synthetic",
            "config": dict(workload=WORKLOAD, l2_between_steps="flushed (512 MiB write)", train_sharding="none" if world == 1 else ("train rows sharded over %d ranks, " % world) + ("top-2 keys all-gathered over NVLink peer memory inside the merge kernel" if r.get("peer_exchange") else "NCCL all-gather of top-2 keys + merge"),
                           launch="CUDA graph replay of the step" if r.get("graphed") else "stream launches"),
            "clocks": r["clocks"],
            "roofline": dict(bound="tensor", achieved=ach, peak=peak, unit="TFLOP/s", frac=ach / peak, traffic=ncu_traffic("l2_tc_rank_kernel"),
                             kernel=kname, kernel_ms=k_ms,
                             peak_src="%s: %s" % (pk["_src"], peak_name), flops_per_launch=flops,
                             step_share={k: (v[0] / max(v[1], 1)) for k, v in kern.items() if v[1]}),
            "e2e": dict(value=pairs / e2e_s, unit="pairs/s", h2d_bytes_per_step=int(bi) * world, d2h_bytes_per_step=int(bo) * world, ms=e2e_s * 1e3,
                        matches=int(n_matches), api=e2e_api),
            "gpu_launches": int(r["launches_timed"]),
        }
        if not args.no_cpu and world == 1:
            cb, cidx, cdist = cpu_l2_sample(r["q"], r["t"])
            n1 = len(cidx)
            gi = r["idx"][:n1, 0]
            differ = np.nonzero(cidx[:, 0] != gi)[0]
            # north_star: float NN indices must agree except for near-ties within 1e-5 relative distance.  For every query
            # whose first neighbour differs from the reference's NMSLIB (built -Ofast: its fp32 sums are not reproducible),
            # the two candidates' exact distances (fp64 on the host) must be within 1e-5 of each other.
            gaps = []
            for qi in differ:
                a = np.sqrt(((r["q"][qi].astype(np.float64) - r["t"][gi[qi]].astype(np.float64)) ** 2).sum())
                b = np.sqrt(((r["q"][qi].astype(np.float64) - r["t"][cidx[qi, 0]].astype(np.float64)) ** 2).sum())
                gaps.append(abs(a - b) / max(a, b, 1e-300))
            cb["nn_agreement_with_gpu"] = float((cidx[:, 0] == gi).mean())
            cb["nn_disagreements"] = int(len(differ))
            cb["nn_disagreement_max_rel_gap"] = float(max(gaps)) if gaps else 0.0
            if gaps and max(gaps) >= 1e-5:
                raise RuntimeError("GPU and reference first neighbours differ beyond a near-tie: relative gap %g" % max(gaps))
            out["cpu_baseline"] = cb
        if world > 1:
            # oracle check of the MERGED sharded result: 512 queries against the concatenation of every rank's shard
            from oracle import knn as oknn
            sel = np.linspace(0, r["nq"] - 1, 512).astype(int)
            full_t = np.concatenate([make_c2(rk)[1] for rk in range(world)], 0)
            oi, od = oknn.knn_l2(r["q"][sel], full_t, 2, 0, threads=os.cpu_count() or 1)
            if not (np.array_equal(r["idx"][sel], oi) and np.array_equal(r["dist"][sel].view(np.int32), od.view(np.int32))):
                raise RuntimeError("sharded merged top-2 differs from the oracle over the concatenated shards")
            out["config"]["oracle_check"] = "512 queries vs oracle over the %d concatenated rows: idx equal, distances bit-exact" % len(full_t)
    parts = {}
    if not args.no_parts:
        # the two multi-GPU workloads of BASELINE.json, at every N (collective: every rank takes part, rank 0 reports)
        c5 = bench_database_c5(g, args.steps, world, rank, dist)
        c4 = bench_pair_batch_c4(g, world, rank, dist)
        if rank == 0:
            parts["database_c5"] = c5
            parts["pair_batch_c4"] = c4
        if world == 1:
            parts["hamming_c1"] = bench_hamming(g, args.steps, args.warmup)
            parts["ransac_c3"] = bench_ransac(g, max(3, args.steps // 4), 3)
            parts["l2_tf32_c2"] = bench_l2_tf32(g, max(5, args.steps // 2), 3)
            parts.update(bench_next_rows(g))
    if rank == 0 and parts:
        out["parts"] = parts
        out["gpu_launches_parts"] = int(sum(v.get("gpu_launches", 0) for v in parts.values() if isinstance(v, dict)))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(out))


if __name__ == "__main__":
    main()
