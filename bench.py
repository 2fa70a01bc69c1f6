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
shard, the queries are replicated, per-rank top-2 keys are all-gathered over NCCL and merged
(weak scaling: pairs grow with N).

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
    p = os.path.join(ROOT, "profiles", "r1_microbench.json")
    if os.path.exists(p):
        d = json.load(open(p))
        d["_src"] = "measured: tools/microbench.cu -> profiles/r1_microbench.json"
        return d
    # CUDA programming guide throughput table x 148 SMs x 1.965 GHz
    return dict(popc_gops=4650.0, dfma_ginst=18600.0, ddiv_rn_gops=1300.0, _src="fallback: programming-guide rates")


def ncu_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed `ncu --set full` capture of this
    bench command (profiles/r1_ncu_traffic.json); None when there is no capture for the kernel."""
    p = os.path.join(ROOT, "profiles", "r1_ncu_traffic.json")
    if os.path.exists(p):
        return json.load(open(p)).get(kernel, {}).get("dram_bytes_per_launch")
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

    ms = g.time_steps(step, steps, warmup)
    lib.mpl_prof_enable(1)
    for _ in range(5):
        g.flush_l2()
        step()
    torch.cuda.synchronize()
    lib.mpl_prof_enable(0)
    k_ms, k_n = g.prof_read(PROF["hamming"])
    g.prof_read(PROF["merge"])
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
    pairs = float(nq) * nt
    step_ms = float(np.mean(ms))
    kern_ms = k_ms / max(k_n, 1)
    pk = peaks()
    pp = pipe_peaks()
    bytes_alg = (nq + nt) * nb + nq * 16
    popc = pairs * (nb // 4)
    return dict(
        workload="C1: ORB 8192x8192x32B Hamming k=2 + ratio (LINEAR)", value=pairs / (step_ms * 1e-3), unit="pairs/s", ms_per_step=step_ms,
        dtype="u32 popc",
        roofline=dict(bound="hbm", achieved=bytes_alg / (kern_ms * 1e-3) / 1e9, peak=pk["hbm_gbs"], unit="GB/s",
                      frac=bytes_alg / (kern_ms * 1e-3) / 1e9 / pk["hbm_gbs"], traffic=ncu_traffic("hamming_top2_kernel"), peak_src=pk["_src"],
                      kernel="hamming_top2_kernel", kernel_ms=kern_ms,
                      note="compute-bound on the integer pipe, not HBM (algorithmic bytes are 0.65 MB): see alu"),
        # SURVEY 8(d) counts ceil(B/4) POPC per pair (the reference's scalar form).  The kernel folds 8 XOR words through 3
        # carry-save adders and issues 5 POPC + 19 logic/add/min-max instructions per pair, so that figure can exceed 1: the
        # binding pipe is the integer ALU (LOP3/IADD3/VIMNMX, 64 lanes/clk/SM = the measured SHF rate).
        alu=dict(bound="int popc issue (algorithmic count); executed: int alu issue", achieved_gpopc_s=popc / (kern_ms * 1e-3) / 1e9,
                 peak_gpopc_s=pp["popc_gops"], frac=popc / (kern_ms * 1e-3) / 1e9 / pp["popc_gops"], popc_per_pair=nb // 4,
                 popc_executed_per_pair=5 * (nb // 32) if nb % 32 == 0 else None,
                 alu_instr_per_pair=HAMMING_ALU_PER_PAIR.get(nb),
                 alu_frac=(pairs * HAMMING_ALU_PER_PAIR[nb] / (kern_ms * 1e-3) / 1e9 / pp["shf_funnel_ginst"]) if nb in HAMMING_ALU_PER_PAIR and "shf_funnel_ginst" in pp else None,
                 peak_src=pp["_src"]),
        e2e=dict(value=pairs / float(np.median(ts)), unit="pairs/s", h2d_bytes_per_step=int(hq.nbytes + ht.nbytes), d2h_bytes_per_step=int(nq * 16),
                 matches=int(len(mt))))


# ------------------------------------------------------------------------------------------------
# C4 unit: ONE image pair, AKAZE-like 61-byte descriptors 5000 x 5000 + per-pair RANSAC (1000 samples), end to end through
# mpl_match_pose_pair (host buffers in, matches / E / mask out; matches stay on the device between the two halves)
# ------------------------------------------------------------------------------------------------
def bench_pair_pipeline(g, reps):
    import datagen, datagen_pose
    from matchinglib_poselib_b200 import poselib
    n, nb = 5000, 61
    sc = datagen_pose.make_scene(n, 0.0, seed=1004)
    rng = np.random.default_rng(1004)
    t = rng.integers(0, 256, (n, nb), dtype=np.uint8)
    q = rng.integers(0, 256, (n, nb), dtype=np.uint8)
    good = rng.random(n) < 0.5                                   # half of the queries are noisy copies of "their" train row
    flips = np.packbits(rng.random((n, nb * 8)) < 0.08, axis=1)
    q[good] = t[good] ^ flips[good]
    th = sc["th"]
    r = None
    ts = []
    for i in range(reps + 2):
        g.torch.cuda.synchronize()
        t0 = time.perf_counter()
        r = poselib.matchAndEstimatePose(q, t, sc["p1"], sc["p2"], "LINEAR", True, th, 1000, 0.999, True, seed=5)
        if i >= 2:
            ts.append(time.perf_counter() - t0)
    dt = float(np.median(ts))
    # a batch of pairs: one host thread per pair in flight (contexts are per thread: own stream and scratch), the GPU
    # overlaps the latency-bound per-pair RANSAC kernels of different pairs
    from concurrent.futures import ThreadPoolExecutor
    n_threads = min(32, 2 * (os.cpu_count() or 1))            # threads mostly wait for the device
    n_pairs = 8 * n_threads

    def one(i):
        return poselib.matchAndEstimatePose(q, t, sc["p1"], sc["p2"], "LINEAR", True, th, 1000, 0.999, True, seed=5 + i)["count"]

    with ThreadPoolExecutor(n_threads) as ex:
        list(ex.map(one, range(2 * n_threads)))                  # warm-up: every thread creates its context
        g.torch.cuda.synchronize()
        t0 = time.perf_counter()
        counts = list(ex.map(one, range(n_pairs)))
        bt = time.perf_counter() - t0
    return dict(workload="C4 unit: one pair, AKAZE-like 61 B x 5000 x 5000 Hamming k=2 + ratio, then RANSAC 5pt (1000 device-drawn samples, refit), "
                         "mpl_match_pose_pair end to end from host buffers",
                value=n_pairs / bt, unit="image pairs/s", ms_per_pair=dt * 1e3, single_thread_pairs_per_s=1.0 / dt,
                batch=dict(pairs=n_pairs, host_threads=n_threads, seconds=bt, min_inliers=int(min(counts))),
                matches=int(len(r["matches"])), inliers=int(r["count"]),
                ransac_iterations=int(r["iters"]), h2d_bytes_per_pair=int(q.nbytes + t.nbytes + 2 * n * 16), d2h_bytes_per_pair=int(len(r["matches"]) * 17 + 72))


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
        kernels=dict(sampson_score_ms=score_ms, fivept_solve_ms=so_ms / max(so_n, 1),
                     score_evals_per_s=evals / (score_ms * 1e-3), score_fp64_tflops=evals * 33 / (score_ms * 1e-3) / 1e12),
        best_inliers=int(d_counts[:n_models].max().item()),
        e2e=dict(value=evals / float(np.median(ts)), unit="evaluations/s", h2d_bytes_per_step=int(p1.nbytes + p2.nbytes + samples.nbytes),
                 d2h_bytes_per_step=int(n_models * 12 + n), inliers=int(cnt.value)))


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
                       (n1, len(q), len(t), dt, cores, "reference NMSLIB seq_search/l2, q%%T striping" if kind == "reference" else "oracle port of cvflann linear L2")), idx


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
    ap.add_argument("--no-parts", action="store_true", help="skip the Hamming (C1) and RANSAC (C3) legs")
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
            cb, cidx = cpu_l2_sample(r["q"], r["t"])
            n1 = len(cidx)
            agree = float((cidx[:, 0] == r["idx"][:n1, 0]).mean())
            cb["nn_agreement_with_gpu"] = agree
            out["cpu_baseline"] = cb
        if not args.no_parts and world == 1:
            out["parts"] = dict(hamming_c1=bench_hamming(g, args.steps, args.warmup), ransac_c3=bench_ransac(g, max(3, args.steps // 4), 3),
                                pair_pipeline_c4=bench_pair_pipeline(g, 10))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(out))


if __name__ == "__main__":
    main()
