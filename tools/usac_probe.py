"""GPU vs reference USAC (oracle/_ref/libusac_ref.so) on synthetic scenes: per-seed statistics of both sides and the wall clock
per call (host buffers in, host results out).  `python tools/usac_probe.py detail` prints every seed of a few scenes."""
import sys, time, json
import numpy as np
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import datagen_pose
from oracle import usac as ousac
from matchinglib_poselib_b200 import poselib, _lib

_lib.check(_lib.load().mpl_init(0), "mpl_init")


def order(sc, seed, mix=0.6):
    rng = np.random.default_rng(seed)
    out = ousac.sampson(sc["E"], sc["p1"], sc["p2"]) >= sc["th"] ** 2
    return np.argsort(out + mix * rng.random(len(out)), kind="stable").astype(np.uint32)


if len(sys.argv) > 1 and sys.argv[1] == "detail":
    keys = ("count", "hyp_count", "model_count", "rejected_sample_count", "rejected_model_count", "total_points_verified", "num_local_optimizations")
    for n, outl, sseed, prosac in [(800, 0.05, 31, False), (2000, 0.55, 21, True), (2000, 0.55, 21, False), (60, 0.1, 7, True)]:
        sc = datagen_pose.make_scene(n, outl, seed=sseed)
        idx = order(sc, sseed) if prosac else None
        truth = int((ousac.sampson(sc["E"], sc["p1"], sc["p2"]) < sc["th"] ** 2).sum())
        print("scene n=%d outl=%.2f prosac=%s truth=%d" % (n, outl, prosac, truth))
        for seed in range(1, 9):
            g = poselib.usac5pt(sc["p1"], sc["p2"], sc["th"], sorted_idx=idx, seed=seed)
            g0 = poselib.usac5pt(sc["p1"], sc["p2"], sc["th"], sorted_idx=idx, seed=seed, lo_mode=0)
            o = ousac.usac(sc["p1"], sc["p2"], sc["th"], sorted_idx=idx, seed=seed)
            print("  seed %d GPU %s sprt (%.4f %.4f)   [lo_mode 0: %s]" % (seed, [g[k] for k in keys], g["sprt_delta"], g["sprt_epsilon"], [g0[k] for k in keys[:2]]))
            print("         REF %s sprt (%.4f %.4f)" % ([o[k] for k in keys], o["sprt_delta"], o["sprt_epsilon"]))
    sys.exit(0)

rows = []
for n, outl, prosac in [(300, 0.2, True), (1000, 0.3, True), (1000, 0.3, False), (3000, 0.5, True), (3000, 0.5, False), (5000, 0.7, True), (10000, 0.3, True)]:
    sc = datagen_pose.make_scene(n, outl, seed=1)
    truth = int((ousac.sampson(sc["E"], sc["p1"], sc["p2"]) < sc["th"] ** 2).sum())
    idx = order(sc, 1) if prosac else None
    poselib.usac5pt(sc["p1"], sc["p2"], sc["th"], sorted_idx=idx, seed=1)
    tg, to, hg, ho, cg, co, bg, lg = [], [], [], [], [], [], [], []
    for seed in range(1, 21):
        t0 = time.perf_counter(); g = poselib.usac5pt(sc["p1"], sc["p2"], sc["th"], sorted_idx=idx, seed=seed); tg.append(time.perf_counter() - t0)
        hg.append(g["hyp_count"]); cg.append(g["count"]); bg.append(g["n_batches"]); lg.append(g["num_local_optimizations"])
        t0 = time.perf_counter(); o = ousac.usac(sc["p1"], sc["p2"], sc["th"], sorted_idx=idx, seed=seed); to.append(time.perf_counter() - t0)
        if o: ho.append(o["hyp_count"]); co.append(o["count"])
    row = dict(n=n, outliers=outl, prosac=prosac, truth=truth, gpu_ms=float(np.median(tg) * 1e3), ref_ms=float(np.median(to) * 1e3) if ho else None,
               gpu_hyp=float(np.median(hg)), ref_hyp=float(np.median(ho)) if ho else None, gpu_count=float(np.median(cg)), ref_count=float(np.median(co)) if co else None,
               gpu_batches=float(np.median(bg)), gpu_lo=float(np.median(lg)))
    rows.append(row); print(json.dumps(row))
json.dump(rows, open("gpurun_out/usac_probe.json", "w"), indent=1)
