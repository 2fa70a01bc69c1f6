"""Experiment (VERDICT r1 item 10 ii): cv::solvePoly's Durand-Kerner loop with an early exit (every update of a sweep below
2^-40 relative) against the reference's 300 sweeps, over >= 10^6 minimal samples of the C3 scene: real-root-set mismatches,
model differences, inlier-count differences, final (count, iterations, mask) of RANSAC runs, and the solve-time gain.
    python tools/solver_early_exit.py > profiles/r2_solver_early_exit.json"""
import ctypes, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import datagen_pose
import matchinglib_poselib_b200 as m
from matchinglib_poselib_b200 import poselib

lib = m._lib.load(); m._lib.check(lib.mpl_init(0), "init")
n, H, CH = 10000, 1_000_000, 100_000
sc = datagen_pose.make_scene(n, 0.3, seed=1003)


def solve_ms():
    ms, k = ctypes.c_double(0), ctypes.c_longlong(0)
    lib.mpl_prof_read(5, ctypes.byref(ms), ctypes.byref(k))
    return ms.value / max(k.value, 1)


tot = dict(samples=0, models_exact=0, rootset_mismatch_samples=0, models_compared=0, max_model_diff=0.0, models_diff_gt_1e12=0,
           count_mismatch_models=0, max_count_diff=0)
t_exact, t_early = [], []
lib.mpl_prof_enable(1)
for c in range(H // CH):
    s = datagen_pose.draw_samples_fast(n, CH, seed=5000 + c)
    lib.mpl_set_solver_early_exit(0); solve_ms()
    a = poselib.ransac5pt(sc["p1"], sc["p2"], s, sc["th"], max_iters=CH, trace=True); t_exact.append(solve_ms())
    lib.mpl_set_solver_early_exit(1)
    b = poselib.ransac5pt(sc["p1"], sc["p2"], s, sc["th"], max_iters=CH, trace=True); t_early.append(solve_ms())
    lib.mpl_set_solver_early_exit(0)
    same = a["n_models"] == b["n_models"]
    tot["samples"] += CH; tot["models_exact"] += int(a["n_models"].sum()); tot["rootset_mismatch_samples"] += int((~same).sum())
    oa = np.concatenate([[0], np.cumsum(a["n_models"])]); ob = np.concatenate([[0], np.cumsum(b["n_models"])])
    # samples with equal root counts: models correspond slot by slot (same root order)
    ia = np.concatenate([np.arange(oa[i], oa[i + 1]) for i in np.nonzero(same)[0]]) if same.any() else np.zeros(0, int)
    ib = np.concatenate([np.arange(ob[i], ob[i + 1]) for i in np.nonzero(same)[0]]) if same.any() else np.zeros(0, int)
    d = np.minimum(np.linalg.norm(a["models"][ia] - b["models"][ib], axis=1), np.linalg.norm(a["models"][ia] + b["models"][ib], axis=1))
    tot["models_compared"] += len(ia); tot["max_model_diff"] = max(tot["max_model_diff"], float(d.max()))
    tot["models_diff_gt_1e12"] += int((d > 1e-12).sum())
    dc = np.abs(a["counts"][ia].astype(int) - b["counts"][ib].astype(int))
    tot["count_mismatch_models"] += int((dc > 0).sum()); tot["max_count_diff"] = max(tot["max_count_diff"], int(dc.max()))
# whole RANSAC runs as estimateEssentialMat drives them (adaptive stop, refit): 200 scenes, exact vs early
runs = dict(scenes=0, count_iters_mask_equal=0, max_E_diff=0.0)
for k in range(200):
    nn = 300 + 41 * k
    s2 = datagen_pose.make_scene(nn, 0.1 + 0.003 * k, seed=7000 + k)
    smp = datagen_pose.draw_samples_fast(nn, 300, seed=k)
    lib.mpl_set_solver_early_exit(0); x = poselib.ransac5pt(s2["p1"], s2["p2"], smp, s2["th"], max_iters=300, refine=True)
    lib.mpl_set_solver_early_exit(1); y = poselib.ransac5pt(s2["p1"], s2["p2"], smp, s2["th"], max_iters=300, refine=True)
    lib.mpl_set_solver_early_exit(0)
    runs["scenes"] += 1
    eq = x["count"] == y["count"] and x["iters"] == y["iters"] and np.array_equal(x["mask"], y["mask"])
    runs["count_iters_mask_equal"] += int(eq)
    if x["ok"] and y["ok"]:
        runs["max_E_diff"] = max(runs["max_E_diff"], float(min(np.linalg.norm(x["E"] - y["E"]), np.linalg.norm(x["E"] + y["E"]))))
print(json.dumps(dict(experiment="Durand-Kerner early exit (update < 2^-40 relative) vs 300 sweeps", totals=tot, ransac_runs=runs,
                      solve_ms_per_100k_samples=dict(exact=float(np.median(t_exact)), early=float(np.median(t_early))),
                      speedup=float(np.median(t_exact) / np.median(t_early)))))
