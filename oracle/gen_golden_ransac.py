"""Generate tests/golden/ransac_golden.npz by running the REFERENCE's 5-point RANSAC arithmetic in this
container (needs /root/reference and cv2; the GPU box needs neither).

Fidelity choices
  * the generated coefficient table of CvEMEstimator::getCoeffMat (five-point.cpp:603-824) and the
    polynomial coefficient lines of run5Point (:416-428) are READ FROM /root/reference AT RUN TIME and
    exec-ed against numpy arrays (they are valid Python once the trailing ';' is dropped) -- nothing is
    copied into this repository;
  * SVD / inverse / solvePoly / solveZ go through the same OpenCV primitives via cv2
    (cv2.SVDecomp(FULL_UV), cv2.invert, cv2.solvePoly) as five-point.cpp:385-455 does in C++;
  * Sampson error follows five-point.cpp:476-503 in the k-ascending scalar order that OpenCV's small gemm
    was verified to produce, then (float) cast, then  err <= th*th  (modelest.cpp:69-83);
  * the hypothesise-and-verify loop, selection and tie-break rules and cvRANSACUpdateNumIters1 follow
    modelest.cpp:86-109, 343-474 with PRE-DRAWN samples instead of std::rand() (getSubset :567-610).
"""
import os
import re
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cv2  # noqa: E402
import datagen_pose  # noqa: E402

REF = "/root/reference/matchinglib_poselib/source/poselib/source/five-point-nister/five-point.cpp"


def _load_tables():
    lines = open(REF).read().split("\n")
    a_lines = [l.strip().rstrip(";") for l in lines[611:812] if re.match(r"\s*A\[\d+\]=", l)]
    c_lines = [l.strip().rstrip(";") for l in lines[415:430] if re.match(r"\s*c\[\d+\]\s*=", l)]
    perm_line = [l for l in lines[805:820] if "int perm[20]" in l][0]
    perm = [int(v) for v in re.search(r"\{(.*)\}", perm_line).group(1).split(",")]
    assert len(a_lines) == 200 and len(c_lines) == 11 and len(perm) == 20
    return compile("\n".join(a_lines), "getCoeffMat", "exec"), compile("\n".join(c_lines), "coeffs", "exec"), perm


A_CODE, C_CODE, PERM = _load_tables()


def get_coeff_mat(e):
    ep2 = e * e
    ep3 = ep2 * e
    A = np.zeros(200)
    exec(A_CODE, {"e": e, "ep2": ep2, "ep3": ep3, "A": A})
    AA = np.zeros(200)
    for i in range(20):
        for j in range(10):
            AA[i + j * 20] = A[PERM[i] + j * 20]
    return AA.reshape(10, 20)


def run5point(q1, q2, debug=None):
    """CvEMEstimator::run5Point (five-point.cpp:366-471).  q1, q2: n x 2 float64.  Returns list of 3x3 E."""
    n = q1.shape[0]
    Q = np.empty((n, 9))
    Q[:, 0] = q1[:, 0] * q2[:, 0]; Q[:, 1] = q1[:, 1] * q2[:, 0]; Q[:, 2] = q2[:, 0]
    Q[:, 3] = q1[:, 0] * q2[:, 1]; Q[:, 4] = q1[:, 1] * q2[:, 1]; Q[:, 5] = q2[:, 1]
    Q[:, 6] = q1[:, 0]; Q[:, 7] = q1[:, 1]; Q[:, 8] = 1.0
    _, _, vt = cv2.SVDecomp(Q, flags=cv2.SVD_FULL_UV)
    EE = vt[5:9].copy()                      # rows = EE.t() of the reference (4 x 9)
    A = get_coeff_mat(EE.reshape(-1))
    ok, inv = cv2.invert(np.ascontiguousarray(A[:, :10]))
    A = cv2.gemm(inv, np.ascontiguousarray(A[:, 10:]), 1.0, None, 0.0)
    b = np.zeros((3, 13))
    for i in range(3):
        r1 = np.zeros(13); r2 = np.zeros(13)
        a1, a2 = A[i * 2 + 4], A[i * 2 + 5]
        r1[1:4] = a1[0:3]; r1[5:8] = a1[3:6]; r1[9:13] = a1[6:10]
        r2[0:3] = a2[0:3]; r2[4:7] = a2[3:6]; r2[8:12] = a2[6:10]
        b[i] = r1 - r2
    bf = b.reshape(-1)
    c = np.zeros(11)
    exec(C_CODE, {"b": bf, "c": c})
    _, roots = cv2.solvePoly(c.reshape(1, 11))
    roots = roots.reshape(-1, 2)
    if debug is not None:
        debug.update(EE=EE, B=b.copy(), c=c.copy(), roots=roots.copy())
    models = []
    for re_, im_ in roots:
        if abs(im_) > 1e-10:
            continue
        z1 = re_; z2 = z1 * z1; z3 = z2 * z1; z4 = z3 * z1
        bz = np.empty((3, 3))
        for j in range(3):
            br = b[j]
            bz[j, 0] = br[0] * z3 + br[1] * z2 + br[2] * z1 + br[3]
            bz[j, 1] = br[4] * z3 + br[5] * z2 + br[6] * z1 + br[7]
            bz[j, 2] = br[8] * z4 + br[9] * z3 + br[10] * z2 + br[11] * z1 + br[12]
        _, _, vt3 = cv2.SVDecomp(bz)
        xy1 = vt3[-1]
        if abs(xy1[2]) < 1e-10:
            continue
        x = xy1[0] / xy1[2]; y = xy1[1] / xy1[2]
        Evec = EE[0] * x + EE[1] * y + EE[2] * z1 + EE[3]
        Evec = Evec / cv2.norm(Evec)
        models.append(Evec.reshape(3, 3).copy())
    return models


def sampson_err(E, p1, p2):
    """computeReprojError3 (five-point.cpp:476-503): float32 error per correspondence."""
    x1, y1, x2, y2 = p1[:, 0], p1[:, 1], p2[:, 0], p2[:, 1]
    Ex1 = [(E[k, 0] * x1 + E[k, 1] * y1) + E[k, 2] * 1.0 for k in range(3)]
    Etx2 = [(E[0, k] * x2 + E[1, k] * y2) + E[2, k] * 1.0 for k in range(2)]
    r = (x2 * Ex1[0] + y2 * Ex1[1]) + 1.0 * Ex1[2]
    a = Ex1[0] * Ex1[0]; b = Ex1[1] * Ex1[1]; c = Etx2[0] * Etx2[0]; d = Etx2[1] * Etx2[1]
    return (r * r / (a + b + c + d)).astype(np.float32)


def update_num_iters(p, ep, model_points, max_iters):
    """cvRANSACUpdateNumIters1 (modelest.cpp:86-109)."""
    p = min(max(p, 0.0), 1.0); ep = min(max(ep, 0.0), 1.0)
    num = max(1.0 - p, np.finfo(float).tiny)
    denom = 1.0 - (1.0 - ep) ** model_points
    if denom < np.finfo(float).tiny:
        return 0
    num = np.log(num); denom = np.log(denom)
    if denom >= 0 or -num >= max_iters * (-denom):
        return max_iters
    return int(np.round(num / denom))      # std::round: half away from zero; num/denom > 0 here


def run_ransac(p1, p2, samples, th, conf=0.999, max_iters=1000, refine=False, trace=None):
    """CvModelEstimator3::runRANSAC (modelest.cpp:343-474) with pre-drawn samples."""
    N = p1.shape[0]
    niters = max_iters
    best_count = 0; best_E = None; best_mask = None; errminsum = np.finfo(float).max
    th2 = th * th
    it = 0
    while it < niters and it < len(samples):
        s = samples[it]
        models = run5point(p1[s], p2[s])
        if trace is not None:
            trace["n_models"].append(len(models))
        for E in models:
            err = sampson_err(E, p1, p2)
            mask = err.astype(np.float64) <= th2
            good = int(mask.sum())
            esum = float(cv2.sumElems(err.reshape(1, -1))[0])
            if trace is not None:
                trace["models"].append(E.copy()); trace["counts"].append(good); trace["errsums"].append(esum)
            if good > max(best_count, 4):
                best_mask = mask; best_E = E.copy(); best_count = good
                niters = update_num_iters(conf, (N - good) / N, 5, niters)
                errminsum = esum
            elif good == max(best_count, 5) and errminsum < np.finfo(float).max and errminsum > esum:
                best_mask = mask; best_E = E.copy(); errminsum = esum
        it += 1
    iters_run = it
    if refine and best_count > 0:
        inl = np.nonzero(best_mask)[0]
        models = run5point(p1[inl], p2[inl])
        for E in models:
            err = sampson_err(E, p1, p2)
            mask = err.astype(np.float64) <= th2
            good = int(mask.sum())
            esum = float(cv2.sumElems(err.reshape(1, -1))[0])
            if good > max(best_count, 4):
                best_mask = mask; best_E = E.copy(); best_count = good; errminsum = esum
            elif good == best_count and errminsum < np.finfo(float).max and errminsum > esum:
                best_mask = mask; best_E = E.copy(); errminsum = esum
    return best_count > 0, best_E, best_mask, best_count, iters_run


def main():
    out = {}
    # (1) solver + scoring trace on a small scene: every model of every sample, counts, error sums
    for name, N, H, outl, seed in [("small", 300, 60, 0.3, 7), ("mid", 2000, 40, 0.45, 8), ("clean", 150, 30, 0.0, 9)]:
        sc = datagen_pose.make_scene(N, outlier_frac=outl, seed=seed)
        samples = datagen_pose.draw_samples(sc["p1"], sc["p2"], H, seed=seed + 100)
        trace = {"n_models": [], "models": [], "counts": [], "errsums": []}
        ok, E, mask, cnt, iters = run_ransac(sc["p1"], sc["p2"], samples, sc["th"], max_iters=H, trace=trace)
        # trace covers only the iterations the adaptive stop allowed; complete it for the remaining samples
        full = {"n_models": [], "models": [], "counts": [], "errsums": []}
        for s in samples:
            ms = run5point(sc["p1"][s], sc["p2"][s])
            full["n_models"].append(len(ms))
            for Em in ms:
                err = sampson_err(Em, sc["p1"], sc["p2"])
                full["models"].append(Em); full["counts"].append(int((err.astype(np.float64) <= sc["th"] ** 2).sum()))
                full["errsums"].append(float(cv2.sumElems(err.reshape(1, -1))[0]))
        out.update({f"{name}_p1": sc["p1"], f"{name}_p2": sc["p2"], f"{name}_th": np.float64(sc["th"]),
                    f"{name}_samples": samples, f"{name}_R": sc["R"], f"{name}_t": sc["t"],
                    f"{name}_n_models": np.array(full["n_models"], np.int32), f"{name}_models": np.array(full["models"]).reshape(-1, 9),
                    f"{name}_counts": np.array(full["counts"], np.int32), f"{name}_errsums": np.array(full["errsums"])})
        for refine in (False, True):
            ok, E, mask, cnt, iters = run_ransac(sc["p1"], sc["p2"], samples, sc["th"], max_iters=1000, refine=refine)
            tag = f"{name}_ref{int(refine)}"
            out.update({tag + "_E": E.reshape(9), tag + "_mask": mask.astype(np.uint8), tag + "_count": np.int32(cnt), tag + "_iters": np.int32(iters)})
            print(name, "refine", refine, "count", cnt, "iters", iters, "models/sample", np.mean(full["n_models"]))
    # (2) solver internals for a handful of samples (B matrix, polynomial, roots) to pin the restated algebra
    sc = datagen_pose.make_scene(100, outlier_frac=0.0, seed=21)
    samples = datagen_pose.draw_samples(sc["p1"], sc["p2"], 8, seed=22)
    Bs, cs, rs = [], [], []
    for s in samples:
        dbg = {}
        run5point(sc["p1"][s], sc["p2"][s], dbg)
        Bs.append(dbg["B"]); cs.append(dbg["c"]); rs.append(dbg["roots"])
    out.update(dbg_p1=sc["p1"], dbg_p2=sc["p2"], dbg_samples=samples, dbg_B=np.array(Bs), dbg_c=np.array(cs), dbg_roots=np.array(rs))
    # (3) cv2.solvePoly known answers (pins the restated Durand-Kerner iteration, incl. root ORDER)
    rng = np.random.default_rng(33)
    polys = rng.standard_normal((20, 11))
    polys[5, 10] = 0.0; polys[6, 9:] = 0.0                 # degenerate leading coefficients
    roots = np.array([cv2.solvePoly(p.reshape(1, 11))[1].reshape(-1, 2) for p in polys])
    out.update(poly_coeffs=polys, poly_roots=roots)
    path = os.path.join(ROOT, "tests", "golden", "ransac_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path))


if __name__ == "__main__":
    main()
