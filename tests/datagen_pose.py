"""Seeded synthetic two-view scenes and pre-drawn minimal samples."""
import numpy as np

FLT_EPSILON = float(np.finfo(np.float32).eps)


def rodrigues(axis, angle):
    axis = axis / np.linalg.norm(axis)
    K = np.array([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
    return np.eye(3) + np.sin(angle) * K + (1 - np.cos(angle)) * (K @ K)


def make_scene(n, outlier_frac=0.3, seed=0, noise_px=0.5, f=800.0, th_px=1.6):
    """Random R (angle <= 15 deg), unit t; n 3-D points with depth U[2,20]; normalised image coordinates in
    both cameras; Gaussian noise noise_px/f; a fraction replaced by uniform outliers in a 1392x512 image.
    th = th_px * 4 / (sqrt(2) * (fx0+fy0+fx1+fy1))  (tests/poselib-test/main.cpp:1597-1599)."""
    rng = np.random.default_rng(seed)
    R = rodrigues(rng.standard_normal(3), np.deg2rad(rng.uniform(2, 15)))
    t = rng.standard_normal(3); t /= np.linalg.norm(t)
    w, h = 1392.0 / f, 512.0 / f
    z = rng.uniform(2, 20, n)
    x1 = np.stack([rng.uniform(-w / 2, w / 2, n), rng.uniform(-h / 2, h / 2, n)], 1)
    X = np.concatenate([x1 * z[:, None], z[:, None]], 1)
    X2 = X @ R.T + t
    x2 = X2[:, :2] / X2[:, 2:3]
    p1 = x1 + rng.standard_normal((n, 2)) * noise_px / f
    p2 = x2 + rng.standard_normal((n, 2)) * noise_px / f
    n_out = int(round(outlier_frac * n))
    if n_out:
        sel = rng.choice(n, n_out, replace=False)
        p2[sel] = np.stack([rng.uniform(-w / 2, w / 2, n_out), rng.uniform(-h / 2, h / 2, n_out)], 1)
    th = th_px * 4.0 / (np.sqrt(2.0) * 4.0 * f)
    tx = np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]])
    E = tx @ R
    return dict(p1=np.ascontiguousarray(p1), p2=np.ascontiguousarray(p2), R=R, t=t, E=E / np.linalg.norm(E), th=float(th))


def _subset_ok(pts):
    """CvModelEstimator3::checkSubset (modelest.cpp:613-650) for the LAST point of the prefix."""
    i = len(pts) - 1
    for j in range(i):
        dx1 = pts[j][0] - pts[i][0]; dy1 = pts[j][1] - pts[i][1]
        for k in range(j):
            dx2 = pts[k][0] - pts[i][0]; dy2 = pts[k][1] - pts[i][1]
            if abs(dx2 * dy1 - dy2 * dx1) <= FLT_EPSILON * (abs(dx1) + abs(dy1) + abs(dx2) + abs(dy2)):
                return False
    return True


def draw_samples(p1, p2, h, seed=0):
    """h rows of 5 distinct indices, each prefix passing the reference's collinearity test in both images
    (getSubset, modelest.cpp:567-610) so that either side would accept every row."""
    rng = np.random.default_rng(seed)
    n = p1.shape[0]
    out = np.empty((h, 5), np.int32)
    for r in range(h):
        while True:
            idx = []
            fail = False
            for _ in range(300):
                c = int(rng.integers(0, n))
                if c in idx:
                    continue
                cand = idx + [c]
                if _subset_ok(p1[cand]) and _subset_ok(p2[cand]):
                    idx = cand
                    if len(idx) == 5:
                        break
            else:
                fail = True
            if not fail and len(idx) == 5:
                break
        out[r] = idx
    return out


def draw_samples_fast(n, h, seed=0):
    """Bulk version for the benchmark sizes (H = 100 000): distinct indices by rejection, no collinearity
    screening (random continuous scenes never trip the FLT_EPSILON test)."""
    rng = np.random.default_rng(seed)
    s = rng.integers(0, n, (h, 5), dtype=np.int64)
    while True:
        srt = np.sort(s, 1)
        bad = (srt[:, 1:] == srt[:, :-1]).any(1)
        if not bad.any():
            break
        s[bad] = rng.integers(0, n, (int(bad.sum()), 5))
    return s.astype(np.int32)
