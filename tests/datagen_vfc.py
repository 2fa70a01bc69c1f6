"""Synthetic match sets for the VFC filter: a smooth motion field (similarity + a gentle quadratic warp) with pixel noise for the
inliers, uniformly random right points for the outliers."""
import numpy as np


def scene(n=3000, outlier_frac=0.4, seed=0, w=1392, h=512, noise_px=0.7, shuffle=True):
    rng = np.random.default_rng(seed)
    kp1 = np.stack([rng.uniform(0, w, n), rng.uniform(0, h, n)], 1)
    ang = np.deg2rad(rng.uniform(-4, 4)); s = rng.uniform(0.95, 1.05)
    c, si = np.cos(ang) * s, np.sin(ang) * s
    d = kp1 - np.array([w / 2, h / 2])
    warp = 2e-5 * np.stack([d[:, 0] * d[:, 1], d[:, 0] ** 2 - d[:, 1] ** 2], 1)
    kp2 = np.stack([c * d[:, 0] - si * d[:, 1], si * d[:, 0] + c * d[:, 1]], 1) + np.array([w / 2, h / 2]) + rng.uniform(-25, 25, 2) + warp
    kp2 += rng.standard_normal((n, 2)) * noise_px
    out = np.zeros(n, bool)
    out[rng.choice(n, int(round(outlier_frac * n)), replace=False)] = True
    kp2[out] = np.stack([rng.uniform(0, w, out.sum()), rng.uniform(0, h, out.sum())], 1)
    perm = rng.permutation(n) if shuffle else np.arange(n)
    kp2s = np.empty_like(kp2); kp2s[perm] = kp2                           # train keypoint perm[i] is the partner of query keypoint i
    matches = np.stack([np.arange(n), perm], 1).astype(np.int32)
    return kp1.astype(np.float32), kp2s.astype(np.float32), matches, ~out
