"""Seeded keypoint / match sets for the GMS filter: a smooth image motion (translation, optional rotation / scale) with
pixel noise carries a fraction of the matches, the rest are random pairings."""
import numpy as np


def scene(n, seed, W=1392, H=512, frac=0.6, rot=0.0, sc=1.0, border=False):
    rng = np.random.default_rng(seed)
    k1 = np.stack([rng.uniform(0, W - 1, n), rng.uniform(0, H - 1, n)], 1).astype(np.float32)
    c, s = np.cos(rot), np.sin(rot)
    ctr = np.array([W / 2, H / 2])
    k2 = ((k1 - ctr) @ np.array([[c, -s], [s, c]]).T) * sc + ctr + np.array([15.0, -7.0]) + rng.normal(0, 1.0, (n, 2))
    k2 = np.clip(k2, 0, [W - 1.001, H - 1.001]).astype(np.float32)
    if border:                                                   # keypoints on the last pixel row / column (shifted grids fall off)
        k1[:20, 0] = W - 1; k1[20:40, 1] = H - 1
    out = rng.random(n) > frac
    perm = rng.permutation(n)
    tr = np.arange(n); tr[out] = perm[out]
    keep = rng.random(n) < 0.8                                   # not every keypoint has a match
    m = np.stack([np.arange(n), tr], 1).astype(np.int32)[keep]
    return k1, k2, m, (W, H)


CASES = {
    "stereo3000": dict(scene=dict(n=3000, seed=0), use_scale=False, use_rotation=False),
    "stereo8000": dict(scene=dict(n=8000, seed=1, border=True), use_scale=False, use_rotation=False),
    "few": dict(scene=dict(n=60, seed=2), use_scale=False, use_rotation=False),
    "rotated": dict(scene=dict(n=4000, seed=3, rot=0.8), use_scale=False, use_rotation=True),
    "scaled": dict(scene=dict(n=4000, seed=4, sc=1.4), use_scale=True, use_rotation=False),
    "both": dict(scene=dict(n=3000, seed=5, rot=0.5, sc=0.7, frac=0.8), use_scale=True, use_rotation=True),
}


def dmatch(m):
    """[n, 2] (queryIdx, trainIdx) -> cv::DMatch records"""
    d = np.zeros(len(m), np.dtype([("queryIdx", "<i4"), ("trainIdx", "<i4"), ("imgIdx", "<i4"), ("distance", "<f4")]))
    d["queryIdx"], d["trainIdx"], d["imgIdx"] = m[:, 0], m[:, 1], -1
    return d
