"""CPU restatement of poselib::getPoseTriangPts.  TEST INFRASTRUCTURE ONLY: imported by tests/, never by the product.

Follows (paths relative to /root/reference/matchinglib_poselib/source/poselib/source):
  getPoseTriangPts         pose_estim.cpp:913-946
  recoverPose              five-point-nister/five-point.cpp:150-338
  decomposeEssentialMat    five-point-nister/five-point.cpp:340-352
  getTfromTransEssential   pose_helper.cpp:422-433
  cv::triangulatePoints    OpenCV 4.2.0 calib3d/src/triangulate.cpp (third party, not vendored): per point the 4x4 matrix
                           [x*P[2]-P[0]; y*P[2]-P[1]] of both cameras, SVD, last row of Vt.

Pinned by oracle/gen_golden_pose.py, which runs the same steps with cv2's own primitives (cv2.SVDecomp,
cv2.triangulatePoints, OpenCV 4.13 python build) and stores tests/golden/pose_golden.npz; the sign of a homogeneous
solution and the labelling of (R1, R2, +-t) are the SVD's free choices, so parity is stated on the sign-free outputs
(masks, counts, Q = X/W, and the chosen R, t).
"""
import numpy as np


def decompose_essential(E, svd=None):
    """five-point.cpp:340-352.  `svd(E) -> (w, U, Vt)` defaults to numpy's."""
    if svd is None:
        U, w, Vt = np.linalg.svd(E)
    else:
        w, U, Vt = svd(E)
    if np.linalg.det(U) < 0:
        U = -U
    if np.linalg.det(Vt) < 0:
        Vt = -Vt
    W = np.array([[0., 1, 0], [-1, 0, 0], [0, 0, 1]])
    return U @ W @ Vt, U @ W.T @ Vt, U[:, 2:3] * 1.0


def triangulate(P0, P1, x1, x2):
    """cv::triangulatePoints on 2 x n arrays -> 4 x n homogeneous points."""
    n = x1.shape[1]
    A = np.empty((n, 4, 4))
    A[:, 0] = x1[0][:, None] * P0[2] - P0[0]
    A[:, 1] = x1[1][:, None] * P0[2] - P0[1]
    A[:, 2] = x2[0][:, None] * P1[2] - P1[0]
    A[:, 3] = x2[1][:, None] * P1[2] - P1[1]
    _, _, Vt = np.linalg.svd(A)
    return Vt[:, 3, :].T.copy()


def t_from_trans_essential(E):
    """pose_helper.cpp:422-433."""
    nz = lambda d: -1e-3 < d < 1e-3
    assert nz(E[0, 1] / E[1, 0] + 1.0) and nz(E[0, 2] / E[2, 0] + 1.0) and nz(E[1, 2] / E[2, 1] + 1.0)
    t = np.array([[E[1, 2]], [E[2, 0]], [E[0, 1]]])
    n = float(np.linalg.norm(t))
    return t / n if abs(n - 1.0) > 1e-3 else t


def recover_pose(E, p1, p2, mask=None, dist=50.0, translatE=False, triang=triangulate, svd=None):
    """five-point.cpp:150-338.  p1, p2: n x 2.  mask: None (not requested), or an n-vector (its content is used; pass
    ones for "requested but empty").  Returns (good, R, t, Q n x 3, mask or None)."""
    x1, x2 = np.asarray(p1, np.float64).T, np.asarray(p2, np.float64).T
    n = x1.shape[1]
    if translatE:
        R1 = np.eye(3); R2 = None; t = t_from_trans_essential(np.asarray(E, np.float64))
    else:
        R1, R2, t = decompose_essential(np.asarray(E, np.float64), svd)
    P0 = np.eye(3, 4)
    combos = [(R1, t), (R2, t), (R1, -t), (R2, -t)]
    masks, Qs = [], []
    for R, tt in combos:
        if R is None:
            masks.append(np.zeros(n, np.uint8)); Qs.append(None)
            continue
        P = np.hstack([R, tt])
        Q = triang(P0, P, x1, x2)
        m = Q[2] * Q[3] > 0
        q = P @ Q
        m = (q[2] * Q[3] > 0) & m
        with np.errstate(divide="ignore", invalid="ignore"):
            Q = Q.copy(); Q[0] /= Q[3]; Q[1] /= Q[3]; Q[2] /= Q[3]
        m = (Q[2] < dist) & m
        masks.append(m.astype(np.uint8) * 255); Qs.append(Q)
    if mask is not None:
        mk = np.asarray(mask, np.uint8).reshape(-1)
        masks = [mk & m for m in masks]
    good = [int(np.count_nonzero(m)) for m in masks]
    g1, g2, g3, g4 = good
    if g1 >= g2 and g1 >= g3 and g1 >= g4:
        k = 0
    elif g2 and g2 >= g1 and g2 >= g3 and g2 >= g4:
        k = 1
    elif g3 >= g1 and g3 >= g2 and g3 >= g4:
        k = 2
    elif g4:
        k = 3
    else:
        return g4, np.eye(3), np.zeros((3, 1)), np.zeros((n, 3)), (np.zeros(n, np.uint8) if mask is not None else None)
    R, tt = combos[k]
    return good[k], R, tt, Qs[k][:3].T.copy(), (masks[k] if mask is not None else None)


# ---- the driver's pixel -> camera conversion in front of estimateEssentialMat (tests/poselib-test/main.cpp:1522-1596) ----
def img_to_cam(points, K):
    """poselib::ImgToCamCoordTrans (pose_helper.cpp:1100-1109): float32 in, float32 out, arithmetic in double."""
    pts = np.asarray(points, np.float32)
    K = np.asarray(K, np.float64)
    x = ((pts[:, 0].astype(np.float64) - K[0, 2]) / K[0, 0]).astype(np.float32)
    y = ((pts[:, 1].astype(np.float64) - K[1, 2]) / K[1, 1]).astype(np.float32)
    return np.stack([x, y], 1)


def lens_dist_oulu(distorted, dist, iters=10):
    """poselib::LensDist_Oulu (pose_helper.cpp:1232-1279), vectorised over points.  Returns (corrected float32 [n,2], ok [n])."""
    d = np.asarray(dist, np.float64).reshape(-1)
    k1, k2, p1, p2, k3, k4, k5, k6 = d[:8]
    dx = distorted[:, 0].astype(np.float64); dy = distorted[:, 1].astype(np.float64)
    cx = distorted[:, 0].astype(np.float32).copy(); cy = distorted[:, 1].astype(np.float32).copy()

    def terms(cx, cy):
        x = cx.astype(np.float64); y = cy.astype(np.float64)
        r2 = x * x + y * y
        _2xy = 2.0 * x * y
        rad = (1.0 + ((k3 * r2 + k2) * r2 + k1) * r2) / (1.0 + ((k6 * r2 + k5) * r2 + k4) * r2)
        d0 = p1 * _2xy + p2 * (r2 + 2.0 * x * x)
        d1 = p1 * (r2 + 2.0 * y * y) + p2 * _2xy
        return x, y, rad, d0, d1
    for _ in range(iters):
        x, y, rad, d0, d1 = terms(cx, cy)
        cx = ((dx - d0) / rad).astype(np.float32)
        cy = ((dy - d1) / rad).astype(np.float32)
    x, y, rad, d0, d1 = terms(cx, cy)
    px = (x * rad + d0 - dx).astype(np.float32); py = (y * rad + d1 - dy).astype(np.float32)
    ok = ~(np.sqrt(px * px + py * py, dtype=np.float32) > np.float32(0.25))
    return np.stack([cx, cy], 1), ok


def remove_lens_dist(points1, points2, dist1, dist2):
    """poselib::Remove_LensDist (pose_helper.cpp:1169-1228).  Returns (ok, points1, points2, keep mask over the input)."""
    near_zero = lambda v: -1e-3 < v < 1e-3                                     # pose_helper.h:82-87
    n = len(points1)
    if near_zero(float(np.sum(dist1))) and near_zero(float(np.sum(dist2))):
        return True, points1, points2, np.ones(n, bool)
    c1, ok1 = lens_dist_oulu(points1, dist1)
    c2, ok2 = lens_dist_oulu(points2, dist2)
    keep = ok1 & ok2
    if int(keep.sum()) < 16:
        return False, points1, points2, keep
    return True, c1[keep], c2[keep], keep
