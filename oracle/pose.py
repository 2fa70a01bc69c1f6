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
