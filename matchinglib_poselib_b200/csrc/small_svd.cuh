// small_svd.cuh -- one-sided Jacobi SVD of small square matrices (host + device), shared by recover_pose.cu and arrsac.cu.
#pragma once
#include <cuda_runtime.h>
#include <math.h>

// One-sided (Hestenes) Jacobi SVD of an N x N matrix, the method cv::SVD::compute uses for small matrices: the rows of
// At (= columns of A) are rotated until mutually orthogonal; V accumulates the rotations.  Returns in `x` the right
// singular vector of the smallest singular value (unit norm; the sign is arbitrary, callers only use sign-free
// combinations).  Works on A^T A's eigen-structure without forming it, so small singular values keep full accuracy.
template <int N>
__host__ __device__ inline void jacobi_null_vector(const double (&A)[N][N], double (&x)[N], double (*Vout)[N] = nullptr, double* wout = nullptr)
{
    double U[N][N], V[N][N];                     // U[j] = column j of A (rotated), V[j] = right singular vector j
    for (int j = 0; j < N; ++j)
        for (int i = 0; i < N; ++i) { U[j][i] = A[i][j]; V[j][i] = (i == j) ? 1.0 : 0.0; }
    for (int sweep = 0; sweep < 40; ++sweep) {
        bool changed = false;
        for (int p = 0; p < N - 1; ++p)
            for (int q = p + 1; q < N; ++q) {
                double a = 0, b = 0, c = 0;
                for (int i = 0; i < N; ++i) { a += U[p][i] * U[p][i]; b += U[q][i] * U[q][i]; c += U[p][i] * U[q][i]; }
                if (fabs(c) <= 2.220446049250313e-16 * sqrt(a * b) || c == 0.0) continue;
                changed = true;
                const double zeta = (b - a) / (2.0 * c);
                const double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                const double cs = 1.0 / sqrt(1.0 + t * t), sn = cs * t;
                for (int i = 0; i < N; ++i) {
                    const double up = U[p][i], uq = U[q][i];
                    U[p][i] = cs * up - sn * uq; U[q][i] = sn * up + cs * uq;
                    const double vp = V[p][i], vq = V[q][i];
                    V[p][i] = cs * vp - sn * vq; V[q][i] = sn * vp + cs * vq;
                }
            }
        if (!changed) break;
    }
    int best = 0; double wbest = 0;
    for (int j = 0; j < N; ++j) {
        double w = 0;
        for (int i = 0; i < N; ++i) w += U[j][i] * U[j][i];
        if (wout) wout[j] = sqrt(w);
        if (j == 0 || w < wbest) { wbest = w; best = j; }
    }
    for (int i = 0; i < N; ++i) x[i] = V[best][i];
    if (Vout) for (int j = 0; j < N; ++j) for (int i = 0; i < N; ++i) Vout[j][i] = V[j][i];
}

