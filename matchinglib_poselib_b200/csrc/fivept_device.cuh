// fivept_device.cuh -- per-thread Nister 5-point solver in fp64 (device functions).
//
// Follows the ALGORITHM of CvEMEstimator::run5Point (five-point.cpp:366-471 of the reference's
// poselib/source/five-point-nister) but is written for one CUDA thread per minimal sample:
//   1. null space of the 5x9 epipolar constraint matrix by Householder QR of its transpose
//      (the reference takes the last four right-singular vectors of an SVD; any orthonormal basis of the
//      same null space yields the same essential matrices),
//   2. the ten cubic constraints det(E) = 0, 2 E E^T E - tr(E E^T) E = 0 expanded by polynomial arithmetic
//      with compile-time monomial tables (the reference evaluates a generated table, five-point.cpp:603-824),
//      in Nister's monomial order,
//   3. Gauss-Jordan elimination of the first ten columns (reference: 10x10 inverse times the rest, :394),
//   4. B(z) (3x3 polynomial matrix) and its determinant, a degree-10 polynomial (:396-428),
//   5. roots by the Durand-Kerner iteration of cv::solvePoly (OpenCV 4.2.0 core/src/mathfuncs.cpp: start
//      roots (1+i)^k, in-place updates, at most 300 sweeps) with an exit as soon as a sweep leaves every
//      root bit-identical -- from that point the remaining sweeps of the reference are no-ops, so the root
//      set, the root order and the |imag| <= 1e-10 acceptance rule are the reference's,
//   6. per real root: null vector of B(z) (cross products instead of SVD::solveZ), reject |w| < 1e-10,
//      E = x*E0 + y*E1 + z*E2 + E3, normalised to unit Frobenius norm (:440-464).
#pragma once
#include <cuda_runtime.h>
#include <math.h>

namespace fivept {

// ---- compile-time monomial tables -----------------------------------------------------------------
// linear   l[4]  : x y z 1
// quadratic q[10]: xx yy zz xy xz yz x y z 1
// cubic    c[20] : Nister order  x3 y3 x2y xy2 x2z x2 y2z y2 xyz xy | xz2 xz x yz2 yz y z3 z2 z 1
struct Exp { int x, y, z; };
__host__ __device__ constexpr Exp lin_exp(int i) { return i == 0 ? Exp{1,0,0} : i == 1 ? Exp{0,1,0} : i == 2 ? Exp{0,0,1} : Exp{0,0,0}; }
__host__ __device__ constexpr Exp quad_exp(int i)
{
    return i == 0 ? Exp{2,0,0} : i == 1 ? Exp{0,2,0} : i == 2 ? Exp{0,0,2} : i == 3 ? Exp{1,1,0} : i == 4 ? Exp{1,0,1} :
           i == 5 ? Exp{0,1,1} : i == 6 ? Exp{1,0,0} : i == 7 ? Exp{0,1,0} : i == 8 ? Exp{0,0,1} : Exp{0,0,0};
}
__host__ __device__ constexpr Exp cub_exp(int i)
{
    return i == 0 ? Exp{3,0,0} : i == 1 ? Exp{0,3,0} : i == 2 ? Exp{2,1,0} : i == 3 ? Exp{1,2,0} : i == 4 ? Exp{2,0,1} :
           i == 5 ? Exp{2,0,0} : i == 6 ? Exp{0,2,1} : i == 7 ? Exp{0,2,0} : i == 8 ? Exp{1,1,1} : i == 9 ? Exp{1,1,0} :
           i == 10 ? Exp{1,0,2} : i == 11 ? Exp{1,0,1} : i == 12 ? Exp{1,0,0} : i == 13 ? Exp{0,1,2} : i == 14 ? Exp{0,1,1} :
           i == 15 ? Exp{0,1,0} : i == 16 ? Exp{0,0,3} : i == 17 ? Exp{0,0,2} : i == 18 ? Exp{0,0,1} : Exp{0,0,0};
}
__host__ __device__ constexpr int quad_index(Exp e)
{
    for (int i = 0; i < 10; ++i) { Exp q = quad_exp(i); if (q.x == e.x && q.y == e.y && q.z == e.z) return i; }
    return -1;
}
__host__ __device__ constexpr int cub_index(Exp e)
{
    for (int i = 0; i < 20; ++i) { Exp q = cub_exp(i); if (q.x == e.x && q.y == e.y && q.z == e.z) return i; }
    return -1;
}
__host__ __device__ constexpr int LL(int i, int j) { return quad_index(Exp{lin_exp(i).x + lin_exp(j).x, lin_exp(i).y + lin_exp(j).y, lin_exp(i).z + lin_exp(j).z}); }
__host__ __device__ constexpr int QL(int i, int j) { return cub_index(Exp{quad_exp(i).x + lin_exp(j).x, quad_exp(i).y + lin_exp(j).y, quad_exp(i).z + lin_exp(j).z}); }

template <int I, int J> struct LLc { static constexpr int v = LL(I, J); };
template <int I, int J> struct QLc { static constexpr int v = QL(I, J); };

// q += s * a * b   (linear x linear -> quadratic), fully unrolled with compile-time target indices
template <int I = 0, int J = 0>
__device__ __forceinline__ void mul_ll(const double* a, const double* b, double s, double* q)
{
    if constexpr (I < 4) {
        q[LLc<I, J>::v] = fma(s * a[I], b[J], q[LLc<I, J>::v]);
        if constexpr (J < 3) mul_ll<I, J + 1>(a, b, s, q); else mul_ll<I + 1, 0>(a, b, s, q);
    }
}
// c += s * q * l   (quadratic x linear -> cubic)
template <int I = 0, int J = 0>
__device__ __forceinline__ void mul_ql(const double* q, const double* l, double s, double* c)
{
    if constexpr (I < 10) {
        c[QLc<I, J>::v] = fma(s * q[I], l[J], c[QLc<I, J>::v]);
        if constexpr (J < 3) mul_ql<I, J + 1>(q, l, s, c); else mul_ql<I + 1, 0>(q, l, s, c);
    }
}

// ---- step 1: null space of the 5x9 constraint matrix ----------------------------------------------
// rows of Q: [x1x2, y1x2, x2, x1y2, y1y2, y2, x1, y1, 1]  (five-point.cpp:375-383)
// ee[b][k]: basis vector b (0..3 <-> x, y, z, 1), component k (row-major E).
__device__ inline void null_space_5x9(const double* x1, const double* y1, const double* x2, const double* y2, double ee[4][9])
{
    double M[9][5];                       // Q^T, columns = constraints
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        M[0][j] = x1[j] * x2[j]; M[1][j] = y1[j] * x2[j]; M[2][j] = x2[j];
        M[3][j] = x1[j] * y2[j]; M[4][j] = y1[j] * y2[j]; M[5][j] = y2[j];
        M[6][j] = x1[j];         M[7][j] = y1[j];         M[8][j] = 1.0;
    }
    double v[5][9], beta[5];              // Householder vectors
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        double nrm2 = 0;
#pragma unroll
        for (int i = 0; i < 9; ++i) if (i >= j) nrm2 += M[i][j] * M[i][j];
        const double nrm = sqrt(nrm2);
        const double alpha = M[j][j] >= 0 ? -nrm : nrm;
        double vn2 = 0;
#pragma unroll
        for (int i = 0; i < 9; ++i) {
            v[j][i] = i < j ? 0.0 : (i == j ? M[i][j] - alpha : M[i][j]);
            vn2 += v[j][i] * v[j][i];
        }
        beta[j] = vn2 > 0 ? 2.0 / vn2 : 0.0;
#pragma unroll
        for (int c = 0; c < 5; ++c) {
            if (c <= j) continue;
            double d = 0;
#pragma unroll
            for (int i = 0; i < 9; ++i) if (i >= j) d += v[j][i] * M[i][c];
            d *= beta[j];
#pragma unroll
            for (int i = 0; i < 9; ++i) if (i >= j) M[i][c] -= d * v[j][i];
        }
    }
    // null basis: H0 H1 H2 H3 H4 e_{5+b}
#pragma unroll
    for (int b = 0; b < 4; ++b) {
        double e[9];
#pragma unroll
        for (int i = 0; i < 9; ++i) e[i] = (i == 5 + b) ? 1.0 : 0.0;
#pragma unroll
        for (int j = 4; j >= 0; --j) {
            double d = 0;
#pragma unroll
            for (int i = 0; i < 9; ++i) d += v[j][i] * e[i];
            d *= beta[j];
#pragma unroll
            for (int i = 0; i < 9; ++i) e[i] -= d * v[j][i];
        }
#pragma unroll
        for (int i = 0; i < 9; ++i) ee[b][i] = e[i];
    }
}

// ---- step 2: 10 x 20 constraint matrix -------------------------------------------------------------
__device__ inline void constraint_matrix(const double ee[4][9], double A[10][20])
{
    double L[9][4];                        // E entries as linear polynomials (x, y, z, 1)
#pragma unroll
    for (int k = 0; k < 9; ++k) { L[k][0] = ee[0][k]; L[k][1] = ee[1][k]; L[k][2] = ee[2][k]; L[k][3] = ee[3][k]; }
#pragma unroll
    for (int r = 0; r < 10; ++r)
#pragma unroll
        for (int c = 0; c < 20; ++c) A[r][c] = 0.0;
    // det(E) = E0 (E4 E8 - E5 E7) - E1 (E3 E8 - E5 E6) + E2 (E3 E7 - E4 E6)
    {
        double m[10];
#pragma unroll
        for (int i = 0; i < 10; ++i) m[i] = 0;
        mul_ll(L[4], L[8], 1.0, m); mul_ll(L[5], L[7], -1.0, m); mul_ql(m, L[0], 1.0, A[0]);
#pragma unroll
        for (int i = 0; i < 10; ++i) m[i] = 0;
        mul_ll(L[3], L[8], 1.0, m); mul_ll(L[5], L[6], -1.0, m); mul_ql(m, L[1], -1.0, A[0]);
#pragma unroll
        for (int i = 0; i < 10; ++i) m[i] = 0;
        mul_ll(L[3], L[7], 1.0, m); mul_ll(L[4], L[6], -1.0, m); mul_ql(m, L[2], 1.0, A[0]);
    }
    // G = E E^T (symmetric, quadratic entries), tr = trace
    double G[6][10];                       // 00 01 02 11 12 22
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
        for (int j = 0; j < 10; ++j) G[i][j] = 0;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        mul_ll(L[0 + k], L[0 + k], 1.0, G[0]); mul_ll(L[0 + k], L[3 + k], 1.0, G[1]); mul_ll(L[0 + k], L[6 + k], 1.0, G[2]);
        mul_ll(L[3 + k], L[3 + k], 1.0, G[3]); mul_ll(L[3 + k], L[6 + k], 1.0, G[4]); mul_ll(L[6 + k], L[6 + k], 1.0, G[5]);
    }
    double tr[10];
#pragma unroll
    for (int j = 0; j < 10; ++j) tr[j] = G[0][j] + G[3][j] + G[5][j];
    // rows 1..9: (2 G - tr I) E
    constexpr int gi[3][3] = {{0, 1, 2}, {1, 3, 4}, {2, 4, 5}};
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            double* row = A[1 + 3 * i + j];
#pragma unroll
            for (int k = 0; k < 3; ++k) mul_ql(G[gi[i][k]], L[3 * k + j], 2.0, row);
            mul_ql(tr, L[3 * i + j], -1.0, row);
        }
}

// ---- step 3: Gauss-Jordan on the first ten columns ---------------------------------------------------
__device__ inline bool eliminate10(double A[10][20])
{
    for (int c = 0; c < 10; ++c) {
        int p = c; double best = fabs(A[c][c]);
        for (int r = c + 1; r < 10; ++r) { const double v = fabs(A[r][c]); if (v > best) { best = v; p = r; } }
        if (!(best > 1e-300)) return false;
        if (p != c) for (int k = c; k < 20; ++k) { const double t = A[c][k]; A[c][k] = A[p][k]; A[p][k] = t; }
        const double inv = 1.0 / A[c][c];
        for (int k = c; k < 20; ++k) A[c][k] *= inv;
        for (int r = 0; r < 10; ++r) {
            if (r == c) continue;
            const double f = A[r][c];
            for (int k = c; k < 20; ++k) A[r][k] = fma(-f, A[c][k], A[r][k]);
        }
    }
    return true;
}

// ---- step 5: cv::solvePoly's Durand-Kerner iteration ------------------------------------------------
// c[0..10] ascending.  re/im[10] out.  Returns the effective degree n (leading |c| <= DBL_EPSILON dropped,
// missing roots reported as (0, 0) like cv::solvePoly does).
__device__ inline int durand_kerner10(const double* c, double* re, double* im)
{
    int n = 10;
    for (; n > 1; --n) if (fabs(c[n]) > 2.220446049250313e-16) break;
    {   // start roots (1+i)^k
        double pr = 1, pi = 0;
        for (int i = 0; i < n; ++i) { re[i] = pr; im[i] = pi; const double t = pr - pi; pi = pr + pi; pr = t; }
    }
    for (int iter = 0; iter < 300; ++iter) {
        bool changed = false;
        double max_diff = 0;
        for (int i = 0; i < n; ++i) {
            const double pr = re[i], pi = im[i];
            double nr = c[n], ni = 0, dr = c[n], di = 0;
            int same = 1;
            for (int j = 0; j < n; ++j) {
                // num = num * p + c[n-j-1]   (no FMA contraction: products rounded like the reference's)
                const double t = __dadd_rn(__dsub_rn(__dmul_rn(nr, pr), __dmul_rn(ni, pi)), c[n - j - 1]);
                ni = __dadd_rn(__dmul_rn(nr, pi), __dmul_rn(ni, pr));
                nr = t;
                if (j != i) {
                    const double er = __dsub_rn(pr, re[j]), ei = __dsub_rn(pi, im[j]);
                    if (er != 0 || ei != 0) {
                        const double t2 = __dsub_rn(__dmul_rn(dr, er), __dmul_rn(di, ei));
                        di = __dadd_rn(__dmul_rn(dr, ei), __dmul_rn(di, er));
                        dr = t2;
                    } else ++same;
                }
            }
            // num /= denom  (cv::Complex division: scale by 1/|denom|^2)
            const double tt = __ddiv_rn(1.0, __dadd_rn(__dmul_rn(dr, dr), __dmul_rn(di, di)));
            double qr = __dmul_rn(__dadd_rn(__dmul_rn(nr, dr), __dmul_rn(ni, di)), tt);
            double qi = __dmul_rn(__dadd_rn(__dmul_rn(-nr, di), __dmul_rn(ni, dr)), tt);
            if (same > 1) {              // coincident iterates (cv::solvePoly's multiple-root branch)
                const double old_re = qr, old_im = qi;
                const int sq = (same % 2 == 0) ? same / 2 : same / 2 - 1;
                for (int k = 0; k < sq; ++k) {
                    qr = old_re * old_re + old_im * old_im; qr = sqrt(qr); qr += old_re;
                    qi = qr - old_re; qr /= 2; qr = sqrt(qr); qi /= 2; qi = sqrt(qi);
                    if (old_re < 0) qi = -qi;
                }
                if (same % 2 != 0) {
                    const double m = pow(sqrt(qr * qr + qi * qi), 1.0 / 3), ang = atan2(qi, qr) / 3;
                    qr = m * cos(ang); qi = m * sin(ang);
                }
            }
            const double nre = __dsub_rn(pr, qr), nim = __dsub_rn(pi, qi);
            changed |= (nre != pr) || (nim != pi);
            re[i] = nre; im[i] = nim;
            max_diff = fmax(max_diff, fabs(qr) + fabs(qi));
        }
        if (!changed || !(max_diff > 0)) break;    // fixed point: every further sweep of the reference is a no-op
    }
    for (int i = 0; i < n; ++i) if (fabs(im[i]) < 1e-100) im[i] = 0;
    for (int i = n; i < 10; ++i) { re[i] = 0; im[i] = 0; }
    return n;
}

// Same iteration for the full degree (the common case), every index a compile-time constant: the ten iterates and the
// coefficients stay in registers instead of dynamically indexed local memory.  300 sweeps are the reference's: the
// iteration practically never becomes bit-stationary (measured on the oracle: 99.5 % of the samples use all sweeps).
// EARLY (off by default, mpl_set_solver_early_exit): leave the loop once a whole sweep moved every root by less than
// 2^-40 of its magnitude.  The iteration converges quadratically, so every later sweep of the reference only re-rounds the
// last bits; the roots then differ from the 300-sweep roots by a few ulp (DESIGN.md section 4.4 has the measured effect).
template <bool EARLY>
__device__ __forceinline__ void durand_kerner_deg10(const double (&c)[11], double (&re)[10], double (&im)[10])
{
    {
        double pr = 1, pi = 0;
#pragma unroll
        for (int i = 0; i < 10; ++i) { re[i] = pr; im[i] = pi; const double t = pr - pi; pi = pr + pi; pr = t; }
    }
#pragma unroll 1
    for (int iter = 0; iter < 300; ++iter) {
        bool changed = false;
        bool settled = true;
        double max_diff = 0;
#pragma unroll
        for (int i = 0; i < 10; ++i) {
            const double pr = re[i], pi = im[i];
            double nr = c[10], ni = 0, dr = c[10], di = 0;
            int same = 1;
#pragma unroll
            for (int j = 0; j < 10; ++j) {
                const double t = __dadd_rn(__dsub_rn(__dmul_rn(nr, pr), __dmul_rn(ni, pi)), c[10 - j - 1]);
                ni = __dadd_rn(__dmul_rn(nr, pi), __dmul_rn(ni, pr));
                nr = t;
                if (j != i) {
                    const double er = __dsub_rn(pr, re[j]), ei = __dsub_rn(pi, im[j]);
                    if (er != 0 || ei != 0) {
                        const double t2 = __dsub_rn(__dmul_rn(dr, er), __dmul_rn(di, ei));
                        di = __dadd_rn(__dmul_rn(dr, ei), __dmul_rn(di, er));
                        dr = t2;
                    } else ++same;
                }
            }
            const double tt = __ddiv_rn(1.0, __dadd_rn(__dmul_rn(dr, dr), __dmul_rn(di, di)));
            double qr = __dmul_rn(__dadd_rn(__dmul_rn(nr, dr), __dmul_rn(ni, di)), tt);
            double qi = __dmul_rn(__dadd_rn(__dmul_rn(-nr, di), __dmul_rn(ni, dr)), tt);
            if (same > 1) {              // coincident iterates (cv::solvePoly's multiple-root branch)
                const double old_re = qr, old_im = qi;
                const int sq = (same % 2 == 0) ? same / 2 : same / 2 - 1;
                for (int k = 0; k < sq; ++k) {
                    qr = old_re * old_re + old_im * old_im; qr = sqrt(qr); qr += old_re;
                    qi = qr - old_re; qr /= 2; qr = sqrt(qr); qi /= 2; qi = sqrt(qi);
                    if (old_re < 0) qi = -qi;
                }
                if (same % 2 != 0) {
                    const double m = pow(sqrt(qr * qr + qi * qi), 1.0 / 3), ang = atan2(qi, qr) / 3;
                    qr = m * cos(ang); qi = m * sin(ang);
                }
            }
            const double nre = __dsub_rn(pr, qr), nim = __dsub_rn(pi, qi);
            changed |= (nre != pr) || (nim != pi);
            re[i] = nre; im[i] = nim;
            max_diff = fmax(max_diff, fabs(qr) + fabs(qi));
            if (EARLY) settled &= (fabs(qr) + fabs(qi)) <= 9.094947017729282e-13 * (fabs(nre) + fabs(nim));   // 2^-40
        }
        if (!changed || !(max_diff > 0)) break;
        if (EARLY && settled) break;
    }
#pragma unroll
    for (int i = 0; i < 10; ++i) if (fabs(im[i]) < 1e-100) im[i] = 0;
}

// ---- steps 3-6 from a given null-space basis -------------------------------------------------------
// models: up to 10 x 9 doubles; returns the count.
template <bool EARLY = false>
__device__ inline int solve_from_basis(const double ee[4][9], double* models)
{
    double A[10][20];
    constraint_matrix(ee, A);
    if (!eliminate10(A)) return 0;
    // B(z) row i = (row 4+2i) - z (row 5+2i) on the last ten columns [xz2 xz x | yz2 yz y | z3 z2 z 1]
    // stored by ascending power of z: bx[i][0..3], by[i][0..3], bw[i][0..4]
    double bx[3][4], by[3][4], bw[3][5];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const double* a1 = &A[4 + 2 * i][10];
        const double* a2 = &A[5 + 2 * i][10];
        bx[i][0] = a1[2]; bx[i][1] = a1[1] - a2[2]; bx[i][2] = a1[0] - a2[1]; bx[i][3] = -a2[0];
        by[i][0] = a1[5]; by[i][1] = a1[4] - a2[5]; by[i][2] = a1[3] - a2[4]; by[i][3] = -a2[3];
        bw[i][0] = a1[9]; bw[i][1] = a1[8] - a2[9]; bw[i][2] = a1[7] - a2[8]; bw[i][3] = a1[6] - a2[7]; bw[i][4] = -a2[6];
    }
    double c[11];
#pragma unroll
    for (int i = 0; i < 11; ++i) c[i] = 0;
    constexpr int perm[6][4] = {{0, 1, 2, 1}, {1, 2, 0, 1}, {2, 0, 1, 1}, {0, 2, 1, -1}, {1, 0, 2, -1}, {2, 1, 0, -1}};
#pragma unroll
    for (int s = 0; s < 6; ++s) {
        double t2[7];
#pragma unroll
        for (int i = 0; i < 7; ++i) t2[i] = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) t2[i + j] = fma(bx[perm[s][0]][i], by[perm[s][1]][j], t2[i + j]);
#pragma unroll
        for (int i = 0; i < 7; ++i)
#pragma unroll
            for (int j = 0; j < 5; ++j) c[i + j] = fma((double)perm[s][3] * t2[i], bw[perm[s][2]][j], c[i + j]);
    }
    double re[10], im[10];
    if (fabs(c[10]) > 2.220446049250313e-16) durand_kerner_deg10<EARLY>(c, re, im);
    else durand_kerner10(c, re, im);
    int count = 0;
    for (int i = 0; i < 10; ++i) {
        if (fabs(im[i]) > 1e-10) continue;
        const double z = re[i];
        double Bz[3][3];
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            Bz[j][0] = ((bx[j][3] * z + bx[j][2]) * z + bx[j][1]) * z + bx[j][0];
            Bz[j][1] = ((by[j][3] * z + by[j][2]) * z + by[j][1]) * z + by[j][0];
            Bz[j][2] = (((bw[j][4] * z + bw[j][3]) * z + bw[j][2]) * z + bw[j][1]) * z + bw[j][0];
        }
        // null vector: the largest of the three row cross products, normalised to unit length
        double best[3] = {0, 0, 0}, bn = -1;
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const int r0 = a == 2 ? 0 : a, r1 = a == 0 ? 1 : 2;   // (0,1) (1,2) (0,2)
            const double cx = Bz[r0][1] * Bz[r1][2] - Bz[r0][2] * Bz[r1][1];
            const double cy = Bz[r0][2] * Bz[r1][0] - Bz[r0][0] * Bz[r1][2];
            const double cz = Bz[r0][0] * Bz[r1][1] - Bz[r0][1] * Bz[r1][0];
            const double nn = cx * cx + cy * cy + cz * cz;
            if (nn > bn) { bn = nn; best[0] = cx; best[1] = cy; best[2] = cz; }
        }
        if (!(bn > 0)) continue;
        const double inv = 1.0 / sqrt(bn);
        const double w = best[2] * inv;
        if (fabs(w) < 1e-10) continue;                       // five-point.cpp:452
        const double x = best[0] / best[2], y = best[1] / best[2];
        double e[9], nrm = 0;
#pragma unroll
        for (int k = 0; k < 9; ++k) { e[k] = ee[0][k] * x + ee[1][k] * y + ee[2][k] * z + ee[3][k]; nrm += e[k] * e[k]; }
        nrm = 1.0 / sqrt(nrm);
#pragma unroll
        for (int k = 0; k < 9; ++k) models[9 * count + k] = e[k] * nrm;
        ++count;
    }
    return count;
}

}  // namespace fivept
