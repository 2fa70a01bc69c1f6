/*
 * ransac5pt_oracle.c -- CPU restatement of the reference's RANSAC 5-point essential-matrix path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under matchinglib_poselib_b200/ may link or call this file; it is the
 * checker for tests/, __graft_entry__.smoke() and the timed CPU baseline of bench.py.
 *
 * Reference (paths relative to /root/reference/matchinglib_poselib/source/poselib/source/five-point-nister):
 *   five-point.cpp:366-471  CvEMEstimator::run5Point   (Nister 5-point: null space -> 10x20 constraint matrix ->
 *                                                       10x10 elimination -> 3x3 polynomial matrix B(z) -> degree-10
 *                                                       polynomial -> cv::solvePoly -> SVD::solveZ -> E)
 *   five-point.cpp:603-824  getCoeffMat                (generated table; here re-DERIVED by polynomial arithmetic from
 *                                                       det(E) = 0 and 2 E E^T E - tr(E E^T) E = 0, same monomial order)
 *   five-point.cpp:476-503  computeReprojError3        (Sampson error, double -> float)
 *   modelest.cpp:69-83      findInliers                (err <= th*th)
 *   modelest.cpp:86-109     cvRANSACUpdateNumIters1
 *   modelest.cpp:343-474    runRANSAC                  (selection / tie-break / adaptive stop / least-squares refit)
 * Third-party semantics restated from OpenCV 4.2.0 (not under /root/reference; pinned by ci/make_opencv.sh:6):
 *   core/src/mathfuncs.cpp  cv::solvePoly  (Durand-Kerner, in-place updates, 300 iterations)
 *   core/src/lapack.cpp     SVD (one-sided Jacobi), Mat::inv(DECOMP_LU)
 *
 * Parity pin: tests/test_oracle_ransac.py checks this file against tests/golden/ransac_golden.npz, which
 * oracle/gen_golden_ransac.py produced by running the reference's own table + cv2's OpenCV primitives here.
 * E is basis independent, so models agree with the reference to rounding (matched per sample by nearest E,
 * sign-free); solvePoly root ORDER is pinned separately on raw polynomials.
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math -shared -fPIC   (no FMA contraction: the reference has none)
 */
#include <math.h>
#include <float.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* ---------------------------------------------------------------------------------------------- */
/* monomials of degree <= 3 in (x, y, z), Nister's order = the reference's column order after its  */
/* permutation (five-point.cpp:813): the first ten are eliminated, the last ten carry B(z).        */
static const int MON[20][3] = {
    {3,0,0},{0,3,0},{2,1,0},{1,2,0},{2,0,1},{2,0,0},{0,2,1},{0,2,0},{1,1,1},{1,1,0},
    {1,0,2},{1,0,1},{1,0,0},{0,1,2},{0,1,1},{0,1,0},{0,0,3},{0,0,2},{0,0,1},{0,0,0}};
static int MON_IDX[4][4][4];
static int mon_ready = 0;
static void mon_init(void)
{
    if (mon_ready) return;
    memset(MON_IDX, -1, sizeof(MON_IDX));
    for (int m = 0; m < 20; ++m) MON_IDX[MON[m][0]][MON[m][1]][MON[m][2]] = m;
    mon_ready = 1;
}
static void pmul_acc(const double *a, const double *b, double s, double *c)
{   /* c += s * a * b  (all 20-coefficient polynomials; terms above degree 3 cannot occur for our inputs) */
    for (int i = 0; i < 20; ++i) {
        if (a[i] == 0.0) continue;
        for (int j = 0; j < 20; ++j) {
            if (b[j] == 0.0) continue;
            int ex = MON[i][0] + MON[j][0], ey = MON[i][1] + MON[j][1], ez = MON[i][2] + MON[j][2];
            if (ex + ey + ez > 3) continue;
            c[MON_IDX[ex][ey][ez]] += s * a[i] * b[j];
        }
    }
}

/* 10 x 20 constraint matrix from the null-space basis ee[4][9] (x, y, z, 1 components). */
static void coeff_mat(const double ee[4][9], double A[10][20])
{
    mon_init();
    double E[9][20];                       /* E entries as linear polynomials */
    memset(E, 0, sizeof(E));
    for (int k = 0; k < 9; ++k) {
        E[k][MON_IDX[1][0][0]] = ee[0][k]; E[k][MON_IDX[0][1][0]] = ee[1][k];
        E[k][MON_IDX[0][0][1]] = ee[2][k]; E[k][MON_IDX[0][0][0]] = ee[3][k];
    }
    memset(A, 0, sizeof(double) * 200);
    /* row 0: det(E) */
    double t[20];
#define MINOR(a, b, c, d, sgn)                                       \
    memset(t, 0, sizeof(t)); pmul_acc(E[a], E[b], 1.0, t); pmul_acc(E[c], E[d], -1.0, t)
    MINOR(4, 8, 5, 7, 1); pmul_acc(E[0], t, 1.0, A[0]);
    MINOR(3, 8, 5, 6, 1); pmul_acc(E[1], t, -1.0, A[0]);
    MINOR(3, 7, 4, 6, 1); pmul_acc(E[2], t, 1.0, A[0]);
    /* EEt (symmetric), trace */
    double EEt[3][3][20], tr[20];
    memset(EEt, 0, sizeof(EEt)); memset(tr, 0, sizeof(tr));
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            for (int k = 0; k < 3; ++k) pmul_acc(E[3 * i + k], E[3 * j + k], 1.0, EEt[i][j]);
    for (int i = 0; i < 3; ++i) for (int m = 0; m < 20; ++m) tr[m] += EEt[i][i][m];
    /* rows 1..9: 2 EEt E - tr E */
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            double *row = A[1 + 3 * i + j];
            for (int k = 0; k < 3; ++k) pmul_acc(EEt[i][k], E[3 * k + j], 2.0, row);
            pmul_acc(tr, E[3 * i + j], -1.0, row);
        }
}

/* ---------------------------------------------------------------------------------------------- */
/* one-sided (Hestenes) Jacobi SVD of an n x 9 matrix: returns V (9x9, columns) sorted by descending */
/* singular value.  Stands in for cv::SVD::compute(Q, W, U, Vt, FULL_UV) -- only Vt rows 5..8 are used. */
static void right_singular_vectors(const double *Q, int n, double V[9][9])
{
    double *A = (double *)malloc(sizeof(double) * n * 9);
    memcpy(A, Q, sizeof(double) * n * 9);
    for (int i = 0; i < 9; ++i) for (int j = 0; j < 9; ++j) V[i][j] = (i == j);
    for (int sweep = 0; sweep < 60; ++sweep) {
        int rotated = 0;
        for (int p = 0; p < 8; ++p)
            for (int q = p + 1; q < 9; ++q) {
                double a = 0, b = 0, c = 0;
                for (int r = 0; r < n; ++r) { double x = A[r * 9 + p], y = A[r * 9 + q]; a += x * x; b += y * y; c += x * y; }
                if (fabs(c) <= 1e-300 || fabs(c) <= DBL_EPSILON * sqrt(a * b)) continue;
                rotated = 1;
                double zeta = (b - a) / (2.0 * c);
                double tt = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                double cs = 1.0 / sqrt(1.0 + tt * tt), sn = cs * tt;
                for (int r = 0; r < n; ++r) {
                    double x = A[r * 9 + p], y = A[r * 9 + q];
                    A[r * 9 + p] = cs * x - sn * y; A[r * 9 + q] = sn * x + cs * y;
                }
                for (int r = 0; r < 9; ++r) {
                    double x = V[r][p], y = V[r][q];
                    V[r][p] = cs * x - sn * y; V[r][q] = sn * x + cs * y;
                }
            }
        if (!rotated) break;
    }
    double w[9]; int ord[9];
    for (int j = 0; j < 9; ++j) { double s = 0; for (int r = 0; r < n; ++r) s += A[r * 9 + j] * A[r * 9 + j]; w[j] = s; ord[j] = j; }
    for (int i = 0; i < 9; ++i) for (int j = i + 1; j < 9; ++j) if (w[ord[j]] > w[ord[i]]) { int tmp = ord[i]; ord[i] = ord[j]; ord[j] = tmp; }
    double Vs[9][9];
    for (int j = 0; j < 9; ++j) for (int r = 0; r < 9; ++r) Vs[r][j] = V[r][ord[j]];
    memcpy(V, Vs, sizeof(Vs));
    free(A);
}

/* A1^{-1} A2 by Gauss-Jordan with partial pivoting (Mat::inv(DECOMP_LU) * A2 of five-point.cpp:394). */
static int eliminate(double A[10][20])
{
    for (int c = 0; c < 10; ++c) {
        int p = c; double best = fabs(A[c][c]);
        for (int r = c + 1; r < 10; ++r) if (fabs(A[r][c]) > best) { best = fabs(A[r][c]); p = r; }
        if (best < 1e-300) return -1;
        if (p != c) for (int k = 0; k < 20; ++k) { double tmp = A[c][k]; A[c][k] = A[p][k]; A[p][k] = tmp; }
        double inv = 1.0 / A[c][c];
        for (int k = 0; k < 20; ++k) A[c][k] *= inv;
        for (int r = 0; r < 10; ++r) {
            if (r == c) continue;
            double f = A[r][c];
            if (f == 0.0) continue;
            for (int k = 0; k < 20; ++k) A[r][k] -= f * A[c][k];
        }
    }
    return 0;
}

/* ---------------------------------------------------------------------------------------------- */
/* cv::solvePoly restated (OpenCV 4.2.0 core/src/mathfuncs.cpp): Durand-Kerner with Gauss-Seidel style  */
/* in-place updates, start roots (1+i)^k, maxIters = 300, stop only when the largest update is exactly 0. */
typedef struct { double re, im; } cplx;
static cplx cmul(cplx a, cplx b) { cplx r = {a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re}; return r; }
static cplx cadd(cplx a, cplx b) { cplx r = {a.re + b.re, a.im + b.im}; return r; }
static cplx csub(cplx a, cplx b) { cplx r = {a.re - b.re, a.im - b.im}; return r; }
static cplx cdivv(cplx a, cplx b)
{   /* cv::Complex operator/ : t = 1/(|b|^2); ((a.re*b.re + a.im*b.im)*t, (-a.re*b.im + a.im*b.re)*t) */
    double t = 1.0 / (b.re * b.re + b.im * b.im);
    cplx r = {(a.re * b.re + a.im * b.im) * t, (-a.re * b.im + a.im * b.re) * t};
    return r;
}
/* coeffs[0..n0] ascending powers; roots[n0] out.  Returns the effective degree. */
int orc_solve_poly(const double *coeffs_in, int n0, double *roots_out /* 2*n0 */, int max_iters)
{
    cplx coeffs[32], roots[32];
    int n = n0;
    for (int i = 0; i <= n0; ++i) { coeffs[i].re = coeffs_in[i]; coeffs[i].im = 0; }
    for (; n > 1; --n) if (fabs(coeffs[n].re) + fabs(coeffs[n].im) > DBL_EPSILON) break;
    cplx p = {1, 0}, r = {1, 1};
    for (int i = 0; i < n; ++i) { roots[i] = p; p = cmul(p, r); }
    if (max_iters <= 0) max_iters = 1000;
    for (int iter = 0; iter < max_iters; ++iter) {
        double maxDiff = 0;
        for (int i = 0; i < n; ++i) {
            p = roots[i];
            cplx num = coeffs[n], denom = coeffs[n];
            int num_same_root = 1;
            for (int j = 0; j < n; ++j) {
                num = cadd(cmul(num, p), coeffs[n - j - 1]);
                if (j != i) {
                    cplx d = csub(p, roots[j]);
                    if (d.re != 0 || d.im != 0) denom = cmul(denom, d);
                    else num_same_root++;
                }
            }
            num = cdivv(num, denom);
            if (num_same_root > 1) {
                double old_re = num.re, old_im = num.im;
                int square_root_times = (num_same_root % 2 == 0) ? num_same_root / 2 : num_same_root / 2 - 1;
                for (int k = 0; k < square_root_times; ++k) {
                    num.re = old_re * old_re + old_im * old_im;
                    num.re = sqrt(num.re); num.re += old_re;
                    num.im = num.re - old_re;
                    num.re /= 2; num.re = sqrt(num.re);
                    num.im /= 2; num.im = sqrt(num.im);
                    if (old_re < 0) num.im = -num.im;
                }
                if (num_same_root % 2 != 0) {
                    double m = pow(sqrt(num.re * num.re + num.im * num.im), 1.0 / 3);
                    double ang = atan2(num.im, num.re) / 3;
                    num.re = m * cos(ang); num.im = m * sin(ang);
                }
            }
            roots[i] = csub(p, num);
            double ad = sqrt(num.re * num.re + num.im * num.im);
            if (ad > maxDiff) maxDiff = ad;
        }
        if (maxDiff <= 0) break;
    }
    for (int i = 0; i < n; ++i) if (fabs(roots[i].im) < 1e-100) roots[i].im = 0;
    for (int i = n; i < n0; ++i) { roots[i].re = 0; roots[i].im = 0; }   /* degree dropped: cv2 4.13 returns (0,0) for the missing roots */
    for (int i = 0; i < n0; ++i) { roots_out[2 * i] = roots[i].re; roots_out[2 * i + 1] = roots[i].im; }
    return n;
}

/* ---------------------------------------------------------------------------------------------- */
/* null vector of a 3x3 matrix (SVD::solveZ in the reference): smallest right singular vector via     */
/* Jacobi eigen-decomposition of M^T M would square the conditioning; use the one-sided Jacobi above  */
/* on the 3x3 instead (same routine family as cv::SVD).                                              */
static void null_vector3(const double M[3][3], double z[3])
{
    double A[3][3], V[3][3] = {{1,0,0},{0,1,0},{0,0,1}};
    memcpy(A, M, sizeof(A));
    for (int sweep = 0; sweep < 60; ++sweep) {
        int rotated = 0;
        for (int p = 0; p < 2; ++p)
            for (int q = p + 1; q < 3; ++q) {
                double a = 0, b = 0, c = 0;
                for (int r = 0; r < 3; ++r) { a += A[r][p] * A[r][p]; b += A[r][q] * A[r][q]; c += A[r][p] * A[r][q]; }
                if (fabs(c) <= 1e-300 || fabs(c) <= DBL_EPSILON * sqrt(a * b)) continue;
                rotated = 1;
                double zeta = (b - a) / (2.0 * c);
                double tt = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                double cs = 1.0 / sqrt(1.0 + tt * tt), sn = cs * tt;
                for (int r = 0; r < 3; ++r) {
                    double x = A[r][p], y = A[r][q]; A[r][p] = cs * x - sn * y; A[r][q] = sn * x + cs * y;
                    x = V[r][p]; y = V[r][q]; V[r][p] = cs * x - sn * y; V[r][q] = sn * x + cs * y;
                }
            }
        if (!rotated) break;
    }
    int m = 0; double best = DBL_MAX;
    for (int j = 0; j < 3; ++j) { double s = 0; for (int r = 0; r < 3; ++r) s += A[r][j] * A[r][j]; if (s < best) { best = s; m = j; } }
    for (int r = 0; r < 3; ++r) z[r] = V[r][m];
}

/* CvEMEstimator::run5Point (five-point.cpp:366-471).  q1, q2: n x 2 doubles.  models: up to 10 x 9. */
int orc_run5point(const double *q1, const double *q2, int n, double *models)
{
    double *Q = (double *)malloc(sizeof(double) * n * 9);
    for (int i = 0; i < n; ++i) {                         /* five-point.cpp:375-383 */
        double x1 = q1[2 * i], y1 = q1[2 * i + 1], x2 = q2[2 * i], y2 = q2[2 * i + 1];
        double *r = Q + 9 * i;
        r[0] = x1 * x2; r[1] = y1 * x2; r[2] = x2; r[3] = x1 * y2; r[4] = y1 * y2; r[5] = y2; r[6] = x1; r[7] = y1; r[8] = 1.0;
    }
    double V[9][9];
    right_singular_vectors(Q, n, V);
    free(Q);
    double ee[4][9];
    for (int b = 0; b < 4; ++b) for (int k = 0; k < 9; ++k) ee[b][k] = V[k][5 + b];     /* Vt rows 5..8 (:388) */
    double A[10][20];
    coeff_mat(ee, A);
    if (eliminate(A)) return 0;
    /* B(z): row i = (eliminated row 4+2i) - z * (eliminated row 5+2i)   (five-point.cpp:396-414) */
    double b[3][13];
    for (int i = 0; i < 3; ++i) {
        const double *a1 = &A[4 + 2 * i][10], *a2 = &A[5 + 2 * i][10];
        double r1[13] = {0}, r2[13] = {0};
        r1[1] = a1[0]; r1[2] = a1[1]; r1[3] = a1[2]; r1[5] = a1[3]; r1[6] = a1[4]; r1[7] = a1[5];
        r1[9] = a1[6]; r1[10] = a1[7]; r1[11] = a1[8]; r1[12] = a1[9];
        r2[0] = a2[0]; r2[1] = a2[1]; r2[2] = a2[2]; r2[4] = a2[3]; r2[5] = a2[4]; r2[6] = a2[5];
        r2[8] = a2[6]; r2[9] = a2[7]; r2[10] = a2[8]; r2[11] = a2[9];
        for (int k = 0; k < 13; ++k) b[i][k] = r1[k] - r2[k];
    }
    /* det B(z) -> c[0..10] ascending (the reference expands the same determinant term by term, :416-428).
       Entry polynomials, ascending powers: col0/col1 degree 3 (b[.][3],b[.][2],b[.][1],b[.][0]), col2 degree 4. */
    double P[3][3][5];
    memset(P, 0, sizeof(P));
    for (int i = 0; i < 3; ++i) {
        for (int k = 0; k < 4; ++k) { P[i][0][k] = b[i][3 - k]; P[i][1][k] = b[i][7 - k]; }
        for (int k = 0; k < 5; ++k) P[i][2][k] = b[i][12 - k];
    }
    double c[11] = {0};
    static const int perm[6][4] = {{0,1,2,1},{1,2,0,1},{2,0,1,1},{0,2,1,-1},{1,0,2,-1},{2,1,0,-1}};
    for (int s = 0; s < 6; ++s) {
        double t2[8] = {0};
        for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) t2[i + j] += P[perm[s][0]][0][i] * P[perm[s][1]][1][j];
        for (int i = 0; i < 7; ++i) for (int j = 0; j < 5; ++j) c[i + j] += perm[s][3] * t2[i] * P[perm[s][2]][2][j];
    }
    double roots[20];
    orc_solve_poly(c, 10, roots, 300);
    int count = 0;
    for (int i = 0; i < 10; ++i) {
        if (fabs(roots[2 * i + 1]) > 1e-10) continue;
        double z1 = roots[2 * i], z2 = z1 * z1, z3 = z2 * z1, z4 = z3 * z1;
        double bz[3][3];
        for (int j = 0; j < 3; ++j) {
            const double *br = b[j];
            bz[j][0] = br[0] * z3 + br[1] * z2 + br[2] * z1 + br[3];
            bz[j][1] = br[4] * z3 + br[5] * z2 + br[6] * z1 + br[7];
            bz[j][2] = br[8] * z4 + br[9] * z3 + br[10] * z2 + br[11] * z1 + br[12];
        }
        double xy1[3];
        null_vector3(bz, xy1);
        if (fabs(xy1[2]) < 1e-10) continue;
        double x = xy1[0] / xy1[2], y = xy1[1] / xy1[2];
        double *e = models + 9 * count, nrm = 0;
        for (int k = 0; k < 9; ++k) { e[k] = ee[0][k] * x + ee[1][k] * y + ee[2][k] * z1 + ee[3][k]; nrm += e[k] * e[k]; }
        nrm = sqrt(nrm);
        for (int k = 0; k < 9; ++k) e[k] /= nrm;
        ++count;
    }
    return count;
}

/* computeReprojError3 + findInliers (five-point.cpp:476-503, modelest.cpp:69-83).  err/mask may be NULL.
 * Returns the inlier count; *errsum = sum of all float errors accumulated in double (cv::sum). */
int orc_sampson(const double *E, const double *p1, const double *p2, int n, double th, float *err, uint8_t *mask,
                double *errsum)
{
    const double th2 = th * th;
    int good = 0; double s = 0;
    for (int i = 0; i < n; ++i) {
        double x1 = p1[2 * i], y1 = p1[2 * i + 1], x2 = p2[2 * i], y2 = p2[2 * i + 1];
        double Ex0 = (E[0] * x1 + E[1] * y1) + E[2] * 1.0;
        double Ex1 = (E[3] * x1 + E[4] * y1) + E[5] * 1.0;
        double Ex2 = (E[6] * x1 + E[7] * y1) + E[8] * 1.0;
        double Et0 = (E[0] * x2 + E[3] * y2) + E[6] * 1.0;
        double Et1 = (E[1] * x2 + E[4] * y2) + E[7] * 1.0;
        double r = (x2 * Ex0 + y2 * Ex1) + 1.0 * Ex2;
        double a = Ex0 * Ex0, b = Ex1 * Ex1, c = Et0 * Et0, d = Et1 * Et1;
        float e = (float)(r * r / (a + b + c + d));
        int in = (double)e <= th2;
        if (err) err[i] = e;
        if (mask) mask[i] = (uint8_t)in;
        good += in; s += (double)e;
    }
    if (errsum) *errsum = s;
    return good;
}

/* cvRANSACUpdateNumIters1 (modelest.cpp:86-109) */
int orc_update_num_iters(double p, double ep, int model_points, int max_iters)
{
    p = p < 0 ? 0 : (p > 1 ? 1 : p); ep = ep < 0 ? 0 : (ep > 1 ? 1 : ep);
    double num = 1.0 - p; if (num < DBL_MIN) num = DBL_MIN;
    double denom = 1.0 - pow(1.0 - ep, model_points);
    if (denom < DBL_MIN) return 0;
    num = log(num); denom = log(denom);
    return denom >= 0 || -num >= (double)max_iters * (-denom) ? max_iters : (int)round(num / denom);
}

/* runRANSAC (modelest.cpp:343-474) over pre-drawn samples[h][5].  Optional trace outputs (per model, in
 * evaluation order, at most 10*h entries): trace_models[.][9], trace_counts, trace_errsums, n_models[h].
 * Returns 1 when a model was found.  *iters_run = samples consumed before the adaptive stop. */
int orc_ransac5pt(const double *p1, const double *p2, int n, const int32_t *samples, int h, double th, double conf,
                  int max_iters, int refine, double *E_out, uint8_t *mask_out, int *n_inl, int *iters_run,
                  double *trace_models, int32_t *trace_counts, double *trace_errsums, int32_t *n_models, int full_trace)
{
    int niters = max_iters, best = 0, nt = 0, it;
    double errminsum = DBL_MAX, models[90], esum;
    uint8_t *mask = (uint8_t *)malloc(n), *bestmask = (uint8_t *)calloc(n, 1);
    for (it = 0; it < h && (full_trace || it < niters); ++it) {
        double s1[10], s2[10];
        for (int j = 0; j < 5; ++j) { int id = samples[5 * it + j]; s1[2*j] = p1[2*id]; s1[2*j+1] = p1[2*id+1]; s2[2*j] = p2[2*id]; s2[2*j+1] = p2[2*id+1]; }
        int nm = orc_run5point(s1, s2, 5, models);
        if (n_models) n_models[it] = nm;
        const int live = it < niters;          /* with full_trace the loop keeps tracing past the adaptive stop */
        for (int m = 0; m < nm; ++m) {
            int good = orc_sampson(models + 9 * m, p1, p2, n, th, NULL, mask, &esum);
            if (trace_models) memcpy(trace_models + 9 * nt, models + 9 * m, 72);
            if (trace_counts) trace_counts[nt] = good;
            if (trace_errsums) trace_errsums[nt] = esum;
            ++nt;
            if (!live) continue;
            if (good > (best > 4 ? best : 4)) {
                memcpy(bestmask, mask, n); memcpy(E_out, models + 9 * m, 72); best = good;
                niters = orc_update_num_iters(conf, (double)(n - good) / n, 5, niters);
                errminsum = esum;
            } else if (good == (best > 5 ? best : 5) && errminsum < DBL_MAX && errminsum > esum) {
                memcpy(bestmask, mask, n); memcpy(E_out, models + 9 * m, 72); errminsum = esum;
            }
        }
        if (live && iters_run) *iters_run = it + 1;
    }
    if (refine && best > 0) {                               /* modelest.cpp:420-464 */
        double *a1 = (double *)malloc(sizeof(double) * 2 * n), *a2 = (double *)malloc(sizeof(double) * 2 * n);
        int k = 0;
        for (int i = 0; i < n; ++i) if (bestmask[i]) { a1[2*k] = p1[2*i]; a1[2*k+1] = p1[2*i+1]; a2[2*k] = p2[2*i]; a2[2*k+1] = p2[2*i+1]; ++k; }
        int nm = orc_run5point(a1, a2, k, models);
        if (nm <= 0) {                                      /* modelest.cpp:443-445: "return result" with result still false */
            free(a1); free(a2); free(mask); free(bestmask);
            if (n_inl) *n_inl = 0;
            return 0;
        }
        for (int m = 0; m < nm; ++m) {
            int good = orc_sampson(models + 9 * m, p1, p2, n, th, NULL, mask, &esum);
            if (good > (best > 4 ? best : 4)) { memcpy(bestmask, mask, n); memcpy(E_out, models + 9 * m, 72); best = good; errminsum = esum; }
            else if (good == best && errminsum < DBL_MAX && errminsum > esum) { memcpy(bestmask, mask, n); memcpy(E_out, models + 9 * m, 72); errminsum = esum; }
        }
        free(a1); free(a2);
    }
    if (mask_out) memcpy(mask_out, bestmask, n);
    if (n_inl) *n_inl = best;
    free(mask); free(bestmask);
    return best > 0;
}

/* runLMeDS (modelest.cpp:483-564) over pre-drawn samples[h][5]; the caller passes h = the reference's fixed iteration count
 * (orc_lmeds_niters).  Errors are floats sorted through their bit patterns as ints (icvSortDistances over (int*)err.data,
 * modelest.cpp:477-480, 539); the median of an even count is the FLOAT sum of the two middle values times 0.5 (:541-542);
 * strict '<' keeps the first minimal model (:544); sigma and the final findInliers as :553-559.  medians (optional): one
 * value per model in evaluation order.  Returns 1 when count >= 5 (result), E_out / mask_out written whenever a model had a
 * finite median (like the reference writes model and mask before it decides result). */
static int cmp_i32(const void *a, const void *b) { int32_t x = *(const int32_t *)a, y = *(const int32_t *)b; return x < y ? -1 : x > y; }
int orc_lmeds_niters(double conf, int max_iters)
{
    int niters = (int)round(log(1. - conf) / log(1. - pow(1. - 0.45, 5.0)));
    if (niters < 3) niters = 3;
    return niters < max_iters ? niters : max_iters;
}
int orc_lmeds5pt(const double *p1, const double *p2, int n, const int32_t *samples, int h, double *E_out, uint8_t *mask_out,
                 int *n_inl, double *min_median, double *sigma_out, double *medians, int32_t *n_models)
{
    double minMedian = DBL_MAX, models[90];
    float *err = (float *)malloc(sizeof(float) * n);
    int nt = 0;
    if (n_inl) *n_inl = 0;
    for (int it = 0; it < h; ++it) {
        double s1[10], s2[10];
        for (int j = 0; j < 5; ++j) { int id = samples[5 * it + j]; s1[2*j] = p1[2*id]; s1[2*j+1] = p1[2*id+1]; s2[2*j] = p2[2*id]; s2[2*j+1] = p2[2*id+1]; }
        int nm = orc_run5point(s1, s2, 5, models);
        if (n_models) n_models[it] = nm;
        for (int m = 0; m < nm; ++m) {
            orc_sampson(models + 9 * m, p1, p2, n, 0.0, err, NULL, NULL);
            qsort(err, n, sizeof(float), cmp_i32);
            double median = n % 2 != 0 ? (double)err[n / 2] : (double)(err[n / 2 - 1] + err[n / 2]) * 0.5;
            if (medians) medians[nt] = median;
            ++nt;
            if (median < minMedian) { minMedian = median; memcpy(E_out, models + 9 * m, 72); }
        }
    }
    free(err);
    if (min_median) *min_median = minMedian;
    if (!(minMedian < DBL_MAX)) return 0;
    double sigma = 2.5 * 1.4826 * (1 + 5. / (n - 5)) * sqrt(minMedian);
    if (!(sigma > 0.001)) sigma = 0.001;                    /* MAX(sigma, 0.001): NaN -> 0.001 like the macro */
    if (sigma_out) *sigma_out = sigma;
    int count = orc_sampson(E_out, p1, p2, n, sigma, NULL, mask_out, NULL);
    if (n_inl) *n_inl = count;
    return count >= 5;
}
