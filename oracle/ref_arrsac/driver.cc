// oracle/ref_arrsac/driver.cc -- the reference's OWN theia::Arrsac (poselib/include/arrsac/arrsac.h, prosac_sampler.h,
// random_sampler.h, sequential_probability_ratio.h + source/arrsac/sequential_probability_ratio.cc, compiled unmodified from
// /root/reference) driven over raw arrays.  TEST INFRASTRUCTURE ONLY.
//
// What this file restates (the pieces of the reference that need OpenCV / Eigen types and therefore cannot be compiled here):
//   EssentialMatEstimatorTheia::EstimateModel / EstimateModelNonminimal / Error   five-point-nister/modelest.cpp:111-195
//   CvEMEstimator::ValidModel                                                      five-point-nister/five-point.cpp:534-601
//   cv::findFundamentalMat(FM_8POINT) (OpenCV 4.2.0 calib3d/src/fundam.cpp run8Point; third party, not under /root/reference:
//     Hartley normalisation, A^T A, eigenvector of the smallest eigenvalue, rank-2 by SVD, denormalise, F(3,3) = 1)
//   the tail of CvModelEstimator3::runARRSAC (findInliers, acceptance rule)        five-point-nister/modelest.cpp:269-277
// run5Point and computeReprojError3 come from oracle/ransac5pt_oracle.c.
#include <cmath>
#include <cfloat>
#include <cstdint>
#include <cstring>
#include <vector>

#include "arrsac/arrsac.h"

extern "C" int orc_run5point(const double* q1, const double* q2, int n, double* models);
extern "C" int orc_sampson(const double* E, const double* p1, const double* p2, int n, double th, float* err, uint8_t* mask, double* errsum);

namespace {

struct E9 { double e[9]; };

// one-sided Jacobi SVD of an n x n matrix (rows of A), V rows = right singular vectors, w = singular values
template <int N> void jacobi(const double A[N][N], double V[N][N], double w[N])
{
    double U[N][N];
    for (int j = 0; j < N; ++j) for (int i = 0; i < N; ++i) { U[j][i] = A[i][j]; V[j][i] = i == j; }
    for (int sweep = 0; sweep < 60; ++sweep) {
        bool ch = false;
        for (int p = 0; p < N - 1; ++p) for (int q = p + 1; q < N; ++q) {
            double a = 0, b = 0, c = 0;
            for (int i = 0; i < N; ++i) { a += U[p][i] * U[p][i]; b += U[q][i] * U[q][i]; c += U[p][i] * U[q][i]; }
            if (std::fabs(c) <= DBL_EPSILON * std::sqrt(a * b) || c == 0.0) continue;
            ch = true;
            const double z = (b - a) / (2 * c), t = (z >= 0 ? 1.0 : -1.0) / (std::fabs(z) + std::sqrt(1 + z * z)), cs = 1 / std::sqrt(1 + t * t), sn = cs * t;
            for (int i = 0; i < N; ++i) {
                const double up = U[p][i], uq = U[q][i]; U[p][i] = cs * up - sn * uq; U[q][i] = sn * up + cs * uq;
                const double vp = V[p][i], vq = V[q][i]; V[p][i] = cs * vp - sn * vq; V[q][i] = sn * vp + cs * vq;
            }
        }
        if (!ch) break;
    }
    for (int j = 0; j < N; ++j) { double s = 0; for (int i = 0; i < N; ++i) s += U[j][i] * U[j][i]; w[j] = std::sqrt(s); }
}

inline bool isZero(double d) { return d < 1e-3 && d > -1e-3; }                 // five-point.hpp:76-81

bool valid_model(const double* Ein, const double* m1, const double* m2, int n)  // five-point.cpp:534-601
{
    double E[9]; std::memcpy(E, Ein, 72);
    double At[3][3], V[3][3], w[3];
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) At[i][j] = E[3 * j + i];
    jacobi<3>(At, V, w);
    int o[3] = {0, 1, 2};
    for (int a = 0; a < 2; ++a) for (int b = a + 1; b < 3; ++b) if (w[o[b]] > w[o[a]]) std::swap(o[a], o[b]);
    if (w[o[0]] / w[o[1]] > 1.2) return false;
    if (!isZero(0.01 * w[o[2]] / w[o[1]])) return false;
    const double e2[3] = {V[o[2]][0], V[o[2]][1], V[o[2]][2]};
    bool emult = false; int fail = 0;
    for (;;) {
        bool again = false;
        for (int i = 0; i < n && !again; ++i) {
            const double x1 = m1[2 * i], y1 = m1[2 * i + 1], x2 = m2[2 * i], y2 = m2[2 * i + 1];
            const double l1[3] = {e2[1] - e2[2] * y2, e2[2] * x2 - e2[0], e2[0] * y2 - e2[1] * x2};
            const double l2[3] = {E[0] * x1 + E[1] * y1 + E[2], E[3] * x1 + E[4] * y1 + E[5], E[6] * x1 + E[7] * y1 + E[8]};
            for (int j = 0; j < 3; ++j) {
                if (isZero(0.1 * l1[j]) || isZero(0.1 * l2[j])) continue;
                if (l1[j] * l2[j] < 0) {
                    if (!emult) { emult = true; fail = 0; for (int k = 0; k < 9; ++k) E[k] = -E[k]; again = true; break; }
                    ++fail; break;
                }
            }
        }
        if (!again) break;
    }
    return !((float)fail / (float)n >= 0.4f);
}

bool eight_point(const double* m1, const double* m2, int n, double* F)          // OpenCV run8Point
{
    double c1x = 0, c1y = 0, c2x = 0, c2y = 0;
    for (int i = 0; i < n; ++i) { c1x += m1[2 * i]; c1y += m1[2 * i + 1]; c2x += m2[2 * i]; c2y += m2[2 * i + 1]; }
    const double t = 1.0 / n; c1x *= t; c1y *= t; c2x *= t; c2y *= t;
    double s1 = 0, s2 = 0;
    for (int i = 0; i < n; ++i) {
        s1 += std::sqrt((m1[2 * i] - c1x) * (m1[2 * i] - c1x) + (m1[2 * i + 1] - c1y) * (m1[2 * i + 1] - c1y));
        s2 += std::sqrt((m2[2 * i] - c2x) * (m2[2 * i] - c2x) + (m2[2 * i + 1] - c2y) * (m2[2 * i + 1] - c2y));
    }
    s1 *= t; s2 *= t;
    if (s1 < FLT_EPSILON || s2 < FLT_EPSILON) return false;
    s1 = std::sqrt(2.) / s1; s2 = std::sqrt(2.) / s2;
    double A[9][9] = {{0}};
    for (int i = 0; i < n; ++i) {
        const double x1 = (m1[2 * i] - c1x) * s1, y1 = (m1[2 * i + 1] - c1y) * s1, x2 = (m2[2 * i] - c2x) * s2, y2 = (m2[2 * i + 1] - c2y) * s2;
        const double r[9] = {x2 * x1, x2 * y1, x2, y2 * x1, y2 * y1, y2, x1, y1, 1};
        for (int a = 0; a < 9; ++a) for (int b = 0; b < 9; ++b) A[a][b] += r[a] * r[b];
    }
    double V[9][9], w[9];
    jacobi<9>(A, V, w);                                                        // symmetric: singular vectors = eigenvectors
    int mn = 0; for (int j = 1; j < 9; ++j) if (w[j] < w[mn]) mn = j;
    double F0[3][3], V3[3][3], w3[3];
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) F0[i][j] = V[mn][3 * i + j];
    jacobi<3>(F0, V3, w3);
    int m3 = 0; for (int j = 1; j < 3; ++j) if (w3[j] < w3[m3]) m3 = j;
    double Fv[3];
    for (int i = 0; i < 3; ++i) Fv[i] = F0[i][0] * V3[m3][0] + F0[i][1] * V3[m3][1] + F0[i][2] * V3[m3][2];
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) F0[i][j] -= Fv[i] * V3[m3][j];
    const double T1[3][3] = {{s1, 0, -s1 * c1x}, {0, s1, -s1 * c1y}, {0, 0, 1}}, T2[3][3] = {{s2, 0, -s2 * c2x}, {0, s2, -s2 * c2y}, {0, 0, 1}};
    double M[3][3], R[3][3];
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) { double a = 0; for (int k = 0; k < 3; ++k) a += T2[k][i] * F0[k][j]; M[i][j] = a; }
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) { double a = 0; for (int k = 0; k < 3; ++k) a += M[i][k] * T1[k][j]; R[i][j] = a; }
    if (std::fabs(R[2][2]) > FLT_EPSILON) { const double inv = 1.0 / R[2][2]; for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) R[i][j] *= inv; }
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) F[3 * i + j] = R[i][j];
    return true;
}

class Esti : public theia::Estimator<size_t, E9> {
public:
    const double* p1; const double* p2;
    void gather(const std::vector<size_t>& d, std::vector<double>& a, std::vector<double>& b) const {
        a.resize(2 * d.size()); b.resize(2 * d.size());
        for (size_t i = 0; i < d.size(); ++i) { a[2 * i] = p1[2 * d[i]]; a[2 * i + 1] = p1[2 * d[i] + 1]; b[2 * i] = p2[2 * d[i]]; b[2 * i + 1] = p2[2 * d[i] + 1]; }
    }
    bool EstimateModel(const std::vector<size_t>& data, std::vector<E9>* model) const override {       // modelest.cpp:111-146
        std::vector<double> a, b; gather(data, a, b);
        double models[90];
        const int nm = orc_run5point(a.data(), b.data(), (int)data.size(), models);
        if (nm <= 0) return false;
        for (int i = 0; i < nm; ++i) {
            if (!valid_model(models + 9 * i, a.data(), b.data(), (int)data.size())) continue;
            E9 e; std::memcpy(e.e, models + 9 * i, 72); model->push_back(e);
        }
        return !model->empty();
    }
    bool EstimateModelNonminimal(const std::vector<size_t>& data, std::vector<E9>* model) const override {   // :149-179
        std::vector<double> a, b; gather(data, a, b);
        E9 e;
        if (!eight_point(a.data(), b.data(), (int)data.size(), e.e)) return false;
        if (!valid_model(e.e, a.data(), b.data(), (int)data.size())) return false;
        model->push_back(e);
        return true;
    }
    double Error(const size_t& d, const E9& m) const override {                                          // :182-195
        float err; orc_sampson(m.e, p1 + 2 * d, p2 + 2 * d, 1, 1.0, &err, nullptr, nullptr);
        return (double)err;
    }
};

}  // namespace

// runARRSAC up to (not including) the refinement: E_out = best model of Arrsac::Estimate, mask / *n_inl = findInliers.
// Returns 1 when the reference's `result` is true at modelest.cpp:278.  The samplers' static generators keep running across calls.
extern "C" int ref_arrsac(const double* p1, const double* p2, int n, double th, double* E_out, uint8_t* mask, int* n_inl)
{
    *n_inl = 0;
    if (n < 5) return 0;
    Esti esti; esti.p1 = p1; esti.p2 = p2;
    std::vector<size_t> input(n);
    for (int i = 0; i < n; ++i) input[i] = i;
    E9 best; std::memset(&best, 0, sizeof(best));
    theia::Arrsac<size_t, E9> arrsac(5, th * th, 500, 100, 14, 8);             // modelest.cpp:270
    if (!arrsac.Estimate(input, esti, &best)) return 0;
    double es;
    const int good = orc_sampson(best.e, p1, p2, n, th, nullptr, mask, &es);   // findInliers (:273)
    if (((good < 50) && (n > 200)) || (good < 15)) return 0;                   // :274-277
    std::memcpy(E_out, best.e, 72);
    *n_inl = good;
    return 1;
}
