// vo_ransac.cpp — see vo_ransac.hpp.  Follows /root/reference/ransac.cpp stage by stage (line references
// at each function); written against vo_linalg.hpp instead of Eigen.
#include "vo_ransac.hpp"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <iterator>
#include <limits>
#include <random>

#if defined(__x86_64__) || defined(__i386__)
#include <immintrin.h>
#define VO_HAVE_AVX2_PATH 1
#endif

namespace vo {

namespace {

// Consensus scoring of one model over the scored correspondences: flag[i] = Sampson error < threshold, returns the count.
// computeSampsonError (ransac.cpp:12-23) inlined, the same operations in the same order, every one rounded on its own
// (the file is compiled with -ffp-contract=off and the AVX2 form uses separate multiplies and adds, never FMA), so the
// scalar loop and the 4-wide loop give identical flags.
int score_scalar(const double* F, const double* X0, const double* X1, const double* P0, const double* P1, int from, int n, double thr,
                 unsigned char* flag) {
    const double f00 = F[0], f01 = F[1], f02 = F[2], f10 = F[3], f11 = F[4], f12 = F[5], f20 = F[6], f21 = F[7], f22 = F[8];
    int count = 0;
    for (int i = from; i < n; ++i) {
        const double x0 = X0[i], x1 = X1[i], p0 = P0[i], p1 = P1[i];
        const double Fx0 = f00 * x0 + f01 * x1 + f02 * 1.0, Fx1 = f10 * x0 + f11 * x1 + f12 * 1.0, Fx2 = f20 * x0 + f21 * x1 + f22 * 1.0;
        const double Ft0 = f00 * p0 + f10 * p1 + f20 * 1.0, Ft1 = f01 * p0 + f11 * p1 + f21 * 1.0;
        const double e = p0 * Fx0 + p1 * Fx1 + 1.0 * Fx2;
        const double den = Fx0 * Fx0 + Fx1 * Fx1 + Ft0 * Ft0 + Ft1 * Ft1;
        const unsigned char in = (den >= 1e-12) & (e * e / den < thr);  // den < 1e-12: error = max
        flag[i] = in;
        count += in;
    }
    return count;
}

#ifdef VO_HAVE_AVX2_PATH
__attribute__((target("avx2"))) int score_avx2(const double* F, const double* X0, const double* X1, const double* P0, const double* P1, int n,
                                               double thr, unsigned char* flag) {
    const __m256d f00 = _mm256_set1_pd(F[0]), f01 = _mm256_set1_pd(F[1]), f02 = _mm256_set1_pd(F[2]), f10 = _mm256_set1_pd(F[3]),
                  f11 = _mm256_set1_pd(F[4]), f12 = _mm256_set1_pd(F[5]), f20 = _mm256_set1_pd(F[6]), f21 = _mm256_set1_pd(F[7]),
                  f22 = _mm256_set1_pd(F[8]);
    const __m256d eps = _mm256_set1_pd(1e-12), vthr = _mm256_set1_pd(thr);
    int count = 0, i = 0;
    for (; i + 4 <= n; i += 4) {
        const __m256d x0 = _mm256_loadu_pd(X0 + i), x1 = _mm256_loadu_pd(X1 + i), p0 = _mm256_loadu_pd(P0 + i), p1 = _mm256_loadu_pd(P1 + i);
        const __m256d Fx0 = _mm256_add_pd(_mm256_add_pd(_mm256_mul_pd(f00, x0), _mm256_mul_pd(f01, x1)), f02);
        const __m256d Fx1 = _mm256_add_pd(_mm256_add_pd(_mm256_mul_pd(f10, x0), _mm256_mul_pd(f11, x1)), f12);
        const __m256d Fx2 = _mm256_add_pd(_mm256_add_pd(_mm256_mul_pd(f20, x0), _mm256_mul_pd(f21, x1)), f22);
        const __m256d Ft0 = _mm256_add_pd(_mm256_add_pd(_mm256_mul_pd(f00, p0), _mm256_mul_pd(f10, p1)), f20);
        const __m256d Ft1 = _mm256_add_pd(_mm256_add_pd(_mm256_mul_pd(f01, p0), _mm256_mul_pd(f11, p1)), f21);
        const __m256d e = _mm256_add_pd(_mm256_add_pd(_mm256_mul_pd(p0, Fx0), _mm256_mul_pd(p1, Fx1)), Fx2);
        const __m256d den = _mm256_add_pd(_mm256_add_pd(_mm256_add_pd(_mm256_mul_pd(Fx0, Fx0), _mm256_mul_pd(Fx1, Fx1)), _mm256_mul_pd(Ft0, Ft0)),
                                          _mm256_mul_pd(Ft1, Ft1));
        const __m256d err = _mm256_div_pd(_mm256_mul_pd(e, e), den);
        const __m256d in = _mm256_and_pd(_mm256_cmp_pd(den, eps, _CMP_GE_OQ), _mm256_cmp_pd(err, vthr, _CMP_LT_OQ));
        const int m = _mm256_movemask_pd(in);
        flag[i] = m & 1; flag[i + 1] = (m >> 1) & 1; flag[i + 2] = (m >> 2) & 1; flag[i + 3] = (m >> 3) & 1;
        count += __builtin_popcount(m);
    }
    return count + score_scalar(F, X0, X1, P0, P1, i, n, thr, flag);
}
#endif

#ifdef VO_HAVE_AVX2_PATH
// the same loop eight correspondences at a time
__attribute__((target("avx512f"))) int score_avx512(const double* F, const double* X0, const double* X1, const double* P0, const double* P1,
                                                    int n, double thr, unsigned char* flag) {
    const __m512d f00 = _mm512_set1_pd(F[0]), f01 = _mm512_set1_pd(F[1]), f02 = _mm512_set1_pd(F[2]), f10 = _mm512_set1_pd(F[3]),
                  f11 = _mm512_set1_pd(F[4]), f12 = _mm512_set1_pd(F[5]), f20 = _mm512_set1_pd(F[6]), f21 = _mm512_set1_pd(F[7]),
                  f22 = _mm512_set1_pd(F[8]);
    const __m512d eps = _mm512_set1_pd(1e-12), vthr = _mm512_set1_pd(thr);
    int count = 0, i = 0;
    for (; i + 8 <= n; i += 8) {
        const __m512d x0 = _mm512_loadu_pd(X0 + i), x1 = _mm512_loadu_pd(X1 + i), p0 = _mm512_loadu_pd(P0 + i), p1 = _mm512_loadu_pd(P1 + i);
        const __m512d Fx0 = _mm512_add_pd(_mm512_add_pd(_mm512_mul_pd(f00, x0), _mm512_mul_pd(f01, x1)), f02);
        const __m512d Fx1 = _mm512_add_pd(_mm512_add_pd(_mm512_mul_pd(f10, x0), _mm512_mul_pd(f11, x1)), f12);
        const __m512d Fx2 = _mm512_add_pd(_mm512_add_pd(_mm512_mul_pd(f20, x0), _mm512_mul_pd(f21, x1)), f22);
        const __m512d Ft0 = _mm512_add_pd(_mm512_add_pd(_mm512_mul_pd(f00, p0), _mm512_mul_pd(f10, p1)), f20);
        const __m512d Ft1 = _mm512_add_pd(_mm512_add_pd(_mm512_mul_pd(f01, p0), _mm512_mul_pd(f11, p1)), f21);
        const __m512d e = _mm512_add_pd(_mm512_add_pd(_mm512_mul_pd(p0, Fx0), _mm512_mul_pd(p1, Fx1)), Fx2);
        const __m512d den = _mm512_add_pd(_mm512_add_pd(_mm512_add_pd(_mm512_mul_pd(Fx0, Fx0), _mm512_mul_pd(Fx1, Fx1)), _mm512_mul_pd(Ft0, Ft0)),
                                          _mm512_mul_pd(Ft1, Ft1));
        const __m512d err = _mm512_div_pd(_mm512_mul_pd(e, e), den);
        const unsigned m = static_cast<unsigned>(_mm512_cmp_pd_mask(den, eps, _CMP_GE_OQ) & _mm512_cmp_pd_mask(err, vthr, _CMP_LT_OQ));
        for (int b = 0; b < 8; ++b) flag[i + b] = (m >> b) & 1u;
        count += __builtin_popcount(m);
    }
    return count + score_scalar(F, X0, X1, P0, P1, i, n, thr, flag);
}
#endif

bool use_avx512() {
#ifdef VO_HAVE_AVX2_PATH
    static const bool ok = __builtin_cpu_supports("avx512f") && !std::getenv("VO_RANSAC_SCALAR") && !std::getenv("VO_RANSAC_NO_AVX512");
    return ok;
#else
    return false;
#endif
}

bool use_avx2() {
#ifdef VO_HAVE_AVX2_PATH
    static const bool ok = __builtin_cpu_supports("avx2") && !std::getenv("VO_RANSAC_SCALAR");
    return ok;
#else
    return false;
#endif
}

}  // namespace

// test hook (vo_host_score_models): both scoring loops on the same model
int score_correspondences(const double F[9], const double* X0, const double* X1, const double* P0, const double* P1, int n, double thr,
                          unsigned char* flag, bool force_scalar) {
#ifdef VO_HAVE_AVX2_PATH
    if (!force_scalar && use_avx512()) return score_avx512(F, X0, X1, P0, P1, n, thr, flag);
    if (!force_scalar && use_avx2()) return score_avx2(F, X0, X1, P0, P1, n, thr, flag);
#endif
    return score_scalar(F, X0, X1, P0, P1, 0, n, thr, flag);
}

// ransac.cpp:12-23: (x'^T F x)^2 / ((Fx)_0^2 + (Fx)_1^2 + (F^T x')_0^2 + (F^T x')_1^2)
double computeSampsonError(const Mat3& F, const Vec3& x, const Vec3& xp) {
    const Vec3 Fx = mul(F, x);
    const Vec3 Ftx = mul(transpose(F), xp);
    const double e = dot(xp, Fx);
    const double numerator = e * e;
    const double denominator = Fx(0) * Fx(0) + Fx(1) * Fx(1) + Ftx(0) * Ftx(0) + Ftx(1) * Ftx(1);
    if (denominator < 1e-12) return std::numeric_limits<double>::max();
    return numerator / denominator;
}

// ransac.cpp:25-61: Hartley normalisation, centroid to the origin and RMS distance sqrt(2)
void FundamentalMatrix::normalizePoints(const std::vector<PointPair>& points, std::vector<Vec3>& n1,
                                        std::vector<Vec3>& n2, Mat3& T1, Mat3& T2) {
    double mx1 = 0, my1 = 0, mx2 = 0, my2 = 0;
    for (const auto& p : points) { mx1 += p.first.x; my1 += p.first.y; mx2 += p.second.x; my2 += p.second.y; }
    const double n = static_cast<double>(points.size());
    mx1 /= n; my1 /= n; mx2 /= n; my2 /= n;
    double s1 = 0, s2 = 0;
    for (const auto& p : points) {
        s1 += (p.first.x - mx1) * (p.first.x - mx1) + (p.first.y - my1) * (p.first.y - my1);
        s2 += (p.second.x - mx2) * (p.second.x - mx2) + (p.second.y - my2) * (p.second.y - my2);
    }
    s1 = std::sqrt(2.0) / std::sqrt(s1 / n);
    s2 = std::sqrt(2.0) / std::sqrt(s2 / n);
    T1 = Mat3::zero(); T2 = Mat3::zero();
    T1(0, 0) = s1; T1(0, 2) = -s1 * mx1; T1(1, 1) = s1; T1(1, 2) = -s1 * my1; T1(2, 2) = 1;
    T2(0, 0) = s2; T2(0, 2) = -s2 * mx2; T2(1, 1) = s2; T2(1, 2) = -s2 * my2; T2(2, 2) = 1;
    n1.clear(); n2.clear();
    n1.reserve(points.size()); n2.reserve(points.size());
    for (const auto& p : points) {
        n1.push_back(mul(T1, Vec3{{p.first.x, p.first.y, 1.0}}));
        n2.push_back(mul(T2, Vec3{{p.second.x, p.second.y, 1.0}}));
    }
}

// ransac.cpp:63-93.  Note the reference's row layout (p1.x*p2.x, p1.x*p2.y, p1.x, ...): with f reshaped
// row-major it encodes p1^T F p2 = 0, and F is then de-normalised as T2^T F T1 - kept as written.
Mat3 FundamentalMatrix::computeFundamentalMatrix(const std::vector<PointPair>& points) {
    Mat3 T1, T2;
    std::vector<Vec3> n1, n2;
    normalizePoints(points, n1, n2, T1, T2);
    const int m = static_cast<int>(points.size());
    std::vector<double> A(static_cast<size_t>(m) * 9);
    for (int i = 0; i < m; ++i) {
        const Vec3 &p1 = n1[i], &p2 = n2[i];
        double* r = &A[static_cast<size_t>(i) * 9];
        r[0] = p1(0) * p2(0); r[1] = p1(0) * p2(1); r[2] = p1(0);
        r[3] = p1(1) * p2(0); r[4] = p1(1) * p2(1); r[5] = p1(1);
        r[6] = p2(0);         r[7] = p2(1);         r[8] = 1.0;
    }
    Mat3 F;
    if (m == 8) {
        null_vector_8x9(A.data(), F.m);  // minimal sample: the null vector, without the Jacobi sweeps
    } else {
        double s[9], V[81];
        jacobi_svd(A.data(), m, 9, s, V);
        for (int i = 0; i < 9; ++i) F.m[i] = V[i * 9 + 8];  // right singular vector of the smallest singular value
    }
    F = mul(mul(transpose(T2), F), T1);
    // rank 2: zero the smallest singular value
    Mat3 U, W;
    Vec3 sv;
    svd3(F, U, sv, W);
    Mat3 D = Mat3::zero();
    D(0, 0) = sv(0); D(1, 1) = sv(1);
    return mul(mul(U, D), transpose(W));
}

void FundamentalMatrix::fit(const std::vector<PointPair>& sample) {
    if (sample.size() < 8) return;
    F = computeFundamentalMatrix(sample);
    inliers = sample;
}

int FundamentalMatrix::countInliers(const std::vector<PointPair>& data, double threshold) const {
    int count = 0;
    for (const auto& pr : data) {
        const Vec3 p1{{pr.first.x, pr.first.y, 1.0}}, p2{{pr.second.x, pr.second.y, 1.0}};
        count += std::fabs(dot(p2, mul(F, p1))) < threshold;
    }
    return count;
}

RansacStats Ransac::run(FundamentalMatrix& model, const std::vector<PointPair>& data, double probability,
                        double sampsonThreshold, int numThreads, std::uint64_t seed) const {
    RansacStats st;
    const int N = static_cast<int>(data.size());
    const int sampleSize = 8;
    double outlierRatio = 0.5;
    int maxIterations = static_cast<int>(std::log(1.0 - probability) / std::log(1.0 - std::pow(1.0 - outlierRatio, sampleSize)));
    int bestInliers = 0;
    std::vector<PointPair> bestInlierSet;
    std::mt19937 rng(seed ? static_cast<std::mt19937::result_type>(seed) : std::random_device{}());
    if (numThreads < 1) numThreads = 1;
    // the reference scores chunks [t*chunk, min(N, (t+1)*chunk)), t < numThreads: the tail N % numThreads is never scored
    const int chunkSize = N / numThreads;
    const int scored = std::min(N, numThreads * chunkSize);
    std::vector<PointPair> sample;
    // scored points as four plain arrays: the scoring loop below has no branch and no store besides the flag
    std::vector<double> X0(scored), X1(scored), P0(scored), P1(scored);
    std::vector<unsigned char> flag(scored);
    for (int i = 0; i < scored; ++i) { X0[i] = data[i].first.x; X1[i] = data[i].first.y; P0[i] = data[i].second.x; P1[i] = data[i].second.y; }
    for (int iter = 0; iter < maxIterations; ++iter) {
        ++st.iterations;
        // ransac.cpp:142: std::sample = a uniformly random 8-subset in data order.  libstdc++ walks all N elements with
        // one draw each; Floyd's algorithm draws the same distribution with 8 draws (the reference's generator is
        // unseeded, so no particular sequence has to be reproduced).
        sample.clear();
        if (N <= sampleSize) {
            sample = data;
        } else {
            int idx[8], cnt = 0;
            for (int j = N - sampleSize; j < N; ++j) {
                const int t = static_cast<int>(std::uniform_int_distribution<int>(0, j)(rng));
                bool seen = false;
                for (int q = 0; q < cnt; ++q) seen |= idx[q] == t;
                idx[cnt++] = seen ? j : t;
            }
            std::sort(idx, idx + sampleSize);
            for (int q = 0; q < sampleSize; ++q) sample.push_back(data[idx[q]]);
        }
        FundamentalMatrix tempModel;
        tempModel.fit(sample);  // fewer than 8 correspondences: F stays zero, every Sampson error is "max"
        const Mat3 F = tempModel.getMatrix();
        const int count = score_correspondences(F.m, X0.data(), X1.data(), P0.data(), P1.data(), scored, sampsonThreshold, flag.data(), false);
        if (count > bestInliers) {  // the consensus set is materialised only when it is the best so far
            bestInliers = count;
            bestInlierSet.clear();
            for (int i = 0; i < scored; ++i)
                if (flag[i]) bestInlierSet.push_back(data[i]);
            outlierRatio = 1.0 - static_cast<double>(bestInliers) / N;
            const double denom = std::log(1.0 - std::pow(1.0 - outlierRatio, sampleSize));
            if (denom != 0.0) {
                // log(1-p)/denom can be +-inf (all inliers: denom = -inf gives 0, fine) or huge: clamp in double first
                const double it = std::log(1.0 - probability) / denom;
                maxIterations = it >= 2000.0 ? 2000 : (it <= 100.0 ? 100 : static_cast<int>(it));
            }
        }
    }
    st.best_inliers = bestInliers;
    model.fit(bestInlierSet);
    return st;
}

}  // namespace vo
