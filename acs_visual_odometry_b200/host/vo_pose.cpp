// vo_pose.cpp — see vo_pose.hpp.
#include "vo_pose.hpp"

#include <array>
#include <cmath>
#include <cstdlib>
#include <limits>
#include <vector>
#include <cstdio>
#include <fstream>
#include <iomanip>
#include <sstream>
#include <stdexcept>

#if defined(__x86_64__)
#include <immintrin.h>
#define VO_HAVE_AVX2_PATH 1
#endif

namespace vo {

const Mat4 flipZ{{1, 0, 0, 0, 0, 1, 0, 0, 0, 0, -1, 0, 0, 0, 0, 1}};

// PoseUpdate.cpp:14-33
Vec3 rotationMatrixToEulerAngles(const Mat3& R) {
    const double sy = std::sqrt(R(0, 0) * R(0, 0) + R(1, 0) * R(1, 0));
    const bool singular = sy < 1e-6;
    double x, y, z;
    if (!singular) {
        x = std::atan2(R(2, 1), R(2, 2));
        y = std::atan2(-R(2, 0), sy);
        z = std::atan2(R(1, 0), R(0, 0));
    } else {
        x = std::atan2(-R(1, 2), R(1, 1));
        y = std::atan2(-R(2, 0), sy);
        z = 0;
    }
    return Vec3{{x, y, z}};
}

// PoseUpdate.cpp:35-42
Mat3 quat_to_rot(double w, double x, double y, double z) {
    return Mat3{{1 - 2 * y * y - 2 * z * z, 2 * x * y - 2 * z * w, 2 * x * z + 2 * y * w,
                 2 * x * y + 2 * z * w, 1 - 2 * x * x - 2 * z * z, 2 * y * z - 2 * x * w,
                 2 * x * z - 2 * y * w, 2 * y * z + 2 * x * w, 1 - 2 * x * x - 2 * y * y}};
}

// PoseUpdate.cpp:43-50
Mat4 readGTLine(const std::string& line) {
    std::stringstream ss(line);
    Mat4 T = Mat4::identity();
    for (int i = 0; i < 12; ++i) {
        double v;
        if (!(ss >> v)) break;  // operator>> leaves the target at 0 on failure in C++11; a short line is malformed either way
        T(i / 4, i % 4) = v;
    }
    return T;
}

bool readGTFile(const std::string& path, std::vector<Mat4>& poses) {
    std::ifstream in(path);
    if (!in.is_open()) return false;
    std::string line;
    while (std::getline(in, line)) poses.push_back(readGTLine(line));
    return true;
}

// PoseUpdate.cpp:52-69
std::string formatPoseCSVLine(const Mat4& T) {
    std::ostringstream os;
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 4; ++c) {
            os << std::setprecision(9) << T(r, c);
            if (!(r == 2 && c == 3)) os << ",";
        }
    os << "\n";
    return os.str();
}

bool writePoseCSV(const std::string& filename, const std::vector<Mat4>& poses) {
    std::ofstream file(filename);
    if (!file.is_open()) return false;
    for (const Mat4& T : poses) file << formatPoseCSVLine(T);
    return static_cast<bool>(file);
}

std::string frameFileName(const std::string& image_dir, size_t index) {
    char name[32];
    std::snprintf(name, sizeof(name), "%06zu.png", index);
    return image_dir + name;
}

void triangulatePoint(const Mat3& R, const Vec3& t, double x1, double y1, double x2, double y2, double X[4]) {
    // P1 = [I | 0], P2 = [R | t]; rows x*P[2] - P[0], y*P[2] - P[1] for both views
    double A[16];
    const double P1[3][4] = {{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}};
    double P2[3][4];
    for (int i = 0; i < 3; ++i) { for (int j = 0; j < 3; ++j) P2[i][j] = R(i, j); P2[i][3] = t(i); }
    for (int j = 0; j < 4; ++j) {
        A[0 * 4 + j] = x1 * P1[2][j] - P1[0][j];
        A[1 * 4 + j] = y1 * P1[2][j] - P1[1][j];
        A[2 * 4 + j] = x2 * P2[2][j] - P2[0][j];
        A[3 * 4 + j] = y2 * P2[2][j] - P2[1][j];
    }
    // cv::triangulatePoints takes the right singular vector of the smallest singular value of A, i.e. the eigenvector of
    // the smallest eigenvalue of M = A^T A.  Inverse iteration on the Cholesky factor of M (shifted by 1e-14 trace so
    // that it stays positive definite when the rays meet exactly) converges by (s4/s3)^2 per step: one or two steps
    // for inlier correspondences, against ~40 Jacobi rotations.  If it has not converged to 1e-11 after six steps
    // (rays nearly parallel to the baseline: s4 ~ s3) the Jacobi SVD decides, so the result is the SVD's in every case.
    double M[4][4];
    for (int i = 0; i < 4; ++i)
        for (int j = i; j < 4; ++j) {
            double acc = 0.0;
            for (int k = 0; k < 4; ++k) acc += A[k * 4 + i] * A[k * 4 + j];
            M[i][j] = M[j][i] = acc;
        }
    const double shift = 1e-14 * (M[0][0] + M[1][1] + M[2][2] + M[3][3]);
    double L[4][4] = {};
    bool ok = shift > 0.0;
    for (int i = 0; i < 4 && ok; ++i)
        for (int j = 0; j <= i; ++j) {
            double acc = M[i][j] + (i == j ? shift : 0.0);
            for (int k = 0; k < j; ++k) acc -= L[i][k] * L[j][k];
            if (i == j) { ok = acc > 0.0; L[i][i] = std::sqrt(acc); }
            else L[i][j] = acc / L[j][j];
        }
    if (ok) {
        const double d[4] = {1.0 / L[0][0], 1.0 / L[1][1], 1.0 / L[2][2], 1.0 / L[3][3]};
        double x[4] = {0.5, 0.5, 0.5, 0.5};
        for (int it = 0; it < 6; ++it) {
            double y[4];
            for (int i = 0; i < 4; ++i) {  // L z = x
                double acc = x[i];
                for (int k = 0; k < i; ++k) acc -= L[i][k] * y[k];
                y[i] = acc * d[i];
            }
            for (int i = 3; i >= 0; --i) {  // L^T y = z
                double acc = y[i];
                for (int k = i + 1; k < 4; ++k) acc -= L[k][i] * y[k];
                y[i] = acc * d[i];
            }
            const double nrm = std::sqrt(y[0] * y[0] + y[1] * y[1] + y[2] * y[2] + y[3] * y[3]);
            if (!(nrm > 0.0) || !std::isfinite(nrm)) break;
            const double dotp = y[0] * x[0] + y[1] * x[1] + y[2] * x[2] + y[3] * x[3];
            const double scale = (dotp < 0.0 ? -1.0 : 1.0) / nrm;
            double change = 0.0;
            for (int i = 0; i < 4; ++i) {
                const double v = scale * y[i];
                change = std::fmax(change, std::fabs(v - x[i]));
                x[i] = v;
            }
            if (it > 0 && change < 1e-11) {  // reaching 1e-11 within six steps needs a ratio (s4/s3)^2 < ~0.02: the remaining error is < 1e-12
                for (int i = 0; i < 4; ++i) X[i] = x[i];
                return;
            }
        }
    }
    double sv[4], V[16];
    jacobi_svd(A, 4, 4, sv, V);
    for (int i = 0; i < 4; ++i) X[i] = V[i * 4 + 3];
}

#ifdef VO_HAVE_AVX2_PATH
// triangulatePoint for four correspondences at once, one per AVX2 lane: the same operations in the same order, separate
// multiplies and adds (no FMA), so every lane that converges returns bit for bit what the scalar routine returns.
// Lanes whose Cholesky factor breaks down or whose inverse iteration has not converged after six steps are reported
// in the returned mask and left to the scalar routine (which then takes its Jacobi fallback).
__attribute__((target("avx2"))) static unsigned triangulate4(const Mat3& R, const Vec3& t, const double* x1, const double* y1, const double* x2,
                                                             const double* y2, double X[4][4]) {
    typedef __m256d V;
#define S(v) _mm256_set1_pd(v)
    const V vx1 = _mm256_loadu_pd(x1), vy1 = _mm256_loadu_pd(y1), vx2 = _mm256_loadu_pd(x2), vy2 = _mm256_loadu_pd(y2);
    const double P1[3][4] = {{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}};
    double P2[3][4];
    for (int i = 0; i < 3; ++i) { for (int j = 0; j < 3; ++j) P2[i][j] = R(i, j); P2[i][3] = t(i); }
    V A[4][4];
    for (int j = 0; j < 4; ++j) {
        A[0][j] = _mm256_sub_pd(_mm256_mul_pd(vx1, S(P1[2][j])), S(P1[0][j]));
        A[1][j] = _mm256_sub_pd(_mm256_mul_pd(vy1, S(P1[2][j])), S(P1[1][j]));
        A[2][j] = _mm256_sub_pd(_mm256_mul_pd(vx2, S(P2[2][j])), S(P2[0][j]));
        A[3][j] = _mm256_sub_pd(_mm256_mul_pd(vy2, S(P2[2][j])), S(P2[1][j]));
    }
    V M[4][4];
    for (int i = 0; i < 4; ++i)
        for (int j = i; j < 4; ++j) {
            V acc = S(0.0);
            for (int k = 0; k < 4; ++k) acc = _mm256_add_pd(acc, _mm256_mul_pd(A[k][i], A[k][j]));
            M[i][j] = M[j][i] = acc;
        }
    const V shift = _mm256_mul_pd(S(1e-14), _mm256_add_pd(_mm256_add_pd(_mm256_add_pd(M[0][0], M[1][1]), M[2][2]), M[3][3]));
    V ok = _mm256_cmp_pd(shift, S(0.0), _CMP_GT_OQ);
    V L[4][4];
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j <= i; ++j) {
            V acc = i == j ? _mm256_add_pd(M[i][j], shift) : _mm256_add_pd(M[i][j], S(0.0));
            for (int k = 0; k < j; ++k) acc = _mm256_sub_pd(acc, _mm256_mul_pd(L[i][k], L[j][k]));
            if (i == j) { ok = _mm256_and_pd(ok, _mm256_cmp_pd(acc, S(0.0), _CMP_GT_OQ)); L[i][i] = _mm256_sqrt_pd(acc); }
            else L[i][j] = _mm256_div_pd(acc, L[j][j]);
        }
    const V d[4] = {_mm256_div_pd(S(1.0), L[0][0]), _mm256_div_pd(S(1.0), L[1][1]), _mm256_div_pd(S(1.0), L[2][2]), _mm256_div_pd(S(1.0), L[3][3])};
    V x[4] = {S(0.5), S(0.5), S(0.5), S(0.5)};
    V active = ok, done = S(0.0);  // lane masks (all-ones / zero)
    const V sign_bit = S(-0.0), inf = S(std::numeric_limits<double>::infinity());
    for (int it = 0; it < 6; ++it) {
        V y[4];
        for (int i = 0; i < 4; ++i) {  // L z = x
            V acc = x[i];
            for (int k = 0; k < i; ++k) acc = _mm256_sub_pd(acc, _mm256_mul_pd(L[i][k], y[k]));
            y[i] = _mm256_mul_pd(acc, d[i]);
        }
        for (int i = 3; i >= 0; --i) {  // L^T y = z
            V acc = y[i];
            for (int k = i + 1; k < 4; ++k) acc = _mm256_sub_pd(acc, _mm256_mul_pd(L[k][i], y[k]));
            y[i] = _mm256_mul_pd(acc, d[i]);
        }
        const V nrm = _mm256_sqrt_pd(_mm256_add_pd(_mm256_add_pd(_mm256_add_pd(_mm256_mul_pd(y[0], y[0]), _mm256_mul_pd(y[1], y[1])),
                                                                 _mm256_mul_pd(y[2], y[2])), _mm256_mul_pd(y[3], y[3])));
        // nrm > 0 and finite, else the scalar routine breaks out (to its fallback): such lanes stop being active
        const V good = _mm256_and_pd(_mm256_cmp_pd(nrm, S(0.0), _CMP_GT_OQ), _mm256_cmp_pd(nrm, inf, _CMP_LT_OQ));
        active = _mm256_and_pd(active, good);
        const V dotp = _mm256_add_pd(_mm256_add_pd(_mm256_add_pd(_mm256_mul_pd(y[0], x[0]), _mm256_mul_pd(y[1], x[1])), _mm256_mul_pd(y[2], x[2])),
                                     _mm256_mul_pd(y[3], x[3]));
        const V one = _mm256_blendv_pd(S(1.0), S(-1.0), _mm256_cmp_pd(dotp, S(0.0), _CMP_LT_OQ));
        const V scale = _mm256_div_pd(one, nrm);
        V change = S(0.0);
        for (int i = 0; i < 4; ++i) {
            const V v = _mm256_mul_pd(scale, y[i]);
            change = _mm256_max_pd(change, _mm256_andnot_pd(sign_bit, _mm256_sub_pd(v, x[i])));
            x[i] = _mm256_blendv_pd(x[i], v, active);  // lanes that are done (or broken) keep what they have
        }
        if (it > 0) {
            const V conv = _mm256_and_pd(active, _mm256_cmp_pd(change, S(1e-11), _CMP_LT_OQ));
            done = _mm256_or_pd(done, conv);
            active = _mm256_andnot_pd(conv, active);
        }
        if (_mm256_movemask_pd(active) == 0) break;
    }
    double xs[4][4];
    for (int i = 0; i < 4; ++i) _mm256_storeu_pd(xs[i], x[i]);
    for (int lane = 0; lane < 4; ++lane)
        for (int i = 0; i < 4; ++i) X[lane][i] = xs[i][lane];
    return ~static_cast<unsigned>(_mm256_movemask_pd(done)) & 0xFu;  // lanes the scalar routine has to do
#undef S
}

static bool pose_use_avx2() {
    static const bool ok = __builtin_cpu_supports("avx2") && !std::getenv("VO_POSE_SCALAR");
    return ok;
}
#endif

// triangulatePoint over n correspondences (normalised float coordinates); X[i] = the unit null vector of point i
void triangulatePoints(const Mat3& R, const Vec3& t, const Point2f* n1, const Point2f* n2, size_t n, double (*X)[4]) {
    size_t i = 0;
#ifdef VO_HAVE_AVX2_PATH
    if (pose_use_avx2()) {
        for (; i + 4 <= n; i += 4) {
            double x1[4], y1[4], x2[4], y2[4];
            for (int l = 0; l < 4; ++l) { x1[l] = n1[i + l].x; y1[l] = n1[i + l].y; x2[l] = n2[i + l].x; y2[l] = n2[i + l].y; }
            const unsigned todo = triangulate4(R, t, x1, y1, x2, y2, X + i);
            for (int l = 0; l < 4; ++l)
                if (todo & (1u << l)) triangulatePoint(R, t, x1[l], y1[l], x2[l], y2[l], X[i + l]);
        }
    }
#endif
    for (; i < n; ++i) triangulatePoint(R, t, n1[i].x, n1[i].y, n2[i].x, n2[i].y, X[i]);
}

// PoseUpdate.hpp:61-179
PoseResult PoseUpdate::getPoseEx(const Mat3& F, const std::vector<Point2f>& points1, const std::vector<Point2f>& points2,
                                 double GT_pos_norm) const {
    Mat3 E = mul(mul(transpose(K), F), K);
    const double en = frobenius(E);
    E = scaled(E, 1.0 / en);
    int nonzero = 0;
    for (double v : E.m) nonzero += v != 0.0;  // cv::countNonZero: NaN counts as non-zero
    if (nonzero < 5) throw std::runtime_error("Degenerate essential matrix");

    Mat3 U, V;
    Vec3 S;
    svd3(E, U, S, V);
    Mat3 Vt = transpose(V);
    if (det(U) < 0) U = scaled(U, -1.0);
    if (det(Vt) < 0) Vt = scaled(Vt, -1.0);
    const Mat3 W{{0, -1, 0, 1, 0, 0, 0, 0, 1}};
    Mat3 R1 = mul(mul(U, W), Vt);
    Mat3 R2 = mul(mul(U, transpose(W)), Vt);
    if (det(R1) < 0) R1 = scaled(R1, -1.0);
    if (det(R2) < 0) R2 = scaled(R2, -1.0);
    const Vec3 t{{U(0, 2), U(1, 2), U(2, 2)}};
    const Vec3 tn{{-t(0), -t(1), -t(2)}};
    const Mat3* Rs[4] = {&R1, &R1, &R2, &R2};
    const Vec3* ts[4] = {&t, &tn, &t, &tn};

    // cv::undistortPoints(points, K, no distortion): (x - cx) / fx, (y - cy) / fy, stored as float
    const size_t n = std::min(points1.size(), points2.size());
    std::vector<Point2f> n1(n), n2(n);
    for (size_t i = 0; i < n; ++i) {
        n1[i] = Point2f{static_cast<float>((points1[i].x - K(0, 2)) / K(0, 0)), static_cast<float>((points1[i].y - K(1, 2)) / K(1, 1))};
        n2[i] = Point2f{static_cast<float>((points2[i].x - K(0, 2)) / K(0, 0)), static_cast<float>((points2[i].y - K(1, 2)) / K(1, 1))};
    }

    // Cheirality votes (PoseUpdate.hpp:98-146).  The reference triangulates every point for each of the four candidates
    // (R1, t), (R1, -t), (R2, t), (R2, -t), counts the points in front of both cameras and keeps the first candidate with
    // the largest count.  Two things make that cheaper without changing the winner or its count:
    //  * (R, -t) needs no triangulation of its own.  Its DLT matrix is the one of (R, t) with the last column negated
    //    (the rows of view 1 have a zero there), so its null vector is (x, y, z, -w): the homogenised point, z1 and z2
    //    all change sign, and its vote is "both depths negative" on the triangulation of (R, t).  That equals a separate
    //    triangulation of [R|-t] - what the reference does - up to the solver's tolerance, not bit for bit: the inverse
    //    iteration starts from a fixed vector, not the mirrored one, so the two null vectors agree to about 1e-11 before
    //    the float cast and a point with |w| at 1e-6 or a depth at 0 can vote differently.  On near-degenerate data
    //    (millimetre baselines, a third of the points near infinity) winner and count still equal four brute-force
    //    triangulations (tests/test_host.py::test_cheirality_shortcut_against_four_triangulations);
    //  * only the argmax and its count are used afterwards, so a rotation is dropped as soon as neither of its candidates
    //    can reach the leader's count even if every remaining point voted for it (its partial counts are then below the
    //    winner's and cannot be chosen).
    int positive[4] = {0, 0, 0, 0};
    std::vector<std::array<double, 4>> scratch;
    auto vote = [&](int c, size_t i0, size_t i1) {  // candidates c = (R, +t) and c + 1 = (R, -t) over points [i0, i1)
        const Mat3& R = *Rs[c];
        const Vec3& tc = *ts[c];
        std::vector<std::array<double, 4>>& XS = scratch;
        XS.resize(i1 - i0);
        triangulatePoints(R, tc, n1.data() + i0, n2.data() + i0, i1 - i0, reinterpret_cast<double(*)[4]>(XS.data()));
        for (size_t i = i0; i < i1; ++i) {
            double* X = XS[i - i0].data();
            for (int q = 0; q < 4; ++q) X[q] = static_cast<double>(static_cast<float>(X[q]));  // triangulatePoints returns CV_32F for float input
            const double w = X[3];
            if (std::fabs(w) < 1e-6) continue;
            const Vec3 P{{X[0] / w, X[1] / w, X[2] / w}};
            const double z1 = P(2);
            const Vec3 P2 = mul(R, P);
            const double z2 = P2(2) + tc(2);
            if (z1 > 0 && z2 > 0) ++positive[c];
            if (z1 < 0 && z2 < 0) ++positive[c + 1];
        }
    };
    // the first block for both rotations, then the leading rotation to the end, then the other one only for as long as
    // one of its candidates could still reach (or tie) the leader's final count
    const size_t block = 64, first = std::min(n, block);
    vote(0, 0, first);
    vote(2, 0, first);
    const int lead_pair = std::max(positive[2], positive[3]) > std::max(positive[0], positive[1]) ? 2 : 0, other = 2 - lead_pair;
    vote(lead_pair, first, n);
    const int lead = std::max(positive[lead_pair], positive[lead_pair + 1]);
    bool alive[4] = {true, true, true, true};
    for (size_t i0 = first; i0 < n; i0 += block) {
        const int remaining = static_cast<int>(n - i0);
        if (std::max(positive[other], positive[other + 1]) + remaining < lead) { alive[other] = alive[other + 1] = false; break; }
        vote(other, i0, std::min(n, i0 + block));
    }
    PoseResult best;
    for (int c = 0; c < 4; ++c)
        if (alive[c] && positive[c] > best.positive_depth) { best.positive_depth = positive[c]; best.R = *Rs[c]; best.t = *ts[c]; best.candidate = c; }
    if (det(best.R) < 0) best.R = scaled(best.R, -1.0);
    const double tnorm = norm(best.t);
    if (tnorm > 1e-6)
        for (int i = 0; i < 3; ++i) best.t(i) *= GT_pos_norm / tnorm;
    return best;
}

}  // namespace vo
