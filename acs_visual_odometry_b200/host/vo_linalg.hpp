// vo_linalg.hpp — the small dense double-precision linear algebra the host side of the pipeline needs
// (the reference uses Eigen in ransac.cpp and OpenCV core/calib3d in PoseUpdate.hpp; neither exists in
// this image, and the host rows of SURVEY.md section 8f ask for an Eigen/OpenCV-free build).
// Header-only, no dependencies beyond <cmath>.
#pragma once
#include <algorithm>
#include <array>
#include <cmath>
#include <vector>

namespace vo {

struct Vec3 {
    double v[3];
    double& operator()(int i) { return v[i]; }
    double operator()(int i) const { return v[i]; }
};

struct Mat3 {
    double m[9];  // row-major
    double& operator()(int r, int c) { return m[3 * r + c]; }
    double operator()(int r, int c) const { return m[3 * r + c]; }
    static Mat3 zero() { Mat3 a; for (double& x : a.m) x = 0.0; return a; }
    static Mat3 identity() { Mat3 a = zero(); a(0, 0) = a(1, 1) = a(2, 2) = 1.0; return a; }
};

struct Mat4 {
    double m[16];  // row-major
    double& operator()(int r, int c) { return m[4 * r + c]; }
    double operator()(int r, int c) const { return m[4 * r + c]; }
    static Mat4 identity() { Mat4 a; for (double& x : a.m) x = 0.0; a(0, 0) = a(1, 1) = a(2, 2) = a(3, 3) = 1.0; return a; }
};

inline Mat3 mul(const Mat3& a, const Mat3& b) {
    Mat3 c = Mat3::zero();
    for (int i = 0; i < 3; ++i)
        for (int k = 0; k < 3; ++k)
            for (int j = 0; j < 3; ++j) c(i, j) += a(i, k) * b(k, j);
    return c;
}
inline Vec3 mul(const Mat3& a, const Vec3& x) {
    Vec3 y;
    for (int i = 0; i < 3; ++i) y(i) = a(i, 0) * x(0) + a(i, 1) * x(1) + a(i, 2) * x(2);
    return y;
}
inline Mat3 transpose(const Mat3& a) {
    Mat3 t;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) t(i, j) = a(j, i);
    return t;
}
inline Mat3 scaled(const Mat3& a, double s) { Mat3 b = a; for (double& x : b.m) x *= s; return b; }
inline double det(const Mat3& a) {
    return a(0, 0) * (a(1, 1) * a(2, 2) - a(1, 2) * a(2, 1)) - a(0, 1) * (a(1, 0) * a(2, 2) - a(1, 2) * a(2, 0)) +
           a(0, 2) * (a(1, 0) * a(2, 1) - a(1, 1) * a(2, 0));
}
inline double frobenius(const Mat3& a) { double s = 0; for (double x : a.m) s += x * x; return std::sqrt(s); }
inline double dot(const Vec3& a, const Vec3& b) { return a(0) * b(0) + a(1) * b(1) + a(2) * b(2); }
inline double norm(const Vec3& a) { return std::sqrt(dot(a, a)); }
inline Vec3 cross(const Vec3& a, const Vec3& b) {
    return Vec3{{a(1) * b(2) - a(2) * b(1), a(2) * b(0) - a(0) * b(2), a(0) * b(1) - a(1) * b(0)}};
}

inline Mat4 mul(const Mat4& a, const Mat4& b) {
    Mat4 c; for (double& x : c.m) x = 0.0;
    for (int i = 0; i < 4; ++i)
        for (int k = 0; k < 4; ++k)
            for (int j = 0; j < 4; ++j) c(i, j) += a(i, k) * b(k, j);
    return c;
}
// inverse of a rigid/affine 4x4 whose last row is (0,0,0,1): general 3x3 inverse of the upper block
inline Mat4 inverse_affine(const Mat4& a) {
    Mat3 r;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) r(i, j) = a(i, j);
    const double d = det(r);
    Mat3 inv;
    inv(0, 0) = (r(1, 1) * r(2, 2) - r(1, 2) * r(2, 1)) / d; inv(0, 1) = (r(0, 2) * r(2, 1) - r(0, 1) * r(2, 2)) / d;
    inv(0, 2) = (r(0, 1) * r(1, 2) - r(0, 2) * r(1, 1)) / d; inv(1, 0) = (r(1, 2) * r(2, 0) - r(1, 0) * r(2, 2)) / d;
    inv(1, 1) = (r(0, 0) * r(2, 2) - r(0, 2) * r(2, 0)) / d; inv(1, 2) = (r(0, 2) * r(1, 0) - r(0, 0) * r(1, 2)) / d;
    inv(2, 0) = (r(1, 0) * r(2, 1) - r(1, 1) * r(2, 0)) / d; inv(2, 1) = (r(0, 1) * r(2, 0) - r(0, 0) * r(2, 1)) / d;
    inv(2, 2) = (r(0, 0) * r(1, 1) - r(0, 1) * r(1, 0)) / d;
    Mat4 out = Mat4::identity();
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) out(i, j) = inv(i, j);
        out(i, 3) = -(inv(i, 0) * a(0, 3) + inv(i, 1) * a(1, 3) + inv(i, 2) * a(2, 3));
    }
    return out;
}

// Unit null vector of an 8 x 9 matrix (rows = the 8-point equations of a minimal sample): Householder QR of A^T, the
// last column of Q is orthogonal to every row.  For a rank-8 sample this is the right singular vector of the zero
// singular value that JacobiSVD returns (up to sign and rounding), at about a tenth of the cost of the Jacobi sweeps;
// RANSAC only scores such models, and the Sampson error does not depend on the sign or scale of F.
inline void null_vector_8x9(const double* A, double f[9]) {
    double M[9][8], v[8][9], vv[8];
    for (int r = 0; r < 8; ++r)
        for (int c = 0; c < 9; ++c) M[c][r] = A[r * 9 + c];
    for (int k = 0; k < 8; ++k) {
        double nrm = 0.0;
        for (int i = k; i < 9; ++i) nrm += M[i][k] * M[i][k];
        nrm = std::sqrt(nrm);
        for (int i = 0; i < 9; ++i) v[k][i] = i < k ? 0.0 : M[i][k];
        v[k][k] += v[k][k] >= 0.0 ? nrm : -nrm;
        vv[k] = 0.0;
        for (int i = k; i < 9; ++i) vv[k] += v[k][i] * v[k][i];
        if (vv[k] == 0.0) continue;  // zero column: nothing to reflect
        for (int j = k + 1; j < 8; ++j) {
            double d = 0.0;
            for (int i = k; i < 9; ++i) d += v[k][i] * M[i][j];
            d = 2.0 * d / vv[k];
            for (int i = k; i < 9; ++i) M[i][j] -= d * v[k][i];
        }
    }
    for (int i = 0; i < 9; ++i) f[i] = i == 8 ? 1.0 : 0.0;
    for (int k = 7; k >= 0; --k) {
        if (vv[k] == 0.0) continue;
        double d = 0.0;
        for (int i = k; i < 9; ++i) d += v[k][i] * f[i];
        d = 2.0 * d / vv[k];
        for (int i = k; i < 9; ++i) f[i] -= d * v[k][i];
    }
}

// One-sided (Hestenes) Jacobi SVD of a row-major m x n matrix, any m >= 1, n <= 16.
//   A = U diag(s) V^T, singular values sorted in descending order.
// Outputs: s[n], V (n x n, row-major, columns = right singular vectors), and, if `Uout` is given, the
// n columns A v_i / s_i (m x n, row-major; columns with s_i == 0 are left as zero; only for m <= n... see below).
// Tall matrices (m > n, the final fit over hundreds of inliers) are first reduced to their n x n triangular
// factor with Householder reflections - the same preconditioning Eigen::JacobiSVD applies - so the sweeps cost
// O(n^3) instead of O(m n^2); R has the singular values and right singular vectors of A.  Uout is not
// available for m > n (never needed here).
// The null vector the 8-point algorithm needs is column n-1 of V (also when m < n: the working columns
// live in R^m, so at least n - m of them converge to zero).
// Dot product with four independent partial sums (a single running sum is a chain of dependent additions: the long
// columns of the final 8-point fit spent most of their time waiting for the adder).
inline double dot4(const double* a, const double* b, int n) {
    double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
    int i = 0;
    for (; i + 4 <= n; i += 4) { s0 += a[i] * b[i]; s1 += a[i + 1] * b[i + 1]; s2 += a[i + 2] * b[i + 2]; s3 += a[i + 3] * b[i + 3]; }
    for (; i < n; ++i) s0 += a[i] * b[i];
    return (s0 + s1) + (s2 + s3);
}

inline void jacobi_svd(const double* A, int m, int n, double* s, double* V, double* Uout = nullptr) {
    std::vector<double> Wheap;
    double Wstack[16 * 16];
    double* W;
    int rows = m;
    if (m > n) {
        // Householder QR in place on a TRANSPOSED copy (column j contiguous: the dot products and updates below walk
        // memory linearly; about a third of the time of a row-major sweep at 1270 x 9);
        // keep R (upper triangle) as the n x n working matrix
        Wheap.resize((size_t)m * n);
        double* Q = Wheap.data();
        for (int i = 0; i < m; ++i)
            for (int j = 0; j < n; ++j) Q[(size_t)j * m + i] = A[(size_t)i * n + j];
        for (int k = 0; k < n; ++k) {
            double* const ck = Q + (size_t)k * m;
            const double nrm2 = dot4(ck + k, ck + k, m - k);
            const double nrm = std::sqrt(nrm2);
            if (nrm == 0.0) continue;
            const double akk = ck[k];
            const double alpha = akk > 0 ? -nrm : nrm;
            // v = x - alpha e1 (stored in column k below the diagonal, v_k separately), H = I - 2 v v^T / (v^T v)
            const double vk = akk - alpha;
            const double vtv = nrm2 - akk * akk + vk * vk;
            if (vtv == 0.0) continue;
            for (int j = k + 1; j < n; ++j) {
                double* const cj = Q + (size_t)j * m;
                const double d = vk * cj[k] + dot4(ck + k + 1, cj + k + 1, m - k - 1);
                const double f = 2.0 * d / vtv;
                cj[k] -= f * vk;
                for (int i = k + 1; i < m; ++i) cj[i] -= f * ck[i];
            }
            ck[k] = alpha;
        }
        W = Wstack;
        for (int i = 0; i < n; ++i)
            for (int j = 0; j < n; ++j) W[i * n + j] = j >= i ? Q[(size_t)j * m + i] : 0.0;
        rows = n;
        Uout = nullptr;
    } else {
        W = Wstack;
        std::copy(A, A + (size_t)m * n, W);  // m <= n <= 16
    }
    for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) V[i * n + j] = i == j ? 1.0 : 0.0;
    const double eps = 1e-15;
    double scale2 = 0;  // squared Frobenius norm: columns below 1e-30 of it are numerically zero
    for (int i = 0; i < rows * n; ++i) scale2 += W[i] * W[i];
    const double zero2 = scale2 * 1e-60;
    for (int sweep = 0; sweep < 60; ++sweep) {
        bool rotated = false;
        for (int p = 0; p < n - 1; ++p)
            for (int q = p + 1; q < n; ++q) {
                double alpha = 0, beta = 0, gamma = 0;
                for (int i = 0; i < rows; ++i) {
                    const double wp = W[i * n + p], wq = W[i * n + q];
                    alpha += wp * wp; beta += wq * wq; gamma += wp * wq;
                }
                if (gamma == 0.0 || alpha <= zero2 || beta <= zero2 || std::fabs(gamma) <= eps * std::sqrt(alpha * beta)) continue;
                rotated = true;
                const double zeta = (beta - alpha) / (2.0 * gamma);
                const double t = (zeta >= 0 ? 1.0 : -1.0) / (std::fabs(zeta) + std::sqrt(1.0 + zeta * zeta));
                const double c = 1.0 / std::sqrt(1.0 + t * t), sn = c * t;
                for (int i = 0; i < rows; ++i) {
                    const double wp = W[i * n + p], wq = W[i * n + q];
                    W[i * n + p] = c * wp - sn * wq;
                    W[i * n + q] = sn * wp + c * wq;
                }
                for (int i = 0; i < n; ++i) {
                    const double vp = V[i * n + p], vq = V[i * n + q];
                    V[i * n + p] = c * vp - sn * vq;
                    V[i * n + q] = sn * vp + c * vq;
                }
            }
        if (!rotated) break;
    }
    std::array<double, 16> nrm{};
    std::array<int, 16> order{};
    for (int j = 0; j < n; ++j) {
        double a = 0;
        for (int i = 0; i < rows; ++i) a += W[i * n + j] * W[i * n + j];
        nrm[j] = std::sqrt(a); order[j] = j;
    }
    std::stable_sort(order.begin(), order.begin() + n, [&](int a, int b) { return nrm[a] > nrm[b]; });
    double Vs[16 * 16];
    for (int j = 0; j < n; ++j) {
        s[j] = nrm[order[j]];
        for (int i = 0; i < n; ++i) Vs[i * n + j] = V[i * n + order[j]];
    }
    std::copy(Vs, Vs + n * n, V);
    if (Uout)
        for (int j = 0; j < n; ++j)
            for (int i = 0; i < rows; ++i) Uout[(size_t)i * n + j] = s[j] > 0 ? W[i * n + order[j]] / s[j] : 0.0;
}

// Full SVD of a 3x3 matrix: A = U diag(s) V^T with orthogonal U and V (columns of U for vanishing singular
// values are completed to an orthonormal basis), s descending.
inline void svd3(const Mat3& A, Mat3& U, Vec3& s, Mat3& V) {
    double u[9], vv[9], sv[3];
    jacobi_svd(A.m, 3, 3, sv, vv, u);
    for (int i = 0; i < 9; ++i) { U.m[i] = u[i]; V.m[i] = vv[i]; }
    for (int i = 0; i < 3; ++i) s(i) = sv[i];
    const double tiny = 1e-14 * (sv[0] > 0 ? sv[0] : 1.0);
    auto col = [&](const Mat3& M, int j) { return Vec3{{M(0, j), M(1, j), M(2, j)}}; };
    auto setcol = [&](Mat3& M, int j, const Vec3& x) { M(0, j) = x(0); M(1, j) = x(1); M(2, j) = x(2); };
    if (sv[0] <= tiny) { U = Mat3::identity(); return; }
    if (sv[1] <= tiny) {  // rank 1: any unit vector orthogonal to u0
        const Vec3 u0 = col(U, 0);
        Vec3 a = std::fabs(u0(0)) < 0.9 ? Vec3{{1, 0, 0}} : Vec3{{0, 1, 0}};
        Vec3 u1 = cross(u0, a);
        const double n1 = norm(u1);
        for (int i = 0; i < 3; ++i) u1(i) /= n1;
        setcol(U, 1, u1);
        setcol(U, 2, cross(u0, u1));
        return;
    }
    if (sv[2] <= tiny) {  // rank 2: u2 = +-(u0 x u1), sign chosen so that det(U) det(V) > 0 is not forced
        setcol(U, 2, cross(col(U, 0), col(U, 1)));
    }
}

}  // namespace vo
