// vo_ransac.hpp — fundamental-matrix RANSAC on the host.
//
// Same interface and algorithm as the reference's ransac.hpp / ransac.cpp (normalised 8-point fit,
// Sampson-error consensus, adaptive iteration count), without Eigen: the two SVDs are one-sided Jacobi
// (vo_linalg.hpp).  Differences, all deliberate and listed in DESIGN.md:
//   * `seed` makes a run reproducible (the reference seeds from std::random_device, ransac.cpp:137);
//     seed == 0 keeps the reference behaviour;
//   * the consensus loop runs on the calling thread: the pipeline parallelises over frame pairs instead
//     (one RANSAC per pair is independent of every other pair).  `numThreads` is kept because it shapes
//     the result: the reference scores only the first numThreads * (N / numThreads) correspondences
//     (ransac.cpp:152-157), and so does this;
//   * inliers are collected in index order (the reference appends per-thread lists in completion order).
#pragma once
#include <cstdint>
#include <utility>
#include <vector>

#include "vo_linalg.hpp"

namespace vo {

struct Point {
    double x, y;
};
using PointPair = std::pair<Point, Point>;

// ransac.cpp:12-23
double computeSampsonError(const Mat3& F, const Vec3& x, const Vec3& x_prime);

class FundamentalMatrix {
    Mat3 F = Mat3::zero();
    std::vector<PointPair> inliers;

public:
    static Mat3 computeFundamentalMatrix(const std::vector<PointPair>& points);            // ransac.cpp:63-93
    static void normalizePoints(const std::vector<PointPair>& points, std::vector<Vec3>& n1,
                                std::vector<Vec3>& n2, Mat3& T1, Mat3& T2);               // ransac.cpp:25-61
    void fit(const std::vector<PointPair>& sample);                                       // ransac.cpp:95-99
    Mat3 getMatrix() const { return F; }
    int countInliers(const std::vector<PointPair>& data, double threshold) const;         // ransac.cpp:105-114
    const std::vector<PointPair>& getInliers() const { return inliers; }
};

// The consensus-scoring loop of Ransac::run on its own (flags + count for one model); AVX2 when the CPU has it, and the
// scalar loop with force_scalar: both must agree bit for bit (tests/test_host.py).
int score_correspondences(const double F[9], const double* X0, const double* X1, const double* P0, const double* P1, int n, double thr,
                          unsigned char* flag, bool force_scalar);

struct RansacStats {
    int iterations = 0;
    int best_inliers = 0;
};

class Ransac {
public:
    // ransac.cpp:120-194.  Returns the statistics of the run; `model` ends up fitted to the best consensus set.
    RansacStats run(FundamentalMatrix& model, const std::vector<PointPair>& data, double probability,
                    double sampsonThreshold, int numThreads, std::uint64_t seed = 0) const;
    // the reference's signature: the pool is accepted and ignored
    template <class Pool>
    RansacStats run(FundamentalMatrix& model, const std::vector<PointPair>& data, double probability,
                    double sampsonThreshold, int numThreads, Pool&) const {
        return run(model, data, probability, sampsonThreshold, numThreads, 0);
    }
};

}  // namespace vo
