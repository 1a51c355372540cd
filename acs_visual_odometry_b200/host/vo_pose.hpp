// vo_pose.hpp — relative pose from the fundamental matrix, and the on-disk formats of the pipeline
// (SURVEY.md section 8f, rows N3 and N4).
//
// Mirrors /root/reference/PoseUpdate.hpp (PoseUpdate::getPose, :61-179) and PoseUpdate.cpp (flipZ :7-11,
// rotationMatrixToEulerAngles :14-33, quat_to_rot :35-42, readGTLine :43-50, writePoseCSV :52-69) without
// OpenCV: E = K^T F K, SVD, the four (R, t) candidates, linear triangulation of the inlier
// correspondences and the cheirality vote, translation scaled to the ground-truth step length.
#pragma once
#include <string>
#include <utility>
#include <vector>

#include "vo_linalg.hpp"

namespace vo {

struct Point2f {
    float x, y;
};

extern const Mat4 flipZ;  // diag(1, 1, -1, 1)

Vec3 rotationMatrixToEulerAngles(const Mat3& R);  // (roll, pitch, yaw)
Mat3 quat_to_rot(double w, double x, double y, double z);

// One ground-truth line: 12 whitespace-separated numbers = the top 3 rows of a 4x4 pose (KITTI format).
// Missing numbers leave the identity's entries, like the reference's stream extraction into Mat::eye.
Mat4 readGTLine(const std::string& line);
bool readGTFile(const std::string& path, std::vector<Mat4>& poses);
// 12 values per line, setprecision(9), comma after every value except the last of the line.
bool writePoseCSV(const std::string& filename, const std::vector<Mat4>& poses);
std::string formatPoseCSVLine(const Mat4& T);
// "%06d.png" under image_dir (VisualOdometry.cpp:65,73-74)
std::string frameFileName(const std::string& image_dir, size_t index);

struct PoseResult {
    Mat3 R;
    Vec3 t;
    int positive_depth = -1;  // cheirality votes of the chosen candidate
    int candidate = -1;       // 0: (R1, t)  1: (R1, -t)  2: (R2, t)  3: (R2, -t)
};

class PoseUpdate {
public:
    // KITTI sequence camera (PoseUpdate.hpp:36-39)
    Mat3 K{{7.188560000000e+02, 0, 6.071928000000e+02, 0, 7.188560000000e+02, 1.852157000000e+02, 0, 0, 1.0}};

    // Throws std::runtime_error("Degenerate essential matrix") like the reference when E has < 5 non-zeros.
    PoseResult getPoseEx(const Mat3& F, const std::vector<Point2f>& points1, const std::vector<Point2f>& points2,
                         double GT_pos_norm) const;
    std::pair<Mat3, Vec3> getPose(const Mat3& F, const std::vector<Point2f>& points1,
                                  const std::vector<Point2f>& points2, double GT_pos_norm) const {
        const PoseResult r = getPoseEx(F, points1, points2, GT_pos_norm);
        return {r.R, r.t};
    }
};

// cv::triangulatePoints for one correspondence: the right singular vector of the smallest singular value
// of the 4x4 DLT system built from P1 = [I | 0], P2 = [R | t] and normalised image points.
void triangulatePoint(const Mat3& R, const Vec3& t, double x1, double y1, double x2, double y2, double X[4]);
// The same for n correspondences; four at a time on AVX2 when the CPU has it (bit-identical to the scalar routine, which
// also serves the lanes that do not converge; VO_POSE_SCALAR=1 in the environment forces the scalar loop).
void triangulatePoints(const Mat3& R, const Vec3& t, const Point2f* n1, const Point2f* n2, size_t n, double (*X)[4]);

}  // namespace vo
