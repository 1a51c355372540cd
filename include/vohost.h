/* vohost.h — C ABI of the host side of the pipeline (libvohost.so): the consumers of the front end's match
 * lists that SURVEY.md section 8f ranks next - the VisualOdometry frame loop (N1), fundamental-matrix
 * RANSAC (N2), the on-disk formats (N3) and PoseUpdate::getPose (N4).  C++ callers use the classes in
 * acs_visual_odometry_b200/host/ directly (same names as the reference's: VisualOdometry, Ransac,
 * FundamentalMatrix, PoseUpdate, readGTLine, writePoseCSV); this header is what tests and other languages bind.
 * Everything here runs on the host except vo_pipeline_run, which drives the GPU front end through vofront.h.
 * Return value: 0 on success, negative on failure (-1 bad argument, -2 I/O, -3 numerical / runtime error);
 * nothing throws across the ABI.  Points are (x1, y1, x2, y2) rows of doubles: frame-1 and frame-2 pixel
 * coordinates of one correspondence.  Matrices are row-major. */
#ifndef VOHOST_H
#define VOHOST_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* FundamentalMatrix::computeFundamentalMatrix (ransac.cpp:63-93): normalised 8-point fit, rank 2 enforced. n >= 8. */
int vo_host_fundamental(const double* pts, int n, double F[9]);
/* FundamentalMatrix::normalizePoints (ransac.cpp:25-61): n1, n2 receive n x 3 normalised homogeneous points. */
int vo_host_normalize_points(const double* pts, int n, double* n1, double* n2, double T1[9], double T2[9]);
/* computeSampsonError (ransac.cpp:12-23) for n correspondences. */
int vo_host_sampson(const double F[9], const double* pts, int n, double* err);
/* The consensus-scoring loop of Ransac::run on its own (ransac.cpp:159-168): flags[i] = Sampson error of correspondence
 * i < threshold, *count their number.  force_scalar 0: the AVX2 loop when the CPU has it; 1: the scalar loop.  Both give
 * the same flags (separate multiplies and adds, no FMA). */
int vo_host_score(const double F[9], const double* pts, int n, double threshold, int force_scalar, uint8_t* flags, int* count);
/* Ransac::run (ransac.cpp:120-194).  seed 0 = std::random_device like the reference.  inliers (may be NULL)
 * receives the consensus set the final fit used, n_inliers its size (capacity n rows), iterations the count run. */
int vo_host_ransac(const double* pts, int n, double probability, double sampson_threshold, int num_threads,
                   uint64_t seed, double F[9], double* inliers, int* n_inliers, int* iterations);
/* PoseUpdate::getPose (PoseUpdate.hpp:61-179) with the KITTI intrinsics; K (may be NULL) overrides them.
 * p1 / p2: n x 2 floats (inlier pixel coordinates).  votes (may be NULL): cheirality count of the winner. */
int vo_host_get_pose(const double F[9], const float* p1, const float* p2, int n, double gt_step, const double* K,
                     double R[9], double t[3], int* votes);
/* cv::triangulatePoints as getPose uses it (PoseUpdate.hpp:114): P1 = [I | 0], P2 = [R | t], n1 / n2 = n x 2 normalised
 * float coordinates; X receives n x 4 unit null vectors.  force_scalar 0: four points at a time on AVX2 when the CPU has
 * it; 1: one at a time.  Both give the same bits. */
int vo_host_triangulate(const double R[9], const double t[3], const float* n1, const float* n2, int n, int force_scalar, double* X);
/* readGTLine (PoseUpdate.cpp:43-50) and one line of writePoseCSV (:52-69); buf receives a NUL-terminated line. */
int vo_host_read_gt_line(const char* line, double T[16]);
int vo_host_format_pose_csv(const double T[16], char* buf, int cap);
/* rotationMatrixToEulerAngles (PoseUpdate.cpp:14-33): roll, pitch, yaw. */
int vo_host_euler(const double R[9], double rpy[3]);
/* cv::imread(path, IMREAD_GRAYSCALE) stand-in: 8-bit PNG (gray / RGB / RGBA) or binary PGM.  With out == NULL
 * only the size is returned. */
int vo_host_read_gray(const char* path, uint8_t* out, size_t cap, int* rows, int* cols);
int vo_host_write_png(const char* path, const uint8_t* gray, int rows, int cols);
int vo_host_write_pgm(const char* path, const uint8_t* gray, int rows, int cols);

/* VisualOdometry::run (VisualOdometry.cpp:38-193) as the pipelined driver of host/vo_pipeline.hpp:
 * image_dir + "%06d.png" frames, KITTI pose file, output CSV.  preset: VO_PRESET_* of vofront.h; batch: frames
 * per chunk (0 = default 64); seed as above (pair i uses seed + i).
 * stats[16] (may be NULL): 0 frames, 1 pairs solved, 2 skipped frames, 3 missing frames, 4 re-matched pairs,
 * 5 wall s, 6 decode s, 7 front-end s, 8 RANSAC+pose s, 9 compose s, 10 RANSAC iterations, 11 mean keypoints,
 * 12 mean matches, 13 mean inliers. */
/* The GPU context a call builds is kept for the next call on the same device with the same preset and batch (one per
 * process), so a second sequence does not pay the device allocations again. */
int vo_pipeline_run(int device, int num_threads, const char* image_dir, int num_images, const char* pose_file,
                    const char* output_csv, int preset, int batch, uint64_t seed, double* stats);
/* The same with flags: VO_PIPE_REUSE_MODEL keeps one FundamentalMatrix across frames like the reference does
 * (VisualOdometry.cpp:49): a pair whose RANSAC run ends without a consensus set of 8 is then processed with the
 * previous pair's F and inliers (ransac.cpp:95-96 returns early) instead of being skipped.  stats[14] counts them;
 * stats[15] counts frames whose NMS candidate list overflowed (treated as frames without features, not as an error). */
#define VO_PIPE_REUSE_MODEL 1
int vo_pipeline_run_ex(int device, int num_threads, const char* image_dir, int num_images, const char* pose_file,
                       const char* output_csv, int preset, int batch, uint64_t seed, int flags, double* stats);
/* Gives back what vo_pipeline_run keeps between calls: the GPU context of the last run (all its device memory) and the
 * pooled page-locked staging buffers.  For long-lived processes that are done with the pipeline for a while; the next
 * vo_pipeline_run simply builds both again.  The pool is bounded in any case: at most VO_PINNED_POOL_MB (environment,
 * default 1024) megabytes stay pooled, and a pooled block is reused only for requests of at least half its size. */
int vo_pipeline_release(void);
/* message of the last failure on this thread */
const char* vo_host_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* VOHOST_H */
