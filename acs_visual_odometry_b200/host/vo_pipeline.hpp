// vo_pipeline.hpp — the frame loop of VisualOdometry as a pipelined driver.
//
// Mirrors /root/reference/VisualOdometry.h / VisualOdometry.cpp: same class name, constructor arguments,
// compute_descriptor_with_key_points / match_descriptors / run, same skip rules (< 8 matches :108,
// < 8 inliers :147, unreadable frame :78-83), same output file (writePoseCSV of flipZ * T_curr per frame).
// What changes is the schedule.  The reference handles one frame at a time: imread -> extract -> match ->
// RANSAC -> pose.  Here frames move through four overlapped stages in chunks of `batch` frames:
//   decode (host threads)  ->  front end on the GPU (vo_sequence_run: H2D, the four kernels, D2H)
//   ->  RANSAC + pose per frame pair (host threads, pairs are independent)  ->  sequential pose composition.
// The GPU stage matches every consecutive pair (i-1, i) speculatively.  The reference advances its query
// frame only when a pair succeeds (VisualOdometry.cpp:164-166); when a pair fails, the next pair is
// (last valid frame, i) instead, and the composition stage recomputes exactly that pair on demand
// (two vo_extract calls + vo_match).  Failures are rare, so the speculation almost always holds.
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <utility>
#include <vector>

#include "../../include/vofront.h"
#include "vo_image_io.hpp"
#include "vo_pose.hpp"
#include "vo_ransac.hpp"

namespace vo {

struct KeyPoint {  // the fields of cv::KeyPoint the pipeline uses
    struct { float x, y; } pt;
    float size = 1.0f;
};

struct RunStats {
    size_t frames = 0, pairs_ok = 0, skipped_frames = 0, missing_frames = 0, respeculated = 0;
    size_t capacity_frames = 0;    // frames whose NMS candidate list overflowed: treated as frames without features
    size_t stale_model_pairs = 0;  // reuse_model only: pairs whose pose came from the previous pair's model
    double wall_s = 0;        // whole run
    double decode_s = 0;      // sum over chunks, overlapped with the rest
    double frontend_s = 0;    // sum of vo_sequence_run calls (H2D + kernels + D2H)
    double host_s = 0;        // sum of RANSAC + pose stage wall time per chunk (parallel over pairs)
    double compose_s = 0;
    long long ransac_iterations = 0;
    double mean_keypoints = 0, mean_matches = 0, mean_inliers = 0;
};

class VisualOdometry {
public:
    // kernel_filename: the reference's OpenCL program binary (VisualOdometry.cpp:8-24); accepted, unused.
    explicit VisualOdometry(const std::string& kernel_filename, std::size_t num_threads, int device = 0);
    ~VisualOdometry();
    VisualOdometry(const VisualOdometry&) = delete;
    VisualOdometry& operator=(const VisualOdometry&) = delete;

    std::pair<std::vector<std::vector<uint8_t>>, std::vector<KeyPoint>> compute_descriptor_with_key_points(const GrayImage& image);
    std::vector<std::pair<int, int>> match_descriptors(const std::vector<std::vector<uint8_t>>& desc1,
                                                       const std::vector<std::vector<uint8_t>>& desc2);
    void run(const std::string image_dir, size_t num_images, const std::string pose_file, const std::string output_csv);

    // knobs the reference hard-codes
    vo_params params;            // preset VO_PRESET_VO_GPU_REF (what VisualOdometry links), ratio 0.75, REF32
    int batch = 64;              // frames per chunk
    int cand_cap = 0;            // NMS candidates per frame the context holds (vo_create_ex; 0 = its default formula)
    std::uint64_t seed = 0;      // 0: std::random_device per pair, like the reference; else pair i uses seed + i
    double ransac_probability = 0.99, sampson_threshold = 1.0;  // VisualOdometry.cpp:125
    // The reference keeps ONE FundamentalMatrix across frames (VisualOdometry.cpp:49); when a RANSAC run ends without a
    // consensus set of 8, FundamentalMatrix::fit returns early (ransac.cpp:95-96) and the frame is processed with the
    // PREVIOUS pair's F and inliers instead of being skipped.  true restates that; false (default) fits a fresh model
    // per pair, so the "< 8 inliers" guard (VisualOdometry.cpp:147) skips the frame.  Both are checked against the
    // sequential restatement oracle/vo_loop_oracle.py (tests/test_gpu_pipeline.py).
    bool reuse_model = false;
    bool verbose = true;
    RunStats stats;
    const std::vector<Mat4>& poses() const { return estimated_poses_; }
    // The GPU context is built on first use and lives as long as the object (the reference builds its OpenCL context in
    // the constructor, VisualOdometry.cpp:9-25).  A caller that runs several sequences through short-lived objects (the C
    // ABI's vo_pipeline_run) can hand a context over instead of paying the device allocations again.
    struct Context { vo_ctx* ctx = nullptr; int rows = 0, cols = 0, max_keypoints = 0, batch = 0, cand_cap = 0, device = 0; };
    void adopt_context(const Context& c);   // takes ownership if it fits this object's parameters, else destroys it
    Context release_context();              // gives ownership away (ctx == nullptr if none was built)

private:
    struct PairResult;
    void ensure_ctx(int rows, int cols);
    void solve_pair(const float* kp1, const float* kp2, const int32_t* pairs, int n_matches, size_t pair_index,
                    PairResult& out) const;

    std::size_t number_of_threads_;
    int device_;
    vo_ctx* ctx_ = nullptr;
    int ctx_rows_ = 0, ctx_cols_ = 0, ctx_batch_ = 0, ctx_max_kp_ = 0;
    std::vector<Mat4> estimated_poses_;
};

// Returns the pooled page-locked staging buffers of finished runs to the driver; the number of bytes freed.
size_t release_pinned_pool();

}  // namespace vo
