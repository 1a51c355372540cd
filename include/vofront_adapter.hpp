// vofront_adapter.hpp — header-only C++17 adapter: the reference's own entry points on top of the C ABI
// (include/vofront.h), so VisualOdometry / main_pipeline / matching.cpp swap the B200 front-end in by
// changing one #include.  Needs only <opencv2/core.hpp> types (cv::Mat, cv::KeyPoint);
// compiles against real OpenCV or any header providing those two types.
//
// Names, argument meaning and error behaviour follow the reference (paths relative to the reference root):
//   feature_extraction_manager_with_points(cv::Mat&, thread_pool&)
//        feature_extraction_parallel/feature_extraction_parallel.h:20, .cpp:215-349   (CPU thread-pool variant:
//        BGR or gray input, threshold 25000, NMS 5, top 1500)
//   feature_extraction_manager_with_points(const cv::Mat&, const GPU_settings&)
//        feature_extraction_parallel_GPU/feature_extraction_parallel_GPU.h:55, .cpp:194-306 (what VisualOdometry
//        links: gray input, threshold 20000, NMS 3, top 2000)
//   matchCustomBinaryDescriptorsThreadPool(desc1, desc2, pool, numThreads, ratioThreshold)
//        feature_matching_parallel/feature_matching_parallel.h:35-41, .cpp:49-113
//   GPU_settings    feature_extraction_parallel_GPU/corner_detection_parallel_GPU.h:48-52 (here: one vo_ctx)
//   CornerDetectionParallel / FREAK_Parallel stage statics, whole-image form
//        feature_extraction_parallel/corner_detection_parallel.h:38-47, FREAK_feature_descriptor_parallel.h:104-110
// Errors: the reference throws std::runtime_error (feature_extraction_parallel_GPU.cpp:91,98); so does this adapter
// for every non-zero C-ABI status.  Empty descriptor sets give an empty match list (feature_matching_parallel.cpp:57).
// `thread_pool&` and `numThreads` are accepted and ignored: CUDA streams replace the pool.
#pragma once
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "vofront.h"

class thread_pool;  // feature_extraction_parallel/threadpool.h:15 — only ever taken by reference

namespace vofront {

inline void check(int rc, vo_ctx* ctx, const char* what) {
    if (rc != VO_OK) throw std::runtime_error(std::string(what) + ": " + vo_last_error(ctx));
}

// One GPU context, created once (the reference builds GPU_settings once in the VisualOdometry ctor,
// VisualOdometry.cpp:14-23).  Copyable like the reference's struct of OpenCL handles.
struct Context {
    std::shared_ptr<vo_ctx> ctx;
    int max_rows = 0, max_cols = 0, max_keypoints = 0;

    Context() = default;
    Context(int device, int rows, int cols, int max_kp = 2000, int n_slots = 4)
        : max_rows(rows), max_cols(cols), max_keypoints(max_kp) {
        vo_ctx* c = nullptr;
        const int rc = vo_create(&c, device, rows, cols, max_kp, n_slots);
        if (rc != VO_OK) throw std::runtime_error(std::string("vo_create: ") + vo_last_error(nullptr));
        ctx.reset(c, [](vo_ctx* p) { vo_destroy(p); });
    }
    vo_ctx* get() const { return ctx.get(); }
    bool fits(int rows, int cols, int kp) const { return ctx && rows <= max_rows && cols <= max_cols && kp <= max_keypoints; }
};

// Process-wide context for the thread_pool overloads (which carry no GPU handle); grown on demand.
inline Context& default_context(int rows, int cols, int max_kp) {
    static Context c;
    if (!c.fits(rows, cols, max_kp))
        c = Context(0, rows > c.max_rows ? rows : c.max_rows, cols > c.max_cols ? cols : c.max_cols,
                    max_kp > c.max_keypoints ? max_kp : c.max_keypoints);
    return c;
}

using Descriptors = std::vector<std::vector<uint8_t>>;  // n x 512 bytes, one 0/1 byte per test (the reference layout)

inline std::pair<Descriptors, std::vector<cv::KeyPoint>> fetch(const Context& c, int slot) {
    int n = 0;
    check(vo_frame_fetch(c.get(), slot, &n, nullptr, nullptr, nullptr), c.get(), "vo_frame_fetch");
    std::vector<float> xy(2 * (size_t)(n > 0 ? n : 1));
    std::vector<uint8_t> flat(512 * (size_t)(n > 0 ? n : 1));
    if (n > 0) check(vo_frame_fetch(c.get(), slot, &n, xy.data(), flat.data(), nullptr), c.get(), "vo_frame_fetch");
    std::pair<Descriptors, std::vector<cv::KeyPoint>> out;
    out.first.reserve(n);
    out.second.reserve(n);
    for (int k = 0; k < n; ++k) {
        out.first.emplace_back(flat.begin() + 512 * (size_t)k, flat.begin() + 512 * (size_t)(k + 1));
        out.second.emplace_back(cv::Point2f(xy[2 * k], xy[2 * k + 1]), 1.0f);  // corner_detection_parallel.cpp:203
    }
    return out;
}

inline std::pair<Descriptors, std::vector<cv::KeyPoint>> extract(const Context& c, const cv::Mat& image, int preset, int slot = 0) {
    vo_params p;
    vo_params_preset(&p, preset);
    if (image.rows < 1 || image.cols < 1 || image.data == nullptr) throw std::runtime_error("empty image");
    const int ch = image.channels();
    if (ch == 3) check(vo_extract_bgr(c.get(), image.data, image.rows, image.cols, image.step, &p, slot), c.get(), "vo_extract_bgr");
    else if (ch == 1) check(vo_extract(c.get(), image.data, image.rows, image.cols, image.step, &p, slot), c.get(), "vo_extract");
    else throw std::runtime_error("image must be 8-bit gray or BGR");
    return fetch(c, slot);
}

inline std::vector<std::pair<int, int>> match(const Context& c, const Descriptors& d1, const Descriptors& d2, float ratio,
                                              int match_bits = VO_MATCH_REF32) {
    if (d1.empty() || d2.empty()) return {};  // feature_matching_parallel.cpp:57
    auto flatten = [](const Descriptors& d) {
        std::vector<uint8_t> f(512 * d.size());
        for (size_t k = 0; k < d.size(); ++k) {
            if (d[k].size() != 512) throw std::runtime_error("descriptor rows must hold 512 bytes");
            std::copy(d[k].begin(), d[k].end(), f.begin() + 512 * k);
        }
        return f;
    };
    const std::vector<uint8_t> f1 = flatten(d1), f2 = flatten(d2);
    std::vector<int32_t> pairs(2 * d1.size());
    int n = 0;
    check(vo_match_host(c.get(), f1.data(), (int)d1.size(), f2.data(), (int)d2.size(), 512, ratio, match_bits, pairs.data(),
                        (int)d1.size(), &n), c.get(), "vo_match_host");
    std::vector<std::pair<int, int>> out(n);
    for (int k = 0; k < n; ++k) out[k] = {pairs[2 * k], pairs[2 * k + 1]};
    return out;
}

// ---- stage entry points, whole-image form of the reference's public statics ------------------------
// cv::GaussianBlur(src, dst, Size(7,7), 0)  (feature_extraction_parallel.cpp:223)
inline cv::Mat gaussian_blur7(const Context& c, const cv::Mat& gray) {
    if (gray.channels() != 1 || gray.step != (size_t)gray.cols) throw std::runtime_error("gaussian_blur7: packed 8-bit gray expected");
    cv::Mat out(gray.rows, gray.cols, CV_8UC1);
    check(vo_blur(c.get(), gray.data, gray.rows, gray.cols, out.data), c.get(), "vo_blur");
    return out;
}
// direction_gradients_worker + shitomasi_corner_detection_worker over the whole image
// (corner_detection_parallel.cpp:30-105): blurred u8 -> thresholded R_array
inline std::vector<std::vector<float>> shitomasi_response(const Context& c, const cv::Mat& blurred, float threshold) {
    if (blurred.channels() != 1 || blurred.step != (size_t)blurred.cols) throw std::runtime_error("shitomasi_response: packed 8-bit gray expected");
    std::vector<float> flat((size_t)blurred.rows * blurred.cols);
    check(vo_response(c.get(), blurred.data, blurred.rows, blurred.cols, threshold, flat.data()), c.get(), "vo_response");
    std::vector<std::vector<float>> R(blurred.rows);
    for (int i = 0; i < blurred.rows; ++i) R[i].assign(flat.begin() + (size_t)i * blurred.cols, flat.begin() + (size_t)(i + 1) * blurred.cols);
    return R;
}
// CornerDetectionParallel::non_maximum_suppression (corner_detection_parallel.cpp:166-208) followed by the
// manager's sort by (y, x) (feature_extraction_parallel.cpp:276-282): the same keypoint SET as the reference's
// call, already in the order the manager gives them to FREAK.
inline std::vector<cv::KeyPoint> non_maximum_suppression(const Context& c, const std::vector<std::vector<float>>& R_values,
                                                         const int& n_rows, const int& n_cols, const int& k, const int& N) {
    std::vector<float> flat((size_t)n_rows * n_cols);
    for (int i = 0; i < n_rows; ++i) std::copy(R_values[i].begin(), R_values[i].begin() + n_cols, flat.begin() + (size_t)i * n_cols);
    std::vector<float> xy(2 * (size_t)(N > 0 ? N : 1));
    int n = 0;
    check(vo_nms(c.get(), flat.data(), n_rows, n_cols, k, N, xy.data(), &n), c.get(), "vo_nms");
    std::vector<cv::KeyPoint> out;
    out.reserve(n);
    for (int q = 0; q < n; ++q) out.emplace_back(cv::Point2f(xy[2 * q], xy[2 * q + 1]), 1.0f);
    return out;
}
// FREAK_Parallel::FREAK_feature_description_worker over all keypoints (FREAK_feature_descriptor_parallel.cpp:110-151)
inline Descriptors FREAK_feature_description(const Context& c, const std::vector<cv::KeyPoint>& key_points, const cv::Mat& blurred,
                                             std::vector<float>* angles = nullptr) {
    const int n = (int)key_points.size();
    Descriptors out;
    if (n == 0) return out;
    std::vector<float> xy(2 * (size_t)n), ang(n);
    for (int q = 0; q < n; ++q) { xy[2 * q] = key_points[q].pt.x; xy[2 * q + 1] = key_points[q].pt.y; }
    std::vector<uint8_t> flat(512 * (size_t)n);
    check(vo_describe(c.get(), blurred.data, blurred.rows, blurred.cols, xy.data(), n, flat.data(), ang.data()), c.get(), "vo_describe");
    out.reserve(n);
    for (int q = 0; q < n; ++q) out.emplace_back(flat.begin() + 512 * (size_t)q, flat.begin() + 512 * (size_t)(q + 1));
    if (angles) *angles = ang;
    return out;
}

}  // namespace vofront

// ---- the reference's names ------------------------------------------------------------------------
#ifndef VOFRONT_NO_REFERENCE_NAMES

// Stands where VisualOdometry.h:19 holds `GPU_settings gpu_settings;`
struct GPU_settings : vofront::Context {
    using vofront::Context::Context;
};

inline std::pair<std::vector<std::vector<uint8_t>>, std::vector<cv::KeyPoint>> feature_extraction_manager_with_points(
    const cv::Mat& image, const GPU_settings& settings) {
    return vofront::extract(settings, image, VO_PRESET_VO_GPU_REF);
}

inline std::pair<std::vector<std::vector<uint8_t>>, std::vector<cv::KeyPoint>> feature_extraction_manager_with_points(
    cv::Mat& image, thread_pool& /*pool: replaced by CUDA streams*/) {
    return vofront::extract(vofront::default_context(image.rows, image.cols, 2000), image, VO_PRESET_CPU_REF);
}

inline std::vector<std::pair<int, int>> matchCustomBinaryDescriptorsThreadPool(
    const std::vector<std::vector<uint8_t>>& desc1, const std::vector<std::vector<uint8_t>>& desc2, thread_pool& /*pool*/,
    int /*numThreads*/, float ratioThreshold = 0.75f) {
    if (desc1.empty() || desc2.empty()) return {};
    const int n = (int)(desc1.size() > desc2.size() ? desc1.size() : desc2.size());
    return vofront::match(vofront::default_context(64, 64, n > 2000 ? n : 2000), desc1, desc2, ratioThreshold);
}
// Same, on an explicit context (what a VisualOdometry built around GPU_settings calls).
inline std::vector<std::pair<int, int>> matchCustomBinaryDescriptorsThreadPool(
    const std::vector<std::vector<uint8_t>>& desc1, const std::vector<std::vector<uint8_t>>& desc2,
    const GPU_settings& settings, float ratioThreshold = 0.75f) {
    return vofront::match(settings, desc1, desc2, ratioThreshold);
}

#endif  // VOFRONT_NO_REFERENCE_NAMES
