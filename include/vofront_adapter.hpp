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
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <memory>
#include <mutex>
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
    // What the adapter remembers about the descriptor sets it handed out: the reference's API returns descriptors as
    // vector<vector<uint8_t>> and takes them back in the matcher, but the packed descriptors of the last n_slots
    // extractions are still on the device.  match() packs the caller's rows on the host (it has to read them anyway),
    // recognises a set this context produced by comparing ALL of its packed words with the copy kept here (so a recycled
    // or edited vector can never alias a slot) and then matches slot against slot; any other set is uploaded in packed
    // form, 64 bytes per descriptor instead of 512.
    struct Resident { std::vector<uint32_t> packed; int n = -1; };  // 16 words per descriptor, as on the device
    struct State { std::vector<Resident> slots; int next = 0; };

    std::shared_ptr<vo_ctx> ctx;
    std::shared_ptr<State> state;
    int max_rows = 0, max_cols = 0, max_keypoints = 0, n_slots = 0;

    Context() = default;
    Context(int device, int rows, int cols, int max_kp = 2000, int slots = 4)
        : state(std::make_shared<State>()), max_rows(rows), max_cols(cols), max_keypoints(max_kp), n_slots(slots) {
        vo_ctx* c = nullptr;
        const int rc = vo_create(&c, device, rows, cols, max_kp, slots);
        if (rc != VO_OK) throw std::runtime_error(std::string("vo_create: ") + vo_last_error(nullptr));
        ctx.reset(c, [](vo_ctx* p) { vo_destroy(p); });
        state->slots.resize(slots);
    }
    vo_ctx* get() const { return ctx.get(); }
    bool fits(int rows, int cols, int kp) const { return ctx && rows <= max_rows && cols <= max_cols && kp <= max_keypoints; }
    int take_slot() const { const int s = state->next; state->next = (s + 1) % n_slots; state->slots[s].n = -1; return s; }
    // slot holding exactly these (packed) descriptors, or -1
    int find(const std::vector<uint32_t>& packed, int n) const {
        if (!state) return -1;
        for (int s = 0; s < n_slots; ++s) {
            const Resident& r = state->slots[s];
            if (r.n == n && n > 0 && std::memcmp(r.packed.data(), packed.data(), 64 * (size_t)n) == 0) return s;
        }
        return -1;
    }
};

// Process-wide context for the thread_pool overloads (which carry no GPU handle); grown on demand.  Callers get a COPY
// (the handles are shared_ptrs): growing the context for another thread's larger image replaces the process-wide one
// but cannot destroy the context a thread is still working with - that one dies with its last copy.  The reference's
// managers are not re-entrant either (they share a thread_pool), and neither is a vo_ctx: two threads that call in with
// the SAME geometry share one context and must serialise their calls; this only makes differing geometries safe.
inline Context default_context(int rows, int cols, int max_kp) {
    static std::mutex m;
    static Context c;
    std::lock_guard<std::mutex> lk(m);
    if (!c.fits(rows, cols, max_kp))
        c = Context(0, rows > c.max_rows ? rows : (c.max_rows > 64 ? c.max_rows : 64), cols > c.max_cols ? cols : (c.max_cols > 64 ? c.max_cols : 64),
                    max_kp > c.max_keypoints ? max_kp : (c.max_keypoints > 2000 ? c.max_keypoints : 2000));
    return c;
}

// What the context-less entry points (thread_pool signatures, static stage workers) hold for the length of one call: the
// process-wide context AND a process-wide lock, because a vo_ctx serves one call at a time.  Calls from several threads
// therefore queue up instead of racing - on one GPU context they could not overlap anyway.  Re-entrant on a thread.
struct DefaultCall {
    std::unique_lock<std::recursive_mutex> lock;
    Context c;
    operator const Context&() const { return c; }
    vo_ctx* get() const { return c.get(); }
};
inline DefaultCall default_call(int rows, int cols, int max_kp) {
    static std::recursive_mutex m;
    std::unique_lock<std::recursive_mutex> lk(m);
    return DefaultCall{std::move(lk), default_context(rows, cols, max_kp)};
}

using Descriptors = std::vector<std::vector<uint8_t>>;  // n x 512 bytes, one 0/1 byte per test (the reference layout)

// 512 bytes of 0/1 <-> 16 little-endian words: bit (j & 31) of word (j >> 5) = test j (the device layout).
// Eight 0/1 bytes become one byte with a multiply: byte i of x lands on bit 56 + i of x * 0x0102040810204080.
inline bool pack_rows(const Descriptors& d, std::vector<uint32_t>& packed) {
    packed.resize(16 * d.size());
    uint8_t* out = reinterpret_cast<uint8_t*>(packed.data());
    uint64_t bad = 0;
    for (size_t k = 0; k < d.size(); ++k) {
        if (d[k].size() != 512) throw std::runtime_error("descriptor rows must hold 512 bytes");
        const uint8_t* row = d[k].data();
        for (int w = 0; w < 64; ++w) {
            uint64_t x;
            std::memcpy(&x, row + 8 * w, 8);
            bad |= x;
            out[64 * k + w] = (uint8_t)((x * 0x0102040810204080ull) >> 56);
        }
    }
    return (bad & 0xFEFEFEFEFEFEFEFEull) == 0;  // false: some byte is neither 0 nor 1
}
inline const uint64_t* expand_table() {
    struct Lut {
        uint64_t v[256];
        Lut() {
            for (int b = 0; b < 256; ++b) {
                v[b] = 0;
                for (int i = 0; i < 8; ++i) v[b] |= (uint64_t)((b >> i) & 1) << (8 * i);
            }
        }
    };
    static const Lut lut;  // function-local static: initialised once, thread-safe since C++11
    return lut.v;
}

// One synchronous fetch of keypoints and PACKED descriptors (64 bytes each instead of 512 over PCIe); the reference's
// byte-per-test rows are expanded on the host.
inline std::pair<Descriptors, std::vector<cv::KeyPoint>> fetch(const Context& c, int slot) {
    int n = 0;
    std::vector<float> xy(2 * (size_t)c.max_keypoints);
    Context::Resident& res = c.state->slots[slot];
    res.n = -1;
    res.packed.resize(16 * (size_t)c.max_keypoints);
    check(vo_frame_fetch(c.get(), slot, &n, xy.data(), nullptr, res.packed.data()), c.get(), "vo_frame_fetch");
    res.n = n;
    const uint64_t* lut = expand_table();
    const uint8_t* pk = reinterpret_cast<const uint8_t*>(res.packed.data());
    std::pair<Descriptors, std::vector<cv::KeyPoint>> out;
    out.first.resize(n);
    out.second.reserve(n);
    for (int k = 0; k < n; ++k) {
        std::vector<uint8_t>& row = out.first[k];
        row.resize(512);
        for (int w = 0; w < 64; ++w) std::memcpy(row.data() + 8 * w, &lut[pk[64 * (size_t)k + w]], 8);
        out.second.emplace_back(cv::Point2f(xy[2 * k], xy[2 * k + 1]), 1.0f);  // corner_detection_parallel.cpp:203
    }
    return out;
}

inline std::pair<Descriptors, std::vector<cv::KeyPoint>> extract(const Context& c, const cv::Mat& image, int preset, int slot = -1) {
    vo_params p;
    vo_params_preset(&p, preset);
    if (slot < 0) slot = c.take_slot();  // round robin: the sets of the last n_slots frames stay matchable on the device
    else c.state->slots[slot].n = -1;
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
    std::vector<uint32_t> p1, p2;
    if (!pack_rows(d1, p1) || !pack_rows(d2, p2)) throw std::runtime_error("descriptor bytes must be 0 or 1");
    std::vector<int32_t> pairs(2 * d1.size());
    int n = 0;
    const int s1 = c.find(p1, (int)d1.size()), s2 = s1 >= 0 ? c.find(p2, (int)d2.size()) : -1;
    if (s1 >= 0 && s2 >= 0)  // both sets came from this context and are still resident: nothing to upload
        check(vo_match(c.get(), s1, s2, ratio, match_bits, pairs.data(), (int)d1.size(), &n), c.get(), "vo_match");
    else
        check(vo_match_host(c.get(), reinterpret_cast<const uint8_t*>(p1.data()), (int)d1.size(), reinterpret_cast<const uint8_t*>(p2.data()),
                            (int)d2.size(), 64, ratio, match_bits, pairs.data(), (int)d1.size(), &n), c.get(), "vo_match_host");
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
    return vofront::extract(vofront::default_call(image.rows, image.cols, 2000), image, VO_PRESET_CPU_REF);
}

inline std::vector<std::pair<int, int>> matchCustomBinaryDescriptorsThreadPool(
    const std::vector<std::vector<uint8_t>>& desc1, const std::vector<std::vector<uint8_t>>& desc2, thread_pool& /*pool*/,
    int /*numThreads*/, float ratioThreshold = 0.75f) {
    if (desc1.empty() || desc2.empty()) return {};
    const int n = (int)(desc1.size() > desc2.size() ? desc1.size() : desc2.size());
    // the SAME process-wide context the thread_pool extraction overload uses (grown only if a set is larger than it holds)
    return vofront::match(vofront::default_call(1, 1, n), desc1, desc2, ratioThreshold);
}
// Same, on an explicit context (what a VisualOdometry built around GPU_settings calls).
inline std::vector<std::pair<int, int>> matchCustomBinaryDescriptorsThreadPool(
    const std::vector<std::vector<uint8_t>>& desc1, const std::vector<std::vector<uint8_t>>& desc2,
    const GPU_settings& settings, float ratioThreshold = 0.75f) {
    return vofront::match(settings, desc1, desc2, ratioThreshold);
}

// Descriptor-only managers (feature_extraction_parallel.h:19, feature_extraction_parallel_GPU.h:54)
inline std::vector<std::vector<uint8_t>> feature_extraction_manager(const cv::Mat& image, const GPU_settings& settings) {
    return feature_extraction_manager_with_points(image, settings).first;
}
inline std::vector<std::vector<uint8_t>> feature_extraction_manager(cv::Mat& image, thread_pool& pool) {
    return feature_extraction_manager_with_points(image, pool).first;
}

// ---- the reference's stage classes, same static signatures ------------------------------------------
// feature_extraction_parallel/corner_detection_parallel.h:14-47.  They carry no GPU handle, so they run on the
// process-wide context (vofront::default_context), like the thread_pool overloads above.
inline size_t NUMBER_OF_THREADS = 4;   // kept for source compatibility; unused (CUDA streams replace the pool)
inline int BLOCK_SIZE = 100;           // unused: the whole image is one pass (SURVEY.md section 0.2 on the strip race)
constexpr float RESPONSE_THRESHOLD = 25000;

struct bound2d { int start{}, end{}; };
struct interval { bound2d cols, rows; };

class CornerDetectionParallel {
public:
    // corner_detection_parallel.cpp:14-28
    static cv::Mat custom_bgr2gray(cv::Mat& picture) {
        if (picture.channels() != 3) throw std::runtime_error("custom_bgr2gray: 8-bit BGR expected");
        const vofront::DefaultCall c = vofront::default_call(picture.rows, picture.cols, 2000);
        cv::Mat out(picture.rows, picture.cols, CV_8UC1);
        vofront::check(vo_bgr2gray(c.get(), picture.data, picture.rows, picture.cols, picture.step, out.data), c.get(), "vo_bgr2gray");
        return out;
    }
    // :30-58.  Jx, Jy, Jxy: CV_32F matrices of the picture's size (the manager allocates them with cv::Mat::zeros,
    // feature_extraction_parallel.cpp:229-231); only rows/columns strictly inside the interval are written.
    static void direction_gradients_worker(const cv::Mat& picture, const interval& iv, cv::Mat& Jx, cv::Mat& Jy, cv::Mat& Jxy) {
        const vofront::DefaultCall c = vofront::default_call(picture.rows, picture.cols, 2000);
        for (const cv::Mat* m : {&Jx, &Jy, &Jxy})
            if (m->rows != picture.rows || m->cols != picture.cols || m->step != (size_t)picture.cols * 4)
                throw std::runtime_error("direction_gradients_worker: packed CV_32F matrices of the picture's size expected");
        vofront::check(vo_gradients(c.get(), picture.data, picture.rows, picture.cols, picture.step, iv.rows.start, iv.rows.end, iv.cols.start,
                                    iv.cols.end, reinterpret_cast<float*>(Jx.data), reinterpret_cast<float*>(Jy.data),
                                    reinterpret_cast<float*>(Jxy.data)), c.get(), "vo_gradients");
    }
    // :61-105.  `k` is unused, as in the reference (its body computes the Shi-Tomasi response and thresholds with the
    // RESPONSE_THRESHOLD constant).
    static void shitomasi_corner_detection_worker(const cv::Mat& Jx, const cv::Mat& Jy, const cv::Mat& Jxy, const interval& iv, const float& /*k*/,
                                                  std::vector<std::vector<float>>& R_array) {
        const int rows = Jx.rows, cols = Jx.cols;
        const vofront::DefaultCall c = vofront::default_call(rows, cols, 2000);
        vo_params p;
        vo_params_preset(&p, VO_PRESET_CPU_REF);
        p.threshold = RESPONSE_THRESHOLD;
        std::vector<float> flat((size_t)rows * cols, 0.0f);
        vofront::check(vo_response_from_gradients(c.get(), reinterpret_cast<const float*>(Jx.data), reinterpret_cast<const float*>(Jy.data),
                                                  reinterpret_cast<const float*>(Jxy.data), rows, cols, iv.rows.start, iv.rows.end,
                                                  iv.cols.start, iv.cols.end, &p, flat.data()), c.get(), "vo_response_from_gradients");
        for (int i = iv.rows.start + 2; i < iv.rows.end - 2; ++i)
            for (int j = iv.cols.start + 2; j < iv.cols.end - 2; ++j) R_array[i][j] = flat[(size_t)i * cols + j];
    }
    // :166-208, keypoints in the reference's own order: it pops a max-heap of (R, i, j) tuples (:167,198-206), i.e.
    // descending response, ties by descending row, then descending column - a caller that takes "the K strongest" from
    // the front gets the reference's points.  The GPU returns the same SET sorted by (y, x) (what the manager sorts to
    // next, feature_extraction_parallel.cpp:276-282); the heap order is restored here from the caller's own R_values.
    static std::vector<cv::KeyPoint> non_maximum_suppression(const std::vector<std::vector<float>>& R_values, const int& n_rows,
                                                             const int& n_cols, const int& k, const int& N) {
        std::vector<cv::KeyPoint> kps =
            vofront::non_maximum_suppression(vofront::default_call(n_rows, n_cols, N > 2000 ? N : 2000), R_values, n_rows, n_cols, k, N);
        std::sort(kps.begin(), kps.end(), [&](const cv::KeyPoint& a, const cv::KeyPoint& b) {
            const int ai = (int)a.pt.y, aj = (int)a.pt.x, bi = (int)b.pt.y, bj = (int)b.pt.x;
            const float ra = R_values[ai][aj], rb = R_values[bi][bj];
            if (ra != rb) return ra > rb;
            if (ai != bi) return ai > bi;
            return aj > bj;
        });
        return kps;
    }
};

// feature_extraction_parallel/FREAK_feature_descriptor_parallel.h:104-110
constexpr size_t KEY_POINTS_PER_TASK = 30;
class FREAK_Parallel {
public:
    // FREAK_feature_descriptor_parallel.cpp:11-55
    static float compute_orientation(const cv::KeyPoint& point, const cv::Mat& image) {
        std::vector<float> ang;
        vofront::FREAK_feature_description(vofront::default_call(image.rows, image.cols, 2000), {point}, image, &ang);
        return ang.at(0);
    }
    // declared at .h:106 (its definition is commented out in the reference, .cpp:56-100: every keypoint, the start index unused)
    static void FREAK_feature_description(const std::vector<cv::KeyPoint>& key_points, const cv::Mat blurred_gray_picture,
                                          const size_t& /*starting_key_point_index*/, std::vector<std::vector<uint8_t>>& descriptor) {
        FREAK_feature_description_worker(key_points, blurred_gray_picture, 0, descriptor, key_points.size(), key_points.size());
    }
    // .cpp:110-151: keypoints [start, min(num_of_keypoints, start + KEYPOINTS_PER_TASK)) into descriptor[i][0..511]
    static void FREAK_feature_description_worker(const std::vector<cv::KeyPoint>& key_points, const cv::Mat& blurred_gray_picture,
                                                 const size_t& starting_key_point_index, std::vector<std::vector<uint8_t>>& descriptor,
                                                 const size_t& num_of_keypoints, const size_t& KEYPOINTS_PER_TASK = KEY_POINTS_PER_TASK) {
        const size_t hi = num_of_keypoints < starting_key_point_index + KEYPOINTS_PER_TASK ? num_of_keypoints : starting_key_point_index + KEYPOINTS_PER_TASK;
        if (hi <= starting_key_point_index) return;
        const std::vector<cv::KeyPoint> part(key_points.begin() + starting_key_point_index, key_points.begin() + hi);
        const int n = (int)part.size();
        const auto d = vofront::FREAK_feature_description(vofront::default_call(blurred_gray_picture.rows, blurred_gray_picture.cols,
                                                                                   n > 2000 ? n : 2000), part, blurred_gray_picture);
        for (size_t q = 0; q < d.size(); ++q) descriptor[starting_key_point_index + q] = d[q];
    }
};

#endif  // VOFRONT_NO_REFERENCE_NAMES
