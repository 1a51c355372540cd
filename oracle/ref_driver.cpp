// ref_driver.cpp — TEST INFRASTRUCTURE ONLY (part of oracle/; never linked into the product).
//
// C entry points around the REFERENCE's own hot-path code, compiled unmodified and in place
// from /root/reference against oracle/shim (see oracle/Makefile) into oracle/_ref/libvoref.so.
// Nothing from the reference is copied into this repository: this file only *calls*
//   feature_extraction_manager_with_points      feature_extraction_parallel/feature_extraction_parallel.cpp:215
//   CornerDetectionParallel::{custom_bgr2gray, direction_gradients_worker,
//       shitomasi_corner_detection_worker, non_maximum_suppression}
//                                               feature_extraction_parallel/corner_detection_parallel.cpp:14,30,61,166
//   FREAK_Parallel::{compute_orientation, FREAK_feature_description_worker}
//                                               feature_extraction_parallel/FREAK_feature_descriptor_parallel.cpp:11,110
//   matchCustomBinaryDescriptorsThreadPool      feature_matching_parallel/feature_matching_parallel.cpp:49
// Uses: (1) pinning the C restatement in oracle/vo_oracle.c, (2) generating tests/golden/*,
// (3) the timed CPU baseline ("kind": "reference") in bench.py.
#include <chrono>
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#include "feature_extraction_parallel/feature_extraction_parallel.h"
#include "feature_matching_parallel/feature_matching_parallel.h"

namespace {

std::unique_ptr<thread_pool> g_pool;
int g_pool_threads = 0;

thread_pool& pool_for(int threads) {
    if (!g_pool || g_pool_threads != threads) {
        g_pool.reset();
        g_pool = std::make_unique<thread_pool>(static_cast<unsigned>(threads));
        g_pool_threads = threads;
    }
    return *g_pool;
}

cv::Mat gray_to_bgr(const uint8_t* gray, int rows, int cols) {
    // cv::imread(path) (default flags) on a grayscale PNG yields three equal channels; the
    // reference's pair driver feeds that to the manager (matching.cpp:110-117).
    cv::Mat bgr(rows, cols, CV_8UC3);
    for (int i = 0; i < rows; ++i) {
        uchar* d = bgr.ptr<uchar>(i);
        const uint8_t* s = gray + static_cast<size_t>(i) * cols;
        for (int j = 0; j < cols; ++j) d[3 * j] = d[3 * j + 1] = d[3 * j + 2] = s[j];
    }
    return bgr;
}

int export_features(const std::pair<std::vector<std::vector<uint8_t>>, std::vector<cv::KeyPoint>>& out, int cap,
                    float* kp_xy, uint8_t* desc512) {
    const int n = static_cast<int>(out.second.size());
    if (n > cap) return -n;
    for (int i = 0; i < n; ++i) {
        if (kp_xy) { kp_xy[2 * i] = out.second[i].pt.x; kp_xy[2 * i + 1] = out.second[i].pt.y; }
        if (desc512) std::memcpy(desc512 + static_cast<size_t>(i) * 512, out.first[i].data(), 512);
    }
    return n;
}

std::vector<std::vector<uint8_t>> rows_of(const uint8_t* d, int n) {
    std::vector<std::vector<uint8_t>> v(n);
    for (int i = 0; i < n; ++i) v[i].assign(d + static_cast<size_t>(i) * 512, d + static_cast<size_t>(i + 1) * 512);
    return v;
}

double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

}  // namespace

extern "C" {

// Whole reference extract on a BGR image.  block_size >= rows gives the deterministic
// single-strip semantics used for parity (the default BLOCK_SIZE=100 run has a strip race).
int voref_extract_bgr(const uint8_t* bgr, int rows, int cols, int threads, int block_size, int cap, float* kp_xy,
                      uint8_t* desc512) {
    NUMBER_OF_THREADS = static_cast<size_t>(threads);
    BLOCK_SIZE = block_size;
    cv::Mat img(rows, cols, CV_8UC3);
    std::memcpy(img.data, bgr, static_cast<size_t>(rows) * cols * 3);
    auto out = feature_extraction_manager_with_points(img, pool_for(threads));
    return export_features(out, cap, kp_xy, desc512);
}

int voref_extract(const uint8_t* gray, int rows, int cols, int threads, int block_size, int cap, float* kp_xy,
                  uint8_t* desc512) {
    NUMBER_OF_THREADS = static_cast<size_t>(threads);
    BLOCK_SIZE = block_size;
    cv::Mat img = gray_to_bgr(gray, rows, cols);
    auto out = feature_extraction_manager_with_points(img, pool_for(threads));
    return export_features(out, cap, kp_xy, desc512);
}

int voref_match(const uint8_t* d1, int n1, const uint8_t* d2, int n2, int threads, float ratio, int32_t* pairs,
                int cap) {
    const auto a = rows_of(d1, n1), b = rows_of(d2, n2);
    const auto m = matchCustomBinaryDescriptorsThreadPool(a, b, pool_for(threads), threads, ratio);
    const int n = static_cast<int>(m.size());
    if (n > cap) return -n;
    for (int i = 0; i < n; ++i) { pairs[2 * i] = m[i].first; pairs[2 * i + 1] = m[i].second; }
    return n;
}

// ---- stage probes (single strip = whole image) --------------------------------------------------
void voref_bgr2gray(const uint8_t* bgr, int rows, int cols, uint8_t* gray) {
    cv::Mat img(rows, cols, CV_8UC3);
    std::memcpy(img.data, bgr, static_cast<size_t>(rows) * cols * 3);
    const cv::Mat g = CornerDetectionParallel::custom_bgr2gray(img);
    std::memcpy(gray, g.data, static_cast<size_t>(rows) * cols);
}

void voref_blur(const uint8_t* gray, int rows, int cols, uint8_t* out) {
    cv::Mat src(rows, cols, CV_8UC1, const_cast<uint8_t*>(gray)), dst;
    cv::GaussianBlur(src, dst, cv::Size(7, 7), 0);  // the call at feature_extraction_parallel.cpp:223
    std::memcpy(out, dst.data, static_cast<size_t>(rows) * cols);
}

void voref_gradients(const uint8_t* blurred, int rows, int cols, float* jx, float* jy, float* jxy) {
    cv::Mat src(rows, cols, CV_8UC1, const_cast<uint8_t*>(blurred));
    cv::Mat Jx = cv::Mat::zeros(rows, cols, CV_32F), Jy = cv::Mat::zeros(rows, cols, CV_32F),
            Jxy = cv::Mat::zeros(rows, cols, CV_32F);
    const interval whole = {{0, cols}, {0, rows}};
    CornerDetectionParallel::direction_gradients_worker(src, whole, Jx, Jy, Jxy);
    const size_t bytes = static_cast<size_t>(rows) * cols * sizeof(float);
    std::memcpy(jx, Jx.data, bytes); std::memcpy(jy, Jy.data, bytes); std::memcpy(jxy, Jxy.data, bytes);
}

void voref_response(const uint8_t* blurred, int rows, int cols, float* R) {
    cv::Mat src(rows, cols, CV_8UC1, const_cast<uint8_t*>(blurred));
    cv::Mat Jx = cv::Mat::zeros(rows, cols, CV_32F), Jy = cv::Mat::zeros(rows, cols, CV_32F),
            Jxy = cv::Mat::zeros(rows, cols, CV_32F);
    std::vector<std::vector<float>> Rv(rows, std::vector<float>(cols, 0));
    const interval whole = {{0, cols}, {0, rows}};
    CornerDetectionParallel::direction_gradients_worker(src, whole, Jx, Jy, Jxy);
    CornerDetectionParallel::shitomasi_corner_detection_worker(Jx, Jy, Jxy, whole, 0.05f, Rv);
    for (int i = 0; i < rows; ++i) std::memcpy(R + static_cast<size_t>(i) * cols, Rv[i].data(), cols * sizeof(float));
}

// non_maximum_suppression(k, N) followed by the manager's (y, x) sort (feature_extraction_parallel.cpp:274-282).
int voref_nms(const float* R, int rows, int cols, int k, int N, float* kp_xy) {
    std::vector<std::vector<float>> Rv(rows);
    for (int i = 0; i < rows; ++i) Rv[i].assign(R + static_cast<size_t>(i) * cols, R + static_cast<size_t>(i + 1) * cols);
    auto kps = CornerDetectionParallel::non_maximum_suppression(Rv, rows, cols, k, N);
    std::sort(kps.begin(), kps.end(), [](const cv::KeyPoint& a, const cv::KeyPoint& b) {
        const int ay = static_cast<int>(a.pt.y), by = static_cast<int>(b.pt.y);
        if (ay == by) return static_cast<int>(a.pt.x) < static_cast<int>(b.pt.x);
        return ay < by;
    });
    for (size_t i = 0; i < kps.size(); ++i) { kp_xy[2 * i] = kps[i].pt.x; kp_xy[2 * i + 1] = kps[i].pt.y; }
    return static_cast<int>(kps.size());
}

// non_maximum_suppression(k, N) as the reference returns it: heap order, strongest first (corner_detection_parallel.cpp:198-206).
int voref_nms_ranked(const float* R, int rows, int cols, int k, int N, float* kp_xy) {
    std::vector<std::vector<float>> Rv(rows);
    for (int i = 0; i < rows; ++i) Rv[i].assign(R + static_cast<size_t>(i) * cols, R + static_cast<size_t>(i + 1) * cols);
    auto kps = CornerDetectionParallel::non_maximum_suppression(Rv, rows, cols, k, N);
    for (size_t i = 0; i < kps.size(); ++i) { kp_xy[2 * i] = kps[i].pt.x; kp_xy[2 * i + 1] = kps[i].pt.y; }
    return static_cast<int>(kps.size());
}

float voref_orientation(const uint8_t* blurred, int rows, int cols, float x, float y) {
    cv::Mat src(rows, cols, CV_8UC1, const_cast<uint8_t*>(blurred));
    return FREAK_Parallel::compute_orientation(cv::KeyPoint(cv::Point2f(x, y), 1.0f), src);
}

void voref_describe(const uint8_t* blurred, int rows, int cols, const float* kp_xy, int n, uint8_t* desc512) {
    cv::Mat src(rows, cols, CV_8UC1, const_cast<uint8_t*>(blurred));
    std::vector<cv::KeyPoint> kps;
    for (int i = 0; i < n; ++i) kps.emplace_back(cv::Point2f(kp_xy[2 * i], kp_xy[2 * i + 1]), 1.0f);
    std::vector<std::vector<uint8_t>> desc(n, std::vector<uint8_t>(DESCRIPTOR_SIZE));
    const size_t count = static_cast<size_t>(n);
    FREAK_Parallel::FREAK_feature_description_worker(kps, src, 0, desc, count, count);
    for (int i = 0; i < n; ++i) std::memcpy(desc512 + static_cast<size_t>(i) * 512, desc[i].data(), 512);
}

// ---- timed baseline: the reference's CPU thread-pool path over a frame sequence ---------------
// gray_frames: n_frames tightly packed rows*cols u8 images.  Extract every frame through
// feature_extraction_manager_with_points (BGR input, as matching.cpp feeds it; the gray->BGR
// replication is NOT timed), match consecutive frames through matchCustomBinaryDescriptorsThreadPool.
// Warm-up: repeat frame 0/1 until warmup_s seconds have passed (idle vCPUs wake slowly).
int voref_bench_sequence(const uint8_t* gray_frames, int n_frames, int rows, int cols, int threads, int block_size,
                         double warmup_s, double* extract_s, double* match_s, int64_t* total_kp,
                         int64_t* total_matches) {
    NUMBER_OF_THREADS = static_cast<size_t>(threads);
    BLOCK_SIZE = block_size;
    thread_pool& pool = pool_for(threads);
    const size_t fsz = static_cast<size_t>(rows) * cols;
    if (n_frames < 2) return -1;
    {
        cv::Mat a = gray_to_bgr(gray_frames, rows, cols), b = gray_to_bgr(gray_frames + fsz, rows, cols);
        const double t0 = now_s();
        do {
            auto fa = feature_extraction_manager_with_points(a, pool);
            auto fb = feature_extraction_manager_with_points(b, pool);
            auto m = matchCustomBinaryDescriptorsThreadPool(fa.first, fb.first, pool, threads, 0.75f);
            (void)m;
        } while (now_s() - t0 < warmup_s);
    }
    double te = 0, tm = 0;
    int64_t nk = 0, nm = 0;
    std::vector<std::vector<uint8_t>> prev;
    for (int f = 0; f < n_frames; ++f) {
        cv::Mat img = gray_to_bgr(gray_frames + fsz * f, rows, cols);
        double t0 = now_s();
        auto cur = feature_extraction_manager_with_points(img, pool);
        te += now_s() - t0;
        nk += static_cast<int64_t>(cur.second.size());
        if (f > 0) {
            t0 = now_s();
            auto m = matchCustomBinaryDescriptorsThreadPool(prev, cur.first, pool, threads, 0.75f);
            tm += now_s() - t0;
            nm += static_cast<int64_t>(m.size());
        }
        prev = std::move(cur.first);
    }
    *extract_s = te; *match_s = tm; *total_kp = nk; *total_matches = nm;
    return 0;
}

void voref_shutdown() { g_pool.reset(); g_pool_threads = 0; }

}  // extern "C"
