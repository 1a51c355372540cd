// adapter_driver.cpp — TEST PROGRAM: drives include/vofront_adapter.hpp exactly the way the reference's
// callers drive their front-end (VisualOdometry.cpp:29-36, matching.cpp:116-117,148), and dumps the results
// for tests/test_cpp_adapter.py to compare with the oracle.  Compiled against oracle/shim's cv:: stand-in
// (no C++ OpenCV in the image); with real OpenCV the same source compiles unchanged.
//
//   adapter_driver <in.raw> <out.bin>
// in.raw : int32 rows, cols, then two gray u8 images of rows*cols bytes.
// out.bin: for each of the sections below  int32 count  followed by the payload.
#include <cstring>
#include <thread>
#include <atomic>
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>

#include <opencv2/core.hpp>

#include "vofront_adapter.hpp"

class thread_pool {};  // the adapter only takes it by reference (the reference's pool: threadpool.h:15)

static void put_i32(std::ofstream& o, int32_t v) { o.write(reinterpret_cast<const char*>(&v), 4); }

static void dump_features(std::ofstream& o, const std::pair<std::vector<std::vector<uint8_t>>, std::vector<cv::KeyPoint>>& f) {
    put_i32(o, (int32_t)f.second.size());
    for (const auto& k : f.second) {
        o.write(reinterpret_cast<const char*>(&k.pt.x), 4);
        o.write(reinterpret_cast<const char*>(&k.pt.y), 4);
    }
    for (const auto& d : f.first) o.write(reinterpret_cast<const char*>(d.data()), 512);
}
static void dump_matches(std::ofstream& o, const std::vector<std::pair<int, int>>& m) {
    put_i32(o, (int32_t)m.size());
    for (const auto& p : m) { put_i32(o, p.first); put_i32(o, p.second); }
}

int main(int argc, char** argv) {
    if (argc != 3) { std::cerr << "usage: adapter_driver in.raw out.bin\n"; return 2; }
    std::ifstream in(argv[1], std::ios::binary);
    int32_t rows = 0, cols = 0;
    in.read(reinterpret_cast<char*>(&rows), 4);
    in.read(reinterpret_cast<char*>(&cols), 4);
    cv::Mat img1(rows, cols, CV_8UC1), img2(rows, cols, CV_8UC1);
    in.read(reinterpret_cast<char*>(img1.data), (std::streamsize)rows * cols);
    in.read(reinterpret_cast<char*>(img2.data), (std::streamsize)rows * cols);
    if (!in) { std::cerr << "short input\n"; return 2; }
    std::ofstream out(argv[2], std::ios::binary);
    try {
        // --- 1. the VisualOdometry path: GPU_settings built once, gray frames, ratio 0.75 (VisualOdometry.cpp:9-36)
        GPU_settings gpu_settings(0, rows, cols, 2000);
        auto f1 = feature_extraction_manager_with_points(img1, gpu_settings);
        auto f2 = feature_extraction_manager_with_points(img2, gpu_settings);
        auto m = matchCustomBinaryDescriptorsThreadPool(f1.first, f2.first, gpu_settings, 0.75f);
        dump_features(out, f1);
        dump_features(out, f2);
        dump_matches(out, m);
        // --- 2. the thread-pool signatures (matching.cpp:116-117,148): BGR input, pool ignored
        thread_pool pool1;
        cv::Mat bgr(rows, cols, CV_8UC3);
        for (int i = 0; i < rows; ++i)
            for (int j = 0; j < cols; ++j)
                for (int c = 0; c < 3; ++c) bgr.data[(size_t)i * bgr.step + 3 * j + c] = img1.data[(size_t)i * img1.step + j];
        auto g1 = feature_extraction_manager_with_points(bgr, pool1);
        auto g2 = feature_extraction_manager_with_points(img2, pool1);
        auto gm = matchCustomBinaryDescriptorsThreadPool(g1.first, g2.first, pool1, 8, 0.75f);
        dump_features(out, g1);
        dump_features(out, g2);
        dump_matches(out, gm);
        // --- 3. stage entry points
        cv::Mat blurred = vofront::gaussian_blur7(gpu_settings, img1);
        auto R = vofront::shitomasi_response(gpu_settings, blurred, 25000.0f);
        auto kps = vofront::non_maximum_suppression(gpu_settings, R, rows, cols, 5, 1500);
        auto desc = vofront::FREAK_feature_description(gpu_settings, kps, blurred);
        dump_features(out, {desc, kps});
        // --- 3b. the reference's stage classes with their static signatures, chained the way the manager chains them
        // (feature_extraction_parallel.cpp:223-300), in TWO row intervals to exercise the interval argument
        {
            cv::Mat gray = CornerDetectionParallel::custom_bgr2gray(bgr);
            cv::Mat bl = vofront::gaussian_blur7(gpu_settings, gray);
            cv::Mat Jx = cv::Mat::zeros(rows, cols, CV_32F), Jy = cv::Mat::zeros(rows, cols, CV_32F), Jxy = cv::Mat::zeros(rows, cols, CV_32F);
            const interval whole{{0, cols}, {0, rows}};
            CornerDetectionParallel::direction_gradients_worker(bl, whole, Jx, Jy, Jxy);
            std::vector<std::vector<float>> Rw(rows, std::vector<float>(cols, 0.0f));
            const interval top{{0, cols}, {0, rows / 2 + 2}}, bottom{{0, cols}, {rows / 2 - 2, rows}};  // overlapping by the 2-row halo
            CornerDetectionParallel::shitomasi_corner_detection_worker(Jx, Jy, Jxy, top, 0.05f, Rw);
            CornerDetectionParallel::shitomasi_corner_detection_worker(Jx, Jy, Jxy, bottom, 0.05f, Rw);
            auto kp2 = CornerDetectionParallel::non_maximum_suppression(Rw, rows, cols, 5, 1500);
            // the facade's own order (the reference's heap order, strongest first), then the manager's sort
            // (feature_extraction_parallel.cpp:276-282) before the descriptors
            put_i32(out, (int32_t)kp2.size());
            for (const auto& kq : kp2) { out.write(reinterpret_cast<const char*>(&kq.pt.x), 4); out.write(reinterpret_cast<const char*>(&kq.pt.y), 4); }
            std::sort(kp2.begin(), kp2.end(), [](const cv::KeyPoint& a, const cv::KeyPoint& b) {
                const int ay = static_cast<int>(a.pt.y), by = static_cast<int>(b.pt.y);
                return ay == by ? static_cast<int>(a.pt.x) < static_cast<int>(b.pt.x) : ay < by;
            });
            std::vector<std::vector<uint8_t>> d2(kp2.size(), std::vector<uint8_t>(512));
            for (size_t st = 0; st < kp2.size(); st += KEY_POINTS_PER_TASK)
                FREAK_Parallel::FREAK_feature_description_worker(kp2, bl, st, d2, kp2.size());
            dump_features(out, {d2, kp2});
            // gradients of a sub-interval only: everything outside stays zero
            cv::Mat Gx = cv::Mat::zeros(rows, cols, CV_32F), Gy = cv::Mat::zeros(rows, cols, CV_32F), Gxy = cv::Mat::zeros(rows, cols, CV_32F);
            const interval sub{{10, cols - 20}, {5, 40}};
            CornerDetectionParallel::direction_gradients_worker(bl, sub, Gx, Gy, Gxy);
            put_i32(out, rows * cols);
            out.write(reinterpret_cast<const char*>(Gx.data), (std::streamsize)rows * cols * 4);
            out.write(reinterpret_cast<const char*>(Gy.data), (std::streamsize)rows * cols * 4);
            out.write(reinterpret_cast<const char*>(Gxy.data), (std::streamsize)rows * cols * 4);
            const float ang = kp2.empty() ? 0.0f : FREAK_Parallel::compute_orientation(kp2[0], bl);
            out.write(reinterpret_cast<const char*>(&ang), 4);
            // descriptor-only managers
            auto donly = feature_extraction_manager(img1, gpu_settings);
            put_i32(out, (int32_t)donly.size());
            put_i32(out, donly == f1.first ? 1 : 0);
            auto donly2 = feature_extraction_manager(bgr, pool1);
            put_i32(out, donly2 == g1.first ? 1 : 0);
            // a set the adapter did not produce (a copy with one bit flipped) must take the upload path and still be right
            auto edited = f1.first;
            edited[0][0] ^= 1;
            dump_matches(out, matchCustomBinaryDescriptorsThreadPool(edited, f2.first, gpu_settings, 0.75f));
        }
        // --- 4. error behaviour: empty input => empty result (feature_matching_parallel.cpp:57); bad input => std::runtime_error
        std::vector<std::vector<uint8_t>> none;
        put_i32(out, (int32_t)matchCustomBinaryDescriptorsThreadPool(none, f2.first, pool1, 8, 0.75f).size());
        int threw = 0;
        try {
            cv::Mat big(rows + 1, cols, CV_8UC1);
            feature_extraction_manager_with_points(big, gpu_settings);
        } catch (const std::runtime_error&) { threw = 1; }
        put_i32(out, threw);
        // --- 5. two threads through the context-less signatures with DIFFERENT image sizes: thread B's larger image makes
        // the process-wide context grow while thread A is calling in.  Every result must equal the single-threaded one.
        {
            cv::Mat tall(rows + 40, cols, CV_8UC1);  // img2 with its last rows repeated: forces a regrowth
            for (int i = 0; i < tall.rows; ++i) std::memcpy(tall.data + (size_t)i * tall.step, img2.data + (size_t)std::min(i, rows - 1) * img2.step, cols);
            const auto want_a = feature_extraction_manager_with_points(bgr, pool1);
            const auto want_b = feature_extraction_manager_with_points(tall, pool1);
            const auto want_ma = matchCustomBinaryDescriptorsThreadPool(want_a.first, want_a.first, pool1, 4, 0.75f);
            const auto want_mb = matchCustomBinaryDescriptorsThreadPool(want_b.first, want_a.first, pool1, 4, 0.75f);
            std::atomic<int> bad{0};
            auto worker = [&](bool b) {
                try {
                    thread_pool my_pool;
                    for (int it = 0; it < 6; ++it) {
                        const auto got = b ? feature_extraction_manager_with_points(tall, my_pool) : feature_extraction_manager_with_points(bgr, my_pool);
                        const auto& want = b ? want_b : want_a;
                        if (got.first != want.first || got.second.size() != want.second.size()) ++bad;
                        if (matchCustomBinaryDescriptorsThreadPool(got.first, want_a.first, my_pool, 4, 0.75f) != (b ? want_mb : want_ma)) ++bad;
                    }
                } catch (const std::exception&) { bad += 100; }
            };
            std::thread ta(worker, false), tb(worker, true);
            ta.join(); tb.join();
            put_i32(out, (int32_t)want_b.second.size());
            put_i32(out, bad.load());
        }
    } catch (const std::exception& e) {
        std::cerr << "adapter_driver: " << e.what() << "\n";
        return 1;
    }
    return 0;
}
