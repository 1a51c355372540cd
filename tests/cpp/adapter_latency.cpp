// adapter_latency.cpp — MEASUREMENT PROGRAM: the reference's own frame loop (VisualOdometry.cpp:64-100, 163-165) written
// against include/vofront_adapter.hpp exactly as a maintainer would leave it after swapping the header: one synchronous
// extraction and one match per frame, descriptors travelling as vector<vector<uint8_t>>.  Prints one JSON line.
//
//   adapter_latency <frames.raw> [repeats]      frames.raw: int32 n, rows, cols, then n gray u8 frames
#include <chrono>
#include <cstdio>
#include <fstream>
#include <iostream>

#include <opencv2/core.hpp>

#include "vofront_adapter.hpp"

int main(int argc, char** argv) {
    if (argc < 2) { std::cerr << "usage: adapter_latency frames.raw [repeats]\n"; return 2; }
    const int repeats = argc > 2 ? std::atoi(argv[2]) : 2;
    std::ifstream in(argv[1], std::ios::binary);
    int32_t n = 0, rows = 0, cols = 0;
    in.read(reinterpret_cast<char*>(&n), 4); in.read(reinterpret_cast<char*>(&rows), 4); in.read(reinterpret_cast<char*>(&cols), 4);
    std::vector<cv::Mat> frames;
    for (int i = 0; i < n; ++i) {
        frames.emplace_back(rows, cols, CV_8UC1);
        in.read(reinterpret_cast<char*>(frames.back().data), (std::streamsize)rows * cols);
    }
    if (!in || n < 2) { std::cerr << "short input\n"; return 2; }
    try {
        GPU_settings gpu_settings(0, rows, cols, 2000);
        double best = 1e30, best_extract = 0, best_match = 0;
        long long matches = 0, keypoints = 0;
        for (int rep = 0; rep < repeats; ++rep) {
            matches = keypoints = 0;
            const auto t0 = std::chrono::steady_clock::now();
            auto [desc1, kpts1] = feature_extraction_manager_with_points(frames[0], gpu_settings);
            double te = 0, tm = 0;
            for (int i = 1; i < n; ++i) {
                const auto a0 = std::chrono::steady_clock::now();
                auto [desc2, kpts2] = feature_extraction_manager_with_points(frames[i], gpu_settings);
                const auto a1 = std::chrono::steady_clock::now();
                const auto m = matchCustomBinaryDescriptorsThreadPool(desc1, desc2, gpu_settings, 0.75f);
                const auto a2 = std::chrono::steady_clock::now();
                te += std::chrono::duration<double>(a1 - a0).count();
                tm += std::chrono::duration<double>(a2 - a1).count();
                matches += (long long)m.size();
                keypoints += (long long)kpts2.size();
                desc1 = std::move(desc2);   // VisualOdometry.cpp:164-165
                kpts1 = std::move(kpts2);
            }
            const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            if (s < best) { best = s; best_extract = te; best_match = tm; }
        }
        std::printf("{\"workload\": \"%d KITTI-shaped %dx%d frames through the reference's own loop on vofront_adapter.hpp (one synchronous "
                    "extraction + one match per frame, vector<vector<uint8_t>> descriptors)\", \"frames_per_s\": %.1f, \"ms_per_frame\": %.4f, "
                    "\"extract_ms_per_frame\": %.4f, \"match_ms_per_pair\": %.4f, \"keypoints_per_frame\": %.1f, \"matches_per_pair\": %.1f}\n",
                    n, cols, rows, n / best, 1e3 * best / n, 1e3 * best_extract / (n - 1), 1e3 * best_match / (n - 1),
                    (double)keypoints / (n - 1), (double)matches / (n - 1));
    } catch (const std::exception& e) {
        std::cerr << "adapter_latency: " << e.what() << "\n";
        return 1;
    }
    return 0;
}
