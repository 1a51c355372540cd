// vo_main.cpp — the command line of the reference's main_pipeline.cpp (:7-43):
//   vo_main <num_threads> <image_dir> <num_images> <pose_file> <output_csv> [seed] [batch]
// The OpenCL kernel binary the reference looks for (../kernels/feature_extraction_kernel_functions.bin) has no
// counterpart: the CUDA kernels are linked into libvofront.so.
#include <cstdio>
#include <cstdlib>
#include <exception>
#include <string>

#include "vo_pipeline.hpp"

int main(int argc, char* argv[]) {
    if (argc < 6 || argc > 8) {
        std::fprintf(stderr, "Usage: %s <num_threads> <image_dir> <num_images> <pose_file> <output_csv> [seed] [batch]\n", argv[0]);
        return EXIT_FAILURE;
    }
    std::size_t num_threads = 0, num_images = 0;
    try {
        num_threads = static_cast<std::size_t>(std::stoul(argv[1]));
    } catch (const std::exception&) {
        std::fprintf(stderr, "Invalid num_threads: %s\n", argv[1]);
        return EXIT_FAILURE;
    }
    try {
        num_images = static_cast<std::size_t>(std::stoul(argv[3]));
    } catch (const std::exception&) {
        std::fprintf(stderr, "Invalid num_images: %s\n", argv[3]);
        return EXIT_FAILURE;
    }
    try {
        vo::VisualOdometry vo("", num_threads);
        if (argc > 6) vo.seed = std::stoull(argv[6]);
        if (argc > 7) vo.batch = std::stoi(argv[7]);
        vo.run(argv[2], num_images, argv[4], argv[5]);
        const vo::RunStats& s = vo.stats;
        if (!s.frames) return EXIT_FAILURE;
        std::printf("frames %zu  solved pairs %zu  skipped %zu  missing %zu  re-matched %zu\n", s.frames, s.pairs_ok, s.skipped_frames,
                    s.missing_frames, s.respeculated);
        std::printf("wall %.3f s (%.1f frames/s)  front end %.3f s (%.1f frames/s)  RANSAC+pose %.3f s  decode %.3f s  compose %.3f s\n",
                    s.wall_s, s.frames / s.wall_s, s.frontend_s, s.frontend_s > 0 ? s.frames / s.frontend_s : 0.0, s.host_s, s.decode_s,
                    s.compose_s);
    } catch (const std::exception& e) {
        std::fprintf(stderr, "error: %s\n", e.what());
        return EXIT_FAILURE;
    }
    return EXIT_SUCCESS;
}
