// vo_image_io.hpp — grayscale frame readers for the pipeline driver (stands in for
// cv::imread(path, cv::IMREAD_GRAYSCALE), VisualOdometry.cpp:65,76; OpenCV is not available here).
// PNG: 8-bit grayscale (KITTI odometry frames), non-interlaced, decoded with zlib.  8-bit RGB / RGBA / gray+alpha
// are reduced the way cv::imread does it (libpng: (R*9797 + G*19234 + B*3737) >> 15); 16-bit and
// palette images are rejected.  PGM: binary P5, maxval <= 255.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace vo {

struct GrayImage {
    int rows = 0, cols = 0;
    std::vector<uint8_t> data;  // rows x cols, tightly packed
    bool empty() const { return data.empty(); }
};

// Returns an empty image on failure (missing file, unsupported format), like cv::imread; `err` says why.
GrayImage readGray(const std::string& path, std::string* err = nullptr);
GrayImage decodePNG(const uint8_t* bytes, size_t n, std::string* err = nullptr);
GrayImage decodePGM(const uint8_t* bytes, size_t n, std::string* err = nullptr);
// Writers used by the tests and tools (8-bit grayscale PNG with stored/deflate blocks, binary PGM).
bool writePNG(const std::string& path, const uint8_t* gray, int rows, int cols);
bool writePGM(const std::string& path, const uint8_t* gray, int rows, int cols);

}  // namespace vo
