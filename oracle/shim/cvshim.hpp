// cvshim.hpp — TEST INFRASTRUCTURE ONLY (part of oracle/; never linked into the product).
//
// A header-only stand-in for the small slice of the OpenCV C++ API that the reference's
// four hot-path translation units touch, so that those TUs can be compiled UNMODIFIED,
// in place from /root/reference, into oracle/_ref/libvoref.so (no C++ OpenCV exists in
// this image; only the Python cv2 wheel).  Symbol surface taken from a grep of
//   feature_extraction_parallel/{corner_detection_parallel,FREAK_feature_descriptor_parallel,
//   feature_extraction_parallel}.cpp and feature_matching_parallel/feature_matching_parallel.cpp.
//
// The only arithmetic in here is
//   * saturate_cast<uchar>(double)   (OpenCV: cvRound = lrint, then clamp), and
//   * GaussianBlur(u8, Size(7,7), sigma=0, BORDER_REFLECT_101), written in OpenCV's exact
//     fixed-point form: out = (sum_y sum_x w[y] w[x] src + 2048) >> 12, w = [2,7,14,18,14,7,2].
//     tests/test_oracle.py checks this form bit-for-bit against Python cv2.GaussianBlur.
// Display / drawing calls are no-ops (no graphics stack on either box).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#define CV_8U 0
#define CV_32F 5
#define CV_8UC1 0
#define CV_8UC3 16
#define CV_32FC1 5

typedef unsigned char uchar;

namespace cv {

template <typename T> struct Point_ {
    T x{}, y{};
    Point_() = default;
    template <typename A, typename B> Point_(A x_, B y_) : x(static_cast<T>(x_)), y(static_cast<T>(y_)) {}
};
typedef Point_<int> Point;
typedef Point_<float> Point2f;

struct Size {
    int width{}, height{};
    Size() = default;
    Size(int w, int h) : width(w), height(h) {}
};

struct Scalar {
    double val[4]{};
    Scalar() = default;
    Scalar(double a, double b = 0, double c = 0, double d = 0) : val{a, b, c, d} {}
};

struct Vec3b {
    uchar val[3]{};
    Vec3b() = default;
    Vec3b(uchar a, uchar b, uchar c) : val{a, b, c} {}
    uchar& operator[](int i) { return val[i]; }
    const uchar& operator[](int i) const { return val[i]; }
};

struct KeyPoint {
    Point2f pt;
    float size{};
    float angle{-1.f};
    float response{};
    int octave{};
    int class_id{-1};
    KeyPoint() = default;
    KeyPoint(Point2f p, float s) : pt(p), size(s) {}
    KeyPoint(float x, float y, float s) : pt(x, y), size(s) {}
};

struct DMatch {
    int queryIdx{-1}, trainIdx{-1}, imgIdx{-1};
    float distance{};
    DMatch() = default;
    DMatch(int q, int t, float d) : queryIdx(q), trainIdx(t), distance(d) {}
};

enum { THRESH_TRUNC = 2, RANSAC = 8, COLOR_GRAY2BGR = 8, IMREAD_GRAYSCALE = 0, IMREAD_COLOR = 1 };

template <typename T> static inline T saturate_cast(double v);
template <> inline uchar saturate_cast<uchar>(double v) {
    const long iv = std::lrint(v);  // cvRound: round-half-to-even
    return static_cast<uchar>(iv < 0 ? 0 : (iv > 255 ? 255 : iv));
}

class Mat {
public:
    int rows{0}, cols{0};
    uchar* data{nullptr};
    size_t step{0};

    Mat() = default;
    Mat(int r, int c, int type) { create(r, c, type); }
    // Wrap caller-owned memory (like cv::Mat(rows, cols, type, void* data)).
    Mat(int r, int c, int type, void* ext, size_t step_ = 0) : rows(r), cols(c), type_(type) {
        data = static_cast<uchar*>(ext);
        step = step_ ? step_ : static_cast<size_t>(c) * elemSize();
    }
    static Mat zeros(int r, int c, int type) { return Mat(r, c, type); }  // create() zero-fills

    void create(int r, int c, int type) {
        rows = r; cols = c; type_ = type;
        step = static_cast<size_t>(c) * elemSize();
        buf_ = std::make_shared<std::vector<uchar>>(step * static_cast<size_t>(r), 0);
        data = buf_->data();
    }
    int type() const { return type_; }
    int channels() const { return (type_ >> 3) + 1; }
    size_t elemSize() const { return static_cast<size_t>(channels()) * ((type_ & 7) == CV_32F ? 4 : 1); }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    Mat clone() const {
        Mat m(rows, cols, type_);
        for (int i = 0; i < rows; ++i) std::memcpy(m.data + i * m.step, data + i * step, m.step);
        return m;
    }

    template <typename T> T* ptr(int i = 0) { return reinterpret_cast<T*>(data + static_cast<size_t>(i) * step); }
    template <typename T> const T* ptr(int i = 0) const { return reinterpret_cast<const T*>(data + static_cast<size_t>(i) * step); }
    template <typename T> T& at(int i, int j) { return ptr<T>(i)[j]; }
    template <typename T> const T& at(int i, int j) const { return ptr<T>(i)[j]; }

    Mat operator/(float d) const {
        Mat m(rows, cols, type_);
        if ((type_ & 7) == CV_32F)
            for (int i = 0; i < rows; ++i)
                for (int j = 0; j < cols * channels(); ++j) m.ptr<float>(i)[j] = ptr<float>(i)[j] / d;
        return m;
    }

private:
    int type_{0};
    std::shared_ptr<std::vector<uchar>> buf_;
};

static inline int reflect101(int p, int n) {
    if (n == 1) return 0;
    while (p < 0 || p >= n) p = p < 0 ? -p : 2 * n - 2 - p;
    return p;
}

// cv::GaussianBlur for CV_8UC1, ksize 7x7, sigma 0, default border: exact fixed-point form.
static inline void GaussianBlur(const Mat& src, Mat& dst, Size ksize, double sigma) {
    if (ksize.width != 7 || ksize.height != 7 || sigma != 0 || src.type() != CV_8UC1)
        throw std::runtime_error("cvshim: only GaussianBlur(u8c1, 7x7, 0) is restated");
    static const int w[7] = {2, 7, 14, 18, 14, 7, 2};
    const int H = src.rows, W = src.cols;
    Mat out(H, W, CV_8UC1);
    std::vector<uint16_t> hs(static_cast<size_t>(H) * W);
    for (int i = 0; i < H; ++i) {
        const uchar* s = src.ptr<uchar>(i);
        uint16_t* h = hs.data() + static_cast<size_t>(i) * W;
        for (int j = 0; j < W; ++j) {
            int acc = 0;
            if (j >= 3 && j + 3 < W) for (int k = 0; k < 7; ++k) acc += w[k] * s[j + k - 3];
            else for (int k = 0; k < 7; ++k) acc += w[k] * s[reflect101(j + k - 3, W)];
            h[j] = static_cast<uint16_t>(acc);
        }
    }
    for (int i = 0; i < H; ++i) {
        const uint16_t* r[7];
        for (int k = 0; k < 7; ++k) r[k] = hs.data() + static_cast<size_t>(reflect101(i + k - 3, H)) * W;
        uchar* d = out.ptr<uchar>(i);
        for (int j = 0; j < W; ++j) {
            uint32_t acc = 2048;
            for (int k = 0; k < 7; ++k) acc += static_cast<uint32_t>(w[k]) * r[k][j];
            d[j] = static_cast<uchar>(acc >> 12);
        }
    }
    dst = out;
}

// ---- display / drawing / misc: no-ops (never reached on the hot path) -------------------------
static inline void imshow(const std::string&, const Mat&) {}
template <class... A> static inline void drawMatches(const A&...) {}
static inline int waitKey(int = 0) { return -1; }
static inline void destroyAllWindows() {}
static inline bool imwrite(const std::string&, const Mat&) { return false; }
static inline Mat imread(const std::string&, int = IMREAD_COLOR) { return Mat(); }
static inline void cvtColor(const Mat&, Mat&, int) {}
static inline void rectangle(Mat&, Point, Point, const Scalar&, int = 1) {}
static inline void circle(Mat&, Point, int, const Scalar&, int = 1) {}
static inline double threshold(const Mat&, Mat&, double, double, int) { return 0; }
static inline void minMaxLoc(const Mat&, double*, double* = nullptr) {}
static inline Mat findHomography(const std::vector<Point2f>& a, const std::vector<Point2f>&, int, double,
                                 std::vector<uchar>& mask) {
    mask.assign(a.size(), 1);
    return Mat();
}

}  // namespace cv
