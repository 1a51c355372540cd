// vo_image_io.cpp — see vo_image_io.hpp.
#include "vo_image_io.hpp"

#include <zlib.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace vo {
namespace {

void set_err(std::string* err, const char* msg) { if (err) *err = msg; }
uint32_t be32(const uint8_t* p) { return (uint32_t(p[0]) << 24) | (uint32_t(p[1]) << 16) | (uint32_t(p[2]) << 8) | p[3]; }
void put_be32(std::vector<uint8_t>& v, uint32_t x) { v.push_back(x >> 24); v.push_back(x >> 16); v.push_back(x >> 8); v.push_back(x); }

int paeth(int a, int b, int c) {
    const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

bool read_file(const std::string& path, std::vector<uint8_t>& out) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END);
    const long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    if (n < 0) { std::fclose(f); return false; }
    out.resize(static_cast<size_t>(n));
    const size_t got = n ? std::fread(out.data(), 1, out.size(), f) : 0;
    std::fclose(f);
    return got == out.size();
}

void write_chunk(std::vector<uint8_t>& png, const char type[4], const std::vector<uint8_t>& body) {
    put_be32(png, static_cast<uint32_t>(body.size()));
    const size_t start = png.size();
    png.insert(png.end(), type, type + 4);
    png.insert(png.end(), body.begin(), body.end());
    put_be32(png, static_cast<uint32_t>(crc32(0L, png.data() + start, static_cast<uInt>(png.size() - start))));
}

}  // namespace

GrayImage decodePNG(const uint8_t* b, size_t n, std::string* err) {
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    GrayImage img;
    if (n < 8 || std::memcmp(b, sig, 8) != 0) { set_err(err, "not a PNG file"); return img; }
    size_t pos = 8;
    uint32_t width = 0, height = 0;
    int depth = 0, ctype = -1, interlace = 0;
    std::vector<uint8_t> idat;
    bool end = false;
    while (!end && pos + 12 <= n) {
        const uint32_t len = be32(b + pos);
        const uint8_t* type = b + pos + 4;
        if (pos + 12 + static_cast<size_t>(len) > n) { set_err(err, "truncated PNG chunk"); return img; }
        const uint8_t* body = b + pos + 8;
        if (!std::memcmp(type, "IHDR", 4) && len >= 13) {
            width = be32(body); height = be32(body + 4); depth = body[8]; ctype = body[9]; interlace = body[12];
        } else if (!std::memcmp(type, "IDAT", 4)) {
            idat.insert(idat.end(), body, body + len);
        } else if (!std::memcmp(type, "IEND", 4)) {
            end = true;
        }
        pos += 12 + len;
    }
    if (!width || !height || width > 65535 || height > 65535) { set_err(err, "bad PNG header"); return img; }
    if (depth != 8 || interlace != 0 || !(ctype == 0 || ctype == 2 || ctype == 4 || ctype == 6)) {
        set_err(err, "unsupported PNG (need 8-bit, non-interlaced gray / gray+alpha / RGB / RGBA)");
        return img;
    }
    const int ch = ctype == 0 ? 1 : ctype == 4 ? 2 : ctype == 2 ? 3 : 4;
    const size_t stride = static_cast<size_t>(width) * ch;
    std::vector<uint8_t> raw((stride + 1) * height);
    uLongf raw_len = static_cast<uLongf>(raw.size());
    if (uncompress(raw.data(), &raw_len, idat.data(), static_cast<uLong>(idat.size())) != Z_OK || raw_len != raw.size()) {
        set_err(err, "PNG inflate failed");
        return img;
    }
    // undo the scanline filters in place
    std::vector<uint8_t> prev(stride, 0);
    img.rows = static_cast<int>(height); img.cols = static_cast<int>(width);
    img.data.resize(static_cast<size_t>(width) * height);
    for (uint32_t y = 0; y < height; ++y) {
        uint8_t* line = raw.data() + y * (stride + 1);
        const int ft = line[0];
        uint8_t* cur = line + 1;
        for (size_t i = 0; i < stride; ++i) {
            const int a = i >= static_cast<size_t>(ch) ? cur[i - ch] : 0, up = prev[i], c = i >= static_cast<size_t>(ch) ? prev[i - ch] : 0;
            int v = cur[i];
            switch (ft) {
                case 0: break;
                case 1: v += a; break;
                case 2: v += up; break;
                case 3: v += (a + up) >> 1; break;
                case 4: v += paeth(a, up, c); break;
                default: set_err(err, "bad PNG filter"); img = GrayImage(); return img;
            }
            cur[i] = static_cast<uint8_t>(v);
        }
        std::memcpy(prev.data(), cur, stride);
        uint8_t* dst = img.data.data() + static_cast<size_t>(y) * width;
        if (ch <= 2) {
            for (uint32_t x = 0; x < width; ++x) dst[x] = cur[x * ch];
        } else {  // cv::imread(GRAYSCALE) lets libpng reduce colour: (R*9797 + G*19234 + B*3737) >> 15 (matches cv2 4.13 bit for bit)
            for (uint32_t x = 0; x < width; ++x) {
                const uint8_t* p = cur + static_cast<size_t>(x) * ch;
                dst[x] = static_cast<uint8_t>((p[0] * 9797 + p[1] * 19234 + p[2] * 3737) >> 15);
            }
        }
    }
    return img;
}

GrayImage decodePGM(const uint8_t* b, size_t n, std::string* err) {
    GrayImage img;
    size_t pos = 0;
    auto token = [&](long& out) {
        while (pos < n) {
            if (b[pos] == '#') { while (pos < n && b[pos] != '\n') ++pos; }
            else if (b[pos] == ' ' || b[pos] == '\t' || b[pos] == '\n' || b[pos] == '\r') ++pos;
            else break;
        }
        if (pos >= n || b[pos] < '0' || b[pos] > '9') return false;
        out = 0;
        while (pos < n && b[pos] >= '0' && b[pos] <= '9') out = out * 10 + (b[pos++] - '0');
        return true;
    };
    if (n < 2 || b[0] != 'P' || b[1] != '5') { set_err(err, "not a binary PGM (P5)"); return img; }
    pos = 2;
    long w = 0, h = 0, maxv = 0;
    if (!token(w) || !token(h) || !token(maxv) || w <= 0 || h <= 0 || maxv <= 0 || maxv > 255) { set_err(err, "bad PGM header"); return img; }
    ++pos;  // the single whitespace after maxval
    if (pos + static_cast<size_t>(w) * h > n) { set_err(err, "truncated PGM"); return img; }
    img.rows = static_cast<int>(h); img.cols = static_cast<int>(w);
    img.data.assign(b + pos, b + pos + static_cast<size_t>(w) * h);
    return img;
}

GrayImage readGray(const std::string& path, std::string* err) {
    std::vector<uint8_t> bytes;
    if (!read_file(path, bytes)) { set_err(err, "cannot read file"); return GrayImage(); }
    if (bytes.size() >= 2 && bytes[0] == 'P' && bytes[1] == '5') return decodePGM(bytes.data(), bytes.size(), err);
    return decodePNG(bytes.data(), bytes.size(), err);
}

bool writePNG(const std::string& path, const uint8_t* gray, int rows, int cols) {
    std::vector<uint8_t> raw(static_cast<size_t>(cols + 1) * rows);
    for (int y = 0; y < rows; ++y) {
        raw[static_cast<size_t>(y) * (cols + 1)] = 0;  // filter: none
        std::memcpy(&raw[static_cast<size_t>(y) * (cols + 1) + 1], gray + static_cast<size_t>(y) * cols, cols);
    }
    uLongf clen = compressBound(static_cast<uLong>(raw.size()));
    std::vector<uint8_t> comp(clen);
    if (compress2(comp.data(), &clen, raw.data(), static_cast<uLong>(raw.size()), 1) != Z_OK) return false;
    comp.resize(clen);
    std::vector<uint8_t> png = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    std::vector<uint8_t> ihdr;
    put_be32(ihdr, static_cast<uint32_t>(cols)); put_be32(ihdr, static_cast<uint32_t>(rows));
    ihdr.push_back(8); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
    write_chunk(png, "IHDR", ihdr);
    write_chunk(png, "IDAT", comp);
    write_chunk(png, "IEND", {});
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) return false;
    const bool ok = std::fwrite(png.data(), 1, png.size(), f) == png.size();
    std::fclose(f);
    return ok;
}

bool writePGM(const std::string& path, const uint8_t* gray, int rows, int cols) {
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) return false;
    std::fprintf(f, "P5\n%d %d\n255\n", cols, rows);
    const bool ok = std::fwrite(gray, 1, static_cast<size_t>(rows) * cols, f) == static_cast<size_t>(rows) * cols;
    std::fclose(f);
    return ok;
}

}  // namespace vo
