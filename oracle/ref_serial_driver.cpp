// ref_serial_driver.cpp — TEST INFRASTRUCTURE ONLY (part of oracle/; never linked into the product).
//
// C entry points around the REFERENCE's own FULL-LENGTH matcher, compiled unmodified and in place:
//   hammingDistance(d1, d2, length)              matching_serial.cpp:24-40   (popcount over all `length` bytes)
//   matchCustomBinaryDescriptorsSerial(d1, d2)   matching_serial.cpp:42-77
// matching_serial.cpp is a stand-alone driver program: its `main` (which needs the serial extractor and a
// display) is renamed to an unused static function so that only the two functions above are emitted.
// It redefines findCommonMatches / convertToDMatch / applyRANSAC, which feature_matching_parallel.cpp
// also defines, so it lives in its own library: oracle/_ref/libvoref_serial.so (see oracle/Makefile).
// Uses: pinning the FULL512 mode of the restatement (oracle/vo_oracle.c) and the FULL512 golden vectors.
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <limits>
#include <vector>

#define main static __attribute__((unused)) ref_matching_serial_main
#include "matching_serial.cpp"  // resolved through -I$(REF): the reference's file, where it lies
#undef main

extern "C" {

int voref_serial_hamming(const uint8_t* d1, const uint8_t* d2, int length) { return hammingDistance(d1, d2, length); }

// d1: n1 x row_bytes, d2: n2 x row_bytes (one 0/1 byte per test); returns the number of (query, train) pairs
int voref_serial_match(const uint8_t* d1, int n1, const uint8_t* d2, int n2, int row_bytes, float ratio, int32_t* pairs, int cap) {
    std::vector<std::vector<uint8_t>> a(n1), b(n2);
    for (int i = 0; i < n1; ++i) a[i].assign(d1 + static_cast<size_t>(i) * row_bytes, d1 + static_cast<size_t>(i + 1) * row_bytes);
    for (int i = 0; i < n2; ++i) b[i].assign(d2 + static_cast<size_t>(i) * row_bytes, d2 + static_cast<size_t>(i + 1) * row_bytes);
    const auto m = matchCustomBinaryDescriptorsSerial(a, b, ratio);
    const int n = static_cast<int>(m.size());
    if (n > cap) return -n;
    for (int i = 0; i < n; ++i) { pairs[2 * i] = m[i].first; pairs[2 * i + 1] = m[i].second; }
    return n;
}

}  // extern "C"
