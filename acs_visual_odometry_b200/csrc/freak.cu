// freak.cu — K3: FREAK-style orientation + 512-test binary descriptor.
//
// Replaces FREAK_Parallel::compute_orientation (feature_extraction_parallel/FREAK_feature_descriptor_parallel.cpp:11-55)
// and FREAK_Parallel::FREAK_feature_description_worker (:110-151), i.e. the reference's simplified FREAK:
// 43 single-pixel samples, orientation from all 903 point pairs, a non-orthogonal "rotation"
//   x' = (int)(kx + px*cos + py*sin),  y' = (int)(ky + px*sin + py*cos),
// and 512 pairwise comparisons.  Reproduced as written.
//
// Exactness: the orientation sums are accumulated in the reference's pair order with one fp32 rounding
// per operation (products ic*dx are exact integers; the division is IEEE), so O_x and O_y are
// bit-identical to the CPU.  The only inexact step is atan2f/cosf/sinf: glibc's are < 1 ulp but not
// correctly rounded; here they are evaluated in fp64 and rounded once to fp32.
#include "freak_tables.h"
#include "vo_internal.cuh"

namespace {

constexpr int FK_NT = 128;

__constant__ float c_pair_dx[VO_FREAK_PAIRS];
__constant__ float c_pair_dy[VO_FREAK_PAIRS];
__constant__ float c_pair_norm[VO_FREAK_PAIRS];
__constant__ signed char c_px[VO_FREAK_POINTS];
__constant__ signed char c_py[VO_FREAK_POINTS];
__constant__ unsigned char c_test_a[VO_FREAK_TESTS];
__constant__ unsigned char c_test_b[VO_FREAK_TESTS];

__global__ void __launch_bounds__(FK_NT) k_freak(const FreakArgs a) {
    __shared__ float s_v[VO_FREAK_POINTS][FK_NT];  // per-thread sample values, bank-conflict free
    const int f = blockIdx.y;
    const int q = blockIdx.x * FK_NT + threadIdx.x;
    const int n = a.nkp[f];
    if (q >= n) return;
    const int t = threadIdx.x;
    const uint8_t* __restrict__ img = a.blur + (size_t)f * a.blur_fstride;
    const float2 kp = a.kp[(size_t)f * a.kp_stride + q];
    const int kx = (int)kp.x, ky = (int)kp.y;

    // unrotated samples (always in bounds: NMS keeps 37 <= x <= W-37, 35 <= y <= H-35)
    for (int p = 0; p < VO_FREAK_POINTS; ++p)
        s_v[p][t] = (float)img[(size_t)(ky + c_py[p]) * a.blur_pitch + (kx + c_px[p])];

    float ox = 0.0f, oy = 0.0f;
    int k = 0;
    for (int p = 0; p < VO_FREAK_POINTS; ++p) {
        const float vp = s_v[p][t];
        for (int r = p + 1; r < VO_FREAK_POINTS; ++r, ++k) {
            const float ic = __fsub_rn(vp, s_v[r][t]);
            ox = __fadd_rn(ox, __fdiv_rn(__fmul_rn(ic, c_pair_dx[k]), c_pair_norm[k]));
            oy = __fadd_rn(oy, __fdiv_rn(__fmul_rn(ic, c_pair_dy[k]), c_pair_norm[k]));
        }
    }
    float angle = 0.0f;
    if (!(isnan(ox) || isnan(oy))) angle = (float)atan2((double)oy, (double)ox);
    a.angles[(size_t)f * a.kp_stride + q] = angle;
    const float cs = (float)cos((double)angle), sn = (float)sin((double)angle);

    // rotated samples: rm = {cos, -sin, sin, cos}; x' = kx + px*rm0 + py*rm2; y' = ky + (-1)*px*rm1 + py*rm3
    const float nsn = -sn;
    for (int p = 0; p < VO_FREAK_POINTS; ++p) {
        const float px = (float)c_px[p], py = (float)c_py[p];
        const float fx = __fadd_rn(__fadd_rn(kp.x, __fmul_rn(px, cs)), __fmul_rn(py, sn));
        const float fy = __fadd_rn(__fadd_rn(kp.y, __fmul_rn(__fmul_rn(-1.0f, px), nsn)), __fmul_rn(py, cs));
        const int sx = (int)fx, sy = (int)fy;  // truncation toward zero
        s_v[p][t] = (float)img[(size_t)sy * a.blur_pitch + sx];
    }
    uint32_t* __restrict__ out = a.desc + ((size_t)f * a.kp_stride + q) * 16;
    for (int w = 0; w < 16; ++w) {
        uint32_t bits = 0;
#pragma unroll 8
        for (int b = 0; b < 32; ++b) {
            const int j = w * 32 + b;
            bits |= (s_v[c_test_a[j]][t] > s_v[c_test_b[j]][t] ? 1u : 0u) << b;
        }
        out[w] = bits;
    }
}

// n x 512 bytes (0/1) -> n x 16 words; one warp per descriptor, one ballot per word.
__global__ void k_pack(const uint8_t* __restrict__ bytes, int n, uint32_t* __restrict__ packed, int* flag) {
    const int q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (q >= n) return;
    bool bad = false;
    for (int w = 0; w < 16; ++w) {
        const uint8_t v = bytes[(size_t)q * 512 + w * 32 + lane];
        bad |= v > 1;
        const uint32_t word = __ballot_sync(0xFFFFFFFFu, v & 1);
        if (lane == 0) packed[(size_t)q * 16 + w] = word;
    }
    if (bad) atomicOr(flag, 2);
}

__global__ void k_unpack(const uint32_t* __restrict__ packed, int n, uint8_t* __restrict__ bytes) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)n * 512) return;
    const size_t q = idx >> 9;
    const int j = (int)(idx & 511);
    bytes[idx] = (uint8_t)((packed[q * 16 + (j >> 5)] >> (j & 31)) & 1u);
}

}  // namespace

cudaError_t freak_upload_tables() {
    static float dx[VO_FREAK_PAIRS], dy[VO_FREAK_PAIRS], nm[VO_FREAK_PAIRS];
    int k = 0;
    for (int p = 0; p < VO_FREAK_POINTS; ++p)
        for (int r = p + 1; r < VO_FREAK_POINTS; ++r, ++k) {
            const float fx = (float)(VO_FREAK_PX[p] - VO_FREAK_PX[r]), fy = (float)(VO_FREAK_PY[p] - VO_FREAK_PY[r]);
            dx[k] = fx; dy[k] = fy;
            nm[k] = sqrtf(fx * fx + fy * fy);  // IEEE sqrt on exact small integers, as the reference computes it
        }
    cudaError_t e;
    if ((e = cudaMemcpyToSymbol(c_pair_dx, dx, sizeof(dx))) != cudaSuccess) return e;
    if ((e = cudaMemcpyToSymbol(c_pair_dy, dy, sizeof(dy))) != cudaSuccess) return e;
    if ((e = cudaMemcpyToSymbol(c_pair_norm, nm, sizeof(nm))) != cudaSuccess) return e;
    if ((e = cudaMemcpyToSymbol(c_px, VO_FREAK_PX, sizeof(VO_FREAK_PX))) != cudaSuccess) return e;
    if ((e = cudaMemcpyToSymbol(c_py, VO_FREAK_PY, sizeof(VO_FREAK_PY))) != cudaSuccess) return e;
    if ((e = cudaMemcpyToSymbol(c_test_a, VO_FREAK_TEST_A, sizeof(VO_FREAK_TEST_A))) != cudaSuccess) return e;
    if ((e = cudaMemcpyToSymbol(c_test_b, VO_FREAK_TEST_B, sizeof(VO_FREAK_TEST_B))) != cudaSuccess) return e;
    return cudaSuccess;
}

cudaError_t launch_freak(const FreakArgs& a, int n_frames, int max_n, cudaStream_t st) {
    if (max_n <= 0) return cudaSuccess;
    const dim3 grid((max_n + FK_NT - 1) / FK_NT, n_frames);
    k_freak<<<grid, FK_NT, 0, st>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_pack_bytes(const uint8_t* bytes, int n, uint32_t* packed, int* flag, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    k_pack<<<(n * 32 + 255) / 256, 256, 0, st>>>(bytes, n, packed, flag);
    return cudaGetLastError();
}

cudaError_t launch_unpack_bytes(const uint32_t* packed, int n, uint8_t* bytes, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    const size_t total = (size_t)n * 512;
    k_unpack<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(packed, n, bytes);
    return cudaGetLastError();
}
