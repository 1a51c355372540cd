// freak.cu — K3: FREAK-style orientation + 512-test binary descriptor.
//
// Replaces FREAK_Parallel::compute_orientation (feature_extraction_parallel/FREAK_feature_descriptor_parallel.cpp:11-55)
// and FREAK_Parallel::FREAK_feature_description_worker (:110-151), i.e. the reference's simplified FREAK:
// 43 single-pixel samples, orientation from all 903 point pairs, a non-orthogonal "rotation"
//   x' = (int)(kx + px*cos + py*sin),  y' = (int)(ky + px*sin + py*cos),
// and 512 pairwise comparisons.  Reproduced as written.
//
// Mapping: ONE THREAD PER KEYPOINT.  The orientation is a 903-term fp32 sum whose rounding depends on
// the order, so it is a dependent chain per keypoint; giving every lane its own chain keeps all 32 lanes
// busy (a warp-per-keypoint layout would idle 30 lanes during the chain).  The 43 unrotated samples sit in
// shared memory ([point][thread], conflict free); the pair constants (dx, dy, |d|, 1/|d|) are one
// broadcast 16-byte load per pair; the 43 rotated samples stay in registers and the 512 tests are fully
// unrolled register compares (freak_tables.h X-macros).
//
// Exactness: the orientation sums are accumulated in the reference's pair order with one fp32 rounding
// per operation.  ic*dx is an exact integer; the IEEE quotient (ic*dx)/|d| is obtained without a divide:
// q = m*rn, r = fma(-q, |d|, m), q' = fma(r, rn, q) with rn = RN(1/|d|) - Markstein's correction, checked
// EXHAUSTIVELY against m/|d| for all 903 pairs x 511 intensity differences x 2 axes
// (tests/test_oracle.py::test_reciprocal_division_is_exact).  So O_x and O_y are bit-identical to the CPU.
// The only inexact step is atan2f/cosf/sinf: glibc's are < 1 ulp but not correctly rounded; here they are
// evaluated in fp64 and rounded once to fp32 (measured: angles differ by at most 1 ulp, descriptor bits
// identical on every fixture).
#include "freak_tables.h"
#include "vo_internal.cuh"

namespace {

constexpr int FK_NT = 64;

__device__ float4 g_pair[VO_FREAK_PAIRS];  // (dx, dy, |d|, RN(1/|d|)) in the reference's lexicographic pair order
__constant__ signed char c_px[VO_FREAK_POINTS];
__constant__ signed char c_py[VO_FREAK_POINTS];

__global__ void __launch_bounds__(FK_NT) k_freak(const FreakArgs a) {
    __shared__ float s_v[VO_FREAK_POINTS][FK_NT];
    const int f = blockIdx.y;
    const int q = blockIdx.x * FK_NT + threadIdx.x;
    const int n = a.nkp[f];
    if (q >= n) return;
    const int t = threadIdx.x;
    const uint8_t* __restrict__ img = a.blur + (size_t)f * a.blur_fstride;
    const float2 kp = a.kp[(size_t)f * a.kp_stride + q];
    const int kx = (int)kp.x, ky = (int)kp.y;
    const int pitch = (int)a.blur_pitch;
    const uint8_t* __restrict__ centre = img + (size_t)ky * a.blur_pitch + kx;

    // unrotated samples (always in bounds: NMS keeps 37 <= x <= W-37, 35 <= y <= H-35)
#pragma unroll
    for (int p = 0; p < VO_FREAK_POINTS; ++p) s_v[p][t] = (float)centre[(int)c_py[p] * pitch + (int)c_px[p]];

    float ox = 0.0f, oy = 0.0f;
    const float4* __restrict__ pc = g_pair;
    for (int p = 0; p < VO_FREAK_POINTS - 1; ++p) {
        const float vp = s_v[p][t];
#pragma unroll 6
        for (int r = p + 1; r < VO_FREAK_POINTS; ++r) {
            const float4 c = __ldg(pc++);  // uniform address: one broadcast load
            const float ic = __fsub_rn(vp, s_v[r][t]);
            const float mx = __fmul_rn(ic, c.x), my = __fmul_rn(ic, c.y);  // exact integers
            const float qx = __fmul_rn(mx, c.w), qy = __fmul_rn(my, c.w);
            const float rx = __fmaf_rn(-qx, c.z, mx), ry = __fmaf_rn(-qy, c.z, my);
            ox = __fadd_rn(ox, __fmaf_rn(rx, c.w, qx));  // ox += RN((ic*dx)/|d|)
            oy = __fadd_rn(oy, __fmaf_rn(ry, c.w, qy));
        }
    }
    float angle = 0.0f;
    if (!(isnan(ox) || isnan(oy))) angle = (float)atan2((double)oy, (double)ox);
    a.angles[(size_t)f * a.kp_stride + q] = angle;
    const float cs = (float)cos((double)angle), sn = (float)sin((double)angle);

    // rotated samples: rm = {cos, -sin, sin, cos}; x' = (int)(kx + px*rm0 + py*rm2); y' = (int)(ky + (-1)*px*rm1 + py*rm3)
    // (-1)*px*rm1 = (-px)*(-sin) = RN(px*sin) exactly, so the products below are the reference's.
    float v[VO_FREAK_POINTS];
#define VO_SAMPLE(i, PX, PY)                                                                             \
    {                                                                                                    \
        const float fx = __fadd_rn(__fadd_rn(kp.x, __fmul_rn((float)(PX), cs)), __fmul_rn((float)(PY), sn)); \
        const float fy = __fadd_rn(__fadd_rn(kp.y, __fmul_rn((float)(PX), sn)), __fmul_rn((float)(PY), cs)); \
        v[i] = (float)img[(size_t)(int)fy * a.blur_pitch + (int)fx]; /* truncation toward zero */           \
    }
    VO_FREAK_POINT_LIST(VO_SAMPLE)
#undef VO_SAMPLE
    uint32_t w[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) w[k] = 0u;
#define VO_TEST(j, A, B) w[(j) >> 5] |= (v[A] > v[B] ? 1u : 0u) << ((j) & 31);
    VO_FREAK_TEST_LIST(VO_TEST)
#undef VO_TEST
    uint4* __restrict__ out = reinterpret_cast<uint4*>(a.desc + ((size_t)f * a.kp_stride + q) * 16);
    out[0] = make_uint4(w[0], w[1], w[2], w[3]);
    out[1] = make_uint4(w[4], w[5], w[6], w[7]);
    out[2] = make_uint4(w[8], w[9], w[10], w[11]);
    out[3] = make_uint4(w[12], w[13], w[14], w[15]);
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
    static float4 pairs[VO_FREAK_PAIRS];
    int k = 0;
    for (int p = 0; p < VO_FREAK_POINTS; ++p)
        for (int r = p + 1; r < VO_FREAK_POINTS; ++r, ++k) {
            const float fx = (float)(VO_FREAK_PX[p] - VO_FREAK_PX[r]), fy = (float)(VO_FREAK_PY[p] - VO_FREAK_PY[r]);
            const float nm = sqrtf(fx * fx + fy * fy);  // IEEE sqrt on exact small integers, as the reference computes it
            pairs[k] = make_float4(fx, fy, nm, 1.0f / nm);  // IEEE reciprocal (host arithmetic, no fast-math)
        }
    cudaError_t e;
    if ((e = cudaMemcpyToSymbol(g_pair, pairs, sizeof(pairs))) != cudaSuccess) return e;
    if ((e = cudaMemcpyToSymbol(c_px, VO_FREAK_PX, sizeof(VO_FREAK_PX))) != cudaSuccess) return e;
    if ((e = cudaMemcpyToSymbol(c_py, VO_FREAK_PY, sizeof(VO_FREAK_PY))) != cudaSuccess) return e;
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
