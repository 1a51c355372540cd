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
// shared memory ([point][thread], conflict free).  Per pair the kernel issues six instructions: one shared
// load, one constant-bank load of the pair's weights through the uniform datapath (no LSU traffic), the
// intensity difference, and three packed-fp32 operations (FMUL2 / FFMA2 / FADD2) that carry the x and y
// components together.  Rows of the pair table are padded to multiples of six with zero-weight entries
// (adding +-0 never changes a sum that started at +0), so the inner loop has no remainder.
//
// Exactness: the orientation sums are accumulated in the reference's pair order with one fp32 rounding
// per term.  The reference's term is RN(RN(ic*d) / |d|) with ic*d an exact integer, i.e. RN(ic * w) for the
// real number w = d/|d|.  Here w is split into wh = RN(w), wl = RN(w - wh) and the term is
// fma(ic, wh, RN(ic*wl)): the argument of the final rounding is within 2^-47 of ic*w, so the result can
// differ only if ic*w lies that close to a rounding boundary - checked EXHAUSTIVELY against the IEEE
// quotient for all 903 pairs x 511 intensity differences x 2 axes: zero mismatches
// (tests/test_oracle.py::test_split_weight_product_is_exact).  So O_x and O_y are bit-identical to the CPU.
// The only inexact step is atan2f/cosf/sinf: glibc's are < 1 ulp but not correctly rounded; here they are
// evaluated in fp64 and rounded once to fp32 (measured: angles differ by at most 1 ulp, descriptor bits
// identical on every fixture).
#include "freak_tables.h"
#include "vo_internal.cuh"

namespace {

constexpr int FK_NT = 64;
constexpr int FK_GROUP = 6;  // pairs per inner-loop trip
// padded pair table: row p holds the pairs (p, r), r = 43 - 6*ceil((42-p)/6) .. 42; entries with r <= p are zero
constexpr int FK_TABLE = 1008;  // sum over p of 6*ceil((42-p)/6)

struct __align__(16) PairW { float whx, why, wlx, wly; };
constexpr int FK_GROUPS = FK_TABLE / FK_GROUP;  // 168 groups of six pairs, in the reference's pair order
__constant__ PairW c_w[FK_TABLE + FK_GROUP];    // + one dummy group so that the look-ahead loads stay in bounds
__constant__ short2 c_grp[FK_GROUPS + 1];       // group -> (first point p, first second point r)
__constant__ signed char c_px[VO_FREAK_POINTS];
__constant__ signed char c_py[VO_FREAK_POINTS];

typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk(u64 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }

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

    // O += RN(ic * d / |d|) over the 903 pairs in the reference's order; (x, y) components packed.
    // The product of the packed multiply feeds only the addend of the FMA, and the FMA only the addend of
    // the sum, so ptxas has nothing to contract.
    // The loop runs over the 168 groups of six pairs, two per trip with ping-pong register sets: the weights
    // (constant bank, uniform datapath) and samples (shared memory) of the next group are requested before the
    // running group is evaluated, so their latency hides behind its 30 instructions; without the look-ahead every
    // group stalled on its own constant loads (ncu: 43 % of the loop's samples on the short scoreboard).
    u64 o2 = 0ull;
    PairW wa[FK_GROUP], wb[FK_GROUP];
    float va[FK_GROUP], vb[FK_GROUP], vpa, vpb;
    auto load = [&](PairW (&w)[FK_GROUP], float (&v)[FK_GROUP], float& vp, int G) {
        const short2 g = c_grp[G];  // uniform
        vp = s_v[g.x][t];
#pragma unroll
        for (int i = 0; i < FK_GROUP; ++i) { w[i] = c_w[G * FK_GROUP + i]; v[i] = s_v[g.y + i][t]; }
    };
    auto eval = [&](const PairW (&w)[FK_GROUP], const float (&v)[FK_GROUP], float vp) {
#pragma unroll
        for (int i = 0; i < FK_GROUP; ++i) {
            const float ic = __fsub_rn(vp, v[i]);
            const u64 ic2 = pk(ic, ic);
            o2 = add2(o2, fma2(ic2, pk(w[i].whx, w[i].why), mul2(ic2, pk(w[i].wlx, w[i].wly))));
        }
    };
    load(wa, va, vpa, 0);
#pragma unroll 1
    for (int G = 0; G < FK_GROUPS; G += 2) {
        load(wb, vb, vpb, G + 1);
        eval(wa, va, vpa);
        load(wa, va, vpa, G + 2);  // G + 2 == FK_GROUPS on the last trip: the dummy group
        eval(wb, vb, vpb);
    }
    float ox, oy;
    upk(o2, ox, oy);
    float angle = 0.0f;
    if (!(isnan(ox) || isnan(oy))) angle = (float)atan2((double)oy, (double)ox);
    a.angles[(size_t)f * a.kp_stride + q] = angle;
    const float cs = (float)cos((double)angle), sn = (float)sin((double)angle);

    // rotated samples: rm = {cos, -sin, sin, cos}; x' = (int)(kx + px*rm0 + py*rm2); y' = (int)(ky + (-1)*px*rm1 + py*rm3)
    // (-1)*px*rm1 = (-px)*(-sin) = RN(px*sin) exactly, so the products below are the reference's.
    int v[VO_FREAK_POINTS];
#define VO_SAMPLE(i, PX, PY)                                                                             \
    {                                                                                                    \
        const float fx = __fadd_rn(__fadd_rn(kp.x, __fmul_rn((float)(PX), cs)), __fmul_rn((float)(PY), sn)); \
        const float fy = __fadd_rn(__fadd_rn(kp.y, __fmul_rn((float)(PX), sn)), __fmul_rn((float)(PY), cs)); \
        v[i] = img[(int)fy * pitch + (int)fx]; /* truncation toward zero */                               \
    }
    VO_FREAK_POINT_LIST(VO_SAMPLE)
#undef VO_SAMPLE
    // test j: bit = v[A] > v[B] = sign of v[B] - v[A]; one subtract and one funnel shift push it into the word
    // (first test ends up in bit 31: the word is bit-reversed once at the end)
    uint32_t w[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) w[i] = 0u;
#define VO_TEST(j, A, B) w[(j) >> 5] = __funnelshift_l(static_cast<uint32_t>(v[B] - v[A]), w[(j) >> 5], 1);
    VO_FREAK_TEST_LIST(VO_TEST)
#undef VO_TEST
#pragma unroll
    for (int i = 0; i < 16; ++i) w[i] = __brev(w[i]);
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
    static PairW tab[FK_TABLE + FK_GROUP];  // the dummy group stays zero
    int k = 0;
    for (int p = 0; p < VO_FREAK_POINTS - 1; ++p) {
        const int groups = (VO_FREAK_POINTS - 1 - p + FK_GROUP - 1) / FK_GROUP;
        for (int r = VO_FREAK_POINTS - FK_GROUP * groups; r < VO_FREAK_POINTS; ++r, ++k) {
            PairW w = {0.f, 0.f, 0.f, 0.f};
            if (r > p) {
                const float fx = (float)(VO_FREAK_PX[p] - VO_FREAK_PX[r]), fy = (float)(VO_FREAK_PY[p] - VO_FREAK_PY[r]);
                const float nm = sqrtf(fx * fx + fy * fy);  // IEEE sqrt on exact small integers, as the reference computes it
                // w = d / |d| split into RN(w) and RN(w - RN(w)); host IEEE double arithmetic, no fast-math
                const double wx = (double)fx / (double)nm, wy = (double)fy / (double)nm;
                w.whx = (float)wx; w.wlx = (float)(wx - (double)w.whx);
                w.why = (float)wy; w.wly = (float)(wy - (double)w.why);
            }
            tab[k] = w;
        }
    }
    if (k != FK_TABLE) return cudaErrorInvalidValue;
    static short2 grp[FK_GROUPS + 1];
    int gi = 0;
    for (int p = 0; p < VO_FREAK_POINTS - 1; ++p) {
        const int groups = (VO_FREAK_POINTS - 1 - p + FK_GROUP - 1) / FK_GROUP;
        for (int g = 0; g < groups; ++g, ++gi) grp[gi] = make_short2((short)p, (short)(VO_FREAK_POINTS - FK_GROUP * groups + FK_GROUP * g));
    }
    if (gi != FK_GROUPS) return cudaErrorInvalidValue;
    grp[FK_GROUPS] = make_short2(0, 0);
    cudaError_t e;
    if ((e = cudaMemcpyToSymbol(c_grp, grp, sizeof(grp))) != cudaSuccess) return e;
    if ((e = cudaMemcpyToSymbol(c_w, tab, sizeof(tab))) != cudaSuccess) return e;
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
