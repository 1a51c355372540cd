// freak.cu — K3: FREAK-style orientation + 512-test binary descriptor.
//
// Replaces FREAK_Parallel::compute_orientation (feature_extraction_parallel/FREAK_feature_descriptor_parallel.cpp:11-55)
// and FREAK_Parallel::FREAK_feature_description_worker (:110-151), i.e. the reference's simplified FREAK:
// 43 single-pixel samples, orientation from all 903 point pairs, a non-orthogonal "rotation"
//   x' = (int)(kx + px*cos + py*sin),  y' = (int)(ky + px*sin + py*cos),
// and 512 pairwise comparisons.  Reproduced as written.
//
// Mapping: ONE THREAD OWNS TWO KEYPOINTS.  The orientation is a 903-term fp32 sum whose rounding depends on
// the order, so it is a dependent chain per keypoint; giving every lane its own chains keeps all 32 lanes
// busy (a warp-per-keypoint layout would idle 30 lanes during the chain).  The two keypoints of a thread ride in
// the two halves of Blackwell's packed fp32 instructions: the 43 unrotated samples of both sit side by side in
// shared memory ([point][thread] float2, conflict free), so per pair of sample points the kernel issues one
// 8-byte shared load, one packed subtract (both intensity differences), and per axis FMUL2 / FFMA2 / FADD2 on
// (keypoint A, keypoint B) with the pair's weight as a broadcast scalar operand (the shared layout was since
// changed to byte windows, see the kernel); the weights come from the constant
// bank through the uniform datapath (no LSU traffic) and are shared by the two keypoints.  That is 5.75 issue
// slots per (pair, keypoint) against 8.6 for the one-keypoint-per-thread form this replaces (the x/y packing of
// that form needed one subtract, three packed operations and two constant loads per keypoint), and two
// independent accumulation chains per thread instead of one.  Rows of the pair table are padded to multiples of
// six with zero-weight entries (adding +-0 never changes a sum that started at +0), so the inner loop has no
// remainder.
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

#ifndef VO_FK_NT
#define VO_FK_NT 128
#endif
#ifndef VO_FK_MINB
#define VO_FK_MINB 4
#endif
constexpr int FK_NT = VO_FK_NT;    // threads per CTA
constexpr int FK_KP = 2;     // keypoints per thread
constexpr int FK_GROUP = 6;  // pairs per inner-loop trip
// padded pair table: row p holds the pairs (p, r), r = 43 - 6*ceil((42-p)/6) .. 42; entries with r <= p are zero
constexpr int FK_TABLE = 1008;  // sum over p of 6*ceil((42-p)/6)

// weights of one pair, each duplicated for the two keypoints of a thread: hi and lo parts of d/|d| per axis
struct __align__(16) PairW { float whx, wlx, why, wly; };  // hi and lo parts of d/|d| per axis
constexpr int FK_GROUPS = FK_TABLE / FK_GROUP;  // 168 groups of six pairs, in the reference's pair order
__constant__ PairW c_w[FK_TABLE + FK_GROUP];    // + one dummy group so that the look-ahead loads stay in bounds
__constant__ short2 c_grp[FK_GROUPS + 1];       // group -> (first point p, window w: second points 1+6w .. 6+6w)
constexpr int FK_WINDOWS = (VO_FREAK_POINTS - 1) / FK_GROUP;  // 7
__constant__ signed char c_px[VO_FREAK_POINTS];
__constant__ signed char c_py[VO_FREAK_POINTS];

typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk(u64 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ u64 sub2(u64 a, u64 b) { u64 d; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ u64 f2_as_u64(float2 v) { return pk(v.x, v.y); }

#ifndef VO_FK_SPLIT
#define VO_FK_SPLIT 0
#endif
#ifndef VO_FK_MINB_ORIENT
#define VO_FK_MINB_ORIENT 8
#endif
#ifndef VO_FK_MINB_DESCRIBE
#define VO_FK_MINB_DESCRIBE 5
#endif

// Rotated samples and the 512 tests of one keypoint (FREAK_feature_descriptor_parallel.cpp:110-151), given its angle.
__device__ __forceinline__ void describe_keypoint(const FreakArgs& a, int f, int q, float2 kp, float angle, const uint8_t* __restrict__ img, int pitch) {
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

// Split form (VO_FK_SPLIT=1, off): the descriptor half as its own kernel, one keypoint per thread, reading the angles the
// orientation kernel wrote - the orientation loop then runs at 64 registers and 32 warps per SM instead of 119 and 16.
// Measured (profiles/r02E_variants_freak_split.jsonl): 1.07 ms per 1000 KITTI frames against 0.648 ms for the single
// kernel, identical outputs - the descriptor phase hides behind other warps' orientation loops in the single kernel and
// is exposed on its own.  Kept as a build option only.
__global__ void __launch_bounds__(FK_NT, VO_FK_MINB_DESCRIBE) k_describe(const FreakArgs a) {
    const int f = blockIdx.y, q = blockIdx.x * FK_NT + threadIdx.x;
    if (q >= a.nkp[f]) return;
    describe_keypoint(a, f, q, a.kp[(size_t)f * a.kp_stride + q], a.angles[(size_t)f * a.kp_stride + q], a.blur + (size_t)f * a.blur_fstride,
                      (int)a.blur_pitch);
}

template <bool SPLIT>
__global__ void __launch_bounds__(FK_NT, SPLIT ? VO_FK_MINB_ORIENT : VO_FK_MINB) k_freak(const FreakArgs a) {
    // Unrotated samples in shared memory, per thread (one conflict-free vector row per warp access):
    //   s_p[p]   = sample p of keypoint 0 (low byte) and of keypoint 1 (high byte);
    //   s_win[w] = the six samples 1+6w .. 6+6w of keypoint 0 (bytes 0-5 of .x/.y) and of keypoint 1 (.z/.w).
    // A byte becomes a float with one PRMT: (byte, 0x00, 0x00, 0x4B) is the number 2^23 + byte, and the difference of
    // two numbers biased like that is the exact intensity difference, so the bias never has to be removed.
    // The second points r of every group of the padded pair table are exactly one such window (rows end at r = 42 and are
    // padded at the front to multiples of six), so a group costs one 16-byte and one 2-byte shared load per thread
    // where float samples cost 7 x 8 bytes (14 shared-memory wavefronts per warp instead of 5).  ncu of the float layout: the L1/shared data pipe was the busiest unit of the kernel (57 %).
    __shared__ unsigned short s_p[VO_FREAK_POINTS][FK_NT];
    __shared__ uint4 s_win[FK_WINDOWS][FK_NT];
    const int f = blockIdx.y;
    const int n = a.nkp[f];
    const int t = threadIdx.x;
    // keypoints of this thread: q0 and q0 + FK_NT (coalesced reads and writes per half); a thread whose second
    // keypoint does not exist carries a copy of its first through the arithmetic and stores nothing for it
    const int q0 = blockIdx.x * (FK_KP * FK_NT) + t;
    if (q0 >= n) return;
    const int q1 = q0 + FK_NT;
    const bool has1 = q1 < n;
    static_assert(FK_KP == 2, "the packed halves carry exactly two keypoints");
    const uint8_t* __restrict__ img = a.blur + (size_t)f * a.blur_fstride;
    const int pitch = (int)a.blur_pitch;
    const float2 kp0 = a.kp[(size_t)f * a.kp_stride + q0], kp1 = a.kp[(size_t)f * a.kp_stride + (has1 ? q1 : q0)];

    // unrotated samples (always in bounds: NMS keeps 37 <= x <= W-37, 35 <= y <= H-35)
    {
        const uint8_t* __restrict__ c0 = img + (size_t)(int)kp0.y * a.blur_pitch + (int)kp0.x;
        const uint8_t* __restrict__ c1 = img + (size_t)(int)kp1.y * a.blur_pitch + (int)kp1.x;
        uint32_t b0[VO_FREAK_POINTS], b1[VO_FREAK_POINTS];  // all 86 gathers in flight before the first is used
#pragma unroll
        for (int p = 0; p < VO_FREAK_POINTS; ++p) {
            const int off = (int)c_py[p] * pitch + (int)c_px[p];
            b0[p] = c0[off]; b1[p] = c1[off];
            s_p[p][t] = (unsigned short)(b0[p] | (b1[p] << 8));
        }
#pragma unroll
        for (int w = 0; w < FK_WINDOWS; ++w) {
            const int r = 1 + FK_GROUP * w;
            s_win[w][t] = make_uint4(b0[r] | (b0[r + 1] << 8) | (b0[r + 2] << 16) | (b0[r + 3] << 24), b0[r + 4] | (b0[r + 5] << 8),
                                     b1[r] | (b1[r + 1] << 8) | (b1[r + 2] << 16) | (b1[r + 3] << 24), b1[r + 4] | (b1[r + 5] << 8));
        }
    }

    // O += RN(ic * d / |d|) over the 903 pairs in the reference's order; the two keypoints of the thread packed.
    // The product of the packed multiply feeds only the addend of the FMA, and the FMA only the addend of
    // the sum, so ptxas has nothing to contract.
    // The loop runs over the 168 groups of six pairs, two per trip with ping-pong register sets: the weights
    // (constant bank, uniform datapath) and samples (shared memory) of the next group are requested before the
    // running group is evaluated, so their latency hides behind its instructions; without the look-ahead every
    // group stalled on its own constant loads (ncu: 43 % of the loop's samples on the short scoreboard).
    u64 ox2 = 0ull, oy2 = 0ull;  // (keypoint 0, keypoint 1)
    // 2^23 in a register the compiler cannot fold (rows is never negative): PRMT has one immediate slot, and it should
    // hold the selector - with a known constant here ptxas spends a MOV per PRMT on materialising selectors instead
    const uint32_t bias = 0x4B000000u | (static_cast<uint32_t>(a.rows) >> 31);
    uint4 va, vb;
    uint32_t vpa, vpb;
    auto load = [&](uint4& v, uint32_t& vp, int G) {
        const short2 g = c_grp[G];  // uniform: (first point p, window of the six second points)
        vp = s_p[g.x][t];
        v = s_win[g.y][t];
    };
    auto eval = [&](const uint4& v, uint32_t vpb2, int G) {
        const u64 vp = pk(__uint_as_float(__byte_perm(vpb2, bias, 0x7540u)), __uint_as_float(__byte_perm(vpb2, bias, 0x7541u)));
#pragma unroll
        for (int i = 0; i < FK_GROUP; ++i) {
            // uniform-datapath operands; pk(w, w) becomes the packed instructions' scalar-broadcast operand form
            // (SASS: FMUL2 R, R.F32x2.HI_LO, UR.F32), so the weights are neither duplicated nor moved
            const PairW w = c_w[G * FK_GROUP + i];
            const uint32_t sel = 0x7540u | (uint32_t)(i & 3);  // (sample byte, 0x00, 0x00, 0x4B) = 2^23 + sample
            const u64 s2 = pk(__uint_as_float(__byte_perm(i < 4 ? v.x : v.y, bias, sel)),
                              __uint_as_float(__byte_perm(i < 4 ? v.z : v.w, bias, sel)));
            const u64 ic2 = sub2(vp, s2);
            ox2 = add2(ox2, fma2(ic2, pk(w.whx, w.whx), mul2(ic2, pk(w.wlx, w.wlx))));
            oy2 = add2(oy2, fma2(ic2, pk(w.why, w.why), mul2(ic2, pk(w.wly, w.wly))));
        }
    };
    load(va, vpa, 0);
#pragma unroll 1
    for (int G = 0; G < FK_GROUPS; G += 2) {
        load(vb, vpb, G + 1);
        eval(va, vpa, G);
        load(va, vpa, G + 2);  // G + 2 == FK_GROUPS on the last trip: the dummy group
        eval(vb, vpb, G + 1);
    }
    float ox0, ox1, oy0, oy1;
    upk(ox2, ox0, ox1);
    upk(oy2, oy0, oy1);

#pragma unroll 1
    for (int k = 0; k < FK_KP; ++k) {
        if (k == 1 && !has1) break;
        const int q = k ? q1 : q0;
        const float2 kp = k ? kp1 : kp0;
        const float ox = k ? ox1 : ox0, oy = k ? oy1 : oy0;
        float angle = 0.0f;
        if (!(isnan(ox) || isnan(oy))) angle = (float)atan2((double)oy, (double)ox);
        a.angles[(size_t)f * a.kp_stride + q] = angle;
        if (!SPLIT) describe_keypoint(a, f, q, kp, angle, img, pitch);
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
        for (int g = 0; g < groups; ++g, ++gi) grp[gi] = make_short2((short)p, (short)((VO_FREAK_POINTS - FK_GROUP * groups + FK_GROUP * g - 1) / FK_GROUP));
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
    const dim3 grid((max_n + FK_KP * FK_NT - 1) / (FK_KP * FK_NT), n_frames);
    if (VO_FK_SPLIT) {
        k_freak<true><<<grid, FK_NT, 0, st>>>(a);
        k_describe<<<dim3((max_n + FK_NT - 1) / FK_NT, n_frames), FK_NT, 0, st>>>(a);
    } else {
        k_freak<false><<<grid, FK_NT, 0, st>>>(a);
    }
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
