// stencil.cu — K1: fused blur -> gradients -> 5x5 structure tensor -> Shi-Tomasi response, ONE pass.
//
// Replaces, for a whole batch of frames per launch:
//   cv::GaussianBlur(u8, 7x7, sigma 0)              feature_extraction_parallel/feature_extraction_parallel.cpp:223
//   CornerDetectionParallel::direction_gradients_worker        corner_detection_parallel.cpp:30-58
//   CornerDetectionParallel::shitomasi_corner_detection_worker corner_detection_parallel.cpp:61-105
// with whole-image (single strip) semantics, bit-exact (see "Exactness" below).
//
// Design (B200).  At 6 B/px the kernel is issue-bound long before it is HBM-bound, so every choice below
// buys instructions:
//   * one WARP owns a 128-pixel window (4 adjacent pixels per lane) that yields 116 output columns (6 halo
//     columns each side) and marches down the rows of its segment, FOUR rows per loop trip.  No block
//     barrier; horizontal neighbours come from warp shuffles, vertical history (raw rows, blurred rows,
//     gradient rows) from per-lane rings in shared memory that only their owner reads back;
//   * the four rows of a trip share their ring reads (10 raw rows feed 4 blurred rows, 10 blurred rows
//     feed 4 gradient and 4 response rows), and because the trip length divides the ring length every
//     ring slot is a compile-time offset from one of two base registers that swap each trip: no address
//     arithmetic in the loop;
//   * pixels are processed two per instruction wherever the arithmetic allows: 16-bit SWAR lanes for
//     the vertical blur taps, dp2a for the horizontal taps (weights pre-scaled so the blurred byte and
//     its fp16 form fall out of one byte permute), half2 for the gradients (integers <= 2040, exact in
//     fp16), and Blackwell's packed fp32 (FFMA2 / FADD2 / FMUL2) for the structure-tensor sums and the
//     eigenvalue;
//   * sum_5x5 Jxy telescopes to E[x-3]+E[x-2]-E[x+2]-E[x+3], E[x] = B[o-3][x]+B[o-2][x]-B[o+2][x]-B[o+3][x]
//     (Jxy is a difference of differences; BORDER_REFLECT_101 makes it vanish on the image border
//     exactly as the reference's zero-initialised border does);
//   * the sums of squares are running sums over the 5-row window in fp32, exact (integers < 2^23), updated with
//     Blackwell's mixed-precision FMA (fp16 factors from either register half, fp32 accumulator: SASS FHFMA);
//   * the eigenvalue is computed for every pixel, branch-free: an earlier version skipped warp rows whose
//     min(a, b) could not reach the threshold, but on textured frames nearly every row of 116 pixels holds
//     a corner-like pixel and the branch only kept the scheduler from overlapping the eigenvalue's
//     dependency chain with the next row's blur (measured: 4 % slower with the skip);
//   * input rows are requested one trip (four rows) ahead; image borders are mirrored with two lane
//     shuffles and a byte permute instead of gathers (edge strips only).
// HBM traffic per pixel: 1 B in, 1 B blurred out, 4 B R out.
//
// Exactness: everything up to a, b, c is integer arithmetic carried in exact containers.  The seven
// rounding operations of the eigenvalue are separate round-to-nearest operations.  ptxas (12.9) contracts
// mul.rn.f32x2 + add.rn.f32x2 into FFMA2 even with -fmad=false, so every product whose rounding matters
// (a*b, c*c, tr*tr) is a scalar operation; packed ops are used only where fusing cannot change the
// result (FMAs by the exact constants -1 / 4 / 0.5, sums, the sqrt refinement's intended FMAs).
#include <cuda_fp16.h>

#include <algorithm>
#include <cstdlib>
#include <type_traits>

#include "vo_internal.cuh"

namespace {

constexpr int SW = 116;    // output columns per warp strip: the 128-pixel window minus 6 halo columns each side
constexpr int HX = 8;      // window start relative to strip*SW (6 halo columns needed; 8 keeps the 4-pixel groups aligned,
                           // so lane 1 stores only its upper two pixels and lane 30 only its lower two)
// Fused-NMS mode (the response never leaves the SM): strips advance by 112 columns and start at strip*112 - 4.
// Window columns 3..124 hold valid blurred pixels, 6..121 valid responses, and the NMS centres a strip owns are its
// lanes 2..29 (window columns 8..119): their k x k windows (k <= 5) stay inside columns 6..121.  Blurred pixels are
// stored by lanes 1..28 (whole 4-pixel groups only); the last strip also stores lanes 29..31 up to the image edge.
constexpr int SWF = 112;
constexpr int HXF = 4;
// One warp per CTA, 20 CTAs per SM (96 registers): warps are independent tasks, so single-warp CTAs let the block
// scheduler spread a partly filled wave evenly over the SMs (550 four-warp CTAs left some SMs with 16 and some with
// 12 warps).  Measured per 1000 KITTI frames at batch 200: 4x4 1.267 ms, 1x16 1.312, 3x6 1.277, 1x18 1.212, 1x20 1.200.
#ifndef VO_NWARP
#define VO_NWARP 1
#endif
constexpr int NWARP = VO_NWARP;   // warps per CTA (independent tasks)
constexpr int RB = 4;      // rows per loop trip; must divide the ring length 8
#ifndef VO_CTAS
#define VO_CTAS 20
#endif
constexpr int CTAS_PER_SM = VO_CTAS;
#ifndef VO_CTAS_FUSED
#define VO_CTAS_FUSED 17
#endif
constexpr int CTAS_PER_SM_FUSED = VO_CTAS_FUSED;  // 12.4 KB of rings per warp (+ 1 KB the system reserves per CTA)
constexpr unsigned FULL = 0xFFFFFFFFu;
typedef unsigned long long u64;

__device__ __noinline__ int reflect101_any(int p, int n) {  // images smaller than the halo: repeated reflection
    if (n == 1) return 0;
    while (p < 0 || p >= n) p = p < 0 ? -p : 2 * n - 2 - p;
    return p;
}
// BORDER_REFLECT_101 index; one reflection per side covers every halo position of images >= 9 px
__device__ __forceinline__ int reflect101(int p, int n) {
    int q = p < 0 ? -p : p;
    q = q >= n ? 2 * n - 2 - q : q;
    if (static_cast<unsigned>(q) >= static_cast<unsigned>(n)) q = reflect101_any(p, n);
    return q;
}

__device__ __forceinline__ __half2 u2h(uint32_t u) { return *reinterpret_cast<__half2*>(&u); }
__device__ __forceinline__ uint32_t h2u(__half2 h) { return *reinterpret_cast<uint32_t*>(&h); }
__device__ __forceinline__ __half2 prmt_h2(__half2 a, __half2 b, uint32_t sel) { return u2h(__byte_perm(h2u(a), h2u(b), sel)); }

// packed fp32 pairs (sm_100: FFMA2 / FADD2 / FMUL2), IEEE round-to-nearest per element
__device__ __forceinline__ u64 pk(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk(u64 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ u64 sub2(u64 a, u64 b) { u64 d; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ u64 shfl_up2(u64 v) {
    float lo, hi; upk(v, lo, hi);
    return pk(__shfl_up_sync(FULL, lo, 1), __shfl_up_sync(FULL, hi, 1));
}
__device__ __forceinline__ u64 shfl_dn2(u64 v) {
    float lo, hi; upk(v, lo, hi);
    return pk(__shfl_down_sync(FULL, lo, 1), __shfl_down_sync(FULL, hi, 1));
}
// Blackwell mixed-precision FMA (FHFMA): fp16 multiplicands taken from either half of a half2, fp32 addend and
// result, ONE rounding to fp32.  The products here are squares of integers <= 2040 and the sums stay below
// 2^24, so every result is exact; what it buys is the fp16 -> fp32 conversions and the operand packing.
__device__ __forceinline__ float fhfma_lo(__half2 v, float c) {
    float d; asm("{.reg .b16 l, h; mov.b32 {l, h}, %1; fma.rn.f32.f16 %0, l, l, %2;}" : "=f"(d) : "r"(h2u(v)), "f"(c)); return d;
}
__device__ __forceinline__ float fhfma_hi(__half2 v, float c) {
    float d; asm("{.reg .b16 l, h; mov.b32 {l, h}, %1; fma.rn.f32.f16 %0, h, h, %2;}" : "=f"(d) : "r"(h2u(v)), "f"(c)); return d;
}
__device__ __forceinline__ float fhfnma_lo(__half2 v, float c) {  // c - v.lo^2
    float d; asm("{.reg .b16 l, h, n; mov.b32 {l, h}, %1; neg.f16 n, l; fma.rn.f32.f16 %0, n, l, %2;}" : "=f"(d) : "r"(h2u(v)), "f"(c)); return d;
}
__device__ __forceinline__ float fhfnma_hi(__half2 v, float c) {  // c - v.hi^2
    float d; asm("{.reg .b16 l, h, n; mov.b32 {l, h}, %1; neg.f16 n, h; fma.rn.f32.f16 %0, n, h, %2;}" : "=f"(d) : "r"(h2u(v)), "f"(c)); return d;
}
__device__ __forceinline__ float rsqrt_approx(float x) { float y; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

// Four pixels of R = tr/2 - 0.5*sqrt(tr^2 - 4 det), det = a*b - c*c  (corner_detection_parallel.cpp:93-98).
// Every operation rounds once to nearest like the reference's scalar code (its final subtraction in
// double followed by narrowing equals one fp32 rounding: 53 >= 2*24+2).  The three rounded products are
// scalar (ptxas would contract a packed mul + add pair); the sums that follow them are FMAs by an exact
// constant (x*(-1), x*4, x*0.5 are exact, so fma(x, k, y) rounds once, like the reference's add), which ptxas
// cannot contract any further, and they run packed, two pixels per instruction.
// Signs: the chain carries nd = -disc, ns = -s, ne = -e, nsq = -sqrt: round-to-nearest is sign-symmetric, so
// each value is exactly the negative of the reference's and no negation is ever issued.
// disc is an integer-valued float: 0, <= -1 or >= 1.  sqrt: the correctly rounded fast path of sqrtf (rsqrt
// seed, one refinement), valid for every disc >= 1 that can occur here (< 2^50).  The seed is taken at
// disc + 1e-30 (== disc for disc >= 1): disc == 0 gives exactly 0, and disc < 0 gives NaN, which fails the
// threshold comparison exactly like the reference's NaN does.
// Pixel order: A/B index k = pixel k; cE = (c0, c2), cO = (c1, c3) as fp16 (integers <= 2040, exact).
// RAW2 (fused NMS): out = tr - sqrt(disc) = 2R exactly (halving is exact, so RN(tr/2 - sqrt/2) = RN(tr - sqrt) / 2), NOT
// thresholded: the response ring then holds doubled raw responses, the non-zero test and the window test compare
// against the doubled threshold, and the half is taken when a candidate's key is written.  That removes a packed
// multiply per pixel pair and a compare + select per pixel from the steady loop.
template <bool RAW2 = false>
__device__ __forceinline__ void shi_tomasi4(const float (&A)[4], const float (&B)[4], __half2 cE, __half2 cO, const float (&thr)[4],
                                            float (&out)[4]) {
    float cc[4], ab[4], t[4], ntt[4], dz[4];
    cc[0] = fhfma_lo(cE, 0.f); cc[1] = fhfma_lo(cO, 0.f); cc[2] = fhfma_hi(cE, 0.f); cc[3] = fhfma_hi(cO, 0.f);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        ab[k] = __fmul_rn(A[k], B[k]);
        t[k] = __fadd_rn(A[k], B[k]);  // exact (integers < 2^24)
        ntt[k] = __fmul_rn(t[k], -t[k]);
    }
    const u64 m1 = pk(-1.f, -1.f), p4 = pk(4.f, 4.f), half = pk(0.5f, 0.5f);
#pragma unroll
    for (int q = 0; q < 4; q += 2) {
        const u64 det = fma2(pk(cc[q], cc[q + 1]), m1, pk(ab[q], ab[q + 1]));
        const u64 nd = fma2(det, p4, pk(ntt[q], ntt[q + 1]));  // -(tr^2 - 4 det)
        float n0, n1; upk(nd, n0, n1);
        dz[q] = __fsub_rn(1e-30f, n0); dz[q + 1] = __fsub_rn(1e-30f, n1);
        const u64 y = pk(rsqrt_approx(dz[q]), rsqrt_approx(dz[q + 1]));
        const u64 ns = mul2(nd, y), h = mul2(y, half);
        const u64 ne = fma2(ns, ns, nd);
        const u64 nsq = fma2(ne, h, ns);
        if (RAW2) {
            const u64 r = add2(pk(t[q], t[q + 1]), nsq);
            upk(r, out[q], out[q + 1]);
        } else {
            const u64 r = fma2(nsq, half, mul2(pk(t[q], t[q + 1]), half));
            float r0, r1; upk(r, r0, r1);
            out[q] = r0 > thr[q] ? r0 : 0.0f;
            out[q + 1] = r1 > thr[q + 1] ? r1 : 0.0f;
        }
    }
}

// Harris response of four pixels: R = det - k * tr^2, det = a*b - c*c, every operation rounded once to fp32 in this order
// (feature_extraction/corner_detection.cpp:196-199 evaluates the same expression in double; see vofront.h).
__device__ __forceinline__ void harris4(const float (&A)[4], const float (&B)[4], __half2 cE, __half2 cO, const float (&thr)[4], float k,
                                        float (&out)[4]) {
    float cc[4];
    cc[0] = fhfma_lo(cE, 0.f); cc[1] = fhfma_lo(cO, 0.f); cc[2] = fhfma_hi(cE, 0.f); cc[3] = fhfma_hi(cO, 0.f);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const float det = __fsub_rn(__fmul_rn(A[q], B[q]), cc[q]);
        const float t = __fadd_rn(A[q], B[q]);
        const float r = __fsub_rn(det, __fmul_rn(__fmul_rn(k, t), t));  // (k * trace) * trace, left to right
        out[q] = r > thr[q] ? r : 0.0f;
    }
}

// Global stores through an opaque address (the row address is computed once for the three store forms; the opaque
// register would otherwise turn them into generic-space stores).
__device__ __forceinline__ void stg_u32(const void* p, uint32_t v) { asm volatile("st.global.b32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ void stg_u16(const void* p, uint32_t v) { asm volatile("st.global.b16 [%0], %1;" ::"l"(p), "h"(static_cast<unsigned short>(v)) : "memory"); }
__device__ __forceinline__ void stg_f2(const void* p, float a, float b) { asm volatile("st.global.v2.f32 [%0], {%1, %2};" ::"l"(p), "f"(a), "f"(b) : "memory"); }
__device__ __forceinline__ void stg_f4(const void* p, float a, float b, float c, float d) {
    asm volatile("st.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// ---- per-lane history rings ------------------------------------------------------------------------------
// Layout per warp (bytes): raw rows [8][32] uint2 at 0, blurred rows [8][32] uint2 at 2048, gradient rows
// [8][32] uint4 at 4096.  Each lane reads back only what it wrote itself; every warp access is one
// conflict-free vector row.  Slot of step t is t & 7; with t = t0 + c, t0 a multiple of 4, that is
// (c & 7) ^ (t0 & 4): slots 0-3 relative to one base pointer and 4-7 relative to the other, the two bases
// swapping every trip.
constexpr int RING_BYTES = 8192;
// Fused-NMS mode adds a ring of the last eight thresholded response rows, [8][32] float4 at 8192 (every lane may read
// every lane's pixels: the window test of a compacted item needs its neighbours' columns), and the item list of a trip
// (one byte per non-zero 4-pixel group: row-in-trip * 32 + lane) at 12288.
constexpr int RRING_OFF = 8192, ITEM_OFF = 12288;
// ... and a staging buffer of CBUF candidate keys at 12416: candidates are appended to the frame's list in batches, one
// global atomic per flush instead of one per trip (the warp waits for the atomic's return value to place its keys: with
// an atomic per trip that wait was 7 % of the kernel's warp-state samples, ncu source page of capture r02o).
constexpr int CBUF = 32, CBUF_OFF = 12416;
constexpr int RING_BYTES_FUSED = 12416 + CBUF * 8;
__device__ __forceinline__ const char* slot_of(const char* pa, const char* pb, int c, int row_bytes) {
    const int s = c & 7;
    return ((s & 4) ? pb : pa) + (s & 3) * row_bytes;
}
__device__ __forceinline__ uint2 lds2(const char* p) { return *reinterpret_cast<const uint2*>(p); }
__device__ __forceinline__ uint4 lds4(const char* p) { return *reinterpret_cast<const uint4*>(p); }
__device__ __forceinline__ void sts2(const char* p, uint2 v) { *reinterpret_cast<uint2*>(const_cast<char*>(p)) = v; }
__device__ __forceinline__ void sts4(const char* p, uint4 v) { *reinterpret_cast<uint4*>(const_cast<char*>(p)) = v; }

// GATHER: every lane fetches its four (reflected) bytes itself, synchronously: any image size, used for
// images too small for the mirror scheme.  ALIGNED: rows start on 4-byte boundaries (one load per row).
// NMSH = 0: the thresholded response is written to a.R (vo_response, VO_DBG_RESPONSE, windows 7 and 9, tiny images).
// NMSH = 1, 2: fused non-maximum suppression with half window NMSH (corner_detection_parallel.cpp:166-208): the response
// stays in the warp's shared-memory ring and only candidate keys leave the SM (2 B/px + 8 B per candidate).
// SRC = 0: gray input; 1: blurred input (vo_response); 2: a thresholded response map is streamed through the fused NMS
// instead of being computed (vo_nms with windows 3 / 5: the suppression code under test is the production code).
template <int SRC, bool GATHER, bool ALIGNED, int NMSH, bool HARRIS = false>
__global__ void __launch_bounds__(32 * NWARP, NMSH ? CTAS_PER_SM_FUSED : CTAS_PER_SM)
k_stencil(const StencilArgs a, int n_strips, int n_segs, int seg_rows, int n_tasks) {
    constexpr bool FUSED = NMSH > 0, FROM_BLURRED = SRC == 1, FROM_R = SRC == 2;
    static_assert(!HARRIS || !FUSED, "the Harris response takes the response-writing form");
    static_assert(!FROM_R || FUSED, "a response map can only feed the fused NMS");
    __shared__ __align__(16) char s_rings[NWARP][FUSED ? RING_BYTES_FUSED : RING_BYTES];
    const int lane = threadIdx.x & 31;
    // tasks = (frame, row segment, strip), numbered across the whole batch so that no warp slot of a CTA idles
    char* const ring = s_rings[threadIdx.x >> 5];
    int gtask = blockIdx.x * NWARP + (threadIdx.x >> 5);
    for (;;) {  // one task per warp, or - persistent form - tasks from a counter until none is left
    if (a.task_counter) {
        int t = 0;
        if (lane == 0) t = atomicAdd(a.task_counter, 1);
        gtask = __shfl_sync(FULL, t, 0);
    }
    if (gtask >= n_tasks) return;  // whole warp
    const int per_frame = n_strips * n_segs;
    const int f = gtask / per_frame, task = gtask - f * per_frame;
    const int strip = task % n_strips, seg = task / n_strips;
    const int H = a.rows, W = a.cols;
    // rows this task owns (the NMS centres it decides) and rows it computes: a fused segment also computes the four
    // response rows either side of its own (its windows reach NMSH <= 2 rows across the border).  Blurred rows are stored
    // for every COMPUTED row (the store guard below is [ys, ye), and every steady trip), so the two tasks either side of
    // a seam both write its 8 halo rows - the same bytes, a benign write-write overlap that costs 8 / seg_rows extra
    // store requests.  The 2 B/px quoted for the fused form holds where it is quoted: 1000 KITTI frames per launch
    // plan ONE segment per strip (launch_stencil: best_segs = 1), so there is no seam; small batches (several segments
    // of >= 48 rows) pay up to 1/6 more blurred stores.
    const int ys_own = seg * seg_rows, ye_own = min(H, ys_own + seg_rows);
    const int ys = FUSED ? max(0, ys_own - RB) : ys_own, ye = FUSED ? min(H, ye_own + RB) : ye_own;
    const int xs = FUSED ? strip * SWF - HXF : strip * SW - HX;
    const int x = xs + 4 * lane;  // this lane's columns: x .. x+3
    // unfused: outputs of this strip are columns xs+6 .. xs+121: lanes 2..29 store their four pixels, lane 1 its upper
    // pair, lane 30 its lower pair.  Fused: blurred pixels only, whole groups, lanes 1..28 (to the image edge on the last strip)
    const bool st_full = FUSED ? (lane >= 1 && (lane <= 28 || strip == n_strips - 1) && x < W) : (lane >= 2 && lane < 30 && x < W);
    const bool st_l1 = !FUSED && lane == 1 && x + 2 >= 0 && x + 2 < W, st_l30 = !FUSED && lane == 30 && x < W;
    const uint8_t* __restrict__ in = a.in + (size_t)f * a.in_fstride;
    uint8_t* __restrict__ blur_out = (FROM_BLURRED || FROM_R) ? nullptr : a.blur + (size_t)f * a.blur_fstride;
    float* __restrict__ R = (FUSED && !FROM_R) ? nullptr : a.R + (size_t)f * a.R_fstride;
    const float thr = a.threshold;
    const float kInf = __int_as_float(0x7f800000);

    // gradients are zero outside 1 <= j <= W-2 (cv::Mat::zeros border); packing: E = (px0, px2), O = (px1, px3)
    const __half2 mE = __floats2half2_rn((x >= 1 && x <= W - 2) ? 1.f : 0.f, (x + 2 >= 1 && x + 2 <= W - 2) ? 1.f : 0.f);
    const __half2 mO = __floats2half2_rn((x + 1 >= 1 && x + 1 <= W - 2) ? 1.f : 0.f, (x + 3 >= 1 && x + 3 <= W - 2) ? 1.f : 0.f);
    const bool edge_strip = (xs < 1) || (xs + 128 > W - 2);  // warp-uniform: the 128-pixel window touches a border
    // response columns < 2 and > W-3 stay zero
    const bool okc0 = x >= 2 && x <= W - 3, okc1 = x + 1 >= 2 && x + 1 <= W - 3, okc2 = x + 2 >= 2 && x + 2 <= W - 3,
               okc3 = x + 3 >= 2 && x + 3 <= W - 3;
    const float thr4[4] = {okc0 ? thr : kInf, okc1 ? thr : kInf, okc2 ? thr : kInf, okc3 ? thr : kInf};  // nothing exceeds +inf

    // ---- input access ----------------------------------------------------------------------------------
    // Lanes whose first pixel is inside the row read the aligned word(s) holding x..x+3 (up to 7 bytes past
    // the row end: every input buffer of the library has that slack) and funnel-shift them a trip later.
    // Pixels outside the row are BORDER_REFLECT_101 copies of pixels held by at most two other lanes of the
    // same warp: two indexed shuffles and one byte permute rebuild them (edge strips only).
    const bool inside = x >= 0 && x < W;
    int srcA = lane, srcB = lane;
    uint32_t msel = 0x3210u;
    int cx0 = 0, cx1 = 0, cx2 = 0, cx3 = 0;
    if (GATHER) {
        cx0 = reflect101(x, W); cx1 = reflect101(x + 1, W); cx2 = reflect101(x + 2, W); cx3 = reflect101(x + 3, W);
    } else if (edge_strip) {
        int q[4], la = 31;
#pragma unroll
        for (int k = 0; k < 4; ++k) {  // pixels more than 6 beyond the border are never used: clamp them
            const int p = x + k;
            int r = p < 0 ? -p : (p >= W ? 2 * W - 2 - p : p);
            r = min(max(r, 0), W - 1);
            q[k] = max(r - xs, 0);  // position inside the strip's 128-pixel window
            la = min(la, min(q[k] >> 2, 31));
        }
        msel = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) msel |= static_cast<uint32_t>(min(max(q[k] - 4 * la, 0), 7)) << (4 * k);
        srcA = la; srcB = min(la + 1, 31);
    }
    uint32_t pa[RB], pb[RB], psh[RB];  // in-flight input rows
    const int xin = inside ? x : 0;
    const long long in_pitch = (long long)a.in_pitch;
    const uint8_t* in_next = in + xin;  // this lane's word of the next row to request (valid while that row is interior)
    auto issue = [&](int row, bool known_interior) {  // rows row .. row+3
        const bool interior = known_interior || (row >= 0 && row + RB <= H);  // warp-uniform: no reflection needed
        const uint8_t* const p0 = known_interior ? in_next : in + (long long)row * in_pitch + xin;
#pragma unroll
        for (int j = 0; j < RB; ++j) {
            const uint8_t* p = p0 + j * in_pitch;
            if (!interior) p = in + (long long)reflect101(row + j, H) * in_pitch + xin;
            if (GATHER) {
                p -= xin;
                pa[j] = p[cx0] | (p[cx1] << 8) | (p[cx2] << 16) | (p[cx3] << 24);
            } else if (ALIGNED) {
                pa[j] = __ldg(reinterpret_cast<const uint32_t*>(p));
            } else {
                const uintptr_t q = reinterpret_cast<uintptr_t>(p);
                const uint32_t* w = reinterpret_cast<const uint32_t*>(q & ~static_cast<uintptr_t>(3));
                pa[j] = __ldg(w); pb[j] = __ldg(w + 1); psh[j] = static_cast<uint32_t>(q & 3) * 8u;
            }
        }
    };

    // ---- ring pointers -----------------------------------------------------------------------------------
    const char* qa = ring + lane * 8;           // raw / blurred rings (blurred at +2048), slots 0-3
    const char* qb = qa + 4 * 256;              //                                         slots 4-7
    const char* ja = ring + 4096 + lane * 16;   // gradient ring
    const char* jb = ja + 4 * 512;
#pragma unroll
    for (int k = 0; k < 4; ++k) {  // rows before the first computed one count as zero
        sts2(qa + k * 256, make_uint2(0u, 0u)); sts2(qb + k * 256, make_uint2(0u, 0u));
        sts2(qa + 2048 + k * 256, make_uint2(0u, 0u)); sts2(qb + 2048 + k * 256, make_uint2(0u, 0u));
        sts4(ja + k * 512, make_uint4(0u, 0u, 0u, 0u)); sts4(jb + k * 512, make_uint4(0u, 0u, 0u, 0u));
    }
    float sa[4] = {0.f, 0.f, 0.f, 0.f}, sb[4] = {0.f, 0.f, 0.f, 0.f};  // running 5-row sums of Jx^2, Jy^2 per pixel
    // fused NMS state: the response ring (slot of step t = t & 7, absolute), the non-zero masks of the last NMSH rows of
    // the previous trip, the centre rows / columns this task decides and the frame's candidate list
    char* const rring = ring + RRING_OFF;
    unsigned char* const s_item = reinterpret_cast<unsigned char*>(ring + ITEM_OFF);
    unsigned long long* __restrict__ cand = FUSED ? a.cand + (size_t)f * a.cand_cap : nullptr;
    unsigned long long* const s_cbuf = reinterpret_cast<unsigned long long*>(ring + CBUF_OFF);
    int cbuf_n = 0;  // keys staged in s_cbuf (warp-uniform)
    auto flush_candidates = [&]() {  // whole warp
        if (cbuf_n == 0) return;
        __syncwarp();
        int pos = 0;
        if (lane == 0) pos = atomicAdd(a.cand_count + f, cbuf_n);
        pos = __shfl_sync(FULL, pos, 0) + lane;
        if (lane < cbuf_n && pos < a.cand_cap) cand[pos] = s_cbuf[lane];
        __syncwarp();
        cbuf_n = 0;
    };
    unsigned nz_prev[2] = {0u, 0u};
    // ring scale and threshold of the window test: computed rows are doubled raw responses against 2 * threshold; a
    // streamed response map (FROM_R) is already thresholded: plain values against 0
    const float nms_thr = FROM_R ? 0.f : 2.0f * thr, key_scale = FROM_R ? 1.0f : 0.5f;
    const int nms_i_lo = max(max(NMSH, 35), ys_own), nms_i_hi = min(min(H - NMSH - 1, H - 35), ye_own - 1);
    const int nms_j_lo = max(NMSH, 37), nms_j_hi = min(W - NMSH - 1, W - 37);
    if (FUSED) {
#pragma unroll
        for (int k = 0; k < 8; ++k) sts4(rring + k * 512 + lane * 16, make_uint4(0u, 0u, 0u, 0u));
    }

    const __half2 hz = __floats2half2_rn(0.f, 0.f), two = __floats2half2_rn(2.f, 2.f), k1024 = __floats2half2_rn(1024.f, 1024.f);

    // Step t handles input row r = ys-6+t (blur mode) or blurred row r-3 (from-blurred mode); after it
    //   B row r-3 is complete, gradient row g = r-4, response row o = r-6.  Ring slot of step t: t & 7.
    const int T = (ye - ys) + 12;
    const int r0 = ys - 6;
    const int load_off = FROM_BLURRED ? -3 : 0;
    if (!FROM_R) issue(r0 + load_off, false);
    in_next += (long long)(r0 + RB + load_off) * in_pitch;  // trip t0 requests rows r0+t0+RB+load_off ..
    // output row pointers of step t0 (this lane's column), advanced by RB rows per trip
    const long long blur_pitch = (long long)a.blur_pitch, R_pitch = (long long)a.R_pitch;
    uint8_t* blur_p = (FROM_BLURRED || FROM_R) ? nullptr : blur_out + (long long)(r0 - 3) * blur_pitch + x;
    float* R_p = FUSED ? nullptr : R + (long long)(r0 - 6) * R_pitch + x;

    // ---- fused NMS: one trip's centre rows -------------------------------------------------------------------
    // Called after the response rows of steps t0 .. t0+3 are in the ring (nz[j] = lanes whose group of row j is not all
    // zero).  The centre rows o0-h .. o0+3-h have complete windows now.  A zero centre can never be a strict maximum, so
    // only non-zero groups are tested: the warp compacts them (ballot masks -> byte list in shared memory) and hands
    // them out one per lane, like k_nms does from global memory.
    auto nms_trip = [&](const int t0, const unsigned (&nz)[RB]) {
        // ---- 5. non-maximum suppression of the centre rows o0-h .. o0+3-h (their windows are complete now) ---
        constexpr int h = NMSH > 0 ? NMSH : 1;
        const int o0 = r0 + t0 - 6;  // response row of step t0
        unsigned M[RB];              // non-zero groups per centre row: lanes 2..29 own the centres
#pragma unroll
        for (int k = 0; k < RB; ++k) {
            const unsigned m = k < h ? nz_prev[2 - h + k] : nz[k - h];
            const int i = o0 - h + k;
            M[k] = (i >= nms_i_lo && i <= nms_i_hi) ? (m & 0x3FFFFFFCu) : 0u;
        }
        nz_prev[0] = nz[2]; nz_prev[1] = nz[3];
        const int total = __popc(M[0]) + __popc(M[1]) + __popc(M[2]) + __popc(M[3]);
        if (total > 0) {  // warp-uniform
            // compact the non-zero groups into the item list, one byte each: centre row k * 32 + lane
            const unsigned lt = (1u << lane) - 1u;
            int cum = 0;
#pragma unroll
            for (int k = 0; k < RB; ++k) {
                if ((M[k] >> lane) & 1u) s_item[cum + __popc(M[k] & lt)] = (unsigned char)(k * 32 + lane);
                cum += __popc(M[k]);
            }
            __syncwarp();  // ring rows and item list written by other lanes are visible from here on
            for (int base = 0; base < total; base += 32) {
                const int e = base + lane;
                const bool live = e < total;
                const int item = live ? (int)s_item[e] : 2;  // idle lanes test a harmless in-range item
                const int k = item >> 5, l = item & 31;
                // window rows = steps t0 + k - 2h .. t0 + k, columns 4l - h .. 4l + 3 + h (contiguous in the ring)
                float m4[4] = {0.f, 0.f, 0.f, 0.f}, cv[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int d = 0; d <= 2 * h; ++d) {
                    const char* rowp = rring + ((t0 + k - 2 * h + d) & 7) * 512 + l * 16;
                    float w[4 + 2 * h];
                    const float4 c4 = *reinterpret_cast<const float4*>(rowp);
                    w[h] = c4.x; w[h + 1] = c4.y; w[h + 2] = c4.z; w[h + 3] = c4.w;
                    if (h == 1) {
                        w[0] = *reinterpret_cast<const float*>(rowp - 4);
                        w[5] = *reinterpret_cast<const float*>(rowp + 16);
                    } else {
                        const float2 l2 = *reinterpret_cast<const float2*>(rowp - 8), r2 = *reinterpret_cast<const float2*>(rowp + 16);
                        w[0] = l2.x; w[1] = l2.y; w[6] = r2.x; w[7] = r2.y;
                    }
                    if (d == h) { cv[0] = c4.x; cv[1] = c4.y; cv[2] = c4.z; cv[3] = c4.w; }
#pragma unroll
                    for (int e4 = 0; e4 < 4; ++e4)
#pragma unroll
                        for (int dm = -h; dm <= h; ++dm)
                            if (!(d == h && dm == 0)) m4[e4] = fmaxf(m4[e4], w[h + e4 + dm]);
                }
                // strict maximum of its window (no neighbour >= centre) inside the FREAK margins; two of the four
                // pixels of a group at most (neighbouring pixels cannot both be strict maxima)
                const int i = o0 - h + k, j0 = xs + 4 * l;
                unsigned long long key[2] = {0ull, 0ull};
                int nc = 0;
#pragma unroll
                for (int e4 = 0; e4 < 4; ++e4) {
                    const int j = j0 + e4;
                    // above the threshold and a strict maximum of its window: neighbours at or below the threshold count as zero
                    // in the reference, and they are below this centre anyway
                    if (live && cv[e4] > nms_thr && cv[e4] > m4[e4] && j >= nms_j_lo && j <= nms_j_hi) {
                        const unsigned long long kk = ((unsigned long long)__float_as_uint(cv[e4] * key_scale) << 32) | ((unsigned long long)i << 16) | (unsigned long long)j;
                        if (nc == 0) key[0] = kk; else key[1] = kk;
                        ++nc;
                    }
                }
                const unsigned b1 = __ballot_sync(FULL, nc >= 1), b2 = __ballot_sync(FULL, nc >= 2);
                const int n_new = __popc(b1) + __popc(b2);
                if (n_new > 0) {  // warp-uniform
                    const int rel = __popc(b1 & lt) + __popc(b2 & lt);
                    if (cbuf_n + n_new > CBUF) flush_candidates();
                    if (n_new > CBUF) {  // a dense round: straight to the frame's list
                        int pos = 0;
                        if (lane == 0) pos = atomicAdd(a.cand_count + f, n_new);
                        pos = __shfl_sync(FULL, pos, 0) + rel;
                        if (nc >= 1 && pos < a.cand_cap) cand[pos] = key[0];
                        if (nc >= 2 && pos + 1 < a.cand_cap) cand[pos + 1] = key[1];
                    } else {
                        if (nc >= 1) s_cbuf[cbuf_n + rel] = key[0];
                        if (nc >= 2) s_cbuf[cbuf_n + rel + 1] = key[1];
                        cbuf_n += n_new;
                    }
                }
            }
            __syncwarp();  // the next trip overwrites ring slots and the item list read above
        }
    };

    // One trip = four steps.  STEADY trips (all four blurred / gradient / response rows inside the segment and the
    // image, a further trip to request) carry no row tests; the first four and the last one or two trips of a
    // segment run the general form.
    auto trip = [&](auto steady_tag, const int t0) {
        constexpr bool STEADY = decltype(steady_tag)::value;
        // (a third body specialised for interior strips was tried: 8 % SLOWER, the two steady bodies together no longer
        // fit the instruction cache; so the edge handling stays predicated on the warp-uniform edge_strip)
        const bool edge = edge_strip;
        unsigned nz[RB] = {0u, 0u, 0u, 0u};                     // fused: lanes whose four responses of row j are not all zero
        char* const rrow = rring + (t0 & 4) * 512 + lane * 16;  // fused: this lane's ring slots of steps t0 .. t0+3
        if (FROM_R) {  // the response rows of this trip come from a.R (zero outside the image)
            if (t0 >= 12) {
#pragma unroll
                for (int j = 0; j < RB; ++j) {
                    const int orow = r0 + t0 + j - 6;
                    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (orow >= 0 && orow < H && x >= 0 && x < W) {
                        v = __ldg(reinterpret_cast<const float4*>(R + (long long)orow * R_pitch + x));
                        if (x + 1 >= W) v.y = 0.f;
                        if (x + 2 >= W) v.z = 0.f;
                        if (x + 3 >= W) v.w = 0.f;
                    }
                    sts4(rrow + j * 512, make_uint4(__float_as_uint(v.x), __float_as_uint(v.y), __float_as_uint(v.z), __float_as_uint(v.w)));
                    nz[j] = __ballot_sync(FULL, fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w)) > 0.f);
                }
                nms_trip(t0, nz);
            }
            return;
        }
        // ---- 0. the four rows requested a trip ago; request the next four -----------------------------------
        uint32_t w[RB];
#pragma unroll
        for (int j = 0; j < RB; ++j) w[j] = (GATHER || ALIGNED) ? pa[j] : __funnelshift_r(pa[j], pb[j], psh[j]);
        if (!GATHER && edge) {
#pragma unroll
            for (int j = 0; j < RB; ++j)
                w[j] = __byte_perm(__shfl_sync(FULL, w[j], srcA), __shfl_sync(FULL, w[j], srcB), msel);
        }
        if (STEADY || t0 + RB < T) issue(r0 + t0 + RB + load_off, STEADY);

        // ---- 1. blurred rows r-3 for the four steps ----------------------------------------------------------
        // index i of the local windows = step offset c + 6, c in [-6, 3]
        __half2 bE[10], bO[10];
#pragma unroll
        for (int c = -6; c < 0; ++c) {
            const uint2 v = lds2(slot_of(qa, qb, c, 256) + 2048);
            bE[c + 6] = u2h(v.x); bO[c + 6] = u2h(v.y);
        }
        if (!FROM_BLURRED) {
            uint32_t rE[10], rO[10];
#pragma unroll
            for (int c = -6; c < 0; ++c) {
                const uint2 v = lds2(slot_of(qa, qb, c, 256));
                rE[c + 6] = v.x; rO[c + 6] = v.y;
            }
#pragma unroll
            for (int j = 0; j < RB; ++j) {
                rE[6 + j] = __byte_perm(w[j], 0u, 0x4240);  // (px0, px2) in 16-bit lanes
                rO[6 + j] = __byte_perm(w[j], 0u, 0x4341);  // (px1, px3)
                sts2(slot_of(qa, qb, j, 256), make_uint2(rE[6 + j], rO[6 + j]));
                // vertical taps 4*[2,7,14,18,14,7,2] on rows r-6..r; 16-bit lanes never exceed 4*64*255 = 65280
                const uint32_t VE = 8u * (rE[j] + rE[j + 6]) + 28u * (rE[j + 1] + rE[j + 5]) + 56u * (rE[j + 2] + rE[j + 4]) + 72u * rE[j + 3];
                const uint32_t VO = 8u * (rO[j] + rO[j + 6]) + 28u * (rO[j + 1] + rO[j + 5]) + 56u * (rO[j + 2] + rO[j + 4]) + 72u * rO[j + 3];
                // VE = (V0, V2), VO = (V1, V3); neighbours' columns through shuffles
                const uint32_t LE = __shfl_up_sync(FULL, VE, 1), LO = __shfl_up_sync(FULL, VO, 1);        // (V-4,V-2) (V-3,V-1)
                const uint32_t RE_ = __shfl_down_sync(FULL, VE, 1), RO_ = __shfl_down_sync(FULL, VO, 1);  // (V4,V6) (V5,V7)
                // horizontal taps 4*[2,7,14,18,14,7,2]: __dp2a_lo(a, b, c) = c + a.lo16*b.byte0 + a.hi16*b.byte1.
                // S = 16*(sum + 2048) + 0x64000000 < 2^24 + 0x64000000: byte 2 is the blurred pixel and
                // bytes 3:2 are the fp16 number 1024 + pixel.
                constexpr uint32_t S_INIT = 16u * 2048u + 0x64000000u;
                uint32_t S0 = S_INIT, S1 = S_INIT, S2 = S_INIT, S3 = S_INIT;
                S0 = __dp2a_lo(LO, 8u | (56u << 8), S0);   S0 = __dp2a_lo(LE, 0u | (28u << 8), S0);
                S0 = __dp2a_lo(VE, 72u | (28u << 8), S0);  S0 = __dp2a_lo(VO, 56u | (8u << 8), S0);
                S1 = __dp2a_lo(LE, 0u | (8u << 8), S1);    S1 = __dp2a_lo(LO, 0u | (28u << 8), S1);
                S1 = __dp2a_lo(VE, 56u | (56u << 8), S1);  S1 = __dp2a_lo(VO, 72u | (28u << 8), S1);
                S1 = __dp2a_lo(RE_, 8u, S1);
                S2 = __dp2a_lo(LO, 0u | (8u << 8), S2);    S2 = __dp2a_lo(VE, 28u | (72u << 8), S2);
                S2 = __dp2a_lo(VO, 56u | (56u << 8), S2);  S2 = __dp2a_lo(RE_, 28u, S2);
                S2 = __dp2a_lo(RO_, 8u, S2);
                S3 = __dp2a_lo(VE, 8u | (56u << 8), S3);   S3 = __dp2a_lo(VO, 28u | (72u << 8), S3);
                S3 = __dp2a_lo(RE_, 56u | (8u << 8), S3);  S3 = __dp2a_lo(RO_, 28u, S3);
                const uint32_t hE = __byte_perm(S0, S2, 0x7632), hO = __byte_perm(S1, S3, 0x7632);  // fp16 1024 + B
                const int rb = r0 + t0 + j - 3;
                if (STEADY || (rb >= ys && rb < ye)) {
                    const uint32_t b4 = __byte_perm(hE, hO, 0x6240);
                    uint8_t* q = blur_p + j * blur_pitch;
                    asm volatile("" : "+l"(q));  // one address for the three store forms
                    if (st_full) stg_u32(q, b4);
                    if (st_l30) stg_u16(q, b4 & 0xFFFFu);
                    if (st_l1) stg_u16(q + 2, b4 >> 16);
                }
                bE[6 + j] = __hsub2(u2h(hE), k1024);
                bO[6 + j] = __hsub2(u2h(hO), k1024);
                sts2(slot_of(qa, qb, j, 256) + 2048, make_uint2(h2u(bE[6 + j]), h2u(bO[6 + j])));
            }
        } else {
#pragma unroll
            for (int j = 0; j < RB; ++j) {
                // u8 -> fp16: 0x6400 | b is the half 1024 + b; subtracting 1024 is exact
                bE[6 + j] = __hsub2(u2h(__byte_perm(w[j], 0x64646464u, 0x4240)), k1024);
                bO[6 + j] = __hsub2(u2h(__byte_perm(w[j], 0x64646464u, 0x4341)), k1024);
                sts2(slot_of(qa, qb, j, 256) + 2048, make_uint2(h2u(bE[6 + j]), h2u(bO[6 + j])));
            }
        }
        if (STEADY || t0 >= 8) {  // gradient rows g = r-4 < ys-2 never enter an output window
#pragma unroll
            for (int j = 0; j < RB; ++j) {
                // ---- 2. gradients of row g from blurred rows g-1, g, g+1 (steps t-2, t-1, t) -----------------
                const int g = r0 + t0 + j - 4;
                __half2 jxE = hz, jxO = hz, jyE = hz, jyO = hz;
                if (STEADY || (g >= 1 && g <= H - 2)) {  // warp-uniform; rows 0 and H-1 stay zero
                    const __half2 pE = bE[4 + j], pO = bO[4 + j], cE = bE[5 + j], cO = bO[5 + j], nE = bE[6 + j], nO = bO[6 + j];
                    const __half2 dE = __hsub2(pE, nE), dO = __hsub2(pO, nO);                                      // [1,0,-1]^T
                    const __half2 sE = __hadd2(__hfma2(cE, two, pE), nE), sO = __hadd2(__hfma2(cO, two, pO), nO);  // [1,2,1]^T
                    // one register each way carries d and s of the boundary pixel
                    const __half2 fromL = u2h(__shfl_up_sync(FULL, __byte_perm(h2u(dO), h2u(sO), 0x7632), 1));    // (d-1, s-1)
                    const __half2 fromR = u2h(__shfl_down_sync(FULL, __byte_perm(h2u(dE), h2u(sE), 0x5410), 1));  // (d4, s4)
                    const __half2 dm1d1 = prmt_h2(fromL, dO, 0x5410), d2d4 = prmt_h2(dE, fromR, 0x5432);
                    const __half2 sm1s1 = prmt_h2(fromL, sO, 0x5432), s2s4 = prmt_h2(sE, fromR, 0x7632);
                    jxE = __hadd2(__hfma2(dE, two, dm1d1), dO);  // (d-1,d1) + 2(d0,d2) + (d1,d3)
                    jxO = __hadd2(__hfma2(dO, two, dE), d2d4);   // (d0,d2) + 2(d1,d3) + (d2,d4)
                    jyE = __hsub2(sm1s1, sO);                    // (s-1,s1) - (s1,s3)
                    jyO = __hsub2(sE, s2s4);                     // (s0,s2) - (s2,s4)
                    if (edge) { jxE = __hmul2(jxE, mE); jxO = __hmul2(jxO, mO); jyE = __hmul2(jyE, mE); jyO = __hmul2(jyO, mO); }
                }
                // ---- 3. running sums over rows g-4..g: add row g, drop row g-5 ---------------------------------
                {
                    const uint4 old = lds4(slot_of(ja, jb, j - 5, 512));  // row g-5 (written at step t-5)
                    sts4(slot_of(ja, jb, j, 512), make_uint4(h2u(jxE), h2u(jxO), h2u(jyE), h2u(jyO)));
                    const __half2 oxE = u2h(old.x), oxO = u2h(old.y), oyE = u2h(old.z), oyO = u2h(old.w);
                    // pixel k of this lane: 0 = E.lo, 1 = O.lo, 2 = E.hi, 3 = O.hi
                    sa[0] = fhfnma_lo(oxE, fhfma_lo(jxE, sa[0])); sb[0] = fhfnma_lo(oyE, fhfma_lo(jyE, sb[0]));
                    sa[1] = fhfnma_lo(oxO, fhfma_lo(jxO, sa[1])); sb[1] = fhfnma_lo(oyO, fhfma_lo(jyO, sb[1]));
                    sa[2] = fhfnma_hi(oxE, fhfma_hi(jxE, sa[2])); sb[2] = fhfnma_hi(oyE, fhfma_hi(jyE, sb[2]));
                    sa[3] = fhfnma_hi(oxO, fhfma_hi(jxO, sa[3])); sb[3] = fhfnma_hi(oyO, fhfma_hi(jyO, sb[3]));
                }
                if (!STEADY && t0 < 12) continue;

                // ---- 4. response row o = r-6 in [ys, ye) -------------------------------------------------------
                const int orow = r0 + t0 + j - 6;
                float out[4] = {0.f, 0.f, 0.f, 0.f};
                if (STEADY || (orow >= 2 && orow <= H - 3)) {  // warp-uniform; other rows stay zero (R_array zero-initialised)
                    // horizontal 5-sums of the running sums (a and b together): columns x-2 .. x+5 needed
                    const u64 s0 = pk(sa[0], sb[0]), s1 = pk(sa[1], sb[1]), s2 = pk(sa[2], sb[2]), s3 = pk(sa[3], sb[3]);
                    const u64 L2 = shfl_up2(s2), L3 = shfl_up2(s3), R0 = shfl_dn2(s0), R1 = shfl_dn2(s1);
                    const u64 S0 = add2(add2(add2(L2, L3), add2(s0, s1)), s2);
                    const u64 S1 = add2(sub2(S0, L2), s3);
                    const u64 S2 = add2(sub2(S1, L3), R0);
                    const u64 S3 = add2(sub2(S2, s0), R1);
                    float A[4], B[4];
                    upk(S0, A[0], B[0]); upk(S1, A[1], B[1]); upk(S2, A[2], B[2]); upk(S3, A[3], B[3]);
                    {
                        // sum_5x5 Jxy, telescoped: E = B[o-3]+B[o-2]-B[o+2]-B[o+3] per column (steps t-6, t-5, t-1, t)
                        const __half2 eE = __hsub2(__hadd2(bE[j], bE[1 + j]), __hadd2(bE[5 + j], bE[6 + j]));
                        const __half2 eO = __hsub2(__hadd2(bO[j], bO[1 + j]), __hadd2(bO[5 + j], bO[6 + j]));
                        const __half2 LeE = u2h(__shfl_up_sync(FULL, h2u(eE), 1)), LeO = u2h(__shfl_up_sync(FULL, h2u(eO), 1));      // (e-4,e-2) (e-3,e-1)
                        const __half2 ReE = u2h(__shfl_down_sync(FULL, h2u(eE), 1)), ReO = u2h(__shfl_down_sync(FULL, h2u(eO), 1));  // (e4,e6)  (e5,e7)
                        const __half2 em2e0 = prmt_h2(LeE, eE, 0x5432), em1e1 = prmt_h2(LeO, eO, 0x5432);
                        const __half2 e2e4 = prmt_h2(eE, ReE, 0x5432), e3e5 = prmt_h2(eO, ReO, 0x5432);
                        const __half2 cE = __hsub2(__hadd2(LeO, em2e0), __hadd2(e2e4, e3e5));   // (c0, c2)
                        const __half2 cO = __hsub2(__hadd2(em2e0, em1e1), __hadd2(e3e5, ReE));  // (c1, c3)
                        if (HARRIS) harris4(A, B, cE, cO, thr4, a.harris_k, out);
                        else shi_tomasi4<FUSED>(A, B, cE, cO, thr4, out);
                    }
                }
                if (FUSED) {
                    // the row goes into the ring as computed: doubled raw responses (zeros for rows outside 2 .. H-3: slots are
                    // reused).  A group is a candidate group when any of its pixels exceeds the doubled threshold (fmaxf drops
                    // the NaN of a negative discriminant).  Columns < 2 and > W-3, which the reference leaves at zero, hold
                    // garbage here: no window of a centre inside the FREAK margins (columns 37 .. W-37) reaches them.
                    sts4(rrow + j * 512, make_uint4(__float_as_uint(out[0]), __float_as_uint(out[1]), __float_as_uint(out[2]), __float_as_uint(out[3])));
                    nz[j] = __ballot_sync(FULL, fmaxf(fmaxf(out[0], out[1]), fmaxf(out[2], out[3])) > nms_thr);
                } else if (STEADY || orow < ye) {
                    float* q = R_p + j * R_pitch;
                    asm volatile("" : "+l"(q));
                    if (st_full) stg_f4(q, out[0], out[1], out[2], out[3]);
                    if (st_l30) stg_f2(q, out[0], out[1]);
                    if (st_l1) stg_f2(q + 2, out[2], out[3]);
                }
            }
            if (FUSED && (STEADY || t0 >= 12)) nms_trip(t0, nz);
        }
        // the ring bases swap: slot (c & 7) ^ 4 of this trip is slot c & 7 of the next
        { const char* t = qa; qa = qb; qb = t; t = ja; ja = jb; jb = t; }
        if (!FROM_BLURRED) blur_p += RB * blur_pitch;
        R_p += RB * R_pitch;
        in_next += RB * in_pitch;
    };
    // steady: t0 >= 16 (response rows >= ys+4 >= 2, all rings filled) and blurred row ys+t0-6 <= ye-1 (which puts
    // the gradient row ys+t0-7 <= H-2, the response row ys+t0-9 <= min(ye-1, H-3), and leaves a further trip)
    // and the rows it requests for the next trip, up to ys+t0+1+load_off, are image rows (no reflection).
    const int t_steady_end = min((ye - ys) + 6, H - ys - 1 - load_off);
#pragma unroll 1
    for (int t0 = 0; t0 < T; t0 += RB) {
        if (t0 >= 16 && t0 < t_steady_end) {
            trip(std::true_type{}, t0);
        } else {
            trip(std::false_type{}, t0);
        }
    }
    if (FUSED) flush_candidates();
    if (!a.task_counter) return;
    __syncwarp();
    }  // next task
}

// custom_bgr2gray: saturate_cast<uchar>(0.114*B + 0.587*G + 0.299*R) in double, round-half-even
// (corner_detection_parallel.cpp:14-28); explicit _rn ops keep nvcc from fusing multiply-adds.
__global__ void k_bgr2gray(const uint8_t* __restrict__ bgr, size_t bgr_pitch, uint8_t* __restrict__ gray,
                           size_t gray_pitch, int rows, int cols) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= cols || y >= rows) return;
    const uint8_t* p = bgr + (size_t)y * bgr_pitch + 3 * x;
    const double v = __dadd_rn(__dadd_rn(__dmul_rn(0.114, (double)p[0]), __dmul_rn(0.587, (double)p[1])),
                               __dmul_rn(0.299, (double)p[2]));
    const int iv = __double2int_rn(v);
    gray[(size_t)y * gray_pitch + x] = (uint8_t)min(max(iv, 0), 255);
}

int sm_count_cached() {
    static int n = 0;
    if (!n) {
        int dev = 0;
        cudaGetDevice(&dev);
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    }
    return n;
}

}  // namespace

bool stencil_can_fuse(int rows, int cols, int nms_window) {
    // fused NMS: windows 3 and 5 (half window <= 2: the ring keeps eight response rows), images with room for a candidate
    // inside the FREAK margins (smaller images have none; they take the unfused path, whose NMS kernel launches nothing)
    const int h = nms_window / 2;
    if (h < 1 || h > 2) return false;
    if (cols < 16 || rows < 8) return false;
    const int i_lo = max(h, 35), i_hi = min(rows - h - 1, rows - 35), j_lo = max(h, 37), j_hi = min(cols - h - 1, cols - 37);
    return i_hi >= i_lo && j_hi >= j_lo;
}

cudaError_t launch_stencil(const StencilArgs& a, int n_frames, int src, cudaStream_t st) {
    const int nms_h = a.nms_h;
    const bool from_blurred = src == STENCIL_SRC_BLURRED, from_R = src == STENCIL_SRC_RESPONSE;
    const bool harris = a.response == VO_RESPONSE_HARRIS;
    if ((a.response != VO_RESPONSE_SHI_TOMASI && !harris) || (harris && nms_h)) return cudaErrorInvalidValue;
    if (src < 0 || src > 2 || nms_h < 0 || nms_h > 2 || (from_R && !nms_h) ||
        (nms_h && (from_blurred || !stencil_can_fuse(a.rows, a.cols, 2 * nms_h + 1) || !a.cand || !a.cand_count)))
        return cudaErrorInvalidValue;
    // blurred and R rows are written with 4-byte / 16-byte vector stores (R read with 16-byte loads) at columns that are multiples of 4
    if (((!nms_h || from_R) && ((a.R_pitch & 3) || (a.R_fstride & 3) || (reinterpret_cast<uintptr_t>(a.R) & 15))) ||
        (!from_blurred && !from_R && ((a.blur_pitch & 3) || (a.blur_fstride & 3))))
        return cudaErrorInvalidValue;
    // unfused: strip s covers columns [s*SW - 2, s*SW + SW - 2).  Fused: strip s starts at s*SWF - 4 and holds valid blurred
    // pixels in window columns 3..124: the last strip must reach column cols-1
    const int n_strips = nms_h ? (max(a.cols - 121, 0) + SWF - 1) / SWF + 1 : (a.cols + 2 + SW - 1) / SW;
    // Segment height.  Every segment recomputes 12 halo rows (fused: 20), but the kernel is latency-bound per warp, so what
    // matters most is that all resident warp slots stay occupied to the end: a sweep over segment counts (KITTI at
    // batch 64/100/200, 4K at batch 1/16/64, profiles/README.md) is flat within 5 % around "three tasks per resident
    // warp slot" and 10-25 % worse for a single under-filled wave; at 1000 KITTI frames per launch 1..4 segments are within
    // 1 % of each other (profiles/r02a_segs.jsonl).  Segments stay at least 48 rows high.
    const int slots = sm_count_cached() * (nms_h ? CTAS_PER_SM_FUSED : CTAS_PER_SM) * NWARP;
    const long long cols_tasks = (long long)n_strips * n_frames;
    int best_segs = (int)((3LL * slots + cols_tasks / 2) / cols_tasks);
    best_segs = max(1, min(best_segs, max(1, a.rows / 48)));
    if (const char* e = getenv("VO_STENCIL_SEGS")) { const int v = atoi(e); if (v >= 1 && v <= (a.rows + 15) / 16) best_segs = v; }  // tuning aid
    const int seg_rows = ((a.rows + best_segs - 1) / best_segs + RB - 1) / RB * RB;
    const int n_segs = (a.rows + seg_rows - 1) / seg_rows;
    const long long n_tasks = (long long)n_strips * n_segs * n_frames;
    if (n_tasks > 0x7FFFFFFFll) return cudaErrorInvalidValue;
    long long n_ctas = (n_tasks + NWARP - 1) / NWARP;
    if (a.task_counter && a.ctas_per_sm_cap > 0) n_ctas = std::min<long long>(n_ctas, (long long)a.ctas_per_sm_cap * sm_count_cached());
    const dim3 grid((unsigned)n_ctas);
    const bool gather = a.cols < 16 || a.rows < 8;
    const bool aligned = ((reinterpret_cast<uintptr_t>(a.in) | a.in_pitch | a.in_fstride) & 3) == 0;
#define VO_LAUNCH(SR, GA, AL, NH) k_stencil<SR, GA, AL, NH><<<grid, 32 * NWARP, 0, st>>>(a, n_strips, n_segs, seg_rows, (int)n_tasks)
    if (from_R) {
        if (nms_h == 1) VO_LAUNCH(2, false, true, 1);
        else VO_LAUNCH(2, false, true, 2);
    } else if (nms_h == 1) {
        if (aligned) VO_LAUNCH(0, false, true, 1);
        else VO_LAUNCH(0, false, false, 1);
    } else if (nms_h == 2) {
        if (aligned) VO_LAUNCH(0, false, true, 2);
        else VO_LAUNCH(0, false, false, 2);
    } else if (harris) {
#define VO_LAUNCH_H(SR, GA, AL) k_stencil<SR, GA, AL, 0, true><<<grid, 32 * NWARP, 0, st>>>(a, n_strips, n_segs, seg_rows, (int)n_tasks)
        if (from_blurred) {
            if (gather) VO_LAUNCH_H(1, true, false);
            else if (aligned) VO_LAUNCH_H(1, false, true);
            else VO_LAUNCH_H(1, false, false);
        } else {
            if (gather) VO_LAUNCH_H(0, true, false);
            else if (aligned) VO_LAUNCH_H(0, false, true);
            else VO_LAUNCH_H(0, false, false);
        }
#undef VO_LAUNCH_H
    } else if (from_blurred) {
        if (gather) VO_LAUNCH(1, true, false, 0);
        else if (aligned) VO_LAUNCH(1, false, true, 0);
        else VO_LAUNCH(1, false, false, 0);
    } else {
        if (gather) VO_LAUNCH(0, true, false, 0);
        else if (aligned) VO_LAUNCH(0, false, true, 0);
        else VO_LAUNCH(0, false, false, 0);
    }
#undef VO_LAUNCH
    return cudaGetLastError();
}

cudaError_t launch_bgr2gray(const uint8_t* bgr, size_t bgr_pitch, uint8_t* gray, size_t gray_pitch, int rows, int cols,
                            cudaStream_t st) {
    const dim3 grid((cols + 255) / 256, rows);
    k_bgr2gray<<<grid, 256, 0, st>>>(bgr, bgr_pitch, gray, gray_pitch, rows, cols);
    return cudaGetLastError();
}
