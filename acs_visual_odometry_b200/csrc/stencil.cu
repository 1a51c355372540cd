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

#include <cstdlib>
#include <type_traits>

#include "vo_internal.cuh"

namespace {

constexpr int SW = 116;    // output columns per warp strip: the 128-pixel window minus 6 halo columns each side
constexpr int HX = 8;      // window start relative to strip*SW (6 halo columns needed; 8 keeps the 4-pixel groups aligned,
                           // so lane 1 stores only its upper two pixels and lane 30 only its lower two)
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
        const u64 r = fma2(nsq, half, mul2(pk(t[q], t[q + 1]), half));
        float r0, r1; upk(r, r0, r1);
        out[q] = r0 > thr[q] ? r0 : 0.0f;
        out[q + 1] = r1 > thr[q + 1] ? r1 : 0.0f;
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
template <bool FROM_BLURRED, bool GATHER, bool ALIGNED>
__global__ void __launch_bounds__(32 * NWARP, CTAS_PER_SM)
k_stencil(const StencilArgs a, int n_strips, int n_segs, int seg_rows, int n_tasks) {
    __shared__ __align__(16) char s_rings[NWARP][RING_BYTES];
    const int lane = threadIdx.x & 31;
    // tasks = (frame, row segment, strip), numbered across the whole batch so that no warp slot of a CTA idles
    const int gtask = blockIdx.x * NWARP + (threadIdx.x >> 5);
    if (gtask >= n_tasks) return;  // whole warp
    char* const ring = s_rings[threadIdx.x >> 5];
    const int per_frame = n_strips * n_segs;
    const int f = gtask / per_frame, task = gtask - f * per_frame;
    const int strip = task % n_strips, seg = task / n_strips;
    const int H = a.rows, W = a.cols;
    const int ys = seg * seg_rows, ye = min(H, ys + seg_rows);
    const int xs = strip * SW - HX;
    const int x = xs + 4 * lane;  // this lane's columns: x .. x+3
    // outputs of this strip: columns xs+6 .. xs+121, i.e. lanes 2..29 whole, lane 1 pixels 2-3, lane 30 pixels 0-1
    // lanes 2..29 store their four pixels, lane 1 its upper pair, lane 30 its lower pair
    const bool st_full = lane >= 2 && lane < 30 && x < W, st_l1 = lane == 1 && x + 2 >= 0 && x + 2 < W, st_l30 = lane == 30 && x < W;
    const uint8_t* __restrict__ in = a.in + (size_t)f * a.in_fstride;
    uint8_t* __restrict__ blur_out = FROM_BLURRED ? nullptr : a.blur + (size_t)f * a.blur_fstride;
    float* __restrict__ R = a.R + (size_t)f * a.R_fstride;
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

    const __half2 hz = __floats2half2_rn(0.f, 0.f), two = __floats2half2_rn(2.f, 2.f), k1024 = __floats2half2_rn(1024.f, 1024.f);

    // Step t handles input row r = ys-6+t (blur mode) or blurred row r-3 (from-blurred mode); after it
    //   B row r-3 is complete, gradient row g = r-4, response row o = r-6.  Ring slot of step t: t & 7.
    const int T = (ye - ys) + 12;
    const int r0 = ys - 6;
    const int load_off = FROM_BLURRED ? -3 : 0;
    issue(r0 + load_off, false);
    in_next += (long long)(r0 + RB + load_off) * in_pitch;  // trip t0 requests rows r0+t0+RB+load_off ..
    // output row pointers of step t0 (this lane's column), advanced by RB rows per trip
    const long long blur_pitch = (long long)a.blur_pitch, R_pitch = (long long)a.R_pitch;
    uint8_t* blur_p = FROM_BLURRED ? nullptr : blur_out + (long long)(r0 - 3) * blur_pitch + x;
    float* R_p = R + (long long)(r0 - 6) * R_pitch + x;

    // One trip = four steps.  STEADY trips (all four blurred / gradient / response rows inside the segment and the
    // image, a further trip to request) carry no row tests; the first four and the last one or two trips of a
    // segment run the general form.
    auto trip = [&](auto steady_tag, const int t0) {
        constexpr bool STEADY = decltype(steady_tag)::value;
        // (a third body specialised for interior strips was tried: 8 % SLOWER, the two steady bodies together no longer
        // fit the instruction cache; so the edge handling stays predicated on the warp-uniform edge_strip)
        const bool edge = edge_strip;
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
                        shi_tomasi4(A, B, cE, cO, thr4, out);
                    }
                }
                if (STEADY || orow < ye) {
                    float* q = R_p + j * R_pitch;
                    asm volatile("" : "+l"(q));
                    if (st_full) stg_f4(q, out[0], out[1], out[2], out[3]);
                    if (st_l30) stg_f2(q, out[0], out[1]);
                    if (st_l1) stg_f2(q + 2, out[2], out[3]);
                }
            }
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

cudaError_t launch_stencil(const StencilArgs& a, int n_frames, bool from_blurred, cudaStream_t st) {
    // blurred and R rows are written with 4-byte / 16-byte vector stores at columns that are multiples of 4
    if ((a.R_pitch & 3) || (a.R_fstride & 3) || (!from_blurred && ((a.blur_pitch & 3) || (a.blur_fstride & 3))))
        return cudaErrorInvalidValue;
    const int n_strips = (a.cols + 2 + SW - 1) / SW;  // strip s covers columns [s*SW - 2, s*SW + SW - 2)
    // Segment height.  Every segment recomputes 12 halo rows, but the kernel is latency-bound per warp, so what
    // matters most is that all resident warp slots stay occupied to the end: a sweep over segment counts (KITTI at
    // batch 64/100/200, 4K at batch 1/16/64, profiles/README.md) is flat within 5 % around "three tasks per resident
    // warp slot" and 10-25 % worse for a single under-filled wave.  Segments stay at least 48 rows high.
    const int slots = sm_count_cached() * CTAS_PER_SM * NWARP;
    const long long cols_tasks = (long long)n_strips * n_frames;
    int best_segs = (int)((3LL * slots + cols_tasks / 2) / cols_tasks);
    best_segs = max(1, min(best_segs, max(1, a.rows / 48)));
    if (const char* e = getenv("VO_STENCIL_SEGS")) { const int v = atoi(e); if (v >= 1 && v <= (a.rows + 15) / 16) best_segs = v; }  // tuning aid
    const int seg_rows = ((a.rows + best_segs - 1) / best_segs + RB - 1) / RB * RB;
    const int n_segs = (a.rows + seg_rows - 1) / seg_rows;
    const long long n_tasks = (long long)n_strips * n_segs * n_frames;
    if (n_tasks > 0x7FFFFFFFll) return cudaErrorInvalidValue;
    const dim3 grid((unsigned)((n_tasks + NWARP - 1) / NWARP));
    const bool gather = a.cols < 16 || a.rows < 8;
    const bool aligned = ((reinterpret_cast<uintptr_t>(a.in) | a.in_pitch | a.in_fstride) & 3) == 0;
#define VO_LAUNCH(FB, GA, AL) k_stencil<FB, GA, AL><<<grid, 32 * NWARP, 0, st>>>(a, n_strips, n_segs, seg_rows, (int)n_tasks)
    if (from_blurred) {
        if (gather) VO_LAUNCH(true, true, false);
        else if (aligned) VO_LAUNCH(true, false, true);
        else VO_LAUNCH(true, false, false);
    } else {
        if (gather) VO_LAUNCH(false, true, false);
        else if (aligned) VO_LAUNCH(false, false, true);
        else VO_LAUNCH(false, false, false);
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
