// stencil.cu — K1: fused blur -> gradients -> 5x5 structure tensor -> Shi-Tomasi response, ONE pass.
//
// Replaces, for a whole batch of frames per launch:
//   cv::GaussianBlur(u8, 7x7, sigma 0)              feature_extraction_parallel/feature_extraction_parallel.cpp:223
//   CornerDetectionParallel::direction_gradients_worker        corner_detection_parallel.cpp:30-58
//   CornerDetectionParallel::shitomasi_corner_detection_worker corner_detection_parallel.cpp:61-105
// with whole-image (single strip) semantics, bit-exact (see "Exactness" below).
//
// Design (B200): the kernel is issue-bound, not HBM-bound (about 60 instructions per pixel against
// 6 B/px of traffic), so it is built to minimise instructions and to touch each pixel's data once:
//   * one WARP owns a strip of 128 columns (4 adjacent pixels per lane, 8-column halo each side ->
//     112 output columns) and marches down the rows of its segment.  There is no block-level barrier
//     and no data ever moves between threads through shared memory: horizontal neighbours come from
//     warp shuffles; vertical neighbours (7 raw rows, 7 blurred rows, 6 gradient rows) live in small
//     per-lane ring buffers in shared memory that only their owner reads back (a register ring would
//     need the row loop unrolled by 8: 90 KB of code that thrashes the instruction cache, and 123
//     registers that cap the SM at 16 warps - measured, see profiles/);
//   * pixels are processed two per instruction: 16-bit SWAR integer lanes for the vertical blur taps,
//     __dp2a for the horizontal taps, half2 arithmetic for the gradients (all values are integers of
//     magnitude <= 2040, exact in fp16);
//   * the sum of the mixed derivative telescopes: sum_5x5 Jxy = E[x-3]+E[x-2]-E[x+2]-E[x+3] with
//     E[x] = B[o-3][x]+B[o-2][x]-B[o+2][x]-B[o+3][x], because Jxy is a difference of differences and the
//     BORDER_REFLECT_101 extension makes it vanish on the image border exactly as the reference's
//     zero-initialised border does;
//   * the sums of squares are running sums over the 5-row window (add the entering row, subtract the
//     leaving one) in fp32, where they are exact integers (< 2^23 for any blurred 8-bit image);
//   * the seven fp32 operations of the eigenvalue run only when some pixel of the warp row can pass the
//     threshold (R <= tr/2 always holds, so tr/2 <= threshold decides R = 0 without computing it);
//   * the next input row is requested one step ahead and only touched (funnel-shifted) a step later.
// HBM traffic per pixel: 1 B in, 1 B blurred out, 4 B R out.
//
// Exactness: everything up to a, b, c is integer arithmetic carried in exact containers; the seven
// rounding operations use __f*_rn intrinsics (no FMA contraction), so R is bit-identical to the
// reference's scalar x86-64 code.
#include <cuda_fp16.h>

#include "vo_internal.cuh"

namespace {

constexpr int SW = 112;    // output columns per warp strip
constexpr int HX = 8;      // column halo on each side (6 needed; 8 keeps the 4-pixel groups aligned)
constexpr int NWARP = 4;   // warps per CTA (independent tasks)
constexpr unsigned FULL = 0xFFFFFFFFu;

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
__device__ __forceinline__ __half2 shfl_up_h2(__half2 v) { return u2h(__shfl_up_sync(FULL, h2u(v), 1)); }
__device__ __forceinline__ __half2 shfl_dn_h2(__half2 v) { return u2h(__shfl_down_sync(FULL, h2u(v), 1)); }
// (hi half of a, lo half of b): the pair that straddles two packed registers
__device__ __forceinline__ __half2 hi_lo(__half2 a, __half2 b) { return u2h(__byte_perm(h2u(a), h2u(b), 0x5432)); }

// R = tr/2 - 0.5*sqrt(tr^2 - 4 det), det = a*b - c*c  (corner_detection_parallel.cpp:93-98); a, b, c exact.
// Every operation rounds once to nearest like the reference's scalar code (its final subtraction in
// double followed by narrowing equals one fp32 rounding: 53 >= 2*24+2).  disc is an integer-valued float
// (all operands are integers), so it is 0, <= -1 or >= 1: clamping the sqrt argument to >= 1 keeps
// __fsqrt_rn on its branch-free path; disc == 0 and disc < 0 (NaN in the reference -> 0) are selected after.
__device__ __forceinline__ float shi_tomasi(float fa, float fb, float fc, float thr) {
    const float det = __fsub_rn(__fmul_rn(fa, fb), __fmul_rn(fc, fc));
    const float tr = __fadd_rn(fa, fb);
    const float disc = __fsub_rn(__fmul_rn(tr, tr), __fmul_rn(4.0f, det));
    float sq = __fsqrt_rn(fmaxf(disc, 1.0f));
    sq = disc == 0.0f ? 0.0f : sq;
    const float r = __fsub_rn(__fmul_rn(tr, 0.5f), __fmul_rn(0.5f, sq));
    return (disc >= 0.0f && r > thr) ? r : 0.0f;
}

// Per-lane history rings.  Each lane reads back only what it wrote itself (no inter-thread traffic, no
// barrier); the [slot][lane] layout makes every warp access one conflict-free vector row.
struct WarpRings {
    uint2 raw[8][32];  // (E, O) u8 pixels in 16-bit lanes, input rows
    uint2 b[8][32];    // (E, O) half2, blurred rows
    uint4 j[8][32];    // (JxE, JxO, JyE, JyO) half2, gradient rows
};

template <bool FROM_BLURRED>
__global__ void __launch_bounds__(32 * NWARP, 6) k_stencil(const StencilArgs a, int n_strips, int n_segs, int seg_rows) {
    __shared__ WarpRings s_rings[NWARP];
    const int lane = threadIdx.x & 31;
    const int task = blockIdx.x * NWARP + (threadIdx.x >> 5);
    if (task >= n_strips * n_segs) return;  // whole warp
    WarpRings& ring = s_rings[threadIdx.x >> 5];
    const int f = blockIdx.y;
    const int strip = task % n_strips, seg = task / n_strips;
    const int H = a.rows, W = a.cols;
    const int ys = seg * seg_rows, ye = min(H, ys + seg_rows);
    const int x = strip * SW - HX + 4 * lane;  // this lane's columns: x .. x+3
    const bool store_lane = lane >= HX / 4 && lane < 32 - HX / 4 && x < W;
    const uint8_t* __restrict__ in = a.in + (size_t)f * a.in_fstride;
    uint8_t* __restrict__ blur_out = FROM_BLURRED ? nullptr : a.blur + (size_t)f * a.blur_fstride;
    float* __restrict__ R = a.R + (size_t)f * a.R_fstride;
    const float thr = a.threshold;

    // gradients are zero outside 1 <= j <= W-2 (cv::Mat::zeros border); packing: E = (px0, px2), O = (px1, px3)
    const __half2 mE = __floats2half2_rn((x >= 1 && x <= W - 2) ? 1.f : 0.f, (x + 2 >= 1 && x + 2 <= W - 2) ? 1.f : 0.f);
    const __half2 mO = __floats2half2_rn((x + 1 >= 1 && x + 1 <= W - 2) ? 1.f : 0.f, (x + 3 >= 1 && x + 3 <= W - 2) ? 1.f : 0.f);
    const bool edge_strip = (strip * SW - HX < 1) || (strip * SW + SW + HX > W - 2);  // warp-uniform
    // response columns < 2 and > W-3 stay zero
    const bool okc0 = x >= 2 && x <= W - 3, okc1 = x + 1 >= 2 && x + 1 <= W - 3, okc2 = x + 2 >= 2 && x + 2 <= W - 3,
               okc3 = x + 3 >= 2 && x + 3 <= W - 3;

    // input access: lanes whose 8-byte window lies inside the row read two aligned words (any row alignment,
    // host layouts such as pitch 1241 are read in place) and funnel-shift them one step later; lanes on the
    // image border gather four reflected bytes.
    const bool fast = x >= 0 && x + 7 < W;
    const int cx0 = reflect101(x, W), cx1 = reflect101(x + 1, W), cx2 = reflect101(x + 2, W), cx3 = reflect101(x + 3, W);
    uint32_t pa = 0, pb = 0, pc = 0, pd = 0;  // in-flight input row
    auto issue = [&](int row) {
        const uint8_t* p = in + (size_t)reflect101(row, H) * a.in_pitch;
        if (fast) {
            const uintptr_t q = reinterpret_cast<uintptr_t>(p + x);
            const uint32_t* w = reinterpret_cast<const uint32_t*>(q & ~static_cast<uintptr_t>(3));
            pa = __ldg(w); pb = __ldg(w + 1); pc = static_cast<uint32_t>(q & 3) * 8u;
        } else {
            pa = p[cx0]; pb = p[cx1]; pc = p[cx2]; pd = p[cx3];
        }
    };

    const __half2 hz = __floats2half2_rn(0.f, 0.f), two = __floats2half2_rn(2.f, 2.f), k1024 = __floats2half2_rn(1024.f, 1024.f);
#pragma unroll
    for (int k = 0; k < 8; ++k) {  // gradient rows before the first computed one count as zero
        ring.j[k][lane] = make_uint4(0u, 0u, 0u, 0u);
        ring.raw[k][lane] = make_uint2(0u, 0u);
        ring.b[k][lane] = make_uint2(0u, 0u);
    }
    float sa0 = 0.f, sa1 = 0.f, sa2 = 0.f, sa3 = 0.f, sb0 = 0.f, sb1 = 0.f, sb2 = 0.f, sb3 = 0.f;  // running 5-row sums of Jx^2, Jy^2

    // Step t handles input row r = ys-6+t (blur mode) or blurred row r-3 (from-blurred mode); after it
    //   B row r-3 is complete, gradient row g = r-4, response row o = r-6.  Ring slot of step t: t & 7.
    const int T = (ye - ys) + 12;
    const int r0 = ys - 6;
    const int load_off = FROM_BLURRED ? -3 : 0;
    issue(r0 + load_off);

#pragma unroll 1
    for (int t = 0; t < T; ++t) {
        const int r = r0 + t;
        const int u = t & 7;
        const uint32_t w = fast ? __funnelshift_r(pa, pb, pc) : (pa | (pb << 8) | (pc << 16) | (pd << 24));
        if (t + 1 < T) issue(r + 1 + load_off);

        // ---- 1. blurred row r-3 -> slot u ---------------------------------------------------------------
        uint32_t BEi, BOi;
        if (!FROM_BLURRED) {
            const uint32_t RE0 = __byte_perm(w, 0u, 0x4240), RO0 = __byte_perm(w, 0u, 0x4341);  // (px0,px2) (px1,px3)
            ring.raw[u][lane] = make_uint2(RE0, RO0);
            const uint2 q1 = ring.raw[(u + 7) & 7][lane], q2 = ring.raw[(u + 6) & 7][lane], q3 = ring.raw[(u + 5) & 7][lane];
            const uint2 q4 = ring.raw[(u + 4) & 7][lane], q5 = ring.raw[(u + 3) & 7][lane], q6 = ring.raw[(u + 2) & 7][lane];
            // vertical taps [2,7,14,18,14,7,2] on rows r-6..r; 16-bit lanes never exceed 64*255
            const uint32_t VE = 2u * (RE0 + q6.x) + 7u * (q1.x + q5.x) + 14u * (q2.x + q4.x) + 18u * q3.x;
            const uint32_t VO = 2u * (RO0 + q6.y) + 7u * (q1.y + q5.y) + 14u * (q2.y + q4.y) + 18u * q3.y;
            // VE = (V0, V2), VO = (V1, V3); neighbours' columns through shuffles
            const uint32_t LE = __shfl_up_sync(FULL, VE, 1), LO = __shfl_up_sync(FULL, VO, 1);        // (V-4,V-2) (V-3,V-1)
            const uint32_t RE_ = __shfl_down_sync(FULL, VE, 1), RO_ = __shfl_down_sync(FULL, VO, 1);  // (V4,V6) (V5,V7)
            // horizontal taps: __dp2a_lo(a, b, c) = c + a.lo16*b.byte0 + a.hi16*b.byte1
            uint32_t S0 = 2048u, S1 = 2048u, S2 = 2048u, S3 = 2048u;
            S0 = __dp2a_lo(LO, 2u | (14u << 8), S0);  S0 = __dp2a_lo(LE, 0u | (7u << 8), S0);
            S0 = __dp2a_lo(VE, 18u | (7u << 8), S0);  S0 = __dp2a_lo(VO, 14u | (2u << 8), S0);
            S1 = __dp2a_lo(LE, 0u | (2u << 8), S1);   S1 = __dp2a_lo(LO, 0u | (7u << 8), S1);
            S1 = __dp2a_lo(VE, 14u | (14u << 8), S1); S1 = __dp2a_lo(VO, 18u | (7u << 8), S1);
            S1 = __dp2a_lo(RE_, 2u, S1);
            S2 = __dp2a_lo(LO, 0u | (2u << 8), S2);   S2 = __dp2a_lo(VE, 7u | (18u << 8), S2);
            S2 = __dp2a_lo(VO, 14u | (14u << 8), S2); S2 = __dp2a_lo(RE_, 7u, S2);
            S2 = __dp2a_lo(RO_, 2u, S2);
            S3 = __dp2a_lo(VE, 2u | (14u << 8), S3);  S3 = __dp2a_lo(VO, 7u | (18u << 8), S3);
            S3 = __dp2a_lo(RE_, 14u | (2u << 8), S3); S3 = __dp2a_lo(RO_, 7u, S3);
            BEi = __byte_perm(S0 >> 12, S2 >> 12, 0x5410);  // (B0, B2)
            BOi = __byte_perm(S1 >> 12, S3 >> 12, 0x5410);  // (B1, B3)
            const int rb = r - 3;
            if (store_lane && rb >= ys && rb < ye)
                *reinterpret_cast<uint32_t*>(blur_out + (size_t)rb * a.blur_pitch + x) = __byte_perm(BEi, BOi, 0x6240);
        } else {
            BEi = __byte_perm(w, 0u, 0x4240);
            BOi = __byte_perm(w, 0u, 0x4341);
        }
        // u16 -> fp16: 0x6400 | b is the half 1024 + b; subtracting 1024 is exact
        const __half2 nE = __hsub2(u2h(BEi | 0x64006400u), k1024), nO = __hsub2(u2h(BOi | 0x64006400u), k1024);
        ring.b[u][lane] = make_uint2(h2u(nE), h2u(nO));
        if (t < 8) continue;  // gradient row g = r-4 < ys-2 never enters an output window

        // ---- 2. gradients of row g = r-4 from blurred rows g-1, g, g+1 ------------------------------------
        const int g = r - 4;
        const uint2 bc = ring.b[(u + 7) & 7][lane];  // row g = r-4; also row o+2 of the response row o = r-6
        __half2 jxE = hz, jxO = hz, jyE = hz, jyO = hz;
        if (g >= 1 && g <= H - 2) {  // warp-uniform; rows 0 and H-1 stay zero
            const uint2 bp = ring.b[(u + 6) & 7][lane];  // row g-1 = r-5
            const __half2 pE = u2h(bp.x), pO = u2h(bp.y), cE = u2h(bc.x), cO = u2h(bc.y);
            const __half2 dE = __hsub2(pE, nE), dO = __hsub2(pO, nO);                                      // [1,0,-1]^T
            const __half2 sE = __hadd2(__hfma2(cE, two, pE), nE), sO = __hadd2(__hfma2(cO, two, pO), nO);  // [1,2,1]^T
            const __half2 LdO = shfl_up_h2(dO), RdE = shfl_dn_h2(dE);  // (d-3,d-1) (d4,d6)
            const __half2 LsO = shfl_up_h2(sO), RsE = shfl_dn_h2(sE);
            const __half2 dm1d1 = hi_lo(LdO, dO), d2d4 = hi_lo(dE, RdE);
            const __half2 sm1s1 = hi_lo(LsO, sO), s2s4 = hi_lo(sE, RsE);
            jxE = __hadd2(__hfma2(dE, two, dm1d1), dO);  // (d-1,d1) + 2(d0,d2) + (d1,d3)
            jxO = __hadd2(__hfma2(dO, two, dE), d2d4);   // (d0,d2) + 2(d1,d3) + (d2,d4)
            jyE = __hsub2(sm1s1, sO);                    // (s-1,s1) - (s1,s3)
            jyO = __hsub2(sE, s2s4);                     // (s0,s2) - (s2,s4)
            if (edge_strip) { jxE = __hmul2(jxE, mE); jxO = __hmul2(jxO, mO); jyE = __hmul2(jyE, mE); jyO = __hmul2(jyO, mO); }
        }
        // ---- 3. running sums over rows g-4..g: add row g, drop row g-5 --------------------------------------
        {
            const uint4 old = ring.j[(u + 3) & 7][lane];  // row g-5 (written at step t-5)
            ring.j[u][lane] = make_uint4(h2u(jxE), h2u(jxO), h2u(jyE), h2u(jyO));
            const __half2 oxE = u2h(old.x), oxO = u2h(old.y), oyE = u2h(old.z), oyO = u2h(old.w);
            float v;
            v = __low2float(jxE);  sa0 = __fmaf_rn(v, v, sa0);  v = __low2float(oxE);  sa0 = __fmaf_rn(-v, v, sa0);
            v = __low2float(jxO);  sa1 = __fmaf_rn(v, v, sa1);  v = __low2float(oxO);  sa1 = __fmaf_rn(-v, v, sa1);
            v = __high2float(jxE); sa2 = __fmaf_rn(v, v, sa2);  v = __high2float(oxE); sa2 = __fmaf_rn(-v, v, sa2);
            v = __high2float(jxO); sa3 = __fmaf_rn(v, v, sa3);  v = __high2float(oxO); sa3 = __fmaf_rn(-v, v, sa3);
            v = __low2float(jyE);  sb0 = __fmaf_rn(v, v, sb0);  v = __low2float(oyE);  sb0 = __fmaf_rn(-v, v, sb0);
            v = __low2float(jyO);  sb1 = __fmaf_rn(v, v, sb1);  v = __low2float(oyO);  sb1 = __fmaf_rn(-v, v, sb1);
            v = __high2float(jyE); sb2 = __fmaf_rn(v, v, sb2);  v = __high2float(oyE); sb2 = __fmaf_rn(-v, v, sb2);
            v = __high2float(jyO); sb3 = __fmaf_rn(v, v, sb3);  v = __high2float(oyO); sb3 = __fmaf_rn(-v, v, sb3);
        }
        if (t < 12) continue;

        // ---- 4. response row o = r-6 in [ys, ye) -------------------------------------------------------------
        const int o = r - 6;
        float out0 = 0.f, out1 = 0.f, out2 = 0.f, out3 = 0.f;
        if (o >= 2 && o <= H - 3) {  // warp-uniform; other rows stay zero (R_array zero-initialised)
            // horizontal 5-sums of the running sums: columns x-2 .. x+5 needed
            const float La2 = __shfl_up_sync(FULL, sa2, 1), La3 = __shfl_up_sync(FULL, sa3, 1);
            const float Ra0 = __shfl_down_sync(FULL, sa0, 1), Ra1 = __shfl_down_sync(FULL, sa1, 1);
            const float Lb2 = __shfl_up_sync(FULL, sb2, 1), Lb3 = __shfl_up_sync(FULL, sb3, 1);
            const float Rb0 = __shfl_down_sync(FULL, sb0, 1), Rb1 = __shfl_down_sync(FULL, sb1, 1);
            const float A0 = (La2 + La3) + (sa0 + sa1) + sa2, A1 = A0 - La2 + sa3, A2 = A1 - La3 + Ra0, A3 = A2 - sa0 + Ra1;
            const float B0 = (Lb2 + Lb3) + (sb0 + sb1) + sb2, B1 = B0 - Lb2 + sb3, B2 = B1 - Lb3 + Rb0, B3 = B2 - sb0 + Rb1;
            const bool need = fmaxf(fmaxf(A0 + B0, A1 + B1), fmaxf(A2 + B2, A3 + B3)) * 0.5f > thr;
            if (__any_sync(FULL, need)) {  // R <= tr/2: rows of flat pixels skip the eigenvalue
                // sum_5x5 Jxy, telescoped: E = B[o-3]+B[o-2]-B[o+2]-B[o+3] per column
                const uint2 b3 = ring.b[(u + 2) & 7][lane], b2 = ring.b[(u + 3) & 7][lane];  // rows o-3 = r-9, o-2 = r-8
                const __half2 eE = __hsub2(__hadd2(u2h(b3.x), u2h(b2.x)), __hadd2(u2h(bc.x), nE));
                const __half2 eO = __hsub2(__hadd2(u2h(b3.y), u2h(b2.y)), __hadd2(u2h(bc.y), nO));
                const __half2 LeE = shfl_up_h2(eE), LeO = shfl_up_h2(eO);  // (e-4,e-2) (e-3,e-1)
                const __half2 ReE = shfl_dn_h2(eE), ReO = shfl_dn_h2(eO);  // (e4,e6)  (e5,e7)
                const __half2 em2e0 = hi_lo(LeE, eE), em1e1 = hi_lo(LeO, eO);
                const __half2 e2e4 = hi_lo(eE, ReE), e3e5 = hi_lo(eO, ReO);
                const __half2 cE = __hsub2(__hadd2(LeO, em2e0), __hadd2(e2e4, e3e5));   // (c0, c2)
                const __half2 cO = __hsub2(__hadd2(em2e0, em1e1), __hadd2(e3e5, ReE));  // (c1, c3)
                out0 = okc0 ? shi_tomasi(A0, B0, __low2float(cE), thr) : 0.f;
                out1 = okc1 ? shi_tomasi(A1, B1, __low2float(cO), thr) : 0.f;
                out2 = okc2 ? shi_tomasi(A2, B2, __high2float(cE), thr) : 0.f;
                out3 = okc3 ? shi_tomasi(A3, B3, __high2float(cO), thr) : 0.f;
            }
        }
        if (store_lane) *reinterpret_cast<float4*>(R + (size_t)o * a.R_pitch + x) = make_float4(out0, out1, out2, out3);
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

}  // namespace

cudaError_t launch_stencil(const StencilArgs& a, int n_frames, bool from_blurred, cudaStream_t st) {
    // blurred and R rows are written with 4-byte / 16-byte vector stores at columns that are multiples of 4
    if ((a.R_pitch & 3) || (a.R_fstride & 3) || (!from_blurred && ((a.blur_pitch & 3) || (a.blur_fstride & 3))))
        return cudaErrorInvalidValue;
    const int n_strips = (a.cols + SW - 1) / SW;
    int seg_rows = 48;
    const int n_segs = (a.rows + seg_rows - 1) / seg_rows;
    const dim3 grid((n_strips * n_segs + NWARP - 1) / NWARP, n_frames);
    if (from_blurred) k_stencil<true><<<grid, 32 * NWARP, 0, st>>>(a, n_strips, n_segs, seg_rows);
    else k_stencil<false><<<grid, 32 * NWARP, 0, st>>>(a, n_strips, n_segs, seg_rows);
    return cudaGetLastError();
}

cudaError_t launch_bgr2gray(const uint8_t* bgr, size_t bgr_pitch, uint8_t* gray, size_t gray_pitch, int rows, int cols,
                            cudaStream_t st) {
    const dim3 grid((cols + 255) / 256, rows);
    k_bgr2gray<<<grid, 256, 0, st>>>(bgr, bgr_pitch, gray, gray_pitch, rows, cols);
    return cudaGetLastError();
}
