// stencil.cu — K1: fused blur -> gradients -> 5x5 structure tensor -> Shi-Tomasi response, one tiled pass.
//
// Replaces, for a whole batch of frames per launch:
//   cv::GaussianBlur(u8, 7x7, sigma 0)              feature_extraction_parallel/feature_extraction_parallel.cpp:223
//   CornerDetectionParallel::direction_gradients_worker        corner_detection_parallel.cpp:30-58
//   CornerDetectionParallel::shitomasi_corner_detection_worker corner_detection_parallel.cpp:61-105
// with whole-image (single strip) semantics.  Everything up to the seven fp32 operations of the
// closed-form eigenvalue is integer and therefore exact; those seven use round-to-nearest intrinsics so
// no FMA contraction can occur, which makes R bit-identical to the reference.
//
// Data flow per CTA (tile TW x TH of R, halo 6): u8 tile -> smem; horizontal 7-tap (u16) -> smem;
// vertical 7-tap -> blurred u8 (smem + HBM); Jx/Jy/Jxy (s16) -> smem; vertical 5-sums (s32) -> smem;
// horizontal 5-sums + response -> HBM.  HBM traffic: 1 B/px in, 1 B/px blurred out, 4 B/px R out.
#include "vo_internal.cuh"

namespace {

constexpr int TW = 64, TH = 32, NT = 256;
constexpr int IN_W = TW + 12, IN_H = TH + 12;   // input tile with halo 6
constexpr int IN_P = IN_W + 4;                  // padded pitch (bytes), keeps rows 4-byte aligned
constexpr int HB_W = TW + 6, HB_H = TH + 12;    // horizontal-blur rows (all input rows), cols -3..TW+2
constexpr int B_W = TW + 6, B_H = TH + 6;       // blurred tile with halo 3
constexpr int B_P = B_W + 2;
constexpr int J_W = TW + 4, J_H = TH + 4;       // gradient tile with halo 2
constexpr int V_W = TW + 4, V_H = TH;           // vertical 5-sums

struct Smem {
    uint8_t in[IN_H][IN_P];
    uint16_t hb[HB_H][HB_W];
    uint8_t b[B_H][B_P];
    short jx[J_H][J_W], jy[J_H][J_W], jxy[J_H][J_W];
    int va[V_H][V_W], vb[V_H][V_W], vc[V_H][V_W];
};

__device__ __forceinline__ int reflect101(int p, int n) {
    if (n == 1) return 0;
    while (p < 0 || p >= n) p = p < 0 ? -p : 2 * n - 2 - p;
    return p;
}

// R = tr/2 - 0.5*sqrt(tr^2 - 4 det), det = a*b - c*c  (corner_detection_parallel.cpp:93-98).
// a, b, c are exact integers < 2^24, so the int->float conversions are exact; every operation below
// rounds once, to nearest, exactly like the scalar x86-64 code of the reference (the reference's final
// subtraction in double followed by narrowing equals one fp32 rounding: 53 >= 2*24+2).
__device__ __forceinline__ float shi_tomasi(int a, int b, int c, float thr) {
    const float fa = (float)a, fb = (float)b, fc = (float)c;
    const float det = __fsub_rn(__fmul_rn(fa, fb), __fmul_rn(fc, fc));
    const float tr = __fadd_rn(fa, fb);
    const float disc = __fsub_rn(__fmul_rn(tr, tr), __fmul_rn(4.0f, det));
    const float r = __fsub_rn(__fmul_rn(tr, 0.5f), __fmul_rn(0.5f, __fsqrt_rn(disc)));
    return r > thr ? r : 0.0f;  // NaN (disc < 0) compares false -> 0, like the reference
}

template <bool FROM_BLURRED>
__global__ void __launch_bounds__(NT) k_stencil(const StencilArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Smem& s = *reinterpret_cast<Smem*>(smem_raw);
    const int f = blockIdx.z;
    const int x0 = blockIdx.x * TW, y0 = blockIdx.y * TH;
    const int H = a.rows, W = a.cols;
    const int tid = threadIdx.x;
    const uint8_t* __restrict__ in = a.in + (size_t)f * a.in_fstride;

    if (!FROM_BLURRED) {
        // A. input tile, BORDER_REFLECT_101 applied on the coordinates
        for (int idx = tid; idx < IN_H * IN_W; idx += NT) {
            const int r = idx / IN_W, c = idx - r * IN_W;
            const int gy = reflect101(y0 - 6 + r, H), gx = reflect101(x0 - 6 + c, W);
            s.in[r][c] = in[(size_t)gy * a.in_pitch + gx];
        }
        __syncthreads();
        // B. horizontal 7-tap [2,7,14,18,14,7,2]; sum <= 64*255 fits u16
        for (int idx = tid; idx < HB_H * HB_W; idx += NT) {
            const int r = idx / HB_W, c = idx - r * HB_W;
            const uint8_t* p = &s.in[r][c];
            const int v = 2 * (p[0] + p[6]) + 7 * (p[1] + p[5]) + 14 * (p[2] + p[4]) + 18 * p[3];
            s.hb[r][c] = (uint16_t)v;
        }
        __syncthreads();
        // C. vertical 7-tap, (S + 2048) >> 12
        uint8_t* __restrict__ bl = a.blur + (size_t)f * a.blur_fstride;
        for (int idx = tid; idx < B_H * B_W; idx += NT) {
            const int r = idx / B_W, c = idx - r * B_W;
            const unsigned v = 2u * (s.hb[r][c] + s.hb[r + 6][c]) + 7u * (s.hb[r + 1][c] + s.hb[r + 5][c]) +
                               14u * (s.hb[r + 2][c] + s.hb[r + 4][c]) + 18u * s.hb[r + 3][c];
            const uint8_t o = (uint8_t)((v + 2048u) >> 12);
            s.b[r][c] = o;
            const int gy = y0 - 3 + r, gx = x0 - 3 + c;
            if (r >= 3 && r < TH + 3 && c >= 3 && c < TW + 3 && gy < H && gx < W) bl[(size_t)gy * a.blur_pitch + gx] = o;
        }
    } else {
        // the input already is the blurred image: only in-image values are ever used (gradients are
        // masked to 1 <= i <= H-2, 1 <= j <= W-2), so clamp the coordinates of the halo.
        for (int idx = tid; idx < B_H * B_W; idx += NT) {
            const int r = idx / B_W, c = idx - r * B_W;
            const int gy = min(max(y0 - 3 + r, 0), H - 1), gx = min(max(x0 - 3 + c, 0), W - 1);
            s.b[r][c] = in[(size_t)gy * a.in_pitch + gx];
        }
    }
    __syncthreads();
    // D. gradients at (y0-2+r, x0-2+c); zero outside 1..H-2 x 1..W-2 (cv::Mat::zeros borders)
    for (int idx = tid; idx < J_H * J_W; idx += NT) {
        const int r = idx / J_W, c = idx - r * J_W;
        const int gy = y0 - 2 + r, gx = x0 - 2 + c;
        int jx = 0, jy = 0, jxy = 0;
        if (gy >= 1 && gy <= H - 2 && gx >= 1 && gx <= W - 2) {
            const uint8_t* p1 = &s.b[r][c];      // row gy-1, cols gx-1..gx+1   (s.b origin is (y0-3, x0-3))
            const uint8_t* p2 = &s.b[r + 1][c];
            const uint8_t* p3 = &s.b[r + 2][c];
            const int sx0 = p1[0] - p3[0], sx1 = p1[1] - p3[1], sx2 = p1[2] - p3[2];
            const int sy0 = p1[0] + 2 * p2[0] + p3[0], sy2 = p1[2] + 2 * p2[2] + p3[2];
            jx = sx0 + 2 * sx1 + sx2;
            jy = sy0 - sy2;
            jxy = sx0 - sx2;
        }
        s.jx[r][c] = (short)jx; s.jy[r][c] = (short)jy; s.jxy[r][c] = (short)jxy;
    }
    __syncthreads();
    // E. vertical 5-sums of Jx^2, Jy^2, Jxy for output rows y0..y0+TH-1, cols x0-2..x0+TW+1
    for (int idx = tid; idx < V_H * V_W; idx += NT) {
        const int r = idx / V_W, c = idx - r * V_W;
        int sa = 0, sb = 0, sc = 0;
#pragma unroll
        for (int m = 0; m < 5; ++m) {
            const int vx = s.jx[r + m][c], vy = s.jy[r + m][c];
            sa += vx * vx; sb += vy * vy; sc += s.jxy[r + m][c];
        }
        s.va[r][c] = sa; s.vb[r][c] = sb; s.vc[r][c] = sc;
    }
    __syncthreads();
    // F. horizontal 5-sums + response; zero outside 2..H-3 x 2..W-3 (R_array zero-initialised)
    float* __restrict__ R = a.R + (size_t)f * a.R_fstride;
    for (int idx = tid; idx < TH * TW; idx += NT) {
        const int r = idx / TW, c = idx - r * TW;
        const int gy = y0 + r, gx = x0 + c;
        if (gy >= H || gx >= W) continue;
        float out = 0.0f;
        if (gy >= 2 && gy <= H - 3 && gx >= 2 && gx <= W - 3) {
            int sa = 0, sb = 0, sc = 0;
#pragma unroll
            for (int n = 0; n < 5; ++n) { sa += s.va[r][c + n]; sb += s.vb[r][c + n]; sc += s.vc[r][c + n]; }
            out = shi_tomasi(sa, sb, sc, a.threshold);
        }
        R[(size_t)gy * a.R_pitch + gx] = out;
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
    static bool configured_dev[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    bool& configured = configured_dev[dev & 63];
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(k_stencil<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(k_stencil<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
        if (e != cudaSuccess) return e;
        configured = true;
    }
    const dim3 grid((a.cols + TW - 1) / TW, (a.rows + TH - 1) / TH, n_frames);
    if (from_blurred) k_stencil<true><<<grid, NT, sizeof(Smem), st>>>(a);
    else k_stencil<false><<<grid, NT, sizeof(Smem), st>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_bgr2gray(const uint8_t* bgr, size_t bgr_pitch, uint8_t* gray, size_t gray_pitch, int rows, int cols,
                            cudaStream_t st) {
    const dim3 grid((cols + 255) / 256, rows);
    k_bgr2gray<<<grid, 256, 0, st>>>(bgr, bgr_pitch, gray, gray_pitch, rows, cols);
    return cudaGetLastError();
}
