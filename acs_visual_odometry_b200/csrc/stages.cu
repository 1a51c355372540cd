// stages.cu — the reference's per-stage worker functions as stand-alone kernels, for the façade classes of
// include/vofront_adapter.hpp (CornerDetectionParallel::direction_gradients_worker / shitomasi_corner_detection_worker
// with their `interval` argument).  The hot path never launches these: extraction runs the fused k_stencil, which
// produces the same numbers without materialising gradients.  One thread per pixel, the reference's operation order,
// every fp32 operation rounded on its own (no contraction), so results equal the CPU code for ANY input, not only for
// gradients that came from an 8-bit image.
#include "vo_internal.cuh"

namespace {

// direction_gradients_worker (corner_detection_parallel.cpp:30-58): rows i in [r0+1, r1-1), columns j in [c0+1, c1-1):
//   sumx[k] = p[i-1][j+k] - p[i+1][j+k], sumy[k] = p[i-1][j+k] + 2 p[i][j+k] + p[i+1][j+k]   (integers, exact in fp32)
//   Jx = sumx[-1] + 2 sumx[0] + sumx[1],  Jy = sumy[-1] - sumy[1],  Jxy = sumx[-1] - sumx[1]
__global__ void k_gradients_ref(const uint8_t* __restrict__ img, size_t pitch, int r0, int r1, int c0, int c1,
                                float* __restrict__ jx, float* __restrict__ jy, float* __restrict__ jxy, int cols) {
    const int j = c0 + 1 + blockIdx.x * blockDim.x + threadIdx.x, i = r0 + 1 + blockIdx.y;
    if (j >= c1 - 1 || i >= r1 - 1) return;
    const uint8_t* a = img + (size_t)(i - 1) * pitch + j;
    const uint8_t* b = img + (size_t)i * pitch + j;
    const uint8_t* c = img + (size_t)(i + 1) * pitch + j;
    int sx[3], sy[3];
#pragma unroll
    for (int k = -1; k <= 1; ++k) {
        sx[k + 1] = (int)a[k] - (int)c[k];
        sy[k + 1] = (int)a[k] + 2 * (int)b[k] + (int)c[k];
    }
    const size_t at = (size_t)i * cols + j;
    jx[at] = (float)(sx[0] + 2 * sx[1] + sx[2]);
    jy[at] = (float)(sy[0] - sy[2]);
    jxy[at] = (float)(sx[0] - sx[2]);
}

// shitomasi_corner_detection_worker (corner_detection_parallel.cpp:61-105): rows [r0+2, r1-2), columns [c0+2, c1-2);
// 5x5 sums accumulated row by row, left to right; det, trace, discriminant and square root in fp32; the final
// subtraction in double, narrowed to float (:96).  response 1: Harris det - k * trace^2 on the same sums (vofront.h).
__global__ void k_response_ref(const float* __restrict__ jx, const float* __restrict__ jy, const float* __restrict__ jxy, int cols,
                               int r0, int r1, int c0, int c1, float threshold, int response, float harris_k, float* __restrict__ R) {
    const int j = c0 + 2 + blockIdx.x * blockDim.x + threadIdx.x, i = r0 + 2 + blockIdx.y;
    if (j >= c1 - 2 || i >= r1 - 2) return;
    float sumjxy = 0.f, jx2 = 0.f, jy2 = 0.f;
    for (int m = -2; m <= 2; ++m)
        for (int n = -2; n <= 2; ++n) {
            const size_t at = (size_t)(i + m) * cols + (j + n);
            const float vx = jx[at], vy = jy[at], vxy = jxy[at];
            sumjxy = __fadd_rn(sumjxy, vxy);
            jx2 = __fadd_rn(jx2, __fmul_rn(vx, vx));
            jy2 = __fadd_rn(jy2, __fmul_rn(vy, vy));
        }
    const float det = __fsub_rn(__fmul_rn(jx2, jy2), __fmul_rn(sumjxy, sumjxy));
    const float trace = __fadd_rn(jx2, jy2);
    float r;
    if (response == VO_RESPONSE_HARRIS) {
        r = __fsub_rn(det, __fmul_rn(__fmul_rn(harris_k, trace), trace));
    } else {
        const float disc = __fsub_rn(__fmul_rn(trace, trace), __fmul_rn(4.0f, det));
        const float sq = __fsqrt_rn(disc);  // NaN for a negative discriminant, like sqrtf
        r = (float)__dsub_rn((double)__fmul_rn(trace, 0.5f), __dmul_rn(0.5, (double)sq));
    }
    R[(size_t)i * cols + j] = r > threshold ? r : 0.0f;
}

}  // namespace

cudaError_t launch_gradients_ref(const uint8_t* img, size_t pitch, int r0, int r1, int c0, int c1, float* jx, float* jy, float* jxy,
                                 int cols, cudaStream_t st) {
    const int w = c1 - c0 - 2, h = r1 - r0 - 2;
    if (w <= 0 || h <= 0) return cudaSuccess;
    k_gradients_ref<<<dim3((w + 127) / 128, h), 128, 0, st>>>(img, pitch, r0, r1, c0, c1, jx, jy, jxy, cols);
    return cudaGetLastError();
}

cudaError_t launch_response_ref(const float* jx, const float* jy, const float* jxy, int cols, int r0, int r1, int c0, int c1,
                                float threshold, int response, float harris_k, float* R, cudaStream_t st) {
    const int w = c1 - c0 - 4, h = r1 - r0 - 4;
    if (w <= 0 || h <= 0) return cudaSuccess;
    k_response_ref<<<dim3((w + 127) / 128, h), 128, 0, st>>>(jx, jy, jxy, cols, r0, r1, c0, c1, threshold, response, harris_k, R);
    return cudaGetLastError();
}
