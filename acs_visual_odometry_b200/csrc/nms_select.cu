// nms_select.cu — K2: non-maximum suppression + candidate compaction, top-N selection, (y, x) ordering.
//
// Replaces CornerDetectionParallel::non_maximum_suppression (corner_detection_parallel.cpp:166-208: strict
// local maximum of the k x k window inside the FREAK safety margins, max-heap, pop up to N => the N
// largest by the tuple (R, i, j)) and the manager's std::sort by (int y, int x)
// (feature_extraction_parallel.cpp:276-282).
//
//   k_nms     one thread per pixel of the margin box; candidates are appended to a per-frame buffer as
//             64-bit keys (float bits of R << 32 | i << 16 | j) with a warp-aggregated atomic.
//             R > 0 on every candidate, so the unsigned order of the keys equals the tuple order.
//   k_select  one CTA per frame: if more than N candidates, an 8-pass MSB radix select finds the N-th
//             largest key (keys are unique, so exactly N survive); survivors are reduced to their
//             position key (i << 16 | j), bitonic-sorted in shared memory and written as (x, y) floats.
#include "vo_internal.cuh"

namespace {

__global__ void __launch_bounds__(256) k_nms(const NmsArgs a) {
    const int f = blockIdx.z;
    const int H = a.rows, W = a.cols, h = a.window / 2;
    // i in [h, H-h) and 35 <= i <= H-35 ; j in [h, W-h) and 37 <= j <= W-37
    const int i_lo = max(h, 35), i_hi = min(H - h - 1, H - 35);
    const int j_lo = max(h, 37), j_hi = min(W - h - 1, W - 37);
    const int i = i_lo + blockIdx.y * blockDim.y + threadIdx.y;
    const int j = j_lo + blockIdx.x * blockDim.x + threadIdx.x;
    if (i > i_hi || j > j_hi) return;
    const float* __restrict__ R = a.R + (size_t)f * a.R_fstride;
    const float c = R[(size_t)i * a.R_pitch + j];
    if (!(c > 0.0f)) return;  // thresholded map: a zero centre can never be strictly greater than its neighbours
    bool is_max = true;
    for (int n = i - h; n <= i + h && is_max; ++n) {
        const float* row = R + (size_t)n * a.R_pitch;
        for (int m = j - h; m <= j + h; ++m)
            if (!(n == i && m == j) && row[m] >= c) { is_max = false; break; }
    }
    if (!is_max) return;
    const int pos = atomicAdd(&a.cand_count[f], 1);
    if (pos < a.cand_cap)
        a.cand[(size_t)f * a.cand_cap + pos] =
            ((unsigned long long)__float_as_uint(c) << 32) | ((unsigned long long)i << 16) | (unsigned long long)j;
}

constexpr int SEL_NT = 1024;

__global__ void __launch_bounds__(SEL_NT) k_select(const SelectArgs a) {
    extern __shared__ uint32_t s_pos[];  // sort_cap entries
    __shared__ int s_hist[256];
    __shared__ unsigned long long s_prefix;
    __shared__ int s_remaining, s_count;
    const int f = blockIdx.x, tid = threadIdx.x;
    const unsigned long long* __restrict__ cand = a.cand + (size_t)f * a.cand_cap;
    int M = a.cand_count[f];
    if (M > a.cand_cap) {
        if (tid == 0) atomicOr(&a.flags[f], 1);  // overflow is reported, never silent
        M = a.cand_cap;
    }
    const int N = a.max_keypoints;
    unsigned long long thresh = 0ull;  // keep keys >= thresh
    if (M > N) {
        if (tid == 0) { s_prefix = 0ull; s_remaining = N; }
        for (int shift = 56; shift >= 0; shift -= 8) {
            if (tid < 256) s_hist[tid] = 0;
            __syncthreads();
            const unsigned long long prefix = s_prefix;
            for (int q = tid; q < M; q += SEL_NT) {
                const unsigned long long key = cand[q];
                const bool in = (shift == 56) || ((key >> (shift + 8)) == (prefix >> (shift + 8)));
                if (in) atomicAdd(&s_hist[(int)((key >> shift) & 255ull)], 1);
            }
            __syncthreads();
            if (tid == 0) {
                int rem = s_remaining, d = 255;
                for (; d > 0; --d) {
                    if (s_hist[d] >= rem) break;
                    rem -= s_hist[d];
                }
                s_remaining = rem;  // rank of the wanted key inside digit d (1-based from the top)
                s_prefix = prefix | ((unsigned long long)d << shift);
            }
            __syncthreads();
        }
        thresh = s_prefix;
    }
    if (tid == 0) s_count = 0;
    for (int q = tid; q < a.sort_cap; q += SEL_NT) s_pos[q] = 0xFFFFFFFFu;
    __syncthreads();
    for (int q = tid; q < M; q += SEL_NT) {
        const unsigned long long key = cand[q];
        if (key >= thresh) {
            const int p = atomicAdd(&s_count, 1);
            if (p < a.sort_cap) s_pos[p] = (uint32_t)(key & 0xFFFFFFFFull);
        }
    }
    __syncthreads();
    const int n = min(s_count, N);
    // bitonic sort ascending over the smallest power of two >= n (padding = 0xFFFFFFFF sorts last)
    int len = 1;
    while (len < n) len <<= 1;
    for (int k = 2; k <= len; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int q = tid; q < len; q += SEL_NT) {
                const int partner = q ^ j;
                if (partner > q) {
                    const uint32_t x = s_pos[q], y = s_pos[partner];
                    const bool up = (q & k) == 0;
                    if ((x > y) == up) { s_pos[q] = y; s_pos[partner] = x; }
                }
            }
            __syncthreads();
        }
    float2* __restrict__ kp = a.kp + (size_t)f * a.kp_stride;
    for (int q = tid; q < n; q += SEL_NT) {
        const uint32_t p = s_pos[q];
        kp[q] = make_float2((float)(p & 0xFFFFu), (float)(p >> 16));  // (x = column j, y = row i)
    }
    if (tid == 0) a.nkp[f] = n;
}

}  // namespace

cudaError_t launch_nms(const NmsArgs& a, int n_frames, cudaStream_t st) {
    const int h = a.window / 2;
    const int ni = min(a.rows - h - 1, a.rows - 35) - max(h, 35) + 1;
    const int nj = min(a.cols - h - 1, a.cols - 37) - max(h, 37) + 1;
    if (ni <= 0 || nj <= 0) return cudaSuccess;  // image smaller than the margins: no candidates
    const dim3 block(64, 4), grid((nj + 63) / 64, (ni + 3) / 4, n_frames);
    k_nms<<<grid, block, 0, st>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_select(const SelectArgs& a, int n_frames, cudaStream_t st) {
    k_select<<<n_frames, SEL_NT, (size_t)a.sort_cap * sizeof(uint32_t), st>>>(a);
    return cudaGetLastError();
}
