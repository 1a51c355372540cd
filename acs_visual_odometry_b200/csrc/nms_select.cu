// nms_select.cu — K2: non-maximum suppression + candidate compaction, top-N selection, (y, x) ordering.
//
// Replaces CornerDetectionParallel::non_maximum_suppression (corner_detection_parallel.cpp:166-208: strict
// local maximum of the k x k window inside the FREAK safety margins, max-heap, pop up to N => the N
// largest by the tuple (R, i, j)) and the manager's std::sort by (int y, int x)
// (feature_extraction_parallel.cpp:276-282).
//
//   k_nms     16-byte loads, four pixels per lane, zero pixels rejected at once, the non-zero groups of a warp
//             compacted and tested one per lane; candidates are appended to a
//             per-frame buffer as 64-bit keys (float bits of R << 32 | i << 16 | j).
//             R > 0 on every candidate, so the unsigned order of the keys equals the tuple order.
//   k_select  one CTA per frame: if more than N candidates, an 8-pass MSB radix select finds the N-th
//             largest key (keys are unique, so exactly N survive); survivors are reduced to their
//             position key (i << 16 | j), bitonic-sorted in shared memory and written as (x, y) floats.
#include "vo_internal.cuh"

namespace {

constexpr int NMS_ROWS = 4;                       // rows per warp (all loads issued before the first test); 2, 6, 8 measured slower
constexpr int NMS_TILE_PX = 8 * NMS_ROWS * 128;   // pixels per CTA: 8 warps x NMS_ROWS rows x 128 columns
constexpr int NMS_STAGE = NMS_TILE_PX / 4;        // strict maxima are never 8-adjacent: at most one per 2x2 cell

// One warp scans NMS_ROWS row segments of 128 pixels with 16-byte loads, four pixels per lane, all loads issued
// before the first test.  A zero centre can never be a strict maximum (its neighbours are >= 0), so groups of zeros
// cost one vector load and a compare.  About one group in twelve holds a non-zero centre, so testing them in place
// would run the window code with two or three lanes active: instead the warp COMPACTS its non-zero groups (ballot +
// prefix popcount into a per-warp byte list in shared memory) and hands them out one per lane.  Every item is tested
// without data-dependent loops: the (2h+1) rows of the window are fetched as three aligned float4 each (L1/L2 hits:
// the lines were just read) and the maximum of the window minus its centre is formed with FMNMX; candidate <=>
// centre > that maximum, which is exactly "no neighbour >= centre" (corner_detection_parallel.cpp:181-190).
// Candidates are staged in shared memory and appended to the frame's list with one global atomic per CTA.
// (Measured slower: 8 rows per warp; barrier-free warps that pipeline several tiles and append with one global
// atomic per warp; fetching all window rows up front with shuffled column neighbours: profiles/README.md.)
template <int HW>
__global__ void __launch_bounds__(256) k_nms(const NmsArgs a) {
    __shared__ unsigned long long s_keys[NMS_STAGE];
    __shared__ unsigned char s_item[8][NMS_ROWS * 32];
    __shared__ int s_n, s_base;
    const int f = blockIdx.z;
    const int H = a.rows, W = a.cols;
    constexpr int h = HW;
    // i in [h, H-h) and 35 <= i <= H-35 ; j in [h, W-h) and 37 <= j <= W-37
    const int i_lo = max(h, 35), i_hi = min(H - h - 1, H - 35);
    const int j_lo = max(h, 37), j_hi = min(W - h - 1, W - 37);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c0 = (j_lo & ~3) + 4 * (blockIdx.x * 32);  // first column of the warp's window (16-byte aligned)
    const int c4 = c0 + 4 * lane;                        // first of this lane's four columns
    const int i0 = i_lo + (blockIdx.y * 8 + warp) * NMS_ROWS;
    if (threadIdx.x == 0) s_n = 0;
    __syncthreads();
    const float* __restrict__ R = a.R + (size_t)f * a.R_fstride;
    if (i0 <= i_hi) {  // warp-uniform
        float4 v[NMS_ROWS];
#pragma unroll
        for (int k = 0; k < NMS_ROWS; ++k)
            v[k] = (c4 <= j_hi && i0 + k <= i_hi) ? __ldg(reinterpret_cast<const float4*>(R + (size_t)(i0 + k) * a.R_pitch + c4))
                                                  : make_float4(0.f, 0.f, 0.f, 0.f);
        int n = 0;
#pragma unroll
        for (int k = 0; k < NMS_ROWS; ++k) {
            const bool nz = v[k].x > 0.f || v[k].y > 0.f || v[k].z > 0.f || v[k].w > 0.f;
            const unsigned b = __ballot_sync(0xFFFFFFFFu, nz);
            if (nz) s_item[warp][n + __popc(b & ((1u << lane) - 1u))] = (unsigned char)(k * 32 + lane);  // NMS_ROWS * 32 <= 256
            n += __popc(b);
        }
        __syncwarp();
        for (int t = lane; t < n; t += 32) {
            const int item = s_item[warp][t];
            const int i = i0 + (item >> 5), cc = c0 + 4 * (item & 31);
            float m[4] = {0.f, 0.f, 0.f, 0.f};  // max over the window without the centre (all values >= 0)
            float cv[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int dn = -h; dn <= h; ++dn) {
                const float4* row = reinterpret_cast<const float4*>(R + (size_t)(i + dn) * a.R_pitch + cc);
                const float4 l = __ldg(row - 1), c = __ldg(row), r = __ldg(row + 1);
                const float w[12] = {l.x, l.y, l.z, l.w, c.x, c.y, c.z, c.w, r.x, r.y, r.z, r.w};  // columns cc-4 .. cc+7
                if (dn == 0) { cv[0] = c.x; cv[1] = c.y; cv[2] = c.z; cv[3] = c.w; }
#pragma unroll
                for (int e = 0; e < 4; ++e)
#pragma unroll
                    for (int dm = -h; dm <= h; ++dm)
                        if (!(dn == 0 && dm == 0)) m[e] = fmaxf(m[e], w[4 + e + dm]);
            }
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int j = cc + e;
                if (cv[e] > m[e] && j >= j_lo && j <= j_hi) {  // cv > m >= 0 implies cv > 0
                    const int sp = atomicAdd(&s_n, 1);  // < NMS_STAGE by the 2x2-cell bound
                    s_keys[sp] = ((unsigned long long)__float_as_uint(cv[e]) << 32) | ((unsigned long long)i << 16) | (unsigned long long)j;
                }
            }
        }
    }
    __syncthreads();
    const int cnt = s_n;
    if (cnt == 0) return;
    if (threadIdx.x == 0) s_base = atomicAdd(&a.cand_count[f], cnt);
    __syncthreads();
    const int base = s_base;
    for (int q = threadIdx.x; q < cnt; q += 256)
        if (base + q < a.cand_cap) a.cand[(size_t)f * a.cand_cap + base + q] = s_keys[q];
}

constexpr int SEL_NT = 1024;

// Whole-warp helpers of k_select (every lane of the warp must call them).
// digit 256 = this lane has nothing to add.  (Aggregating equal digits of a warp with __match_any_sync before the atomic
// was measured: 12 % slower than letting the shared-memory atomics serialise.)
__device__ __forceinline__ void hist_add(int* hist, int digit) {
    if (digit < 256) atomicAdd(&hist[digit], 1);
}
// Position of this lane's element in a list whose length lives in shared memory: one atomic per warp.
__device__ __forceinline__ int warp_append(int* counter, bool has) {
    const unsigned m = __ballot_sync(0xFFFFFFFFu, has);
    int base = 0;
    if (m != 0u && (threadIdx.x & 31) == __ffs(m) - 1) base = atomicAdd(counter, __popc(m));
    base = __shfl_sync(0xFFFFFFFFu, base, m ? __ffs(m) - 1 : 0);
    return base + __popc(m & ((1u << (threadIdx.x & 31)) - 1u));
}

__global__ void __launch_bounds__(SEL_NT) k_select(const SelectArgs a) {
    extern __shared__ uint32_t s_pos[];  // sort_cap entries
    __shared__ int s_hist[256];
    __shared__ unsigned long long s_prefix;
    __shared__ int s_remaining, s_count, s_in_prefix, s_list_n;
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
        // Radix select of the N-th largest key, one byte per pass from the top.  The first passes sweep all M candidates
        // from global memory (L2); as soon as the keys that still match the chosen prefix fit into the sort buffer
        // (sort_cap / 2 keys of 8 bytes, unused until the selection is over) the sweep that follows copies them there
        // while it counts, and every later pass reads that short list from shared memory: three global sweeps instead
        // of eight at 8000 keypoints per frame (the prefix narrows about a hundredfold per byte).
        unsigned long long* const s_keys = reinterpret_cast<unsigned long long*>(s_pos);
        const int list_cap = a.sort_cap / 2;
        if (tid == 0) { s_prefix = 0ull; s_remaining = N; s_in_prefix = M; s_list_n = 0; }
        bool in_smem = false;
        int list_n = 0;
        for (int shift = 56; shift >= 0; shift -= 8) {
            if (tid < 256) s_hist[tid] = 0;
            __syncthreads();
            const unsigned long long prefix = s_prefix;
            if (in_smem) {
                for (int base = 0; base < list_n; base += SEL_NT) {
                    const int q = base + tid;
                    const unsigned long long key = q < list_n ? s_keys[q] : 0ull;
                    const bool in = q < list_n && (key >> (shift + 8)) == (prefix >> (shift + 8));
                    hist_add(s_hist, in ? (int)((key >> shift) & 255ull) : 256);
                }
            } else {
                const bool compact = shift < 56 && s_in_prefix <= list_cap;  // uniform: read after the barrier
                // four keys in flight per thread: the sweep is a chain of L2 round trips otherwise
                for (int base = 0; base < M; base += 4 * SEL_NT) {  // uniform trip count: the warp votes inside
                    const int q0 = base + tid;
                    unsigned long long key[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) key[u] = q0 + u * SEL_NT < M ? cand[q0 + u * SEL_NT] : 0ull;
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const bool in = q0 + u * SEL_NT < M && ((shift == 56) || ((key[u] >> (shift + 8)) == (prefix >> (shift + 8))));
                        hist_add(s_hist, in ? (int)((key[u] >> shift) & 255ull) : 256);
                        if (compact) {
                            const int pos = warp_append(&s_list_n, in);
                            if (in) s_keys[pos] = key[u];
                        }
                    }
                }
                if (compact) { in_smem = true; list_n = s_in_prefix; }
            }
            __syncthreads();
            if (tid < 32) {
                // the digit holding the rem-th largest key: walk the histogram from the top.  Lane l owns bins 8l .. 8l+7;
                // a suffix scan over the lanes finds the owner, which walks its own eight bins (a single thread walking
                // all 256 bins eight times over was a third of the kernel at 8000 keypoints per frame)
                int h8[8], own = 0;
#pragma unroll
                for (int b = 0; b < 8; ++b) { h8[b] = s_hist[8 * tid + b]; own += h8[b]; }
                int incl = own;
#pragma unroll
                for (int off = 1; off < 32; off <<= 1) {
                    const int t = __shfl_down_sync(0xFFFFFFFFu, incl, off);
                    if (tid + off < 32) incl += t;
                }
                const int above = incl - own;  // keys in bins above this lane's
                const int rem = s_remaining;
                const bool mine = above < rem && rem <= incl;
                const unsigned who = __ballot_sync(0xFFFFFFFFu, mine);
                if (who == 0u) {  // fewer than rem keys under this prefix cannot happen while M > N; keep the serial walk's answer
                    if (tid == 0) { s_remaining = rem - (incl - h8[0]); s_prefix = prefix; s_in_prefix = h8[0]; }
                } else if (mine) {
                    int r = rem - above, d = 7;
                    for (; d > 0; --d) {
                        if (h8[d] >= r) break;
                        r -= h8[d];
                    }
                    s_remaining = r;  // rank of the wanted key inside digit d (1-based from the top)
                    s_prefix = prefix | ((unsigned long long)(8 * tid + d) << shift);
                    s_in_prefix = h8[d];  // keys that match the prefix extended by this digit
                }
            }
            __syncthreads();
        }
        thresh = s_prefix;
    }
    if (tid == 0) s_count = 0;
    for (int q = tid; q < a.sort_cap; q += SEL_NT) s_pos[q] = 0xFFFFFFFFu;
    __syncthreads();
    for (int base = 0; base < M; base += 4 * SEL_NT) {
        unsigned long long key[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) key[u] = base + tid + u * SEL_NT < M ? cand[base + tid + u * SEL_NT] : 0ull;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const bool keep = base + tid + u * SEL_NT < M && key[u] >= thresh;
            const int p = warp_append(&s_count, keep);
            if (keep && p < a.sort_cap) s_pos[p] = (uint32_t)(key[u] & 0xFFFFFFFFull);
        }
    }
    __syncthreads();
    const int n = min(s_count, N);
    // Bitonic sort ascending over the smallest power of two >= max(n, 64) (padding = 0xFFFFFFFF sorts last; sort_cap >= 64).
    // A thread holds two neighbouring keys, a warp 64: compare distances 1..32 run in registers (the pair itself, then
    // shuffles), only distances >= 64 go through shared memory with a CTA barrier: 15 barrier stages instead of 66 for
    // 2048 keys.
    int len = 64;
    while (len < n) len <<= 1;
    auto warp_stages = [&](uint32_t& x0, uint32_t& x1, int q0, int k, int j_from) {  // distances j_from, j_from/2, .., 1 of phase k
        const bool up = (q0 & k) == 0;                    // k >= 2: both keys of the thread lie in the same run
        for (int j = j_from; j >= 2; j >>= 1) {
            const uint32_t y0 = __shfl_xor_sync(0xFFFFFFFFu, x0, j >> 1), y1 = __shfl_xor_sync(0xFFFFFFFFu, x1, j >> 1);
            const bool take_min = ((q0 & j) == 0) == up;  // the lower index of the pair keeps the minimum in an ascending run
            x0 = take_min ? min(x0, y0) : max(x0, y0);
            x1 = take_min ? min(x1, y1) : max(x1, y1);
        }
        const uint32_t lo = min(x0, x1), hi = max(x0, x1);
        x0 = up ? lo : hi; x1 = up ? hi : lo;
    };
    for (int q0 = 2 * tid; q0 < len; q0 += 2 * SEL_NT) {  // warp-uniform: len is a multiple of 64; phases k = 2 .. 64
        uint2 x = *reinterpret_cast<const uint2*>(s_pos + q0);  // q0 even: 8-byte aligned
        for (int k = 2; k <= 64; k <<= 1) warp_stages(x.x, x.y, q0, k, k >> 1);
        *reinterpret_cast<uint2*>(s_pos + q0) = x;
    }
    __syncthreads();
    for (int k = 128; k <= len; k <<= 1) {
        int j = k >> 1;
        // two distances per barrier: a thread holds the four keys q, q + j/2, q + j, q + 3j/2 and does the two
        // compare-exchanges of distance j and the two of distance j/2 on them (all four lie in the same run of phase k)
        for (; j >= 128; j >>= 2) {
            const int h = j >> 1;
            for (int qd = tid; qd < (len >> 2); qd += SEL_NT) {
                const int q = ((qd & ~(h - 1)) << 2) | (qd & (h - 1));
                uint32_t a0 = s_pos[q], a1 = s_pos[q + h], a2 = s_pos[q + j], a3 = s_pos[q + j + h];
                const bool up = (q & k) == 0;
                uint32_t t;
                if ((a0 > a2) == up) { t = a0; a0 = a2; a2 = t; }
                if ((a1 > a3) == up) { t = a1; a1 = a3; a3 = t; }
                if ((a0 > a1) == up) { t = a0; a0 = a1; a1 = t; }
                if ((a2 > a3) == up) { t = a2; a2 = a3; a3 = t; }
                s_pos[q] = a0; s_pos[q + h] = a1; s_pos[q + j] = a2; s_pos[q + j + h] = a3;
            }
            __syncthreads();
        }
        for (; j >= 64; j >>= 1) {
            // one compare-exchange per thread and trip: pair p of this distance is (q, q + j) with q = p's bits above log2(j)
            // shifted up by one (every thread works; walking q and skipping the upper partners idled half of them)
            for (int pr = tid; pr < (len >> 1); pr += SEL_NT) {
                const int q = ((pr & ~(j - 1)) << 1) | (pr & (j - 1)), partner = q + j;
                const uint32_t x = s_pos[q], y = s_pos[partner];
                const bool up = (q & k) == 0;
                if ((x > y) == up) { s_pos[q] = y; s_pos[partner] = x; }
            }
            __syncthreads();
        }
        for (int q0 = 2 * tid; q0 < len; q0 += 2 * SEL_NT) {
            uint2 x = *reinterpret_cast<const uint2*>(s_pos + q0);
            warp_stages(x.x, x.y, q0, k, 32);
            *reinterpret_cast<uint2*>(s_pos + q0) = x;
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
    const int i_lo = max(h, 35), i_hi = min(a.rows - h - 1, a.rows - 35);
    const int j_lo = max(h, 37), j_hi = min(a.cols - h - 1, a.cols - 37);
    if (i_hi < i_lo || j_hi < j_lo) return cudaSuccess;  // image smaller than the margins: no candidates
    if (a.R_pitch & 3) return cudaErrorInvalidValue;     // rows are read with 16-byte loads
    const int ncol = j_hi - (j_lo & ~3) + 1, nrow = i_hi - i_lo + 1;
    const dim3 grid((ncol + 127) / 128, (nrow + 8 * NMS_ROWS - 1) / (8 * NMS_ROWS), n_frames);
    switch (h) {
        case 1: k_nms<1><<<grid, 256, 0, st>>>(a); break;
        case 2: k_nms<2><<<grid, 256, 0, st>>>(a); break;
        case 3: k_nms<3><<<grid, 256, 0, st>>>(a); break;
        case 4: k_nms<4><<<grid, 256, 0, st>>>(a); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

cudaError_t launch_select(const SelectArgs& a, int n_frames, cudaStream_t st) {
    k_select<<<n_frames, SEL_NT, (size_t)a.sort_cap * sizeof(uint32_t), st>>>(a);
    return cudaGetLastError();
}
