// match.cu — K4: brute-force Hamming 2-NN matcher with Lowe ratio test.
//
// Replaces matchCustomBinaryDescriptorsThreadPool + hammingDistance
// (feature_matching_parallel/feature_matching_parallel.cpp:39-113).  Two distance definitions:
//   REF32   the reference thread-pool matcher popcounts only the first 32 descriptor bytes, i.e. tests
//           0..31 (:39-47) -> word 0 of the packed descriptor;
//   FULL512 all 512 tests (matching_serial.cpp:24-40) -> all 16 words.
// Per query the reference keeps the two smallest distances of the multiset and the FIRST index that
// attains the minimum (:82-92).  Both follow from the two smallest packed keys (dist << 16 | index):
// keys are unique, their order is (dist, index) lexicographic, so min1 carries (d1, first index) and
// min2's distance is the second smallest of the multiset.
//
//   k_match     grid (query tiles, VO_MATCH_SPLITS train slices, pairs).  One query per thread, query
//               words register-resident; train descriptors staged in shared memory and read with
//               broadcast loads; XOR + POPC on the integer pipe; partial (min1, min2) per slice.
//   k_finalize  one CTA per pair: merges the slices, applies `float(d1) < ratio * float(d2)` in fp32
//               (:95-97) and compacts accepted (query, train) pairs in query order with a block scan.
#include "vo_internal.cuh"

namespace {

constexpr int MT_NT = 128;    // queries per CTA
constexpr int MT_TILE = 128;  // train descriptors per shared-memory tile

__device__ __forceinline__ void push_key(uint32_t key, uint32_t& m1, uint32_t& m2) {
    m2 = min(m2, max(m1, key));
    m1 = min(m1, key);
}

template <int WORDS>
__global__ void __launch_bounds__(MT_NT) k_match(const MatchArgs a) {
    __shared__ __align__(16) uint32_t s_t[MT_TILE * WORDS];
    const int pair = blockIdx.z, split = blockIdx.y;
    const int qs = a.qslot[pair], ts = a.tslot[pair];
    const int n1 = a.nkp[qs], n2 = a.nkp[ts];
    if ((int)blockIdx.x * MT_NT >= n1) return;
    const int qi = blockIdx.x * MT_NT + threadIdx.x;
    const uint32_t* __restrict__ qd = a.desc + (size_t)qs * a.kp_stride * 16;
    const uint32_t* __restrict__ td = a.desc + (size_t)ts * a.kp_stride * 16;

    uint32_t q[WORDS];
#pragma unroll
    for (int w = 0; w < WORDS; ++w) q[w] = qi < n1 ? qd[(size_t)qi * 16 + w] : 0u;

    const int chunk = (n2 + VO_MATCH_SPLITS - 1) / VO_MATCH_SPLITS;
    const int j_begin = split * chunk, j_end = min(n2, j_begin + chunk);
    uint32_t m1 = 0xFFFFFFFFu, m2 = 0xFFFFFFFFu;
    for (int j0 = j_begin; j0 < j_end; j0 += MT_TILE) {
        const int cnt = min(MT_TILE, j_end - j0);
        __syncthreads();
        for (int idx = threadIdx.x; idx < cnt * WORDS; idx += MT_NT) {
            const int r = idx / WORDS, w = idx - r * WORDS;
            s_t[idx] = td[(size_t)(j0 + r) * 16 + w];
        }
        __syncthreads();
        for (int r = 0; r < cnt; ++r) {
            int d = 0;
#pragma unroll
            for (int w = 0; w < WORDS; ++w) d += __popc(q[w] ^ s_t[r * WORDS + w]);
            push_key(((uint32_t)d << 16) | (uint32_t)(j0 + r), m1, m2);
        }
    }
    if (qi < n1) a.partial[((size_t)pair * VO_MATCH_SPLITS + split) * a.max_kp + qi] = make_uint2(m1, m2);
}

// REF32 specialisation.  A compare is XOR + POPC + key + three min/max; the integer ALU (16 lanes per clock and
// scheduler) would be the limit long before the POPC unit, so the bookkeeping runs two compares per instruction on
// 16-bit keys: inside a tile of 128 train descriptors the key is dist << 7 | row (dist <= 32: 13 bits), even rows in
// the low and odd rows in the high half of a register, tracked with VIMNMX.U16x2; the products that build the keys
// run on the FMA pipe (IMAD).  After each tile the four 16-bit survivors are widened to the global dist << 16 | index
// keys.  Same (dist, index) order as the 32-bit keys, so the result is identical.
__device__ __forceinline__ uint32_t min2u(uint32_t a, uint32_t b) { uint32_t d; asm("min.u16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ uint32_t max2u(uint32_t a, uint32_t b) { uint32_t d; asm("max.u16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ void push_key2(uint32_t k2, uint32_t& m1, uint32_t& m2) {
    m2 = min2u(m2, max2u(m1, k2));
    m1 = min2u(m1, k2);
}
__device__ __forceinline__ void widen_key(uint32_t k16, int j0, uint32_t& m1, uint32_t& m2) {
    const uint32_t d = k16 >> 7;
    if (d <= 32u) push_key((d << 16) | (uint32_t)(j0 + (int)(k16 & 127u)), m1, m2);  // d > 32: empty slot or padding row
}

__global__ void __launch_bounds__(MT_NT) k_match_ref32(const MatchArgs a) {
    __shared__ __align__(16) uint32_t s_t[MT_TILE];
    const int pair = blockIdx.z, split = blockIdx.y;
    const int qs = a.qslot[pair], ts = a.tslot[pair];
    const int n1 = a.nkp[qs], n2 = a.nkp[ts];
    if ((int)blockIdx.x * MT_NT >= n1) return;
    const int qi = blockIdx.x * MT_NT + threadIdx.x;
    const uint32_t* __restrict__ td = a.desc + (size_t)ts * a.kp_stride * 16;
    const uint32_t q = qi < n1 ? a.desc[((size_t)qs * a.kp_stride + qi) * 16] : 0u;

    const int chunk = (n2 + VO_MATCH_SPLITS - 1) / VO_MATCH_SPLITS;
    const int j_begin = split * chunk, j_end = min(n2, j_begin + chunk);
    uint32_t m1 = 0xFFFFFFFFu, m2 = 0xFFFFFFFFu;
    for (int j0 = j_begin; j0 < j_end; j0 += MT_TILE) {
        const int cnt = min(MT_TILE, j_end - j0);
        __syncthreads();
        if ((int)threadIdx.x < cnt) s_t[threadIdx.x] = td[(size_t)(j0 + threadIdx.x) * 16];
        __syncthreads();
        uint32_t t1 = 0xFFFFFFFFu, t2 = 0xFFFFFFFFu;  // per-tile trackers: (even rows | odd rows << 16)
        uint32_t rc = 0x00010000u;                    // rows (r, r+1) of the running group
        const int full = cnt & ~3;
        const uint4* s4 = reinterpret_cast<const uint4*>(s_t);
#pragma unroll 4
        for (int r = 0; r < full; r += 4) {
            const uint4 v = s4[r >> 2];
            const uint32_t d0 = __popc(q ^ v.x), d1 = __popc(q ^ v.y), d2 = __popc(q ^ v.z), d3 = __popc(q ^ v.w);
            push_key2(d1 * 0x800000u + (d0 * 128u + rc), t1, t2);
            push_key2(d3 * 0x800000u + (d2 * 128u + (rc + 0x00020002u)), t1, t2);
            rc += 0x00040004u;
        }
        if (full < cnt) {  // last, partial group: rows beyond cnt get distance 63 and are dropped when widening
            const uint4 v = s4[full >> 2];
            const uint32_t d0 = __popc(q ^ v.x);
            const uint32_t d1 = full + 1 < cnt ? __popc(q ^ v.y) : 63u, d2 = full + 2 < cnt ? __popc(q ^ v.z) : 63u;
            push_key2(d1 * 0x800000u + (d0 * 128u + rc), t1, t2);
            push_key2(63u * 0x800000u + (d2 * 128u + (rc + 0x00020002u)), t1, t2);
        }
        widen_key(t1 & 0xFFFFu, j0, m1, m2); widen_key(t1 >> 16, j0, m1, m2);
        widen_key(t2 & 0xFFFFu, j0, m1, m2); widen_key(t2 >> 16, j0, m1, m2);
    }
    if (qi < n1) a.partial[((size_t)pair * VO_MATCH_SPLITS + split) * a.max_kp + qi] = make_uint2(m1, m2);
}

constexpr int FN_NT = 1024;

__global__ void __launch_bounds__(FN_NT) k_finalize(const MatchArgs a) {
    __shared__ int s_warp[FN_NT / 32];
    __shared__ int s_base;
    const int pair = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n1 = a.nkp[a.qslot[pair]], n2 = a.nkp[a.tslot[pair]];
    int2* __restrict__ out = a.pairs + (size_t)pair * a.max_kp;
    int* __restrict__ raw = a.raw + (size_t)pair * a.max_kp * 3;
    if (tid == 0) s_base = 0;
    __syncthreads();
    for (int q0 = 0; q0 < n1; q0 += FN_NT) {
        const int qi = q0 + tid;
        bool accept = false;
        int best_idx = -1;
        if (qi < n1) {
            uint32_t m1 = 0xFFFFFFFFu, m2 = 0xFFFFFFFFu;
            if (n2 > 0) {
#pragma unroll
                for (int s = 0; s < VO_MATCH_SPLITS; ++s) {
                    const uint2 p = a.partial[((size_t)pair * VO_MATCH_SPLITS + s) * a.max_kp + qi];
                    push_key(p.x, m1, m2);
                    push_key(p.y, m1, m2);
                }
            }
            const bool has1 = m1 != 0xFFFFFFFFu, has2 = m2 != 0xFFFFFFFFu;
            const int d1 = has1 ? (int)(m1 >> 16) : 0x7FFFFFFF, d2 = has2 ? (int)(m2 >> 16) : 0x7FFFFFFF;
            best_idx = has1 ? (int)(m1 & 0xFFFFu) : -1;
            raw[(size_t)qi * 3 + 0] = d1; raw[(size_t)qi * 3 + 1] = best_idx; raw[(size_t)qi * 3 + 2] = d2;
            // bestIdx != -1 && secondBestIdx != -1 && bestDist < ratioThreshold * secondBestDist   (int < float*int)
            accept = has1 && has2 && ((float)d1 < __fmul_rn(a.ratio, (float)d2));
        }
        const unsigned bal = __ballot_sync(0xFFFFFFFFu, accept);
        if (lane == 0) s_warp[warp] = __popc(bal);
        __syncthreads();
        int before = 0, total = 0;
        for (int w = 0; w < FN_NT / 32; ++w) {
            const int c = s_warp[w];
            before += w < warp ? c : 0;
            total += c;
        }
        const int base = s_base;
        if (accept) out[base + before + __popc(bal & ((1u << lane) - 1u))] = make_int2(qi, best_idx);
        __syncthreads();
        if (tid == 0) s_base = base + total;
        __syncthreads();
    }
    if (tid == 0) a.nmatch[pair] = s_base;
}

// POPC32 issue rate of this GPU: the denominator of the matcher's roofline.  Every thread runs
// eight independent chains.  WITH_XOR_ADD = the matcher's own instruction mix per compare (XOR, POPC, two adds).
template <bool WITH_XOR_ADD>
__global__ void k_popc_peak(unsigned* out, unsigned seed, int iters) {
    unsigned a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = seed * (threadIdx.x + 1) + i * 0x9E3779B9u + blockIdx.x;
    unsigned acc = 0;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (WITH_XOR_ADD) { acc += __popc(a[i] ^ acc); a[i] += 0x01010101u; }
            else a[i] = __popc(a[i]) + a[i];
        }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) acc += a[i];
    if (acc == 0x12345678u) out[0] = acc;  // never true in practice: keeps the chains alive
}

}  // namespace

cudaError_t launch_popc_peak(int mode, unsigned* scratch, int sms, int iters, cudaStream_t st, double* popc_per_launch) {
    const int threads = 256, blocks = sms * 8;
    if (mode == 0) k_popc_peak<false><<<blocks, threads, 0, st>>>(scratch, 12345u, iters);
    else k_popc_peak<true><<<blocks, threads, 0, st>>>(scratch, 12345u, iters);
    *popc_per_launch = (double)blocks * threads * iters * 8;
    return cudaGetLastError();
}

cudaError_t launch_match(const MatchArgs& a, int n_pairs, int max_n1, cudaStream_t st) {
    if (n_pairs <= 0 || max_n1 <= 0) return cudaSuccess;
    const dim3 grid((max_n1 + MT_NT - 1) / MT_NT, VO_MATCH_SPLITS, n_pairs);
    if (a.match_bits == VO_MATCH_REF32) k_match_ref32<<<grid, MT_NT, 0, st>>>(a);
    else k_match<16><<<grid, MT_NT, 0, st>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_match_finalize(const MatchArgs& a, int n_pairs, cudaStream_t st) {
    if (n_pairs <= 0) return cudaSuccess;
    k_finalize<<<n_pairs, FN_NT, 0, st>>>(a);
    return cudaGetLastError();
}
