// match_static.cu — vo_match_static: the Hamming 2-NN ratio matcher for callers that must not allocate GPU memory.
//
// Every other entry point of libvofront.so works on a vo_ctx, and vo_create() sizes that context's pools with
// cudaMalloc (several hundred megabytes at KITTI shape, the matcher's per-slice scratch among them).  Some users of
// this library cannot do that:
//   - a process whose device memory is handed out in full by one arena at start-up (a robot's perception stack with
//     a fixed memory budget per module, an inference server whose allocator owns the GPU): a second allocator inside
//     a library either fails or eats into somebody else's budget;
//   - a control loop with a latency bound: cudaMalloc and cudaFree synchronise the whole device and take an
//     unpredictable time, so they are banned from the loop and often from the whole steady state;
//   - a plug-in that is handed device pointers and a stream by its host application and is not allowed to create
//     state of its own on the device.
// Such a caller already owns device buffers for its descriptors and its match list; what it lacks is the matcher's
// scratch (the two smallest keys of every query in each of the 8 train slices, written by the first kernel and merged
// by the second).  Here that scratch is a fixed array in the library's own device image, so the call needs nothing
// from any allocator.  The price: the array is sized for the maximum (8192 queries, 512 KB), every GPU that loads
// the library carries it until the process exits, and all calls on one GPU share it.
//
// The kernels are variants of k_match<WORDS> / k_finalize of match.cu (same keys, same merge, same fp32 ratio test,
// hence the same match list bit for bit) that take plain pointers instead of slot pools.  Nothing in match.cu, api.cu
// or any existing path is touched.  Reference behaviour: feature_matching_parallel/feature_matching_parallel.cpp:39-113
// (REF32) and matching_serial.cpp:24-77 (FULL512); this entry point itself has no counterpart in the reference.
#include <mutex>
#include <string>

#include "vo_internal.cuh"

namespace {

constexpr int SM_NT = 128;     // queries per CTA
constexpr int SM_TILE = 128;   // train descriptors per shared-memory tile
constexpr int SM_SPLITS = 8;   // train slices per query tile
constexpr int SM_FN_NT = 1024;
constexpr int SM_MAX_DEVICES = 64;

// The scratch: [SM_SPLITS][VO_STATIC_MATCH_MAX] pairs of packed keys (dist << 16 | train index), 512 KB.
// In HOST code the name g_static_partial is only a handle for cudaGetSymbolAddress - it is not a device address and
// must never be passed to a kernel.  scratch_for_current_device() below turns it into the real address, per GPU.
static __device__ __align__(256) uint2 g_static_partial[SM_SPLITS * VO_STATIC_MATCH_MAX];

__device__ __forceinline__ void push_key(uint32_t key, uint32_t& m1, uint32_t& m2) {
    m2 = min(m2, max(m1, key));
    m1 = min(m1, key);
}

// grid (query tiles, SM_SPLITS).  One query per thread with its words in registers; the train slice passes through
// shared memory in tiles.  Every (tile, slice) CTA writes its queries' two keys even when its slice is empty, so the
// finalize kernel never reads a slot left over from an earlier call: nothing in the scratch has to start at zero.
template <int WORDS>
__global__ void __launch_bounds__(SM_NT) k_match_static(const uint32_t* __restrict__ qd, int n1, const uint32_t* __restrict__ td,
                                                        int n2, uint2* __restrict__ partial) {
    __shared__ __align__(16) uint32_t s_t[SM_TILE * WORDS];
    const int split = blockIdx.y;
    const int qi = blockIdx.x * SM_NT + threadIdx.x;
    uint32_t q[WORDS];
#pragma unroll
    for (int w = 0; w < WORDS; ++w) q[w] = qi < n1 ? qd[(size_t)qi * 16 + w] : 0u;

    const int chunk = (n2 + SM_SPLITS - 1) / SM_SPLITS;
    const int j_begin = split * chunk, j_end = min(n2, j_begin + chunk);
    uint32_t m1 = 0xFFFFFFFFu, m2 = 0xFFFFFFFFu;
    for (int j0 = j_begin; j0 < j_end; j0 += SM_TILE) {
        const int cnt = min(SM_TILE, j_end - j0);
        __syncthreads();
        for (int idx = threadIdx.x; idx < cnt * WORDS; idx += SM_NT) {
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
    if (qi < n1) partial[(size_t)split * VO_STATIC_MATCH_MAX + qi] = make_uint2(m1, m2);
}

// One CTA: merge the slices, `float(d1) < ratio * float(d2)` in fp32 (feature_matching_parallel.cpp:95-97), compact
// the accepted (query, train) pairs in query order, write the count.
__global__ void __launch_bounds__(SM_FN_NT) k_finalize_static(const uint2* __restrict__ partial, int n1, int n2, float ratio,
                                                              int2* __restrict__ out, int* __restrict__ n_out) {
    __shared__ int s_warp[SM_FN_NT / 32];
    __shared__ int s_base;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_base = 0;
    __syncthreads();
    for (int q0 = 0; q0 < n1; q0 += SM_FN_NT) {
        const int qi = q0 + tid;
        bool accept = false;
        int best_idx = -1;
        if (qi < n1 && n2 > 0) {
            uint32_t m1 = 0xFFFFFFFFu, m2 = 0xFFFFFFFFu;
#pragma unroll
            for (int s = 0; s < SM_SPLITS; ++s) {
                const uint2 p = partial[(size_t)s * VO_STATIC_MATCH_MAX + qi];
                push_key(p.x, m1, m2);
                push_key(p.y, m1, m2);
            }
            const bool has1 = m1 != 0xFFFFFFFFu, has2 = m2 != 0xFFFFFFFFu;
            best_idx = has1 ? (int)(m1 & 0xFFFFu) : -1;
            accept = has1 && has2 && ((float)(int)(m1 >> 16) < __fmul_rn(ratio, (float)(int)(m2 >> 16)));
        }
        const unsigned bal = __ballot_sync(0xFFFFFFFFu, accept);
        if (lane == 0) s_warp[warp] = __popc(bal);
        __syncthreads();
        int before = 0, total = 0;
        for (int w = 0; w < SM_FN_NT / 32; ++w) {
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
    if (tid == 0) n_out[0] = s_base;
}

int fail_static(int code, const std::string& msg) {
    vo_internal_set_thread_error("vo_match_static: " + msg);
    return code;
}

// Set-up, once per GPU, with that GPU current: ask the runtime where THIS device's copy of the array lives.
std::mutex g_setup_mu;
uint2* g_scratch_addr[SM_MAX_DEVICES] = {};

cudaError_t scratch_for_current_device(int dev, uint2** out) {
    std::lock_guard<std::mutex> lk(g_setup_mu);
    if (!g_scratch_addr[dev]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_partial);
        if (e != cudaSuccess) return e;
        g_scratch_addr[dev] = static_cast<uint2*>(p);
    }
    *out = g_scratch_addr[dev];
    return cudaSuccess;
}

// A pointer this call may hand to a kernel running on `dev`: device (or managed) memory of that GPU.
bool on_device(const void* p, int dev) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return (at.type == cudaMemoryTypeDevice && at.device == dev) || at.type == cudaMemoryTypeManaged;
}

}  // namespace

int vo_match_static(const uint32_t* d_query, int n1, const uint32_t* d_train, int n2, float ratio, int match_bits,
                    int32_t* d_pairs, int32_t* d_n_out, void* stream) {
    if (match_bits != VO_MATCH_REF32 && match_bits != VO_MATCH_FULL512) return fail_static(VO_ERR_BAD_ARG, "match_bits must be 32 or 512");
    if (n1 < 0 || n2 < 0) return fail_static(VO_ERR_BAD_ARG, "negative descriptor count");
    if (n1 > VO_STATIC_MATCH_MAX || n2 > VO_STATIC_MATCH_MAX)
        return fail_static(VO_ERR_BAD_ARG, "descriptor count " + std::to_string(n1 > VO_STATIC_MATCH_MAX ? n1 : n2) +
                                               " exceeds VO_STATIC_MATCH_MAX (" + std::to_string(VO_STATIC_MATCH_MAX) +
                                               "), the size the built-in scratch array was compiled for; use vo_match_host on a context");
    if (!d_n_out || (n1 > 0 && (!d_query || !d_pairs)) || (n2 > 0 && !d_train)) return fail_static(VO_ERR_BAD_ARG, "NULL buffer");
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail_static(VO_ERR_NO_DEVICE, std::string("cudaGetDevice: ") + cudaGetErrorString(e));
    if (dev < 0 || dev >= SM_MAX_DEVICES) return fail_static(VO_ERR_BAD_ARG, "device ordinal out of range");
    uint2* scratch = nullptr;
    if ((e = scratch_for_current_device(dev, &scratch)) != cudaSuccess)
        return fail_static(VO_ERR_CUDA, std::string("cudaGetSymbolAddress(scratch): ") + cudaGetErrorString(e));
    if (!on_device(d_n_out, dev) || (n1 > 0 && (!on_device(d_query, dev) || !on_device(d_pairs, dev))) || (n2 > 0 && !on_device(d_train, dev)))
        return fail_static(VO_ERR_BAD_ARG, "every buffer must be device memory of the current GPU (host pointers are not accepted: "
                                           "staging them would need an allocation)");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (n1 > 0) {
        const dim3 grid((n1 + SM_NT - 1) / SM_NT, SM_SPLITS);
        if (match_bits == VO_MATCH_REF32) k_match_static<1><<<grid, SM_NT, 0, st>>>(d_query, n1, d_train, n2, scratch);
        else k_match_static<16><<<grid, SM_NT, 0, st>>>(d_query, n1, d_train, n2, scratch);
        if ((e = cudaGetLastError()) != cudaSuccess) return fail_static(VO_ERR_CUDA, std::string("k_match_static: ") + cudaGetErrorString(e));
    }
    k_finalize_static<<<1, SM_FN_NT, 0, st>>>(scratch, n1, n2, ratio, reinterpret_cast<int2*>(d_pairs), d_n_out);
    if ((e = cudaGetLastError()) != cudaSuccess) return fail_static(VO_ERR_CUDA, std::string("k_finalize_static: ") + cudaGetErrorString(e));
    return VO_OK;
}
