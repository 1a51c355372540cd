// api.cu — the C ABI of libvofront.so (include/vofront.h): context, buffers, streams and the launch
// sequences.  No arithmetic of the hot path lives here; it only orchestrates the kernels of
// stencil.cu / nms_select.cu / freak.cu / match.cu.  There is no CPU fallback anywhere: every entry
// point either runs the CUDA kernels or returns an error code.
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include <algorithm>

#include <nvtx3/nvToolsExt.h>

#include "vo_internal.cuh"

namespace {

thread_local std::string g_create_err;

int fail(vo_ctx* c, int code, const std::string& msg) {
    if (c) c->err = msg; else g_create_err = msg;
    return code;
}
int fail_cuda(vo_ctx* c, cudaError_t e, const char* what) {
    return fail(c, VO_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}

#define CK(c, expr)                                                  \
    do {                                                             \
        cudaError_t e_ = (expr);                                     \
        if (e_ != cudaSuccess) return fail_cuda((c), e_, #expr);     \
    } while (0)

// kernel launch with launch counting and optional per-family event timing
// NVTX ranges around every launch, named by kernel family, when VO_NVTX is set in the environment (header-only NVTX3:
// nothing links; a tool that understands NVTX - ncu --nvtx --nvtx-include "stencil/" - can then select a family).
static const char* const kFamilyNames[] = {"stencil", "nms", "select", "freak", "match", "finalize", "misc", "other"};
static bool nvtx_on() {
    static const bool on = getenv("VO_NVTX") != nullptr;
    return on;
}

#define RUN(c, fam, expr)                                            \
    do {                                                             \
        ProfRec pr_{(fam), nullptr, nullptr};                        \
        const bool nvtx_ = nvtx_on();                                \
        if (nvtx_) nvtxRangePushA(kFamilyNames[(fam) & 7]);          \
        if ((c)->profiling) {                                        \
            cudaEventCreate(&pr_.a); cudaEventCreate(&pr_.b);        \
            cudaEventRecord(pr_.a, (c)->s_cur);                  \
        }                                                            \
        cudaError_t e_ = (expr);                                     \
        if (nvtx_) nvtxRangePop();                                   \
        if ((c)->profiling) {                                        \
            cudaEventRecord(pr_.b, (c)->s_cur);                  \
            (c)->prof.push_back(pr_);                                \
        }                                                            \
        (c)->launches++;                                             \
        if (e_ != cudaSuccess) return fail_cuda((c), e_, #expr);     \
    } while (0)

size_t round_up(size_t v, size_t m) { return (v + m - 1) / m * m; }

int check_params(vo_ctx* c, const vo_params* p) {
    if (!p) return fail(c, VO_ERR_BAD_ARG, "params is NULL");
    if (p->nms_window < 3 || p->nms_window > 9 || (p->nms_window & 1) == 0)
        return fail(c, VO_ERR_BAD_ARG, "nms_window must be odd and in 3..9");
    if (p->max_keypoints < 1 || p->max_keypoints > c->max_kp)
        return fail(c, VO_ERR_BAD_ARG, "max_keypoints out of range for this context");
    if (p->match_bits != VO_MATCH_REF32 && p->match_bits != VO_MATCH_FULL512)
        return fail(c, VO_ERR_BAD_ARG, "match_bits must be 32 or 512");
    if (!(p->threshold >= 0.0f)) return fail(c, VO_ERR_BAD_ARG, "threshold must be >= 0");
    if (p->response != VO_RESPONSE_SHI_TOMASI && p->response != VO_RESPONSE_HARRIS)
        return fail(c, VO_ERR_BAD_ARG, "response must be VO_RESPONSE_SHI_TOMASI or VO_RESPONSE_HARRIS");
    if (p->response == VO_RESPONSE_HARRIS && !(p->harris_k >= 0.0f && p->harris_k <= 1.0f))
        return fail(c, VO_ERR_BAD_ARG, "harris_k must be in [0, 1]");
    return VO_OK;
}
int check_geom(vo_ctx* c, int rows, int cols, size_t pitch, size_t min_pitch) {
    if (rows < 1 || cols < 1 || rows > c->max_rows || cols > c->max_cols)
        return fail(c, VO_ERR_BAD_ARG, "image size out of range for this context");
    if (pitch < min_pitch) return fail(c, VO_ERR_BAD_ARG, "pitch smaller than a row");
    return VO_OK;
}

// The response pool [total_slots][max_rows][r_pitch] exists only for the paths that materialise R: the stage entry
// points vo_response / vo_nms, VO_DBG_RESPONSE, and extraction with NMS windows 7 / 9 or on images too small for a
// candidate.  The fused extraction path never writes R, so the pool is allocated on first use.
int ensure_R(vo_ctx* c) {
    if (c->d_R) return VO_OK;
    CK(c, cudaMalloc(&c->d_R, (size_t)c->total_slots * c->r_fstride * sizeof(float)));
    return VO_OK;
}

// All stages for n frames whose gray images are at in_base + k*in_fstride; results go to slots
// first_slot .. first_slot+n-1.  Asynchronous on s_compute.
int run_extract(vo_ctx* c, const uint8_t* in_base, size_t in_fstride, size_t in_pitch, int n, int rows, int cols,
                const vo_params* p, int first_slot, bool from_blurred = false, bool stop_after_response = false) {
    const size_t so = (size_t)first_slot;
    CK(c, cudaMemsetAsync(c->d_cand_count + so, 0, sizeof(int) * n, c->s_cur));
    CK(c, cudaMemsetAsync(c->d_flags + so, 0, sizeof(int) * n, c->s_cur));
    // Fused path: the stencil kernel keeps the response on chip and appends the NMS candidates itself (no R round trip).
    const bool fused = !from_blurred && !stop_after_response && !c->force_unfused && p->response == VO_RESPONSE_SHI_TOMASI &&
                       stencil_can_fuse(rows, cols, p->nms_window);
    int rc_;
    if (!fused && (rc_ = ensure_R(c))) return rc_;
    StencilArgs sa;
    sa.in = in_base; sa.in_fstride = in_fstride; sa.in_pitch = in_pitch;
    sa.blur = c->d_blur + so * c->img_fstride; sa.blur_fstride = c->img_fstride; sa.blur_pitch = c->img_pitch;
    sa.R = fused ? nullptr : c->d_R + so * c->r_fstride; sa.R_fstride = c->r_fstride; sa.R_pitch = c->r_pitch;
    sa.rows = rows; sa.cols = cols; sa.threshold = p->threshold;
    sa.response = p->response; sa.harris_k = p->harris_k;
    sa.nms_h = fused ? p->nms_window / 2 : 0;
    sa.cand = c->d_cand + so * c->cand_cap; sa.cand_cap = c->cand_cap; sa.cand_count = c->d_cand_count + so;
    if (c->stencil_cap > 0) {
        sa.task_counter = c->d_task_counter + (c->task_counter_next++ & 7); sa.ctas_per_sm_cap = c->stencil_cap;
        CK(c, cudaMemsetAsync(sa.task_counter, 0, sizeof(int), c->s_cur));
    }
    RUN(c, FAM_STENCIL, launch_stencil(sa, n, from_blurred ? STENCIL_SRC_BLURRED : STENCIL_SRC_GRAY, c->s_cur));
    for (int k = 0; k < n; ++k) {
        c->slot_rows[first_slot + k] = rows; c->slot_cols[first_slot + k] = cols;
        c->slot_thr[first_slot + k] = p->threshold; c->slot_has_R[first_slot + k] = !fused;
    }
    if (stop_after_response) return VO_OK;
    NmsArgs na;
    na.R = sa.R; na.R_fstride = c->r_fstride; na.R_pitch = c->r_pitch;
    na.cand = sa.cand; na.cand_cap = c->cand_cap; na.cand_count = sa.cand_count;
    na.rows = rows; na.cols = cols; na.window = p->nms_window;
    if (!fused) RUN(c, FAM_NMS, launch_nms(na, n, c->s_cur));
    if (c->wait_before_select) {  // resident two-stream schedule: the previous chunk still reads this slot set's boundary frame
        CK(c, cudaStreamWaitEvent(c->s_cur, c->wait_before_select, 0));
        c->wait_before_select = nullptr;
    }
    SelectArgs se;
    se.cand = na.cand; se.cand_cap = c->cand_cap; se.cand_count = na.cand_count;
    se.kp = c->d_kp + so * c->max_kp; se.kp_stride = c->max_kp; se.nkp = c->d_nkp + so; se.flags = c->d_flags + so;
    se.max_keypoints = p->max_keypoints; se.sort_cap = c->sort_cap;
    RUN(c, FAM_SELECT, launch_select(se, n, c->s_cur));
    FreakArgs fa;
    fa.blur = from_blurred ? in_base : sa.blur;
    fa.blur_fstride = from_blurred ? in_fstride : c->img_fstride;
    fa.blur_pitch = from_blurred ? in_pitch : c->img_pitch;
    fa.kp = se.kp; fa.kp_stride = c->max_kp; fa.nkp = se.nkp;
    fa.desc = c->d_desc + so * c->max_kp * 16; fa.angles = c->d_angles + so * c->max_kp;
    fa.rows = rows; fa.cols = cols;
    RUN(c, FAM_FREAK, launch_freak(fa, n, p->max_keypoints, c->s_cur));
    return VO_OK;
}

// n_pairs (query slot, train slot) pairs -> pairs_out/nmatch_out (device), raw triples in c->d_raw.
int run_match(vo_ctx* c, const int* qslots, const int* tslots, int n_pairs, float ratio, int match_bits,
              int max_n1, int2* pairs_out, int* nmatch_out) {
    MatchArgs ma;
    ma.desc = c->d_desc; ma.kp_stride = c->max_kp; ma.nkp = c->d_nkp;
    ma.partial = c->d_partial; ma.raw = c->d_raw; ma.pairs = pairs_out; ma.nmatch = nmatch_out;
    ma.max_kp = c->max_kp; ma.ratio = ratio; ma.match_bits = match_bits;
    for (int k = 0; k < n_pairs; ++k) { ma.qslot[k] = (short)qslots[k]; ma.tslot[k] = (short)tslots[k]; }
    RUN(c, FAM_MATCH, launch_match(ma, n_pairs, max_n1, c->s_cur));
    RUN(c, FAM_FINALIZE, launch_match_finalize(ma, n_pairs, c->s_cur));
    return VO_OK;
}

int slot_ok(vo_ctx* c, int slot, int n = 1) {
    if (slot < 0 || n < 1 || slot + n > c->n_slots) return fail(c, VO_ERR_BAD_ARG, "slot range out of bounds");
    return VO_OK;
}

int check_flags(vo_ctx* c, const int* flags, int n) {
    for (int k = 0; k < n; ++k) {
        if (flags[k] & 1) return fail(c, VO_ERR_CAPACITY, "NMS candidate buffer overflow (raise threshold or max_keypoints)");
        if (flags[k] & 2) return fail(c, VO_ERR_BAD_ARG, "descriptor bytes must be 0 or 1");
    }
    return VO_OK;
}

int ensure_seq_results(vo_ctx* c, int n_frames) {
    if (n_frames <= c->seq_res_cap) return VO_OK;
    cudaFree(c->d_seq_nkp); cudaFree(c->d_seq_kp); cudaFree(c->d_seq_nmatch); cudaFree(c->d_seq_pairs); cudaFree(c->d_seq_flags);
    c->d_seq_nkp = nullptr; c->d_seq_kp = nullptr; c->d_seq_nmatch = nullptr; c->d_seq_pairs = nullptr; c->d_seq_flags = nullptr;
    c->seq_res_cap = 0;
    CK(c, cudaMalloc(&c->d_seq_nkp, sizeof(int) * n_frames));
    CK(c, cudaMalloc(&c->d_seq_flags, sizeof(int) * n_frames));
    CK(c, cudaMalloc(&c->d_seq_kp, sizeof(float2) * (size_t)n_frames * c->max_kp));
    CK(c, cudaMalloc(&c->d_seq_nmatch, sizeof(int) * n_frames));
    CK(c, cudaMalloc(&c->d_seq_pairs, sizeof(int2) * (size_t)n_frames * c->max_kp));
    c->seq_res_cap = n_frames;
    return VO_OK;
}

int ensure_host_flags(vo_ctx* c, int n, int** out) {
    if (n > c->h_flags_cap) {
        if (c->h_flags) cudaFreeHost(c->h_flags);
        c->h_flags = nullptr; c->h_flags_cap = 0;
        CK(c, cudaMallocHost(&c->h_flags, sizeof(int) * (size_t)n));
        c->h_flags_cap = n;
    }
    *out = c->h_flags;
    return VO_OK;
}

// One chunk of a sequence: frames [g0, g0+cnt) already in (or readable through) in_base; extract into the
// slot set (chunk parity), match pairs (g-1, g), and park the per-frame results in the sequence arrays.
int run_seq_chunk(vo_ctx* c, const uint8_t* in_base, size_t in_fstride, size_t in_pitch, int g0, int cnt, int batch,
                  int rows, int cols, const vo_params* p) {
    const int set = (g0 / batch) & 1, base = set * batch;
    int rc = run_extract(c, in_base, in_fstride, in_pitch, cnt, rows, cols, p, base);
    if (rc) return rc;
    CK(c, cudaMemcpyAsync(c->d_seq_nkp + g0, c->d_nkp + base, sizeof(int) * cnt, cudaMemcpyDeviceToDevice, c->s_cur));
    CK(c, cudaMemcpyAsync(c->d_seq_flags + g0, c->d_flags + base, sizeof(int) * cnt, cudaMemcpyDeviceToDevice, c->s_cur));
    CK(c, cudaMemcpyAsync(c->d_seq_kp + (size_t)g0 * c->max_kp, c->d_kp + (size_t)base * c->max_kp,
                          sizeof(float2) * (size_t)cnt * c->max_kp, cudaMemcpyDeviceToDevice, c->s_cur));
    int qs[VO_MAX_BATCH], ts[VO_MAX_BATCH], np = 0, first_pair = -1;
    for (int g = g0; g < g0 + cnt; ++g) {
        if (g == 0) continue;
        if (first_pair < 0) first_pair = g - 1;
        const int gq = g - 1;
        qs[np] = ((gq / batch) & 1) * batch + gq % batch;
        ts[np] = ((g / batch) & 1) * batch + g % batch;
        ++np;
    }
    if (np > 0) {
        rc = run_match(c, qs, ts, np, p->ratio, p->match_bits, p->max_keypoints,
                       c->d_seq_pairs + (size_t)first_pair * c->max_kp, c->d_seq_nmatch + first_pair);
        if (rc) return rc;
    }
    return VO_OK;
}

}  // namespace

void vo_internal_set_thread_error(const std::string& msg) { g_create_err = msg; }

extern "C" {

int vo_params_preset(vo_params* p, int preset) {
    if (!p) return VO_ERR_BAD_ARG;
    p->ratio = 0.75f;
    p->match_bits = VO_MATCH_REF32;
    p->response = VO_RESPONSE_SHI_TOMASI;
    p->harris_k = 0.04f;  // the usual Harris constant; read only when response == VO_RESPONSE_HARRIS
    if (preset == VO_PRESET_CPU_REF) { p->threshold = 25000.0f; p->nms_window = 5; p->max_keypoints = 1500; }
    else if (preset == VO_PRESET_VO_GPU_REF) { p->threshold = 20000.0f; p->nms_window = 3; p->max_keypoints = 2000; }
    else return VO_ERR_BAD_ARG;
    return VO_OK;
}

int vo_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

int vo_create(vo_ctx** out, int device, int max_rows, int max_cols, int max_keypoints, int n_slots) {
    return vo_create_ex(out, device, max_rows, max_cols, max_keypoints, n_slots, 0);
}

int vo_create_ex(vo_ctx** out, int device, int max_rows, int max_cols, int max_keypoints, int n_slots, int cand_cap) {
    if (!out) return fail(nullptr, VO_ERR_BAD_ARG, "out is NULL");
    if (cand_cap < 0 || (cand_cap > 0 && cand_cap < max_keypoints)) return fail(nullptr, VO_ERR_BAD_ARG, "cand_cap must be 0 (default) or >= max_keypoints");
    *out = nullptr;
    if (max_rows < 1 || max_cols < 1 || max_rows > 32768 || max_cols > 32768)
        return fail(nullptr, VO_ERR_BAD_ARG, "max_rows/max_cols out of range (1..32768)");
    if (max_keypoints < 1 || max_keypoints > 8192) return fail(nullptr, VO_ERR_BAD_ARG, "max_keypoints out of range (1..8192)");
    if (n_slots < 2 || n_slots > 2 * VO_MAX_BATCH) return fail(nullptr, VO_ERR_BAD_ARG, "n_slots out of range (2..2*VO_MAX_BATCH = 2048)");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(nullptr, VO_ERR_NO_DEVICE, "no CUDA device: libvofront has no CPU fallback");
    if (device < 0 || device >= ndev) return fail(nullptr, VO_ERR_BAD_ARG, "device index out of range");
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return fail(nullptr, VO_ERR_NO_DEVICE, "cannot query device");
    if (prop.major != 10) return fail(nullptr, VO_ERR_NO_DEVICE, "device is not sm_100 (kernels are built for sm_100a only)");

    vo_ctx* c = new vo_ctx();
    c->device = device; c->max_rows = max_rows; c->max_cols = max_cols; c->max_kp = max_keypoints;
    c->n_slots = n_slots; c->total_slots = n_slots + 2;  // two hidden slots: host-descriptor matching / stage probes
    c->sm_count = prop.multiProcessorCount;
    c->img_pitch = round_up((size_t)max_cols, 16);
    c->img_fstride = c->img_pitch * (size_t)max_rows;
    c->r_pitch = round_up((size_t)max_cols, 4);
    c->r_fstride = c->r_pitch * (size_t)max_rows;
    // NMS survivors per frame: at least 8 candidates per kept keypoint, and one per 32 pixels for large frames with a
    // small top-N (a 4K frame of the synthetic generator has ~60 k survivors); overflow is reported, never silent
    c->cand_cap = VO_CAND_CAP_MIN > 8 * max_keypoints ? VO_CAND_CAP_MIN : 8 * max_keypoints;
    const long long by_area = (long long)max_rows * max_cols / 32;
    if (by_area > c->cand_cap) c->cand_cap = (int)by_area;
    if (cand_cap > 0) c->cand_cap = cand_cap;  // vo_create_ex: the caller knows its images (dense textures at a low threshold)
    c->sort_cap = 64;  // k_select sorts whole warps' worth (64 keys) in registers
    while (c->sort_cap < max_keypoints) c->sort_cap <<= 1;
    const size_t ts = (size_t)c->total_slots;
    const int pairs_cap = c->total_slots < VO_MAX_BATCH ? c->total_slots : VO_MAX_BATCH;
    c->bytes_cap = 2 * (size_t)max_keypoints * 512;
    const size_t bgr_bytes = (size_t)max_rows * max_cols * 3;
    if (bgr_bytes > c->bytes_cap) c->bytes_cap = bgr_bytes;

#define CKC(expr)                                                                \
    do {                                                                         \
        cudaError_t e_ = (expr);                                                 \
        if (e_ != cudaSuccess) {                                                 \
            const int rc_ = fail_cuda(nullptr, e_, #expr);                       \
            vo_destroy(c);                                                       \
            return rc_;                                                          \
        }                                                                        \
    } while (0)
    CKC(cudaSetDevice(device));
    CKC(cudaMalloc(&c->d_img, ts * c->img_fstride + 16));  // + the words the stencil reads past an unpadded last row
    CKC(cudaMalloc(&c->d_blur, ts * c->img_fstride));
    CKC(cudaMalloc(&c->d_cand, ts * c->cand_cap * sizeof(unsigned long long)));
    CKC(cudaMalloc(&c->d_cand_count, ts * sizeof(int)));
    CKC(cudaMalloc(&c->d_kp, ts * max_keypoints * sizeof(float2)));
    CKC(cudaMalloc(&c->d_desc, ts * max_keypoints * 16 * sizeof(uint32_t)));
    CKC(cudaMalloc(&c->d_angles, ts * max_keypoints * sizeof(float)));
    CKC(cudaMalloc(&c->d_nkp, ts * sizeof(int)));
    CKC(cudaMalloc(&c->d_flags, ts * sizeof(int)));
    CKC(cudaMalloc(&c->d_partial, (size_t)pairs_cap * VO_MATCH_SPLITS * max_keypoints * sizeof(uint2)));
    CKC(cudaMalloc(&c->d_raw, (size_t)pairs_cap * max_keypoints * 3 * sizeof(int)));
    CKC(cudaMalloc(&c->d_pairs, (size_t)pairs_cap * max_keypoints * sizeof(int2)));
    CKC(cudaMalloc(&c->d_nmatch, (size_t)pairs_cap * sizeof(int)));
    CKC(cudaMalloc(&c->d_bytes, c->bytes_cap));
    CKC(cudaMemset(c->d_nkp, 0, ts * sizeof(int)));
    CKC(cudaMemset(c->d_flags, 0, ts * sizeof(int)));
    CKC(cudaMemset(c->d_cand_count, 0, ts * sizeof(int)));
    CKC(cudaMalloc(&c->d_task_counter, 8 * sizeof(int)));
    if (const char* e = getenv("VO_STENCIL_CAP")) c->stencil_cap = atoi(e);  // tuning: persistent stencil CTAs per SM (0 = one CTA per task)
    CKC(cudaMallocHost(&c->h_small, sizeof(int) * 4 * (2 * VO_MAX_BATCH + 8)));
    c->slot_rows = new int[c->total_slots]();
    c->slot_cols = new int[c->total_slots]();
    c->slot_thr = new float[c->total_slots]();
    c->slot_has_R = new bool[c->total_slots]();
    if (const char* e = getenv("VO_FORCE_UNFUSED")) c->force_unfused = e[0] == '1';  // test / measurement aid: R round trip + k_nms
    CKC(cudaStreamCreateWithFlags(&c->s_compute, cudaStreamNonBlocking));
    CKC(cudaStreamCreateWithFlags(&c->s_compute2, cudaStreamNonBlocking));
    c->s_cur = c->s_compute;
    for (int k = 0; k < 2; ++k) CKC(cudaEventCreateWithFlags(&c->ev_chunk[k], cudaEventDisableTiming));
    CKC(cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming));
    CKC(cudaStreamCreateWithFlags(&c->s_h2d, cudaStreamNonBlocking));
    CKC(cudaStreamCreateWithFlags(&c->s_d2h, cudaStreamNonBlocking));
    for (int k = 0; k < 2; ++k) {
        CKC(cudaEventCreateWithFlags(&c->ev_h2d[k], cudaEventDisableTiming));
        CKC(cudaEventCreateWithFlags(&c->ev_done[k], cudaEventDisableTiming));
    }
    CKC(cudaEventCreate(&c->ev_t0));
    CKC(cudaEventCreate(&c->ev_t1));
    CKC(freak_upload_tables());
    CKC(cudaDeviceSynchronize());
#undef CKC
    *out = c;
    return VO_OK;
}

void vo_destroy(vo_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->s_compute) cudaStreamSynchronize(c->s_compute);
    if (c->s_h2d) cudaStreamSynchronize(c->s_h2d);
    if (c->s_d2h) cudaStreamSynchronize(c->s_d2h);
    for (auto& r : c->prof) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
    cudaFree(c->d_img); cudaFree(c->d_blur); cudaFree(c->d_R); cudaFree(c->d_cand); cudaFree(c->d_cand_count); cudaFree(c->d_task_counter);
    cudaFree(c->d_kp); cudaFree(c->d_desc); cudaFree(c->d_angles); cudaFree(c->d_nkp); cudaFree(c->d_flags);
    cudaFree(c->d_partial); cudaFree(c->d_raw); cudaFree(c->d_pairs); cudaFree(c->d_nmatch); cudaFree(c->d_bytes);
    cudaFree(c->d_seq); cudaFree(c->d_seq_nkp); cudaFree(c->d_seq_kp); cudaFree(c->d_seq_nmatch); cudaFree(c->d_seq_pairs);
    cudaFree(c->d_seq_flags); cudaFree(c->d_stage);
    if (c->h_small) cudaFreeHost(c->h_small);
    if (c->h_flags) cudaFreeHost(c->h_flags);
    delete[] c->slot_rows; delete[] c->slot_cols; delete[] c->slot_thr; delete[] c->slot_has_R;
    for (int k = 0; k < 2; ++k) {
        if (c->ev_h2d[k]) cudaEventDestroy(c->ev_h2d[k]);
        if (c->ev_done[k]) cudaEventDestroy(c->ev_done[k]);
    }
    if (c->ev_t0) cudaEventDestroy(c->ev_t0);
    if (c->ev_t1) cudaEventDestroy(c->ev_t1);
    if (c->s_compute) cudaStreamDestroy(c->s_compute);
    if (c->s_compute2) cudaStreamDestroy(c->s_compute2);
    for (int k = 0; k < 2; ++k) if (c->ev_chunk[k]) cudaEventDestroy(c->ev_chunk[k]);
    if (c->ev_join) cudaEventDestroy(c->ev_join);
    if (c->s_h2d) cudaStreamDestroy(c->s_h2d);
    if (c->s_d2h) cudaStreamDestroy(c->s_d2h);
    delete c;
}

const char* vo_last_error(const vo_ctx* c) { return c ? c->err.c_str() : g_create_err.c_str(); }

int vo_sync(vo_ctx* c) {
    if (!c) return VO_ERR_BAD_ARG;
    CK(c, cudaSetDevice(c->device));
    CK(c, cudaStreamSynchronize(c->s_h2d));
    CK(c, cudaStreamSynchronize(c->s_compute));
    CK(c, cudaStreamSynchronize(c->s_d2h));
    return VO_OK;
}

void* vo_host_alloc(size_t bytes) {
    // VO_HOST_ALLOC=wc asks for write-combined pages (the host only ever WRITES frames into these buffers; the copy
    // engine reads them without snooping the CPU caches) - a tuning aid for the multi-GPU end-to-end rate, off by default
    // because reading such memory from the CPU is very slow.
    unsigned flags = cudaHostAllocDefault;
    if (const char* e = getenv("VO_HOST_ALLOC")) {
        if (strstr(e, "wc")) flags |= cudaHostAllocWriteCombined;
        if (strstr(e, "portable")) flags |= cudaHostAllocPortable;
    }
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes, flags) != cudaSuccess) return nullptr;
    return p;
}
void vo_host_free(void* p) { if (p) cudaFreeHost(p); }

// ---- extraction ------------------------------------------------------------------------------------
int vo_extract_batch(vo_ctx* c, const uint8_t* gray, int n, int rows, int cols, size_t pitch, size_t frame_stride,
                     const vo_params* p, int first_slot) {
    if (!c) return VO_ERR_BAD_ARG;
    if (!gray) return fail(c, VO_ERR_BAD_ARG, "image pointer is NULL");
    int rc;
    if ((rc = check_params(c, p)) || (rc = check_geom(c, rows, cols, pitch, (size_t)cols)) || (rc = slot_ok(c, first_slot, n))) return rc;
    if (n > VO_MAX_BATCH) return fail(c, VO_ERR_BAD_ARG, "batch larger than VO_MAX_BATCH");
    if (n > 1 && frame_stride < pitch * (size_t)rows) return fail(c, VO_ERR_BAD_ARG, "frame_stride smaller than a frame");
    CK(c, cudaSetDevice(c->device));
    uint8_t* dst = c->d_img + (size_t)first_slot * c->img_fstride;
    if (pitch == (size_t)cols) {
        // unpadded rows: one linear copy per frame, kept in the host layout (the stencil reads any pitch); a strided
        // 2-D copy of 1241-byte rows costs several times the linear one, most of all from pageable memory
        for (int k = 0; k < n; ++k)
            CK(c, cudaMemcpyAsync(dst + (size_t)k * c->img_fstride, gray + (size_t)k * frame_stride, (size_t)rows * cols,
                                  cudaMemcpyHostToDevice, c->s_compute));
        return run_extract(c, dst, c->img_fstride, (size_t)cols, n, rows, cols, p, first_slot);
    }
    for (int k = 0; k < n; ++k)
        CK(c, cudaMemcpy2DAsync(dst + (size_t)k * c->img_fstride, c->img_pitch, gray + (size_t)k * frame_stride, pitch,
                                (size_t)cols, (size_t)rows, cudaMemcpyHostToDevice, c->s_compute));
    return run_extract(c, dst, c->img_fstride, c->img_pitch, n, rows, cols, p, first_slot);
}

int vo_extract(vo_ctx* c, const uint8_t* gray, int rows, int cols, size_t pitch, const vo_params* p, int slot) {
    return vo_extract_batch(c, gray, 1, rows, cols, pitch, pitch * (size_t)(rows > 0 ? rows : 0), p, slot);
}

int vo_extract_bgr(vo_ctx* c, const uint8_t* bgr, int rows, int cols, size_t pitch, const vo_params* p, int slot) {
    if (!c) return VO_ERR_BAD_ARG;
    if (!bgr) return fail(c, VO_ERR_BAD_ARG, "image pointer is NULL");
    int rc;
    if ((rc = check_params(c, p)) || (rc = check_geom(c, rows, cols, pitch, (size_t)cols * 3)) || (rc = slot_ok(c, slot))) return rc;
    CK(c, cudaSetDevice(c->device));
    CK(c, cudaMemcpy2DAsync(c->d_bytes, (size_t)cols * 3, bgr, pitch, (size_t)cols * 3, (size_t)rows, cudaMemcpyHostToDevice,
                            c->s_compute));
    uint8_t* dst = c->d_img + (size_t)slot * c->img_fstride;
    RUN(c, FAM_MISC, launch_bgr2gray(c->d_bytes, (size_t)cols * 3, dst, c->img_pitch, rows, cols, c->s_compute));
    return run_extract(c, dst, c->img_fstride, c->img_pitch, 1, rows, cols, p, slot);
}

int vo_frame_fetch(vo_ctx* c, int slot, int* n_out, float* kp_xy, uint8_t* desc512, uint32_t* desc_packed) {
    if (!c) return VO_ERR_BAD_ARG;
    int rc;
    if ((rc = slot_ok(c, slot))) return rc;
    if (c->slot_rows[slot] == 0) return fail(c, VO_ERR_STATE, "slot holds no frame");
    CK(c, cudaSetDevice(c->device));
    CK(c, cudaMemcpyAsync(c->h_small, c->d_nkp + slot, sizeof(int), cudaMemcpyDeviceToHost, c->s_compute));
    CK(c, cudaMemcpyAsync(c->h_small + 1, c->d_flags + slot, sizeof(int), cudaMemcpyDeviceToHost, c->s_compute));
    CK(c, cudaStreamSynchronize(c->s_compute));
    const int n = c->h_small[0];
    if ((rc = check_flags(c, c->h_small + 1, 1))) return rc;
    if (n_out) *n_out = n;
    if (n > 0) {
        if (kp_xy) CK(c, cudaMemcpyAsync(kp_xy, c->d_kp + (size_t)slot * c->max_kp, sizeof(float2) * n, cudaMemcpyDeviceToHost, c->s_compute));
        if (desc_packed)
            CK(c, cudaMemcpyAsync(desc_packed, c->d_desc + (size_t)slot * c->max_kp * 16, sizeof(uint32_t) * 16 * n,
                                  cudaMemcpyDeviceToHost, c->s_compute));
        if (desc512) {
            RUN(c, FAM_MISC, launch_unpack_bytes(c->d_desc + (size_t)slot * c->max_kp * 16, n, c->d_bytes, c->s_compute));
            CK(c, cudaMemcpyAsync(desc512, c->d_bytes, (size_t)n * 512, cudaMemcpyDeviceToHost, c->s_compute));
        }
        CK(c, cudaStreamSynchronize(c->s_compute));
    }
    return VO_OK;
}

int vo_debug_fetch(vo_ctx* c, int slot, int what, void* dst) {
    if (!c) return VO_ERR_BAD_ARG;
    if (!dst) return fail(c, VO_ERR_BAD_ARG, "dst is NULL");
    if (slot < 0 || slot >= c->total_slots) return fail(c, VO_ERR_BAD_ARG, "slot out of range");
    const int rows = c->slot_rows[slot], cols = c->slot_cols[slot];
    if (rows == 0) return fail(c, VO_ERR_STATE, "slot holds no frame");
    CK(c, cudaSetDevice(c->device));
    if (what == VO_DBG_BLURRED) {
        CK(c, cudaMemcpy2DAsync(dst, (size_t)cols, c->d_blur + (size_t)slot * c->img_fstride, c->img_pitch, (size_t)cols,
                                (size_t)rows, cudaMemcpyDeviceToHost, c->s_compute));
    } else if (what == VO_DBG_RESPONSE) {
        int rc;
        if ((rc = ensure_R(c))) return rc;
        if (!c->slot_has_R[slot]) {
            // the fused extraction never materialised R: recompute it from the slot's blurred image with the writing form
            // of the same kernel (same arithmetic), at the threshold the slot was extracted with
            StencilArgs sa;
            sa.in = c->d_blur + (size_t)slot * c->img_fstride; sa.in_fstride = c->img_fstride; sa.in_pitch = c->img_pitch;
            sa.blur = nullptr; sa.blur_fstride = c->img_fstride; sa.blur_pitch = c->img_pitch;
            sa.R = c->d_R + (size_t)slot * c->r_fstride; sa.R_fstride = c->r_fstride; sa.R_pitch = c->r_pitch;
            sa.rows = rows; sa.cols = cols; sa.threshold = c->slot_thr[slot];
            sa.nms_h = 0; sa.cand = nullptr; sa.cand_cap = 0; sa.cand_count = nullptr;
            RUN(c, FAM_STENCIL, launch_stencil(sa, 1, STENCIL_SRC_BLURRED, c->s_compute));
            c->slot_has_R[slot] = true;
        }
        CK(c, cudaMemcpy2DAsync(dst, (size_t)cols * 4, c->d_R + (size_t)slot * c->r_fstride, c->r_pitch * 4, (size_t)cols * 4,
                                (size_t)rows, cudaMemcpyDeviceToHost, c->s_compute));
    } else if (what == VO_DBG_ANGLES) {
        CK(c, cudaMemcpyAsync(c->h_small, c->d_nkp + slot, sizeof(int), cudaMemcpyDeviceToHost, c->s_compute));
        CK(c, cudaStreamSynchronize(c->s_compute));
        const int n = c->h_small[0];
        if (n > 0) CK(c, cudaMemcpyAsync(dst, c->d_angles + (size_t)slot * c->max_kp, sizeof(float) * n, cudaMemcpyDeviceToHost, c->s_compute));
    } else {
        return fail(c, VO_ERR_BAD_ARG, "unknown debug item");
    }
    CK(c, cudaStreamSynchronize(c->s_compute));
    return VO_OK;
}

// ---- stage entry points (host in, host out) use the hidden slot n_slots as scratch ------------------
int vo_blur(vo_ctx* c, const uint8_t* gray, int rows, int cols, uint8_t* blurred) {
    if (!c) return VO_ERR_BAD_ARG;
    if (!gray || !blurred) return fail(c, VO_ERR_BAD_ARG, "NULL buffer");
    int rc;
    if ((rc = check_geom(c, rows, cols, (size_t)cols, (size_t)cols))) return rc;
    CK(c, cudaSetDevice(c->device));
    const int slot = c->n_slots;
    uint8_t* dst = c->d_img + (size_t)slot * c->img_fstride;
    CK(c, cudaMemcpy2DAsync(dst, c->img_pitch, gray, (size_t)cols, (size_t)cols, (size_t)rows, cudaMemcpyHostToDevice, c->s_compute));
    vo_params p; vo_params_preset(&p, VO_PRESET_CPU_REF); p.max_keypoints = 1;
    if ((rc = run_extract(c, dst, c->img_fstride, c->img_pitch, 1, rows, cols, &p, slot, false, true))) return rc;
    return vo_debug_fetch(c, slot, VO_DBG_BLURRED, blurred);
}

int vo_response(vo_ctx* c, const uint8_t* blurred, int rows, int cols, float threshold, float* R) {
    vo_params p; vo_params_preset(&p, VO_PRESET_CPU_REF); p.threshold = threshold;
    return vo_response_p(c, blurred, rows, cols, &p, R);
}

int vo_response_p(vo_ctx* c, const uint8_t* blurred, int rows, int cols, const vo_params* pp, float* R) {
    if (!c) return VO_ERR_BAD_ARG;
    if (!blurred || !R || !pp) return fail(c, VO_ERR_BAD_ARG, "NULL buffer");
    int rc;
    if ((rc = check_geom(c, rows, cols, (size_t)cols, (size_t)cols))) return rc;
    vo_params p = *pp; p.max_keypoints = 1; p.nms_window = 5; p.match_bits = VO_MATCH_REF32;
    if ((rc = check_params(c, &p))) return rc;
    CK(c, cudaSetDevice(c->device));
    const int slot = c->n_slots;
    uint8_t* dst = c->d_img + (size_t)slot * c->img_fstride;
    CK(c, cudaMemcpy2DAsync(dst, c->img_pitch, blurred, (size_t)cols, (size_t)cols, (size_t)rows, cudaMemcpyHostToDevice, c->s_compute));
    if ((rc = run_extract(c, dst, c->img_fstride, c->img_pitch, 1, rows, cols, &p, slot, true, true))) return rc;
    return vo_debug_fetch(c, slot, VO_DBG_RESPONSE, R);
}

// ---- per-stage workers with the reference's interval semantics (façade classes of vofront_adapter.hpp) --------------
namespace {
struct DevBuf {  // stage calls are not on the hot path: they allocate what they need and free it on return
    void* p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
};
bool interval_ok(int rows, int cols, int r0, int r1, int c0, int c1) {
    return r0 >= 0 && c0 >= 0 && r1 <= rows && c1 <= cols && r0 <= r1 && c0 <= c1;
}
}  // namespace

int vo_bgr2gray(vo_ctx* c, const uint8_t* bgr, int rows, int cols, size_t pitch, uint8_t* gray) {
    if (!c) return VO_ERR_BAD_ARG;
    if (!bgr || !gray) return fail(c, VO_ERR_BAD_ARG, "NULL buffer");
    int rc;
    if ((rc = check_geom(c, rows, cols, pitch, (size_t)cols * 3))) return rc;
    CK(c, cudaSetDevice(c->device));
    CK(c, cudaMemcpy2DAsync(c->d_bytes, (size_t)cols * 3, bgr, pitch, (size_t)cols * 3, (size_t)rows, cudaMemcpyHostToDevice, c->s_compute));
    uint8_t* dst = c->d_img + (size_t)c->n_slots * c->img_fstride;
    RUN(c, FAM_MISC, launch_bgr2gray(c->d_bytes, (size_t)cols * 3, dst, c->img_pitch, rows, cols, c->s_compute));
    CK(c, cudaMemcpy2DAsync(gray, (size_t)cols, dst, c->img_pitch, (size_t)cols, (size_t)rows, cudaMemcpyDeviceToHost, c->s_compute));
    CK(c, cudaStreamSynchronize(c->s_compute));
    return VO_OK;
}

int vo_gradients(vo_ctx* c, const uint8_t* picture, int rows, int cols, size_t pitch, int r0, int r1, int c0, int c1, float* Jx, float* Jy,
                 float* Jxy) {
    if (!c) return VO_ERR_BAD_ARG;
    if (!picture || !Jx || !Jy || !Jxy) return fail(c, VO_ERR_BAD_ARG, "NULL buffer");
    int rc;
    if ((rc = check_geom(c, rows, cols, pitch, (size_t)cols))) return rc;
    if (!interval_ok(rows, cols, r0, r1, c0, c1)) return fail(c, VO_ERR_BAD_ARG, "interval outside the image");
    const int w = c1 - c0 - 2, h = r1 - r0 - 2;
    if (w <= 0 || h <= 0) return VO_OK;  // the reference's loops run zero times
    CK(c, cudaSetDevice(c->device));
    uint8_t* dimg = c->d_img + (size_t)c->n_slots * c->img_fstride;
    CK(c, cudaMemcpy2DAsync(dimg, c->img_pitch, picture, pitch, (size_t)cols, (size_t)rows, cudaMemcpyHostToDevice, c->s_compute));
    DevBuf g;
    const size_t plane = (size_t)rows * cols;
    CK(c, cudaMalloc(&g.p, 3 * plane * sizeof(float)));
    float* d = static_cast<float*>(g.p);
    RUN(c, FAM_MISC, launch_gradients_ref(dimg, c->img_pitch, r0, r1, c0, c1, d, d + plane, d + 2 * plane, cols, c->s_compute));
    // only the rectangle the worker writes goes back: the rest of the caller's matrices stays as it is
    float* dst[3] = {Jx, Jy, Jxy};
    const size_t off = (size_t)(r0 + 1) * cols + (c0 + 1);
    for (int k = 0; k < 3; ++k)
        CK(c, cudaMemcpy2DAsync(dst[k] + off, (size_t)cols * 4, d + k * plane + off, (size_t)cols * 4, (size_t)w * 4, (size_t)h,
                                cudaMemcpyDeviceToHost, c->s_compute));
    CK(c, cudaStreamSynchronize(c->s_compute));
    return VO_OK;
}

int vo_response_from_gradients(vo_ctx* c, const float* Jx, const float* Jy, const float* Jxy, int rows, int cols, int r0, int r1, int c0,
                               int c1, const vo_params* pp, float* R) {
    if (!c) return VO_ERR_BAD_ARG;
    if (!Jx || !Jy || !Jxy || !R || !pp) return fail(c, VO_ERR_BAD_ARG, "NULL buffer");
    if (rows < 1 || cols < 1) return fail(c, VO_ERR_BAD_ARG, "empty image");
    vo_params p = *pp; p.max_keypoints = 1; p.nms_window = 5; p.match_bits = VO_MATCH_REF32;
    int rc;
    if ((rc = check_params(c, &p))) return rc;
    if (!interval_ok(rows, cols, r0, r1, c0, c1)) return fail(c, VO_ERR_BAD_ARG, "interval outside the image");
    const int w = c1 - c0 - 4, h = r1 - r0 - 4;
    if (w <= 0 || h <= 0) return VO_OK;
    CK(c, cudaSetDevice(c->device));
    DevBuf g;
    const size_t plane = (size_t)rows * cols;
    CK(c, cudaMalloc(&g.p, 4 * plane * sizeof(float)));
    float* d = static_cast<float*>(g.p);
    const float* src[3] = {Jx, Jy, Jxy};
    for (int k = 0; k < 3; ++k) CK(c, cudaMemcpyAsync(d + k * plane, src[k], plane * sizeof(float), cudaMemcpyHostToDevice, c->s_compute));
    RUN(c, FAM_MISC, launch_response_ref(d, d + plane, d + 2 * plane, cols, r0, r1, c0, c1, p.threshold, p.response, p.harris_k, d + 3 * plane,
                                         c->s_compute));
    const size_t off = (size_t)(r0 + 2) * cols + (c0 + 2);
    CK(c, cudaMemcpy2DAsync(R + off, (size_t)cols * 4, d + 3 * plane + off, (size_t)cols * 4, (size_t)w * 4, (size_t)h, cudaMemcpyDeviceToHost,
                            c->s_compute));
    CK(c, cudaStreamSynchronize(c->s_compute));
    return VO_OK;
}

int vo_nms(vo_ctx* c, const float* R, int rows, int cols, int nms_window, int max_keypoints, float* kp_xy, int* n_out) {
    if (!c) return VO_ERR_BAD_ARG;
    if (!R || !kp_xy || !n_out) return fail(c, VO_ERR_BAD_ARG, "NULL buffer");
    vo_params p; vo_params_preset(&p, VO_PRESET_CPU_REF); p.nms_window = nms_window; p.max_keypoints = max_keypoints;
    int rc;
    if ((rc = check_params(c, &p)) || (rc = check_geom(c, rows, cols, (size_t)cols, (size_t)cols))) return rc;
    CK(c, cudaSetDevice(c->device));
    if ((rc = ensure_R(c))) return rc;
    const int slot = c->n_slots;
    const size_t so = (size_t)slot;
    CK(c, cudaMemcpy2DAsync(c->d_R + so * c->r_fstride, c->r_pitch * 4, R, (size_t)cols * 4, (size_t)cols * 4, (size_t)rows,
                            cudaMemcpyHostToDevice, c->s_compute));
    CK(c, cudaMemsetAsync(c->d_cand_count + so, 0, sizeof(int), c->s_compute));
    CK(c, cudaMemsetAsync(c->d_flags + so, 0, sizeof(int), c->s_compute));
    NmsArgs na;
    na.R = c->d_R + so * c->r_fstride; na.R_fstride = c->r_fstride; na.R_pitch = c->r_pitch;
    na.cand = c->d_cand + so * c->cand_cap; na.cand_cap = c->cand_cap; na.cand_count = c->d_cand_count + so;
    na.rows = rows; na.cols = cols; na.window = nms_window;
    if (!c->force_unfused && stencil_can_fuse(rows, cols, nms_window)) {
        // windows 3 and 5: the map is streamed through the suppression code of the fused stencil kernel, i.e. the code
        // the extraction path runs (the stand-alone k_nms serves windows 7 / 9 and VO_FORCE_UNFUSED)
        StencilArgs sa;
        sa.in = nullptr; sa.in_fstride = 0; sa.in_pitch = 0;
        sa.blur = nullptr; sa.blur_fstride = c->img_fstride; sa.blur_pitch = c->img_pitch;
        sa.R = c->d_R + so * c->r_fstride; sa.R_fstride = c->r_fstride; sa.R_pitch = c->r_pitch;
        sa.rows = rows; sa.cols = cols; sa.threshold = 0.f;
        sa.nms_h = nms_window / 2; sa.cand = na.cand; sa.cand_cap = c->cand_cap; sa.cand_count = na.cand_count;
        RUN(c, FAM_STENCIL, launch_stencil(sa, 1, STENCIL_SRC_RESPONSE, c->s_compute));
    } else {
        RUN(c, FAM_NMS, launch_nms(na, 1, c->s_compute));
    }
    SelectArgs se;
    se.cand = na.cand; se.cand_cap = c->cand_cap; se.cand_count = na.cand_count;
    se.kp = c->d_kp + so * c->max_kp; se.kp_stride = c->max_kp; se.nkp = c->d_nkp + so; se.flags = c->d_flags + so;
    se.max_keypoints = max_keypoints; se.sort_cap = c->sort_cap;
    RUN(c, FAM_SELECT, launch_select(se, 1, c->s_compute));
    CK(c, cudaMemcpyAsync(c->h_small, c->d_nkp + slot, sizeof(int), cudaMemcpyDeviceToHost, c->s_compute));
    CK(c, cudaMemcpyAsync(c->h_small + 1, c->d_flags + slot, sizeof(int), cudaMemcpyDeviceToHost, c->s_compute));
    CK(c, cudaStreamSynchronize(c->s_compute));
    if ((rc = check_flags(c, c->h_small + 1, 1))) return rc;
    const int n = c->h_small[0];
    *n_out = n;
    if (n > 0) {
        CK(c, cudaMemcpyAsync(kp_xy, c->d_kp + so * c->max_kp, sizeof(float2) * n, cudaMemcpyDeviceToHost, c->s_compute));
        CK(c, cudaStreamSynchronize(c->s_compute));
    }
    return VO_OK;
}

int vo_describe(vo_ctx* c, const uint8_t* blurred, int rows, int cols, const float* kp_xy, int n, uint8_t* desc512,
                float* angles) {
    if (!c) return VO_ERR_BAD_ARG;
    if (!blurred || (n > 0 && !kp_xy)) return fail(c, VO_ERR_BAD_ARG, "NULL buffer");
    int rc;
    if ((rc = check_geom(c, rows, cols, (size_t)cols, (size_t)cols))) return rc;
    if (n < 0 || n > c->max_kp) return fail(c, VO_ERR_BAD_ARG, "keypoint count out of range");
    for (int k = 0; k < n; ++k) {  // the reference relies on the NMS margins for in-bounds sampling; enforce them here
        const float x = kp_xy[2 * k], y = kp_xy[2 * k + 1];
        if (!(x >= 37 && x <= cols - 37 && y >= 35 && y <= rows - 35))
            return fail(c, VO_ERR_BAD_ARG, "keypoint outside the FREAK safety margins (37 <= x <= cols-37, 35 <= y <= rows-35)");
    }
    if (n == 0) return VO_OK;
    CK(c, cudaSetDevice(c->device));
    const int slot = c->n_slots;
    const size_t so = (size_t)slot;
    CK(c, cudaMemcpy2DAsync(c->d_blur + so * c->img_fstride, c->img_pitch, blurred, (size_t)cols, (size_t)cols, (size_t)rows,
                            cudaMemcpyHostToDevice, c->s_compute));
    CK(c, cudaMemcpyAsync(c->d_kp + so * c->max_kp, kp_xy, sizeof(float2) * n, cudaMemcpyHostToDevice, c->s_compute));
    CK(c, cudaMemcpyAsync(c->d_nkp + slot, &n, sizeof(int), cudaMemcpyHostToDevice, c->s_compute));
    FreakArgs fa;
    fa.blur = c->d_blur + so * c->img_fstride; fa.blur_fstride = c->img_fstride; fa.blur_pitch = c->img_pitch;
    fa.kp = c->d_kp + so * c->max_kp; fa.kp_stride = c->max_kp; fa.nkp = c->d_nkp + slot;
    fa.desc = c->d_desc + so * c->max_kp * 16; fa.angles = c->d_angles + so * c->max_kp;
    fa.rows = rows; fa.cols = cols;
    RUN(c, FAM_FREAK, launch_freak(fa, 1, n, c->s_compute));
    if (desc512) {
        RUN(c, FAM_MISC, launch_unpack_bytes(fa.desc, n, c->d_bytes, c->s_compute));
        CK(c, cudaMemcpyAsync(desc512, c->d_bytes, (size_t)n * 512, cudaMemcpyDeviceToHost, c->s_compute));
    }
    if (angles) CK(c, cudaMemcpyAsync(angles, fa.angles, sizeof(float) * n, cudaMemcpyDeviceToHost, c->s_compute));
    CK(c, cudaStreamSynchronize(c->s_compute));
    return VO_OK;
}

// ---- matching ----------------------------------------------------------------------------------------
static int match_common(vo_ctx* c, int qslot, int tslot, int max_n1, float ratio, int match_bits, int32_t* pairs, int cap,
                        int* n_out) {
    if (match_bits != VO_MATCH_REF32 && match_bits != VO_MATCH_FULL512) return fail(c, VO_ERR_BAD_ARG, "match_bits must be 32 or 512");
    if (!n_out || cap < 0 || (cap > 0 && !pairs)) return fail(c, VO_ERR_BAD_ARG, "bad output buffer");
    int rc;
    if ((rc = run_match(c, &qslot, &tslot, 1, ratio, match_bits, max_n1, c->d_pairs, c->d_nmatch))) return rc;
    CK(c, cudaMemcpyAsync(c->h_small, c->d_nmatch, sizeof(int), cudaMemcpyDeviceToHost, c->s_compute));
    CK(c, cudaMemcpyAsync(c->h_small + 1, c->d_nkp + qslot, sizeof(int), cudaMemcpyDeviceToHost, c->s_compute));
    CK(c, cudaMemcpyAsync(c->h_small + 2, c->d_flags + qslot, sizeof(int), cudaMemcpyDeviceToHost, c->s_compute));
    CK(c, cudaMemcpyAsync(c->h_small + 3, c->d_flags + tslot, sizeof(int), cudaMemcpyDeviceToHost, c->s_compute));
    CK(c, cudaStreamSynchronize(c->s_compute));
    if ((rc = check_flags(c, c->h_small + 2, 2))) return rc;
    const int n = c->h_small[0];
    c->last_match_n1 = c->h_small[1];
    *n_out = n;
    if (n > cap) return fail(c, VO_ERR_CAPACITY, "match output buffer too small");
    if (n > 0) {
        CK(c, cudaMemcpyAsync(pairs, c->d_pairs, sizeof(int2) * n, cudaMemcpyDeviceToHost, c->s_compute));
        CK(c, cudaStreamSynchronize(c->s_compute));
    }
    return VO_OK;
}

int vo_match(vo_ctx* c, int query_slot, int train_slot, float ratio, int match_bits, int32_t* pairs, int cap, int* n_out) {
    if (!c) return VO_ERR_BAD_ARG;
    int rc;
    if ((rc = slot_ok(c, query_slot)) || (rc = slot_ok(c, train_slot))) return rc;
    if (c->slot_rows[query_slot] == 0 || c->slot_rows[train_slot] == 0) return fail(c, VO_ERR_STATE, "slot holds no frame");
    CK(c, cudaSetDevice(c->device));
    return match_common(c, query_slot, train_slot, c->max_kp, ratio, match_bits, pairs, cap, n_out);
}

int vo_match_host(vo_ctx* c, const uint8_t* d1, int n1, const uint8_t* d2, int n2, int row_bytes, float ratio,
                  int match_bits, int32_t* pairs, int cap, int* n_out) {
    if (!c) return VO_ERR_BAD_ARG;
    if (row_bytes != 512 && row_bytes != 64) return fail(c, VO_ERR_BAD_ARG, "row_bytes must be 512 (unpacked) or 64 (packed)");
    if (n1 < 0 || n2 < 0 || n1 > c->max_kp || n2 > c->max_kp) return fail(c, VO_ERR_BAD_ARG, "descriptor count out of range");
    if ((n1 > 0 && !d1) || (n2 > 0 && !d2)) return fail(c, VO_ERR_BAD_ARG, "NULL descriptors");
    CK(c, cudaSetDevice(c->device));
    const int qslot = c->n_slots, tslot = c->n_slots + 1;
    const uint8_t* src[2] = {d1, d2};
    const int cnt[2] = {n1, n2};
    CK(c, cudaMemsetAsync(c->d_flags + qslot, 0, 2 * sizeof(int), c->s_compute));
    for (int k = 0; k < 2; ++k) {
        const int slot = c->n_slots + k;
        uint32_t* dd = c->d_desc + (size_t)slot * c->max_kp * 16;
        CK(c, cudaMemcpyAsync(c->d_nkp + slot, &cnt[k], sizeof(int), cudaMemcpyHostToDevice, c->s_compute));
        if (cnt[k] == 0) continue;
        if (row_bytes == 64) {
            CK(c, cudaMemcpyAsync(dd, src[k], (size_t)cnt[k] * 64, cudaMemcpyHostToDevice, c->s_compute));
        } else {
            uint8_t* stage = c->d_bytes + (size_t)k * c->max_kp * 512;
            CK(c, cudaMemcpyAsync(stage, src[k], (size_t)cnt[k] * 512, cudaMemcpyHostToDevice, c->s_compute));
            RUN(c, FAM_MISC, launch_pack_bytes(stage, cnt[k], dd, c->d_flags + slot, c->s_compute));
        }
    }
    return match_common(c, qslot, tslot, n1, ratio, match_bits, pairs, cap, n_out);
}

int vo_match_raw(vo_ctx* c, int32_t* raw, int n1) {
    if (!c) return VO_ERR_BAD_ARG;
    if (!raw || n1 < 0 || n1 > c->last_match_n1) return fail(c, VO_ERR_BAD_ARG, "n1 exceeds the last match's query count");
    if (n1 == 0) return VO_OK;
    CK(c, cudaSetDevice(c->device));
    CK(c, cudaMemcpyAsync(raw, c->d_raw, sizeof(int) * 3 * n1, cudaMemcpyDeviceToHost, c->s_compute));
    CK(c, cudaStreamSynchronize(c->s_compute));
    return VO_OK;
}

// ---- sequences -----------------------------------------------------------------------------------------
int vo_sequence_upload(vo_ctx* c, const uint8_t* frames, int n_frames, int rows, int cols, size_t pitch, size_t frame_stride) {
    if (!c) return VO_ERR_BAD_ARG;
    if (!frames || n_frames < 1) return fail(c, VO_ERR_BAD_ARG, "no frames");
    int rc;
    if ((rc = check_geom(c, rows, cols, pitch, (size_t)cols))) return rc;
    if (n_frames > 1 && frame_stride < pitch * (size_t)rows) return fail(c, VO_ERR_BAD_ARG, "frame_stride smaller than a frame");
    CK(c, cudaSetDevice(c->device));
    const size_t dp = round_up((size_t)cols, 16), fs = dp * (size_t)rows, need = fs * (size_t)n_frames;
    if (need > c->seq_cap) {
        cudaFree(c->d_seq); c->d_seq = nullptr; c->seq_cap = 0;
        CK(c, cudaMalloc(&c->d_seq, need));
        c->seq_cap = need;
    }
    if ((rc = ensure_seq_results(c, n_frames))) return rc;
    if (frame_stride == pitch * (size_t)rows) {
        CK(c, cudaMemcpy2DAsync(c->d_seq, dp, frames, pitch, (size_t)cols, (size_t)rows * n_frames, cudaMemcpyHostToDevice, c->s_compute));
    } else {
        for (int k = 0; k < n_frames; ++k)
            CK(c, cudaMemcpy2DAsync(c->d_seq + (size_t)k * fs, dp, frames + (size_t)k * frame_stride, pitch, (size_t)cols, (size_t)rows,
                                    cudaMemcpyHostToDevice, c->s_compute));
    }
    CK(c, cudaStreamSynchronize(c->s_compute));
    c->seq_frames = n_frames; c->seq_rows = rows; c->seq_cols = cols; c->seq_pitch = dp; c->seq_fstride = fs;
    return VO_OK;
}

int vo_sequence_run_resident(vo_ctx* c, const vo_params* p, int batch) {
    if (!c) return VO_ERR_BAD_ARG;
    if (c->seq_frames < 1) return fail(c, VO_ERR_STATE, "no sequence uploaded");
    int rc;
    if ((rc = check_params(c, p))) return rc;
    if (batch < 1 || 2 * batch > c->n_slots || batch > VO_MAX_BATCH) return fail(c, VO_ERR_BAD_ARG, "batch must be in 1..n_slots/2");
    CK(c, cudaSetDevice(c->device));
    // Chunks alternate between two streams so that the low-occupancy kernels of one chunk (select, finalize, every
    // kernel's tail) share the SMs with the stencil / NMS kernels of the next.  Chunk k uses slot set k&1 like chunk
    // k-2 on the same stream; the one cross-stream hazard is the boundary frame: select(k+1) and freak(k+1) overwrite
    // the slot set whose last frame match(k) / finalize(k) still read, so select(k+1) waits for chunk k's last
    // kernel (that also orders the shared matcher scratch).  With per-kernel profiling on, everything stays on one
    // stream so that the event pairs measure kernels, not overlaps.
    const bool two = !c->profiling && c->seq_frames > batch;
    if (two) {  // stream 2 starts after everything already queued on the main stream
        CK(c, cudaEventRecord(c->ev_join, c->s_compute));
        CK(c, cudaStreamWaitEvent(c->s_compute2, c->ev_join, 0));
    }
    int chunk = 0;
    rc = VO_OK;
    for (int g0 = 0; g0 < c->seq_frames && rc == VO_OK; g0 += batch, ++chunk) {
        const int cnt = c->seq_frames - g0 < batch ? c->seq_frames - g0 : batch;
        c->s_cur = (two && (chunk & 1)) ? c->s_compute2 : c->s_compute;
        c->wait_before_select = (two && chunk > 0) ? c->ev_chunk[(chunk - 1) & 1] : nullptr;
        rc = run_seq_chunk(c, c->d_seq + (size_t)g0 * c->seq_fstride, c->seq_fstride, c->seq_pitch, g0, cnt, batch,
                           c->seq_rows, c->seq_cols, p);
        if (rc == VO_OK && two && cudaEventRecord(c->ev_chunk[chunk & 1], c->s_cur) != cudaSuccess) rc = fail(c, VO_ERR_CUDA, "cudaEventRecord");
    }
    c->s_cur = c->s_compute;
    c->wait_before_select = nullptr;
    if (two) {  // join: whatever follows on the main stream (timer, fetch, the next pass) sees both streams' work
        if (cudaEventRecord(c->ev_join, c->s_compute2) != cudaSuccess || cudaStreamWaitEvent(c->s_compute, c->ev_join, 0) != cudaSuccess)
            return fail(c, VO_ERR_CUDA, "stream join");
    }
    return rc;
}

int vo_sequence_fetch(vo_ctx* c, vo_seq_result* out) {
    if (!c) return VO_ERR_BAD_ARG;
    if (!out) return fail(c, VO_ERR_BAD_ARG, "out is NULL");
    const int n = c->seq_frames;
    if (n < 1) return fail(c, VO_ERR_STATE, "no sequence uploaded");
    CK(c, cudaSetDevice(c->device));
    std::vector<int> flags(n);
    CK(c, cudaMemcpyAsync(flags.data(), c->d_seq_flags, sizeof(int) * n, cudaMemcpyDeviceToHost, c->s_compute));
    if (out->n_keypoints) CK(c, cudaMemcpyAsync(out->n_keypoints, c->d_seq_nkp, sizeof(int) * n, cudaMemcpyDeviceToHost, c->s_compute));
    if (out->kp_xy) CK(c, cudaMemcpyAsync(out->kp_xy, c->d_seq_kp, sizeof(float2) * (size_t)n * c->max_kp, cudaMemcpyDeviceToHost, c->s_compute));
    if (n > 1) {
        if (out->n_matches) CK(c, cudaMemcpyAsync(out->n_matches, c->d_seq_nmatch, sizeof(int) * (n - 1), cudaMemcpyDeviceToHost, c->s_compute));
        if (out->pairs) CK(c, cudaMemcpyAsync(out->pairs, c->d_seq_pairs, sizeof(int2) * (size_t)(n - 1) * c->max_kp, cudaMemcpyDeviceToHost, c->s_compute));
    }
    CK(c, cudaStreamSynchronize(c->s_compute));
    c->last_seq_flags = flags;
    return check_flags(c, flags.data(), n);
}

int vo_sequence_flags(vo_ctx* c, int32_t* flags, int n) {
    if (!c) return VO_ERR_BAD_ARG;
    if (!flags || n < 0 || n > (int)c->last_seq_flags.size()) return fail(c, VO_ERR_BAD_ARG, "n exceeds the last sequence's frame count");
    for (int k = 0; k < n; ++k) flags[k] = c->last_seq_flags[k];
    return VO_OK;
}

int vo_sequence_run(vo_ctx* c, const uint8_t* frames, int n_frames, int rows, int cols, size_t pitch, size_t frame_stride,
                    const vo_params* p, int batch, vo_seq_result* out) {
    if (!c) return VO_ERR_BAD_ARG;
    if (!frames || n_frames < 1 || !out) return fail(c, VO_ERR_BAD_ARG, "no frames / no result struct");
    int rc;
    if ((rc = check_params(c, p)) || (rc = check_geom(c, rows, cols, pitch, (size_t)cols))) return rc;
    if (batch < 1 || 2 * batch > c->n_slots || batch > VO_MAX_BATCH) return fail(c, VO_ERR_BAD_ARG, "batch must be in 1..n_slots/2");
    if (n_frames > 1 && frame_stride < pitch * (size_t)rows) return fail(c, VO_ERR_BAD_ARG, "frame_stride smaller than a frame");
    CK(c, cudaSetDevice(c->device));
    if ((rc = ensure_seq_results(c, n_frames))) return rc;
    int* flags = nullptr;
    if ((rc = ensure_host_flags(c, n_frames, &flags))) return rc;  // pinned: a pageable D2H target would block the host
    // Frames that lie back to back in host memory go up with ONE linear copy per chunk into a staging
    // area that keeps the host layout (the stencil kernel reads any pitch); a strided 2-D copy of
    // 1241-byte rows runs at a fraction of PCIe bandwidth.
    const bool contiguous = frame_stride == pitch * (size_t)rows;
    if (contiguous) {
        const size_t need = 2 * (size_t)batch * frame_stride;
        if (need > c->stage_cap) {
            CK(c, cudaStreamSynchronize(c->s_compute));
            cudaFree(c->d_stage); c->d_stage = nullptr; c->stage_cap = 0;
            CK(c, cudaMalloc(&c->d_stage, need + 16));  // the stencil reads whole words: up to 7 bytes past a row
            c->stage_cap = need;
        }
    }
    int chunk = 0;
    for (int g0 = 0; g0 < n_frames; g0 += batch, ++chunk) {
        const int cnt = n_frames - g0 < batch ? n_frames - g0 : batch;
        const int set = chunk & 1, base = set * batch;
        uint8_t* dst = contiguous ? c->d_stage + (size_t)set * batch * frame_stride : c->d_img + (size_t)base * c->img_fstride;
        const size_t dst_fstride = contiguous ? frame_stride : c->img_fstride, dst_pitch = contiguous ? pitch : c->img_pitch;
        // H2D of chunk `chunk` may start once the kernels of chunk-2 (same slot set) have finished
        if (chunk >= 2) CK(c, cudaStreamWaitEvent(c->s_h2d, c->ev_done[set], 0));
        if (contiguous) {
            CK(c, cudaMemcpyAsync(dst, frames + (size_t)g0 * frame_stride, (size_t)cnt * frame_stride, cudaMemcpyHostToDevice, c->s_h2d));
        } else {
            for (int k = 0; k < cnt; ++k)
                CK(c, cudaMemcpy2DAsync(dst + (size_t)k * c->img_fstride, c->img_pitch, frames + (size_t)(g0 + k) * frame_stride, pitch,
                                        (size_t)cols, (size_t)rows, cudaMemcpyHostToDevice, c->s_h2d));
        }
        CK(c, cudaEventRecord(c->ev_h2d[set], c->s_h2d));
        CK(c, cudaStreamWaitEvent(c->s_compute, c->ev_h2d[set], 0));
        if ((rc = run_seq_chunk(c, dst, dst_fstride, dst_pitch, g0, cnt, batch, rows, cols, p))) return rc;
        CK(c, cudaEventRecord(c->ev_done[set], c->s_compute));
        // D2H of this chunk's results overlaps the next chunk's kernels
        CK(c, cudaStreamWaitEvent(c->s_d2h, c->ev_done[set], 0));
        CK(c, cudaMemcpyAsync(flags + g0, c->d_seq_flags + g0, sizeof(int) * cnt, cudaMemcpyDeviceToHost, c->s_d2h));
        if (out->n_keypoints) CK(c, cudaMemcpyAsync(out->n_keypoints + g0, c->d_seq_nkp + g0, sizeof(int) * cnt, cudaMemcpyDeviceToHost, c->s_d2h));
        if (out->kp_xy)
            CK(c, cudaMemcpyAsync(out->kp_xy + (size_t)g0 * c->max_kp * 2, c->d_seq_kp + (size_t)g0 * c->max_kp,
                                  sizeof(float2) * (size_t)cnt * c->max_kp, cudaMemcpyDeviceToHost, c->s_d2h));
        const int p0 = g0 == 0 ? 0 : g0 - 1, np = g0 == 0 ? cnt - 1 : cnt;
        if (np > 0) {
            if (out->n_matches) CK(c, cudaMemcpyAsync(out->n_matches + p0, c->d_seq_nmatch + p0, sizeof(int) * np, cudaMemcpyDeviceToHost, c->s_d2h));
            if (out->pairs)
                CK(c, cudaMemcpyAsync(out->pairs + (size_t)p0 * c->max_kp * 2, c->d_seq_pairs + (size_t)p0 * c->max_kp,
                                      sizeof(int2) * (size_t)np * c->max_kp, cudaMemcpyDeviceToHost, c->s_d2h));
        }
    }
    CK(c, cudaStreamSynchronize(c->s_h2d));
    CK(c, cudaStreamSynchronize(c->s_compute));
    CK(c, cudaStreamSynchronize(c->s_d2h));
    c->last_seq_flags.assign(flags, flags + n_frames);
    return check_flags(c, flags, n_frames);
}

// ---- measurement ---------------------------------------------------------------------------------------
int vo_timer_start(vo_ctx* c) {
    if (!c) return VO_ERR_BAD_ARG;
    CK(c, cudaSetDevice(c->device));
    CK(c, cudaEventRecord(c->ev_t0, c->s_compute));
    return VO_OK;
}
int vo_timer_stop(vo_ctx* c, float* ms) {
    if (!c) return VO_ERR_BAD_ARG;
    if (!ms) return fail(c, VO_ERR_BAD_ARG, "ms is NULL");
    CK(c, cudaSetDevice(c->device));
    CK(c, cudaEventRecord(c->ev_t1, c->s_compute));
    CK(c, cudaEventSynchronize(c->ev_t1));
    CK(c, cudaEventElapsedTime(ms, c->ev_t0, c->ev_t1));
    return VO_OK;
}
int vo_bench_stencil(vo_ctx* c, const vo_params* p, int batch, int mode, int reps, float* ms_per_launch) {
    if (!c) return VO_ERR_BAD_ARG;
    if (!ms_per_launch || reps < 1) return fail(c, VO_ERR_BAD_ARG, "bad arguments");
    if (c->seq_frames < 1) return fail(c, VO_ERR_STATE, "no sequence uploaded");
    int rc;
    if ((rc = check_params(c, p))) return rc;
    if (batch < 1 || batch > c->n_slots || batch > VO_MAX_BATCH || batch > c->seq_frames)
        return fail(c, VO_ERR_BAD_ARG, "batch must be in 1..min(n_slots, frames uploaded)");
    const bool fused = mode == 1;
    if (mode != 0 && mode != 1) return fail(c, VO_ERR_BAD_ARG, "mode must be 0 (write R) or 1 (fused NMS)");
    if (fused && !stencil_can_fuse(c->seq_rows, c->seq_cols, p->nms_window)) return fail(c, VO_ERR_BAD_ARG, "this geometry / window has no fused form");
    CK(c, cudaSetDevice(c->device));
    if (!fused && (rc = ensure_R(c))) return rc;
    StencilArgs sa;
    sa.in = c->d_seq; sa.in_fstride = c->seq_fstride; sa.in_pitch = c->seq_pitch;
    sa.blur = c->d_blur; sa.blur_fstride = c->img_fstride; sa.blur_pitch = c->img_pitch;
    sa.R = fused ? nullptr : c->d_R; sa.R_fstride = c->r_fstride; sa.R_pitch = c->r_pitch;
    sa.rows = c->seq_rows; sa.cols = c->seq_cols; sa.threshold = p->threshold;
    sa.nms_h = fused ? p->nms_window / 2 : 0; sa.cand = c->d_cand; sa.cand_cap = c->cand_cap; sa.cand_count = c->d_cand_count;
    float total = 0.f;
    for (int r = 0; r < reps; ++r) {
        CK(c, cudaMemsetAsync(c->d_cand_count, 0, sizeof(int) * batch, c->s_compute));
        if (c->stencil_cap > 0) {
            sa.task_counter = c->d_task_counter + (c->task_counter_next++ & 7); sa.ctas_per_sm_cap = c->stencil_cap;
            CK(c, cudaMemsetAsync(sa.task_counter, 0, sizeof(int), c->s_compute));
        }
        CK(c, cudaEventRecord(c->ev_t0, c->s_compute));
        cudaError_t e = launch_stencil(sa, batch, STENCIL_SRC_GRAY, c->s_compute);
        c->launches++;
        if (e != cudaSuccess) return fail_cuda(c, e, "launch_stencil");
        CK(c, cudaEventRecord(c->ev_t1, c->s_compute));
        CK(c, cudaEventSynchronize(c->ev_t1));
        float ms = 0.f;
        CK(c, cudaEventElapsedTime(&ms, c->ev_t0, c->ev_t1));
        total += ms;
    }
    for (int k = 0; k < batch; ++k) { c->slot_rows[k] = 0; c->slot_has_R[k] = false; }  // the slots hold no complete frame record
    *ms_per_launch = total / reps;
    return VO_OK;
}

int vo_bench_candidates(vo_ctx* c, int batch, long long* total) {
    if (!c) return VO_ERR_BAD_ARG;
    if (!total || batch < 1 || batch > c->n_slots || batch > VO_MAX_BATCH) return fail(c, VO_ERR_BAD_ARG, "bad arguments");
    CK(c, cudaSetDevice(c->device));
    std::vector<int> h(batch);
    CK(c, cudaMemcpyAsync(h.data(), c->d_cand_count, sizeof(int) * batch, cudaMemcpyDeviceToHost, c->s_compute));
    CK(c, cudaStreamSynchronize(c->s_compute));
    long long sum = 0;
    for (int v : h) sum += v;
    *total = sum;
    return VO_OK;
}

int vo_bench_popc(vo_ctx* c, int mode, int reps, double* popc32_per_s) {
    if (!c) return VO_ERR_BAD_ARG;
    if (!popc32_per_s || reps < 1 || (mode != 0 && mode != 1)) return fail(c, VO_ERR_BAD_ARG, "bad arguments");
    CK(c, cudaSetDevice(c->device));
    int sms = 0;
    CK(c, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device));
    unsigned* scratch = reinterpret_cast<unsigned*>(c->d_cand_count);  // one word that is never written in practice
    double best = 0.0;
    for (int r = 0; r < reps + 1; ++r) {  // first launch is the warm-up
        double n = 0;
        CK(c, cudaEventRecord(c->ev_t0, c->s_compute));
        cudaError_t e = launch_popc_peak(mode, scratch, sms, 4096, c->s_compute, &n);
        c->launches++;
        if (e != cudaSuccess) return fail_cuda(c, e, "launch_popc_peak");
        CK(c, cudaEventRecord(c->ev_t1, c->s_compute));
        CK(c, cudaEventSynchronize(c->ev_t1));
        float ms = 0.f;
        CK(c, cudaEventElapsedTime(&ms, c->ev_t0, c->ev_t1));
        if (r > 0 && ms > 0.f) best = std::max(best, n / (ms * 1e-3));
    }
    *popc32_per_s = best;
    return VO_OK;
}

long long vo_launch_count(vo_ctx* c, int reset) {
    if (!c) return -1;
    const long long v = c->launches;
    if (reset) c->launches = 0;
    return v;
}
int vo_profile_enable(vo_ctx* c, int on) {
    if (!c) return VO_ERR_BAD_ARG;
    c->profiling = on != 0;
    return VO_OK;
}
int vo_profile_read(vo_ctx* c, float ms[8], long long launches[8], int reset) {
    if (!c) return VO_ERR_BAD_ARG;
    CK(c, cudaSetDevice(c->device));
    CK(c, cudaStreamSynchronize(c->s_compute));
    for (auto& r : c->prof) {
        float t = 0.f;
        if (cudaEventElapsedTime(&t, r.a, r.b) == cudaSuccess) { c->prof_ms[r.fam] += t; c->prof_n[r.fam] += 1; }
        cudaEventDestroy(r.a); cudaEventDestroy(r.b);
    }
    c->prof.clear();
    for (int k = 0; k < 8; ++k) {
        if (ms) ms[k] = k < FAM_COUNT ? c->prof_ms[k] : 0.f;
        if (launches) launches[k] = k < FAM_COUNT ? c->prof_n[k] : 0;
    }
    if (reset) for (int k = 0; k < FAM_COUNT; ++k) { c->prof_ms[k] = 0.f; c->prof_n[k] = 0; }
    return VO_OK;
}

}  // extern "C"
