// vo_internal.cuh — context layout and kernel launch prototypes shared by the translation units of
// libvofront.so.  Not part of the public ABI (that is include/vofront.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/vofront.h"

#define VO_MAX_BATCH 1024       // frames (or pairs) per batched launch
#define VO_MATCH_SPLITS 8      // train-set slices per query tile (merged by the finalize kernel)
#define VO_CAND_CAP_MIN 32768  // NMS candidate slots per frame (>= 4 * max_keypoints)

enum { FAM_STENCIL = 0, FAM_NMS, FAM_SELECT, FAM_FREAK, FAM_MATCH, FAM_FINALIZE, FAM_MISC, FAM_COUNT };

struct ProfRec { int fam; cudaEvent_t a, b; };

// One frame record ("slot") = a stride into each of these pools.
struct vo_ctx {
    int device = 0, max_rows = 0, max_cols = 0, max_kp = 0, n_slots = 0, total_slots = 0;
    int sm_count = 0;
    size_t img_pitch = 0, img_fstride = 0;  // bytes
    size_t r_pitch = 0, r_fstride = 0;      // floats
    int cand_cap = 0, sort_cap = 0;         // per frame; sort_cap = pow2 >= max_kp

    uint8_t* d_img = nullptr;    // [total_slots][max_rows][img_pitch]  input gray
    uint8_t* d_blur = nullptr;   // same layout                         blurred gray (FREAK samples it)
    float* d_R = nullptr;        // [total_slots][max_rows][r_pitch]    thresholded response; allocated on first use (ensure_R)
    unsigned long long* d_cand = nullptr;  // [total_slots][cand_cap]   (Rbits<<32 | i<<16 | j)
    int* d_cand_count = nullptr; // [total_slots]
    int* d_task_counter = nullptr; int task_counter_next = 0, stencil_cap = 0;  // persistent stencil form (VO_STENCIL_CAP)
    float2* d_kp = nullptr;      // [total_slots][max_kp]               (x, y)
    uint32_t* d_desc = nullptr;  // [total_slots][max_kp][16]           packed descriptors
    float* d_angles = nullptr;   // [total_slots][max_kp]
    int* d_nkp = nullptr;        // [total_slots]
    int* d_flags = nullptr;      // [total_slots]  bit0: candidate overflow, bit1: bad descriptor byte
    int* slot_rows = nullptr;    // host: geometry of the frame held by each slot (0 = empty)
    int* slot_cols = nullptr;
    float* slot_thr = nullptr;   // host: response threshold the slot was extracted with
    bool* slot_has_R = nullptr;  // host: d_R holds the slot's response (unfused extraction / recomputed for VO_DBG_RESPONSE)
    bool force_unfused = false;  // VO_FORCE_UNFUSED=1: extraction writes R and runs k_nms (A/B measurements, parity of both paths)

    // matcher scratch, per pair of a batch
    uint2* d_partial = nullptr;  // [VO_MAX_BATCH][VO_MATCH_SPLITS][max_kp]  two smallest packed keys
    int* d_raw = nullptr;        // [VO_MAX_BATCH][max_kp][3]
    int2* d_pairs = nullptr;     // [VO_MAX_BATCH][max_kp]
    int* d_nmatch = nullptr;     // [VO_MAX_BATCH]
    int last_match_n1 = 0;
    std::vector<int> last_seq_flags;  // per-frame status bits of the last sequence call (vo_sequence_flags)
    int cand_cap_request = 0;         // vo_create_ex: candidates per frame asked for (0 = the default formula)

    uint8_t* d_bytes = nullptr;  // scratch: 2 x max_kp x 512 unpacked descriptor bytes / BGR staging
    size_t bytes_cap = 0;

    uint8_t* d_stage = nullptr; size_t stage_cap = 0;  // 2 x batch frames in host layout (vo_sequence_run)

    // resident sequence
    uint8_t* d_seq = nullptr; size_t seq_cap = 0;
    int seq_frames = 0, seq_rows = 0, seq_cols = 0; size_t seq_pitch = 0, seq_fstride = 0;
    int* d_seq_nkp = nullptr; float2* d_seq_kp = nullptr; int* d_seq_nmatch = nullptr; int2* d_seq_pairs = nullptr;
    int* d_seq_flags = nullptr;
    int seq_res_cap = 0;

    cudaStream_t s_compute = nullptr, s_h2d = nullptr, s_d2h = nullptr;
    // Resident sequences alternate their chunks between s_compute and s_compute2 (see vo_sequence_run_resident);
    // s_cur is the stream run_extract / run_match / run_seq_chunk launch on (s_compute everywhere else).
    cudaStream_t s_compute2 = nullptr, s_cur = nullptr;
    cudaEvent_t ev_chunk[2] = {nullptr, nullptr}, ev_join = nullptr;  // chunk k's kernels done (on stream k&1); pass join
    cudaEvent_t wait_before_select = nullptr;                         // set by the resident path for one run_extract call
    cudaEvent_t ev_h2d[2] = {nullptr, nullptr}, ev_done[2] = {nullptr, nullptr}, ev_t0 = nullptr, ev_t1 = nullptr;
    int* h_small = nullptr;  // pinned scratch for counts/flags
    int* h_flags = nullptr; int h_flags_cap = 0;  // pinned per-frame flags of vo_sequence_run

    long long launches = 0;
    bool profiling = false;
    std::vector<ProfRec> prof;
    float prof_ms[FAM_COUNT] = {0};
    long long prof_n[FAM_COUNT] = {0};
    std::string err;
};

// Message behind vo_last_error(NULL) for entry points that work without a context (match_static.cu).
void vo_internal_set_thread_error(const std::string& msg);

// ---- kernel launchers (each TU owns its kernels) ------------------------------------------------
struct StencilArgs {
    const uint8_t* in; size_t in_fstride, in_pitch;     // gray (or blurred when from_blurred)
    uint8_t* blur; size_t blur_fstride, blur_pitch;     // may be null when from_blurred
    float* R; size_t R_fstride, R_pitch;                // floats; unused (may be null) when nms_h > 0
    int rows, cols; float threshold;
    int response = 0; float harris_k = 0.f;             // VO_RESPONSE_*; HARRIS only with nms_h == 0
    // persistent form (tuning, VO_STENCIL_CAP): ctas_per_sm_cap x SMs single-warp CTAs fetch tasks from *task_counter
    // (zeroed by the caller) instead of one CTA per task, so that the kernel leaves room on every SM for other kernels
    int* task_counter = nullptr; int ctas_per_sm_cap = 0;
    // fused non-maximum suppression: half window 1 or 2 (0 = write R instead); candidates are appended to
    // cand[frame * cand_cap ...] with cand_count[frame] (zeroed by the caller) exactly as k_nms does
    int nms_h; unsigned long long* cand; int cand_cap; int* cand_count;
};
enum { STENCIL_SRC_GRAY = 0, STENCIL_SRC_BLURRED = 1, STENCIL_SRC_RESPONSE = 2 };  // what StencilArgs::in / ::R hold on entry
cudaError_t launch_stencil(const StencilArgs& a, int n_frames, int src, cudaStream_t s);
bool stencil_can_fuse(int rows, int cols, int nms_window);
cudaError_t launch_bgr2gray(const uint8_t* bgr, size_t bgr_pitch, uint8_t* gray, size_t gray_pitch, int rows, int cols,
                            cudaStream_t s);

struct NmsArgs {
    const float* R; size_t R_fstride, R_pitch;
    unsigned long long* cand; int cand_cap; int* cand_count;  // per-frame strides: cand_cap / 1
    int rows, cols, window;
};
cudaError_t launch_nms(const NmsArgs& a, int n_frames, cudaStream_t s);

struct SelectArgs {
    const unsigned long long* cand; int cand_cap; int* cand_count;
    float2* kp; int kp_stride; int* nkp; int* flags;
    int max_keypoints, sort_cap;
};
cudaError_t launch_select(const SelectArgs& a, int n_frames, cudaStream_t s);

struct FreakArgs {
    const uint8_t* blur; size_t blur_fstride, blur_pitch;
    const float2* kp; int kp_stride; const int* nkp;
    uint32_t* desc; float* angles;  // strides kp_stride*16 / kp_stride
    int rows, cols;
};
cudaError_t launch_freak(const FreakArgs& a, int n_frames, int max_n, cudaStream_t s);
cudaError_t freak_upload_tables();

struct MatchArgs {
    const uint32_t* desc; int kp_stride; const int* nkp;  // slot pools
    uint2* partial; int* raw; int2* pairs; int* nmatch;    // per pair of the batch
    int max_kp; float ratio; int match_bits;
    short qslot[VO_MAX_BATCH]; short tslot[VO_MAX_BATCH];
};
cudaError_t launch_match(const MatchArgs& a, int n_pairs, int max_n1, cudaStream_t s_);
cudaError_t launch_match_finalize(const MatchArgs& a, int n_pairs, cudaStream_t s);
cudaError_t launch_gradients_ref(const uint8_t* img, size_t pitch, int r0, int r1, int c0, int c1, float* jx, float* jy, float* jxy,
                                 int cols, cudaStream_t st);
cudaError_t launch_response_ref(const float* jx, const float* jy, const float* jxy, int cols, int r0, int r1, int c0, int c1,
                                float threshold, int response, float harris_k, float* R, cudaStream_t st);
cudaError_t launch_popc_peak(int mode, unsigned* scratch, int sms, int iters, cudaStream_t st, double* popc_per_launch);

cudaError_t launch_pack_bytes(const uint8_t* bytes, int n, uint32_t* packed, int* flag, cudaStream_t s);
cudaError_t launch_unpack_bytes(const uint32_t* packed, int n, uint8_t* bytes, cudaStream_t s);
