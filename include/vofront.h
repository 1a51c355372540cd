/* vofront.h — C ABI of libvofront.so: the B200-native feature front-end
 * (7x7 Gaussian blur -> gradients -> 5x5 structure tensor -> Shi-Tomasi response -> NMS/top-N ->
 * FREAK-style 512-test descriptors -> brute-force Hamming 2-NN ratio matching).
 *
 * This is the drop-in boundary for the hot path of Bohdanok/ACS_Visual_Odometry.  Every entry
 * point names the reference interface it replaces (paths relative to the reference root).
 * Plain pointers and sizes only; no C++ or torch types cross it.  All work is done by hand-written
 * sm_100a CUDA kernels; there is no CPU fallback: without a CUDA device vo_create() fails.
 *
 * Conventions
 *   - every function returning int returns VO_OK (0) or a negative VO_ERR_* code;
 *     vo_last_error(ctx) gives a human-readable message for the last failure on that context
 *     (ctx == NULL: last vo_create() or vo_match_static() failure on this thread).  Nothing throws across the ABI.
 *   - a vo_ctx belongs to one GPU; it is not thread-safe, independent contexts are fully concurrent.
 *   - images are 8-bit grayscale, row-major, `pitch` bytes between rows (pitch >= cols).
 *   - keypoints are (x = column, y = row) float pairs, sorted by (y, x) ascending, exactly the order
 *     of the reference (feature_extraction_parallel/feature_extraction_parallel.cpp:276-282); match
 *     indices refer to that order.
 *   - descriptors: "desc512" = n x 512 bytes, one byte (0/1) per test, the reference's layout
 *     (feature_extraction_parallel.cpp:324); "packed" = n x 16 little-endian uint32 words,
 *     bit (j & 31) of word (j >> 5) = test j.
 */
#ifndef VOFRONT_H
#define VOFRONT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VO_OK 0
#define VO_ERR_BAD_ARG (-1)   /* null pointer, size out of range, slot out of range, bad descriptor bytes */
#define VO_ERR_CUDA (-2)      /* a CUDA runtime call or kernel failed; see vo_last_error */
#define VO_ERR_CAPACITY (-3)  /* NMS candidates or matches exceeded a buffer; never truncated silently */
#define VO_ERR_NO_DEVICE (-4) /* no usable sm_100 device */
#define VO_ERR_STATE (-5)     /* slot holds no frame / sequence not uploaded */

#define VO_MATCH_REF32 32     /* hammingDistance of feature_matching_parallel/feature_matching_parallel.cpp:39-47:
                                 only the first 32 tests are compared (what VisualOdometry runs) */
#define VO_MATCH_FULL512 512  /* all 512 tests, matching_serial.cpp:24-40 */

#define VO_PRESET_CPU_REF 0    /* thread-pool path: threshold 25000 (corner_detection_parallel.h:16),
                                  NMS 5, top-N 1500 (feature_extraction_parallel.cpp:274) */
#define VO_PRESET_VO_GPU_REF 1 /* the variant VisualOdometry links: threshold 20000
                                  (corner_detection_parallel_GPU.h:25), NMS 3, top-N 2000
                                  (feature_extraction_parallel_GPU.cpp:235) */

#define VO_DBG_BLURRED 1  /* rows x cols uint8, tightly packed */
#define VO_DBG_RESPONSE 2 /* rows x cols float (thresholded R_array), tightly packed; extraction keeps the response on
                             chip (NMS fused behind the stencil), so this recomputes it from the slot's blurred image */
#define VO_DBG_ANGLES 3   /* n float, FREAK orientation per keypoint */

typedef struct vo_ctx vo_ctx;

/* The literals the reference scatters over globals, constexprs and call sites. */
typedef struct vo_params {
    float threshold;   /* RESPONSE_THRESHOLD: responses <= threshold become 0 */
    int nms_window;    /* k of non_maximum_suppression (odd, 3..9) */
    int max_keypoints; /* N of non_maximum_suppression; <= the context's max_keypoints */
    float ratio;       /* Lowe ratio, 0.75f at every reference call site */
    int match_bits;    /* VO_MATCH_REF32 or VO_MATCH_FULL512 */
    int response;      /* VO_RESPONSE_SHI_TOMASI (every preset) or VO_RESPONSE_HARRIS */
    float harris_k;    /* k of R = det(M) - k * trace(M)^2 (feature_extraction/corner_detection.cpp:199); used by HARRIS only */
} vo_params;

/* Corner response.  SHI_TOMASI: the smaller eigenvalue, what the thread-pool and OpenCL paths compute
 * (corner_detection_parallel.cpp:93-98).  HARRIS: det - k * trace^2, the formula of the reference's serial variant
 * (feature_extraction/corner_detection.cpp:199), evaluated here in fp32 on the same structure tensor with the same
 * absolute threshold semantics (values <= threshold become 0); the serial variant itself works in fp64 with a threshold
 * relative to the image maximum, so this mode is pinned to the restatement in oracle/ only ("parity unpinned").
 * HARRIS frames take the response-writing form of the stencil kernel followed by the NMS kernel. */
#define VO_RESPONSE_SHI_TOMASI 0
#define VO_RESPONSE_HARRIS 1

int vo_params_preset(vo_params* p, int preset);

/* ---- context --------------------------------------------------------------------------------
 * Replaces the per-process state of the reference: `thread_pool` (feature_extraction_parallel/
 * threadpool.h:15) and `GPU_settings{cl::Program, cl::Device, cl::Context}`
 * (feature_extraction_parallel_GPU/corner_detection_parallel_GPU.h:48-52, built in
 * VisualOdometry.cpp:14-23).  All device buffers are sized here; nothing is allocated per frame.
 * n_slots = number of device-resident frame records (image, blurred image, response, candidates,
 * keypoints, packed descriptors), 2 .. 2048; batched calls use consecutive slots modulo n_slots and take at most
 * 1024 frames (or pairs) per call or chunk. */
int vo_create(vo_ctx** out, int device, int max_rows, int max_cols, int max_keypoints, int n_slots);
/* The same with the capacity of the per-frame NMS candidate list chosen by the caller (0 = the default: max(32768,
 * 8 * max_keypoints, max_rows * max_cols / 32)).  The reference's heap has no limit (corner_detection_parallel.cpp:
 * 170-196); here a frame with more candidates than this is reported with VO_ERR_CAPACITY, never truncated silently. */
int vo_create_ex(vo_ctx** out, int device, int max_rows, int max_cols, int max_keypoints, int n_slots, int cand_cap);
void vo_destroy(vo_ctx* ctx);
const char* vo_last_error(const vo_ctx* ctx);
int vo_device_count(void);
int vo_sync(vo_ctx* ctx);

/* Pinned host memory for callers that want truly asynchronous copies (optional). */
void* vo_host_alloc(size_t bytes);
void vo_host_free(void* p);

/* ---- extraction --------------------------------------------------------------------------------
 * Replaces feature_extraction_manager_with_points
 *   (cv::Mat&, thread_pool&)            feature_extraction_parallel/feature_extraction_parallel.cpp:215
 *   (const cv::Mat&, const GPU_settings&) feature_extraction_parallel_GPU/feature_extraction_parallel_GPU.cpp:194
 * i.e. GaussianBlur 7x7 -> direction_gradients_worker -> shitomasi_corner_detection_worker ->
 * non_maximum_suppression -> sort -> FREAK_feature_description_worker.  The result stays on the device
 * in `slot`; fetch it with vo_frame_fetch and/or match it with vo_match. */
int vo_extract(vo_ctx* ctx, const uint8_t* gray, int rows, int cols, size_t pitch, const vo_params* p, int slot);
/* Same, from an interleaved 8-bit BGR image (custom_bgr2gray, corner_detection_parallel.cpp:14-28). */
int vo_extract_bgr(vo_ctx* ctx, const uint8_t* bgr, int rows, int cols, size_t pitch, const vo_params* p, int slot);
/* n frames, `frame_stride` bytes apart, into slots first_slot .. first_slot+n-1 (mod n_slots), one launch
 * per stage for the whole batch. n <= n_slots. */
int vo_extract_batch(vo_ctx* ctx, const uint8_t* gray, int n, int rows, int cols, size_t pitch, size_t frame_stride,
                     const vo_params* p, int first_slot);

/* Copy a slot's results to the host.  Any output pointer may be NULL.  *n receives the keypoint count;
 * kp_xy needs 2*n floats, desc512 512*n bytes, desc_packed 16*n words (size them with max_keypoints). */
int vo_frame_fetch(vo_ctx* ctx, int slot, int* n, float* kp_xy, uint8_t* desc512, uint32_t* desc_packed);
/* Stage probes for parity tests (VO_DBG_*): the blurred image / response map / angles of the frame that
 * was last extracted into `slot`. */
int vo_debug_fetch(vo_ctx* ctx, int slot, int what, void* dst);

/* Stage entry points mirroring the reference's public statics, host buffers in and out:
 *   vo_blur       cv::GaussianBlur(src, dst, Size(7,7), 0)               feature_extraction_parallel.cpp:223
 *   vo_response   direction_gradients_worker + shitomasi_corner_detection_worker over the whole image
 *                 (corner_detection_parallel.cpp:30-105): blurred u8 -> thresholded R
 *   vo_nms        CornerDetectionParallel::non_maximum_suppression + the manager's sort; R must be a thresholded
 *                 (non-negative) response map, as produced by vo_response
 *                 (corner_detection_parallel.cpp:166-208, feature_extraction_parallel.cpp:276-282)
 *   vo_describe   FREAK_Parallel::FREAK_feature_description_worker      FREAK_feature_descriptor_parallel.cpp:110
 */
int vo_blur(vo_ctx* ctx, const uint8_t* gray, int rows, int cols, uint8_t* blurred);
int vo_response(vo_ctx* ctx, const uint8_t* blurred, int rows, int cols, float threshold, float* R);
/* The same with the response selector of p (p->threshold, p->response, p->harris_k). */
int vo_response_p(vo_ctx* ctx, const uint8_t* blurred, int rows, int cols, const vo_params* p, float* R);
/* The reference's per-stage workers with their `interval` argument (rows [r0, r1), columns [c0, c1)), for callers that
 * drive the stages themselves (vofront_adapter.hpp: class CornerDetectionParallel).  Matrices are packed rows x cols;
 * only the rectangle the reference's worker writes is written, the rest of the output stays untouched.
 *   vo_bgr2gray                  CornerDetectionParallel::custom_bgr2gray            corner_detection_parallel.cpp:14-28
 *   vo_gradients                 ::direction_gradients_worker: rows r0+1 .. r1-2, columns c0+1 .. c1-2     :30-58
 *   vo_response_from_gradients   ::shitomasi_corner_detection_worker: rows r0+2 .. r1-3, columns c0+2 .. c1-3, from
 *                                ANY fp32 gradient matrices, the reference's summation order (p->threshold, p->response,
 *                                p->harris_k; the reference's own `k` argument is unused, :61)            :61-105
 * These run small one-thread-per-pixel kernels and synchronise; extraction never calls them (it runs the fused kernel). */
int vo_bgr2gray(vo_ctx* ctx, const uint8_t* bgr, int rows, int cols, size_t pitch, uint8_t* gray);
int vo_gradients(vo_ctx* ctx, const uint8_t* picture, int rows, int cols, size_t pitch, int r0, int r1, int c0, int c1, float* Jx,
                 float* Jy, float* Jxy);
int vo_response_from_gradients(vo_ctx* ctx, const float* Jx, const float* Jy, const float* Jxy, int rows, int cols, int r0, int r1,
                               int c0, int c1, const vo_params* p, float* R);
int vo_nms(vo_ctx* ctx, const float* R, int rows, int cols, int nms_window, int max_keypoints, float* kp_xy,
           int* n_out);
int vo_describe(vo_ctx* ctx, const uint8_t* blurred, int rows, int cols, const float* kp_xy, int n,
                uint8_t* desc512, float* angles);

/* ---- matching ------------------------------------------------------------------------------------
 * Replaces matchCustomBinaryDescriptorsThreadPool(desc1, desc2, pool, numThreads, ratioThreshold)
 *   feature_matching_parallel/feature_matching_parallel.cpp:49-113
 * pairs receives (queryIdx, trainIdx) int32 pairs ascending in queryIdx; cap = capacity in pairs. */
int vo_match(vo_ctx* ctx, int query_slot, int train_slot, float ratio, int match_bits, int32_t* pairs, int cap,
             int* n_out);
/* Host descriptors in the reference's layout (row_bytes 512: one 0/1 byte per test) or packed
 * (row_bytes 64).  Empty inputs give zero matches, like the reference (:57). */
int vo_match_host(vo_ctx* ctx, const uint8_t* d1, int n1, const uint8_t* d2, int n2, int row_bytes, float ratio,
                  int match_bits, int32_t* pairs, int cap, int* n_out);
/* Per-query (best distance, best index, second-best distance) triples of the last vo_match /
 * vo_match_host call: n1 x 3 int32; INT32_MAX / -1 where undefined (fewer than 1 / 2 train rows). */
int vo_match_raw(vo_ctx* ctx, int32_t* raw, int n1);

/* ---- matching without any allocation -----------------------------------------------------------
 * No counterpart in the reference: this entry point exists only because of how this library manages GPU memory.
 * Everything above needs a vo_ctx, and vo_create() obtains the context's pools - the matcher's scratch among them -
 * with cudaMalloc.  vo_match_static is for callers that cannot do that: processes whose device memory is handed out
 * entirely by an arena of their own (fixed per-module budgets on a robot, an inference server's allocator), loops with
 * a latency bound (cudaMalloc / cudaFree synchronise the device and take an unpredictable time), and plug-ins that get
 * device pointers and a stream from a host application and may not create device state of their own.  It needs no
 * context and calls no cudaMalloc, cudaFree or other allocating routine, on any call.
 *
 * Same result as vo_match_host on the same descriptors (same keys, same merge, same fp32 ratio test): the accepted
 * (queryIdx, trainIdx) pairs ascending in queryIdx.  The caller passes inputs and outputs, all DEVICE memory of the
 * current GPU (host pointers are refused with VO_ERR_BAD_ARG):
 *   d_query, d_train   packed descriptors, n x 16 uint32 (may be NULL when the count is 0)
 *   d_pairs            room for n1 pairs (2 * n1 int32): a query yields at most one match
 *   d_n_out            one int32, the number of pairs written
 *   stream             a cudaStream_t (NULL = the default stream).  The call only enqueues two kernels; synchronise the
 *                      stream before reading the outputs.
 * The scratch between the two kernels (8 x n1 key pairs) lives in an array compiled into the library, 512 KB, sized for
 * VO_STATIC_MATCH_MAX descriptors per side; larger counts are refused with VO_ERR_BAD_ARG (use a context for those).
 * Every GPU that loads the library holds that array until the process exits.
 *   - The first call on each GPU looks up the array's device address (cudaGetSymbolAddress, which also makes the
 *     runtime load the library's device code on that GPU).  Make that first call during start-up, before any stream
 *     capture and outside the latency-critical part; later calls on that GPU only launch.
 *   - All calls on one GPU share the array, and it keeps the contents of the last call.  Nothing in it has to start at
 *     zero (every entry the second kernel reads is written by the first kernel of the same call), but two calls that
 *     overlap on the GPU - different streams, or different host threads on different streams - race on it and give
 *     wrong matches.  Keep all calls on one stream, or order them with events.
 * Errors: vo_last_error(NULL) on the calling thread. */
#define VO_STATIC_MATCH_MAX 8192
int vo_match_static(const uint32_t* d_query, int n1, const uint32_t* d_train, int n2, float ratio, int match_bits,
                    int32_t* d_pairs, int32_t* d_n_out, void* stream);

/* ---- frame sequences (the batched, pipelined form of VisualOdometry::run's front-end half) -------
 * VisualOdometry.cpp:64-100: extract every frame, match frame i-1 (query) against frame i (train).
 * Results per frame f: keypoint count, keypoints; per pair f (= frames f, f+1): match count, pairs. */
typedef struct vo_seq_result {
    int32_t* n_keypoints; /* [n_frames] */
    float* kp_xy;         /* [n_frames][max_keypoints][2] or NULL */
    int32_t* n_matches;   /* [n_frames-1] */
    int32_t* pairs;       /* [n_frames-1][max_keypoints][2] or NULL */
} vo_seq_result;

/* Host frames in, host results out; H2D, kernels and D2H are pipelined over `batch`-frame chunks on
 * separate streams (batch <= n_slots/2). This is the end-to-end path.  The call is bound by the frame
 * copy when `frames` and the result arrays are pinned (vo_host_alloc): 1000 KITTI frames take 9.0 ms
 * against 8.4 ms for the bare copy.  Pageable buffers work but serialise the copies with the host
 * (measured 14.5 ms with pageable result arrays). */
int vo_sequence_run(vo_ctx* ctx, const uint8_t* frames, int n_frames, int rows, int cols, size_t pitch,
                    size_t frame_stride, const vo_params* p, int batch, vo_seq_result* out);
/* Device-resident variant: upload once, then run the kernels only (results stay on the device;
 * vo_sequence_fetch copies them out).  Used to measure kernel throughput with inputs already in HBM. */
int vo_sequence_upload(vo_ctx* ctx, const uint8_t* frames, int n_frames, int rows, int cols, size_t pitch,
                       size_t frame_stride);
int vo_sequence_run_resident(vo_ctx* ctx, const vo_params* p, int batch);
int vo_sequence_fetch(vo_ctx* ctx, vo_seq_result* out);
/* Per-frame status bits of the last vo_sequence_run / vo_sequence_fetch call (bit 0: that frame's candidate list
 * overflowed, the call returned VO_ERR_CAPACITY; all other frames' results are complete and were copied out), so that a
 * caller can treat the frame like a failed one instead of giving up the sequence. */
int vo_sequence_flags(vo_ctx* ctx, int32_t* flags, int n);

/* ---- measurement hooks ------------------------------------------------------------------------ */
/* CUDA-event timing on the context's own compute stream. */
int vo_timer_start(vo_ctx* ctx);
int vo_timer_stop(vo_ctx* ctx, float* ms); /* synchronises; elapsed since vo_timer_start */
/* The stencil kernel alone over the first `batch` frames of the uploaded sequence: mean milliseconds per launch over
 * `reps` launches (CUDA events on the launch stream, input already in HBM, one synchronisation per launch).
 *   mode 0  the response-writing form: u8 in -> blurred u8 + fp32 R out = 6 B/px (what vo_response / VO_DBG_RESPONSE run)
 *   mode 1  the fused form of p->nms_window (3 or 5): u8 in -> blurred u8 out + 8 B per NMS candidate = 2 B/px
 * The two roofline figures of bench.py come from here; they are never mixed.  Slots 0..batch-1 are overwritten. */
int vo_bench_stencil(vo_ctx* ctx, const vo_params* p, int batch, int mode, int reps, float* ms_per_launch);
/* Sum over the first `batch` frames of the candidate counts the last fused stencil launch appended (the 8 B per
 * candidate term of the fused kernel's algorithmic bytes).  Call right after vo_bench_stencil(mode 1). */
int vo_bench_candidates(vo_ctx* ctx, int batch, long long* total);
/* The POPC32 issue rate of this GPU, measured by a small kernel of independent chains on the context's stream (best of
 * `reps` launches, CUDA events): the denominator of the matcher's roofline (16 POPC32 per 512-bit compare, 1 in REF32
 * mode).  mode 0: POPC + one add per chain step; mode 1: the matcher's own mix per compare (XOR, POPC, two adds). */
int vo_bench_popc(vo_ctx* ctx, int mode, int reps, double* popc32_per_s);
/* Number of kernel launches issued by this context since creation (or since the last reset). */
long long vo_launch_count(vo_ctx* ctx, int reset);
/* Accumulated device time (ms, CUDA events) per kernel family since the last reset, when profiling is
 * enabled with vo_profile_enable(ctx, 1): [0] stencil, [1] nms, [2] select/sort, [3] freak, [4] match,
 * [5] match-finalize.  Enabling it inserts events around every launch; leave it off for throughput runs. */
int vo_profile_enable(vo_ctx* ctx, int on);
int vo_profile_read(vo_ctx* ctx, float ms[8], long long launches[8], int reset);

#ifdef __cplusplus
}
#endif
#endif /* VOFRONT_H */
