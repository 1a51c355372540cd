/* vo_oracle.c — TEST INFRASTRUCTURE ONLY (see vo_oracle.h for the rules and the parity status).
 *
 * Plain-C restatement of the reference's CPU thread-pool feature front-end with single-strip
 * (whole-image) semantics: the reference's default multi-strip run has a data race between
 * neighbouring strips, BLOCK_SIZE >= rows is the deterministic setting.
 * Build: gcc -O2 -ffp-contract=off (no FMA contraction; no -march=native).
 * All citations are relative to /root/reference.
 */
#include "vo_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "freak_tables.h"

static int reflect101(int p, int n) {
    if (n == 1) return 0;
    while (p < 0 || p >= n) p = p < 0 ? -p : 2 * n - 2 - p;
    return p;
}

/* CornerDetectionParallel::custom_bgr2gray — feature_extraction_parallel/corner_detection_parallel.cpp:14-28
 * cv::saturate_cast<uchar>(double) = clamp(lrint(v)). */
void vo_oracle_bgr2gray(const uint8_t* bgr, int rows, int cols, uint8_t* gray) {
    for (int i = 0; i < rows; ++i)
        for (int j = 0; j < cols; ++j) {
            const uint8_t* p = bgr + ((size_t)i * cols + j) * 3;
            const double v = 0.114 * p[0] + 0.587 * p[1] + 0.299 * p[2];
            long iv = lrint(v);
            gray[(size_t)i * cols + j] = (uint8_t)(iv < 0 ? 0 : (iv > 255 ? 255 : iv));
        }
}

/* cv::GaussianBlur(src, dst, Size(7,7), 0) — call site feature_extraction_parallel.cpp:223.
 * Third-party arithmetic (OpenCV, version unpinned by the reference): for CV_8U, 7x7, sigma=0
 * OpenCV uses the fixed kernel [2,7,14,18,14,7,2]/64 in fixed point with BORDER_REFLECT_101;
 * the result equals (sum_y sum_x w w src + 2048) >> 12 (checked against cv2 4.13.0 in tests). */
void vo_oracle_blur(const uint8_t* gray, int rows, int cols, uint8_t* blurred) {
    static const int w[7] = {2, 7, 14, 18, 14, 7, 2};
    uint16_t* hs = (uint16_t*)malloc((size_t)rows * cols * sizeof(uint16_t));
    for (int i = 0; i < rows; ++i)
        for (int j = 0; j < cols; ++j) {
            int acc = 0;
            for (int k = 0; k < 7; ++k) acc += w[k] * gray[(size_t)i * cols + reflect101(j + k - 3, cols)];
            hs[(size_t)i * cols + j] = (uint16_t)acc;
        }
    for (int i = 0; i < rows; ++i)
        for (int j = 0; j < cols; ++j) {
            uint32_t acc = 2048;
            for (int k = 0; k < 7; ++k) acc += (uint32_t)w[k] * hs[(size_t)reflect101(i + k - 3, rows) * cols + j];
            blurred[(size_t)i * cols + j] = (uint8_t)(acc >> 12);
        }
    free(hs);
}

/* CornerDetectionParallel::direction_gradients_worker — corner_detection_parallel.cpp:30-58
 * over the whole image; rows/cols 0 and n-1 stay zero (cv::Mat::zeros, feature_extraction_parallel.cpp:231-233). */
void vo_oracle_gradients(const uint8_t* b, int rows, int cols, float* jx, float* jy, float* jxy) {
    const size_t n = (size_t)rows * cols;
    memset(jx, 0, n * sizeof(float)); memset(jy, 0, n * sizeof(float)); memset(jxy, 0, n * sizeof(float));
    for (int i = 1; i < rows - 1; ++i) {
        const uint8_t *r1 = b + (size_t)(i - 1) * cols, *r2 = b + (size_t)i * cols, *r3 = b + (size_t)(i + 1) * cols;
        for (int j = 1; j < cols - 1; ++j) {
            float sumx[3], sumy[3];
            for (int k = -1; k <= 1; ++k) {
                sumx[k + 1] = (float)(r1[j + k] - r3[j + k]);                   /* [1,0,-1]^T */
                sumy[k + 1] = (float)(r1[j + k] + 2 * r2[j + k] + r3[j + k]);   /* [1,2,1]^T  */
            }
            jx[(size_t)i * cols + j] = sumx[0] + 2 * sumx[1] + sumx[2];
            jy[(size_t)i * cols + j] = sumy[0] - sumy[2];
            jxy[(size_t)i * cols + j] = sumx[0] - sumx[2];
        }
    }
}

/* CornerDetectionParallel::shitomasi_corner_detection_worker — corner_detection_parallel.cpp:61-105.
 * The off-diagonal is the SUM of the mixed derivative (not sum of Jx*Jy), as written at :86,:93.
 * The final subtraction is evaluated in double and narrowed to float (:96); R_array is
 * zero-initialised (feature_extraction_parallel.cpp:235) so rows/cols < 2 and >= n-2 stay zero. */
void vo_oracle_response(const uint8_t* b, int rows, int cols, float threshold, float* R) {
    const size_t n = (size_t)rows * cols;
    float* jx = (float*)malloc(n * sizeof(float));
    float* jy = (float*)malloc(n * sizeof(float));
    float* jxy = (float*)malloc(n * sizeof(float));
    vo_oracle_gradients(b, rows, cols, jx, jy, jxy);
    memset(R, 0, n * sizeof(float));
    for (int i = 2; i < rows - 2; ++i)
        for (int j = 2; j < cols - 2; ++j) {
            float sumjxy = 0, jx2 = 0, jy2 = 0;
            for (int m = -2; m <= 2; ++m)
                for (int q = -2; q <= 2; ++q) {
                    const size_t at = (size_t)(i + m) * cols + (j + q);
                    const float vx = jx[at], vy = jy[at], vxy = jxy[at];
                    sumjxy += vxy;
                    jx2 += vx * vx;
                    jy2 += vy * vy;
                }
            const float det = (jx2 * jy2) - (sumjxy * sumjxy);
            const float trace = jx2 + jy2;
            const float r = (float)((double)(trace / 2) - (0.5 * (double)sqrtf(trace * trace - 4 * det)));
            R[(size_t)i * cols + j] = r > threshold ? r : 0;
        }
    free(jx); free(jy); free(jxy);
}

/* Harris response on the same structure tensor: R = det - k * trace^2 in fp32, one rounding per operation, left to
 * right ((k * trace) * trace).  The formula is the reference's serial variant (feature_extraction/corner_detection.cpp:
 * 196-199), which evaluates it in double with a threshold relative to the image maximum; no fp32 / absolute-threshold
 * Harris exists in the reference, so this function has nothing to be pinned against: PARITY UNPINNED (restatement only). */
void vo_oracle_response_harris(const uint8_t* b, int rows, int cols, float threshold, float k, float* R) {
    const size_t n = (size_t)rows * cols;
    float* jx = (float*)malloc(n * sizeof(float));
    float* jy = (float*)malloc(n * sizeof(float));
    float* jxy = (float*)malloc(n * sizeof(float));
    vo_oracle_gradients(b, rows, cols, jx, jy, jxy);
    memset(R, 0, n * sizeof(float));
    for (int i = 2; i < rows - 2; ++i)
        for (int j = 2; j < cols - 2; ++j) {
            float sumjxy = 0, jx2 = 0, jy2 = 0;
            for (int m = -2; m <= 2; ++m)
                for (int q = -2; q <= 2; ++q) {
                    const size_t at = (size_t)(i + m) * cols + (j + q);
                    const float vx = jx[at], vy = jy[at], vxy = jxy[at];
                    sumjxy += vxy;
                    jx2 += vx * vx;
                    jy2 += vy * vy;
                }
            const volatile float ab = jx2 * jy2, cc = sumjxy * sumjxy;  /* volatile: no contraction whatever the flags */
            const volatile float det = ab - cc;
            const volatile float trace = jx2 + jy2;
            const volatile float kt = k * trace;
            const volatile float ktt = kt * trace;
            const float r = det - ktt;
            R[(size_t)i * cols + j] = r > threshold ? r : 0;
        }
    free(jx); free(jy); free(jxy);
}

typedef struct { float r; int i, j; } cand_t;

static int cand_desc(const void* pa, const void* pb) { /* (R, i, j) descending: std::priority_queue<tuple<float,int,int>> */
    const cand_t *a = (const cand_t*)pa, *b = (const cand_t*)pb;
    if (a->r != b->r) return a->r > b->r ? -1 : 1;
    if (a->i != b->i) return a->i > b->i ? -1 : 1;
    return a->j > b->j ? -1 : (a->j < b->j ? 1 : 0);
}
static int cand_yx(const void* pa, const void* pb) { /* feature_extraction_parallel.cpp:276-282 */
    const cand_t *a = (const cand_t*)pa, *b = (const cand_t*)pb;
    if (a->i != b->i) return a->i < b->i ? -1 : 1;
    return a->j < b->j ? -1 : (a->j > b->j ? 1 : 0);
}

/* CornerDetectionParallel::non_maximum_suppression — corner_detection_parallel.cpp:166-208, then the
 * manager's sort by (int y, int x) — feature_extraction_parallel.cpp:276-282. */
int vo_oracle_nms(const float* R, int rows, int cols, int k, int N, float* kp_xy) {
    const int h = k / 2;
    size_t cap = 1024, cnt = 0;
    cand_t* c = (cand_t*)malloc(cap * sizeof(cand_t));
    for (int i = h; i < rows - h; ++i)
        for (int j = h; j < cols - h; ++j) {
            if (!((j >= 37) && (j <= cols - 37) && (i >= 35) && (i <= rows - 35))) continue;
            const float center = R[(size_t)i * cols + j];
            int is_max = 1;
            for (int n = i - h; n <= i + h && is_max; ++n)
                for (int m = j - h; m <= j + h; ++m)
                    if (!(n == i && m == j) && R[(size_t)n * cols + m] >= center) { is_max = 0; break; }
            if (!is_max) continue;
            if (cnt == cap) { cap *= 2; c = (cand_t*)realloc(c, cap * sizeof(cand_t)); }
            c[cnt].r = center; c[cnt].i = i; c[cnt].j = j; ++cnt;
        }
    qsort(c, cnt, sizeof(cand_t), cand_desc);
    const int n_out = cnt < (size_t)N ? (int)cnt : N;
    qsort(c, (size_t)n_out, sizeof(cand_t), cand_yx);
    for (int q = 0; q < n_out; ++q) { kp_xy[2 * q] = (float)c[q].j; kp_xy[2 * q + 1] = (float)c[q].i; }
    free(c);
    return n_out;
}

/* FREAK_Parallel::compute_orientation — FREAK_feature_descriptor_parallel.cpp:11-55.
 * All 903 pairs in lexicographic order, sequential fp32 accumulation, each term (ic * d) / norm. */
float vo_oracle_orientation(const uint8_t* b, int rows, int cols, float x, float y) {
    (void)rows;
    float ox = 0, oy = 0;
    for (int p = 0; p < VO_FREAK_POINTS; ++p) {
        const float ip = (float)b[(size_t)(int)(VO_FREAK_PY[p] + y) * cols + (int)(VO_FREAK_PX[p] + x)];
        for (int q = p + 1; q < VO_FREAK_POINTS; ++q) {
            const float ic = ip - (float)b[(size_t)(int)(VO_FREAK_PY[q] + y) * cols + (int)(VO_FREAK_PX[q] + x)];
            const float dx = (float)(VO_FREAK_PX[p] - VO_FREAK_PX[q]);
            const float dy = (float)(VO_FREAK_PY[p] - VO_FREAK_PY[q]);
            const float norm = sqrtf(dx * dx + dy * dy);
            if (norm == 0.0f) continue;
            ox += ic * (float)(VO_FREAK_PX[p] - VO_FREAK_PX[q]) / norm;
            oy += ic * (float)(VO_FREAK_PY[p] - VO_FREAK_PY[q]) / norm;
        }
    }
    if (isnan(ox) || isnan(oy)) return 0;
    return atan2f(oy, ox);
}

/* FREAK_Parallel::FREAK_feature_description_worker — FREAK_feature_descriptor_parallel.cpp:110-151.
 * rotation_matrix = {cos, -sin, sin, cos}; x' = (int)(kx + px*rm0 + py*rm2), y' = (int)(ky + (-1)*px*rm1 + py*rm3)
 * (not a true rotation; reproduced as written). */
void vo_oracle_describe(const uint8_t* b, int rows, int cols, const float* kp_xy, int n, uint8_t* desc512,
                        float* angles) {
    for (int q = 0; q < n; ++q) {
        const float kx = kp_xy[2 * q], ky = kp_xy[2 * q + 1];
        const float angle = vo_oracle_orientation(b, rows, cols, kx, ky);
        if (angles) angles[q] = angle;
        const float rm[4] = {cosf(angle), -1 * sinf(angle), sinf(angle), cosf(angle)};
        int sx[VO_FREAK_POINTS], sy[VO_FREAK_POINTS];
        for (int p = 0; p < VO_FREAK_POINTS; ++p) {
            const float px = (float)VO_FREAK_PX[p], py = (float)VO_FREAK_PY[p];
            sx[p] = (int)(kx + px * rm[0] + py * rm[2]);
            sy[p] = (int)(ky + (-1) * px * rm[1] + py * rm[3]);
        }
        for (int t = 0; t < VO_FREAK_TESTS; ++t) {
            const int a = VO_FREAK_TEST_A[t], c = VO_FREAK_TEST_B[t];
            desc512[(size_t)q * 512 + t] =
                b[(size_t)sy[a] * cols + sx[a]] > b[(size_t)sy[c] * cols + sx[c]] ? 1 : 0;
        }
    }
}

/* feature_extraction_manager_with_points — feature_extraction_parallel.cpp:215-349 from the gray image on
 * (custom_bgr2gray is vo_oracle_bgr2gray; identity on gray-replicated BGR input). */
int vo_oracle_extract(const uint8_t* gray, int rows, int cols, float threshold, int nms_window, int max_keypoints,
                      float* kp_xy, uint8_t* desc512, float* angles, uint8_t* blurred_out, float* R_out) {
    const size_t n = (size_t)rows * cols;
    uint8_t* b = blurred_out ? blurred_out : (uint8_t*)malloc(n);
    float* R = R_out ? R_out : (float*)malloc(n * sizeof(float));
    vo_oracle_blur(gray, rows, cols, b);
    vo_oracle_response(b, rows, cols, threshold, R);
    const int cnt = vo_oracle_nms(R, rows, cols, nms_window, max_keypoints, kp_xy);
    if (desc512) vo_oracle_describe(b, rows, cols, kp_xy, cnt, desc512, angles);
    if (!blurred_out) free(b);
    if (!R_out) free(R);
    return cnt;
}

static int popcount_bytes(const uint8_t* a, const uint8_t* b, int len) {
    int d = 0;
    for (int i = 0; i < len; ++i) d += __builtin_popcount((unsigned)(a[i] ^ b[i]));
    return d;
}

/* matchCustomBinaryDescriptorsThreadPool — feature_matching_parallel/feature_matching_parallel.cpp:49-113 with
 * hammingDistance (:39-47) = popcount over the FIRST 32 BYTES (match_bits 32); match_bits 512 follows
 * matching_serial.cpp:24-77 (all 512 bytes).  Chunking over threads does not change the result. */
int vo_oracle_match(const uint8_t* d1, int n1, const uint8_t* d2, int n2, int match_bits, float ratio,
                    int32_t* pairs, int32_t* raw) {
    int n_out = 0;
    if (n1 <= 0 || n2 <= 0) return 0;
    const int len = match_bits == 32 ? 32 : 512;
    for (int i = 0; i < n1; ++i) {
        int best_idx = -1, second_idx = -1, best = INT32_MAX, second = INT32_MAX;
        for (int j = 0; j < n2; ++j) {
            const int dist = popcount_bytes(d1 + (size_t)i * 512, d2 + (size_t)j * 512, len);
            if (dist >= second) continue;
            if (dist < best) { second = best; second_idx = best_idx; best = dist; best_idx = j; }
            else if (dist < second) { second = dist; second_idx = j; }
        }
        if (raw) { raw[3 * i] = best; raw[3 * i + 1] = best_idx; raw[3 * i + 2] = second; }
        if (best_idx != -1 && second_idx != -1 && best < ratio * second) {
            if (pairs) { pairs[2 * n_out] = i; pairs[2 * n_out + 1] = best_idx; }
            ++n_out;
        }
    }
    return n_out;
}

void vo_oracle_pack(const uint8_t* desc512, int n, uint32_t* packed) {
    for (int q = 0; q < n; ++q)
        for (int w = 0; w < 16; ++w) {
            uint32_t v = 0;
            for (int k = 0; k < 32; ++k) v |= (uint32_t)(desc512[(size_t)q * 512 + w * 32 + k] & 1u) << k;
            packed[(size_t)q * 16 + w] = v;
        }
}

/* Test helper for the CUDA FREAK kernel's divide-free quotient (acs_visual_odometry_b200/csrc/freak.cu):
 * q = m*rn; r = fma(-q, n, m); q' = fma(r, rn, q) with rn = RN(1/n).  Compares q' with the IEEE quotient m/n
 * that compute_orientation (FREAK_feature_descriptor_parallel.cpp:39-40) evaluates, for EVERY pair of
 * pattern points, every intensity difference -255..255 and both axes.  Returns the number of mismatches. */
long vo_oracle_check_recip_division(long* total_out) {
    long bad = 0, total = 0;
    for (int p = 0; p < VO_FREAK_POINTS; ++p)
        for (int r = p + 1; r < VO_FREAK_POINTS; ++r) {
            const float dx = (float)(VO_FREAK_PX[p] - VO_FREAK_PX[r]), dy = (float)(VO_FREAK_PY[p] - VO_FREAK_PY[r]);
            const float n = sqrtf(dx * dx + dy * dy);
            const float rn = 1.0f / n;
            for (int ic = -255; ic <= 255; ++ic)
                for (int ax = 0; ax < 2; ++ax) {
                    const float m = (float)ic * (ax ? dy : dx);
                    const float q = m * rn;
                    const float got = fmaf(fmaf(-q, n, m), rn, q);
                    ++total;
                    if (got != m / n) ++bad;
                }
        }
    if (total_out) *total_out = total;
    return bad;
}

/* ---- exhaustive proof of the FREAK kernel's split-weight product ------------------------------------------
 * TEST INFRASTRUCTURE.  csrc/freak.cu evaluates the orientation term RN(RN(ic*d)/n) (n = sqrtf(dx^2+dy^2),
 * FREAK_feature_descriptor_parallel.cpp:39-40) as fma(ic, wh, RN(ic*wl)) with w = (double)d/(double)n,
 * wh = (float)w, wl = (float)(w - wh).  Compares the two for EVERY pair of pattern points, every intensity
 * difference -255..255 and both axes.  Returns the number of mismatches. */
long vo_oracle_check_split_weight(long* total_out) {
    long bad = 0, total = 0;
    for (int p = 0; p < VO_FREAK_POINTS; ++p)
        for (int r = p + 1; r < VO_FREAK_POINTS; ++r) {
            const float dx = (float)(VO_FREAK_PX[p] - VO_FREAK_PX[r]), dy = (float)(VO_FREAK_PY[p] - VO_FREAK_PY[r]);
            const float n = sqrtf(dx * dx + dy * dy);
            for (int ax = 0; ax < 2; ++ax) {
                const float d = ax ? dy : dx;
                const double w = (double)d / (double)n;
                const float wh = (float)w;
                const float wl = (float)(w - (double)wh);
                for (int ic = -255; ic <= 255; ++ic) {
                    const float fic = (float)ic;
                    const float m = fic * d;
                    const float tl = fic * wl;
                    const float got = fmaf(fic, wh, tl);
                    ++total;
                    if (got != m / n) ++bad;
                }
            }
        }
    if (total_out) *total_out = total;
    return bad;
}
