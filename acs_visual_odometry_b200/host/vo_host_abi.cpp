// vo_host_abi.cpp — the C ABI of include/vohost.h over the host classes.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <exception>
#include <string>

#include "../../include/vohost.h"
#include "vo_image_io.hpp"
#include "vo_pipeline.hpp"
#include "vo_pose.hpp"
#include "vo_ransac.hpp"

namespace {

thread_local std::string g_err;
int fail(int code, const std::string& msg) { g_err = msg; return code; }

std::vector<vo::PointPair> to_pairs(const double* pts, int n) {
    std::vector<vo::PointPair> v(static_cast<size_t>(n));
    for (int i = 0; i < n; ++i) v[i] = {vo::Point{pts[4 * i], pts[4 * i + 1]}, vo::Point{pts[4 * i + 2], pts[4 * i + 3]}};
    return v;
}
vo::Mat3 to_mat3(const double m[9]) { vo::Mat3 a; std::memcpy(a.m, m, sizeof(a.m)); return a; }

template <class Fn>
int guarded(Fn fn) {
    try { return fn(); }
    catch (const std::exception& e) { return fail(-3, e.what()); }
    catch (...) { return fail(-3, "unknown exception"); }
}

}  // namespace

namespace {
std::mutex g_ctx_mutex;
vo::VisualOdometry::Context g_ctx;  // the context vo_pipeline_run built last (kept until the process ends or it is replaced)
}  // namespace

extern "C" {

const char* vo_host_last_error(void) { return g_err.c_str(); }

int vo_host_fundamental(const double* pts, int n, double F[9]) {
    if (!pts || !F || n < 8) return fail(-1, "need at least 8 correspondences");
    return guarded([&] {
        const vo::Mat3 f = vo::FundamentalMatrix::computeFundamentalMatrix(to_pairs(pts, n));
        std::memcpy(F, f.m, sizeof(f.m));
        return 0;
    });
}

int vo_host_normalize_points(const double* pts, int n, double* n1, double* n2, double T1[9], double T2[9]) {
    if (!pts || !n1 || !n2 || !T1 || !T2 || n < 1) return fail(-1, "null argument");
    return guarded([&] {
        std::vector<vo::Vec3> a, b;
        vo::Mat3 t1, t2;
        vo::FundamentalMatrix::normalizePoints(to_pairs(pts, n), a, b, t1, t2);
        for (int i = 0; i < n; ++i)
            for (int k = 0; k < 3; ++k) { n1[3 * i + k] = a[i](k); n2[3 * i + k] = b[i](k); }
        std::memcpy(T1, t1.m, sizeof(t1.m));
        std::memcpy(T2, t2.m, sizeof(t2.m));
        return 0;
    });
}

int vo_host_triangulate(const double R[9], const double t[3], const float* n1, const float* n2, int n, int force_scalar, double* X) {
    if (!R || !t || !n1 || !n2 || !X || n < 0) return fail(-1, "null argument");
    return guarded([&] {
        const vo::Mat3 r = to_mat3(R);
        const vo::Vec3 tv{{t[0], t[1], t[2]}};
        std::vector<vo::Point2f> a(n), b(n);
        for (int i = 0; i < n; ++i) { a[i] = vo::Point2f{n1[2 * i], n1[2 * i + 1]}; b[i] = vo::Point2f{n2[2 * i], n2[2 * i + 1]}; }
        double(*out)[4] = reinterpret_cast<double(*)[4]>(X);
        if (force_scalar) for (int i = 0; i < n; ++i) vo::triangulatePoint(r, tv, a[i].x, a[i].y, b[i].x, b[i].y, out[i]);
        else vo::triangulatePoints(r, tv, a.data(), b.data(), static_cast<size_t>(n), out);
        return 0;
    });
}

int vo_host_score(const double F[9], const double* pts, int n, double threshold, int force_scalar, uint8_t* flags, int* count) {
    if (!F || !pts || !flags || !count || n < 0) return fail(-1, "null argument");
    return guarded([&] {
        std::vector<double> X0(n), X1(n), P0(n), P1(n);
        for (int i = 0; i < n; ++i) { X0[i] = pts[4 * i]; X1[i] = pts[4 * i + 1]; P0[i] = pts[4 * i + 2]; P1[i] = pts[4 * i + 3]; }
        *count = vo::score_correspondences(F, X0.data(), X1.data(), P0.data(), P1.data(), n, threshold, flags, force_scalar != 0);
        return 0;
    });
}

int vo_host_sampson(const double F[9], const double* pts, int n, double* err) {
    if (!F || !pts || !err || n < 0) return fail(-1, "null argument");
    const vo::Mat3 f = to_mat3(F);
    for (int i = 0; i < n; ++i)
        err[i] = vo::computeSampsonError(f, vo::Vec3{{pts[4 * i], pts[4 * i + 1], 1.0}}, vo::Vec3{{pts[4 * i + 2], pts[4 * i + 3], 1.0}});
    return 0;
}

int vo_host_ransac(const double* pts, int n, double probability, double sampson_threshold, int num_threads,
                   uint64_t seed, double F[9], double* inliers, int* n_inliers, int* iterations) {
    if (!pts || !F || n < 0) return fail(-1, "null argument");
    return guarded([&] {
        vo::FundamentalMatrix model;
        const vo::RansacStats st = vo::Ransac().run(model, to_pairs(pts, n), probability, sampson_threshold, num_threads, seed);
        const vo::Mat3 f = model.getMatrix();
        std::memcpy(F, f.m, sizeof(f.m));
        const auto& inl = model.getInliers();
        if (n_inliers) *n_inliers = static_cast<int>(inl.size());
        if (iterations) *iterations = st.iterations;
        if (inliers)
            for (size_t i = 0; i < inl.size(); ++i) {
                inliers[4 * i] = inl[i].first.x; inliers[4 * i + 1] = inl[i].first.y;
                inliers[4 * i + 2] = inl[i].second.x; inliers[4 * i + 3] = inl[i].second.y;
            }
        return 0;
    });
}

int vo_host_get_pose(const double F[9], const float* p1, const float* p2, int n, double gt_step, const double* K,
                     double R[9], double t[3], int* votes) {
    if (!F || !p1 || !p2 || !R || !t || n < 0) return fail(-1, "null argument");
    return guarded([&] {
        vo::PoseUpdate est;
        if (K) est.K = to_mat3(K);
        std::vector<vo::Point2f> a(static_cast<size_t>(n)), b(static_cast<size_t>(n));
        for (int i = 0; i < n; ++i) { a[i] = vo::Point2f{p1[2 * i], p1[2 * i + 1]}; b[i] = vo::Point2f{p2[2 * i], p2[2 * i + 1]}; }
        const vo::PoseResult r = est.getPoseEx(to_mat3(F), a, b, gt_step);
        std::memcpy(R, r.R.m, sizeof(r.R.m));
        for (int i = 0; i < 3; ++i) t[i] = r.t(i);
        if (votes) *votes = r.positive_depth;
        return 0;
    });
}

int vo_host_read_gt_line(const char* line, double T[16]) {
    if (!line || !T) return fail(-1, "null argument");
    const vo::Mat4 m = vo::readGTLine(line);
    std::memcpy(T, m.m, sizeof(m.m));
    return 0;
}

int vo_host_format_pose_csv(const double T[16], char* buf, int cap) {
    if (!T || !buf || cap < 1) return fail(-1, "null argument");
    vo::Mat4 m;
    std::memcpy(m.m, T, sizeof(m.m));
    const std::string s = vo::formatPoseCSVLine(m);
    if (static_cast<int>(s.size()) + 1 > cap) return fail(-1, "buffer too small");
    std::memcpy(buf, s.c_str(), s.size() + 1);
    return 0;
}

int vo_host_euler(const double R[9], double rpy[3]) {
    if (!R || !rpy) return fail(-1, "null argument");
    const vo::Vec3 e = vo::rotationMatrixToEulerAngles(to_mat3(R));
    for (int i = 0; i < 3; ++i) rpy[i] = e(i);
    return 0;
}

int vo_host_read_gray(const char* path, uint8_t* out, size_t cap, int* rows, int* cols) {
    if (!path || !rows || !cols) return fail(-1, "null argument");
    std::string err;
    const vo::GrayImage g = vo::readGray(path, &err);
    if (g.empty()) return fail(-2, err);
    *rows = g.rows; *cols = g.cols;
    if (out) {
        if (cap < g.data.size()) return fail(-1, "buffer too small");
        std::memcpy(out, g.data.data(), g.data.size());
    }
    return 0;
}

int vo_host_write_png(const char* path, const uint8_t* gray, int rows, int cols) {
    if (!path || !gray || rows < 1 || cols < 1) return fail(-1, "bad argument");
    return vo::writePNG(path, gray, rows, cols) ? 0 : fail(-2, "cannot write file");
}

int vo_host_write_pgm(const char* path, const uint8_t* gray, int rows, int cols) {
    if (!path || !gray || rows < 1 || cols < 1) return fail(-1, "bad argument");
    return vo::writePGM(path, gray, rows, cols) ? 0 : fail(-2, "cannot write file");
}

int vo_pipeline_run(int device, int num_threads, const char* image_dir, int num_images, const char* pose_file,
                    const char* output_csv, int preset, int batch, uint64_t seed, double* stats) {
    return vo_pipeline_run_ex(device, num_threads, image_dir, num_images, pose_file, output_csv, preset, batch, seed, 0, stats);
}

int vo_pipeline_run_ex(int device, int num_threads, const char* image_dir, int num_images, const char* pose_file,
                       const char* output_csv, int preset, int batch, uint64_t seed, int flags, double* stats) {
    if (!image_dir || !pose_file || !output_csv || num_images < 1 || num_threads < 1) return fail(-1, "bad argument");
    return guarded([&] {
        vo::VisualOdometry vo_("", static_cast<size_t>(num_threads), device);
        if (vo_params_preset(&vo_.params, preset) != VO_OK) return fail(-1, "unknown preset");
        if (batch > 0) vo_.batch = batch;
        vo_.seed = seed;
        vo_.reuse_model = (flags & VO_PIPE_REUSE_MODEL) != 0;
        if (const char* e = std::getenv("VO_CAND_CAP")) vo_.cand_cap = std::atoi(e);  // tuning / test aid: candidates per frame
        // the context of the previous call on this device is kept (one per process): a second sequence does not pay the
        // device allocations again, like a long-lived VisualOdometry object
        {
            std::lock_guard<std::mutex> lk(g_ctx_mutex);
            if (g_ctx.ctx && g_ctx.device == device) { vo_.adopt_context(g_ctx); g_ctx = vo::VisualOdometry::Context(); }
        }
        struct Keep {
            vo::VisualOdometry& v;
            ~Keep() {
                vo::VisualOdometry::Context c = v.release_context();
                std::lock_guard<std::mutex> lk(g_ctx_mutex);
                if (g_ctx.ctx) vo_destroy(g_ctx.ctx);
                g_ctx = c;
            }
        } keep{vo_};
        vo_.verbose = false;
        vo_.run(image_dir, static_cast<size_t>(num_images), pose_file, output_csv);
        const vo::RunStats& s = vo_.stats;
        if (s.frames == 0) return fail(-2, "pipeline did not run (pose file or first frame unreadable)");
        if (stats) {
            const double v[14] = {double(s.frames), double(s.pairs_ok), double(s.skipped_frames), double(s.missing_frames),
                                  double(s.respeculated), s.wall_s, s.decode_s, s.frontend_s, s.host_s, s.compose_s,
                                  double(s.ransac_iterations), s.mean_keypoints, s.mean_matches, s.mean_inliers};
            std::memcpy(stats, v, sizeof(v));
            stats[14] = double(s.stale_model_pairs);
            stats[15] = double(s.capacity_frames);
        }
        return 0;
    });
}

int vo_pipeline_release(void) {
    return guarded([&] {
        {
            std::lock_guard<std::mutex> lk(g_ctx_mutex);
            if (g_ctx.ctx) vo_destroy(g_ctx.ctx);
            g_ctx = vo::VisualOdometry::Context();
        }
        vo::release_pinned_pool();
        return 0;
    });
}

}  // extern "C"
