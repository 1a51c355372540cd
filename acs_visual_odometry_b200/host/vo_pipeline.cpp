// vo_pipeline.cpp — see vo_pipeline.hpp.  Calls the GPU front end only through the C ABI (include/vofront.h).
#include "vo_pipeline.hpp"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <future>
#include <mutex>
#include <stdexcept>
#include <thread>

namespace vo {
namespace {

double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// fn(i) for i in [0, n) on up to `threads` host threads (dynamic distribution)
template <class Fn>
void parallel_for(size_t n, size_t threads, Fn fn) {
    if (threads < 1) threads = 1;
    if (threads > n) threads = n;
    if (threads <= 1) { for (size_t i = 0; i < n; ++i) fn(i); return; }
    std::atomic<size_t> next{0};
    std::vector<std::thread> pool;
    pool.reserve(threads);
    for (size_t t = 0; t < threads; ++t)
        pool.emplace_back([&] { for (size_t i = next++; i < n; i = next++) fn(i); });
    for (auto& th : pool) th.join();
}

// Page-locked staging buffers are expensive to create (the driver pins every page) and a run needs about 100 MB of
// them: they are kept in a process-wide pool between runs, like the GPU context (a second sequence through
// vo_pipeline_run pays neither again).  The pool is bounded: a block serves a request only when it is at most twice
// the size asked for (a frame-sized block is never spent on a 260-byte count array), blocks coming back beyond
// VO_PINNED_POOL_MB (default 1024) of pooled memory go back to the driver, oldest first, and release_pinned_pool()
// (vo_pipeline_release in the C ABI) returns everything.
struct PinnedPool {
    std::mutex m;
    std::vector<std::pair<void*, size_t>> free_blocks;  // oldest first
    size_t pooled = 0, cap = 0;
    PinnedPool() {
        const char* e = std::getenv("VO_PINNED_POOL_MB");
        cap = static_cast<size_t>(e ? std::max(0L, std::atol(e)) : 1024L) << 20;
    }
    void* acquire(size_t bytes, size_t* got) {
        {
            std::lock_guard<std::mutex> lk(m);
            size_t best = free_blocks.size();  // best fit: the smallest block that is large enough, and not wastefully so
            for (size_t i = 0; i < free_blocks.size(); ++i) {
                const size_t sz = free_blocks[i].second;
                if (sz >= bytes && sz / 2 <= bytes && (best == free_blocks.size() || sz < free_blocks[best].second)) best = i;
            }
            if (best < free_blocks.size()) {
                void* p = free_blocks[best].first;
                *got = free_blocks[best].second;
                pooled -= *got;
                free_blocks.erase(free_blocks.begin() + best);
                return p;
            }
        }
        *got = bytes;
        return vo_host_alloc(bytes);
    }
    void release(void* p, size_t bytes) {
        if (!p) return;
        std::vector<void*> drop;
        {
            std::lock_guard<std::mutex> lk(m);
            free_blocks.emplace_back(p, bytes);
            pooled += bytes;
            while (pooled > cap && !free_blocks.empty()) {
                drop.push_back(free_blocks.front().first);
                pooled -= free_blocks.front().second;
                free_blocks.erase(free_blocks.begin());
            }
        }
        for (void* q : drop) vo_host_free(q);
    }
    size_t trim() {
        std::vector<std::pair<void*, size_t>> all;
        {
            std::lock_guard<std::mutex> lk(m);
            all.swap(free_blocks);
            pooled = 0;
        }
        size_t bytes = 0;
        for (auto& b : all) { vo_host_free(b.first); bytes += b.second; }
        return bytes;
    }
};
PinnedPool& pinned_pool() {
    static PinnedPool* pool = new PinnedPool();  // leaked on purpose: no destruction-order games with the CUDA runtime at exit
    return *pool;
}

void check(vo_ctx* ctx, int rc, const char* what) {
    if (rc == VO_OK) return;
    throw std::runtime_error(std::string(what) + ": " + vo_last_error(ctx));
}

}  // namespace

size_t release_pinned_pool() { return pinned_pool().trim(); }

struct VisualOdometry::PairResult {
    enum Status { OK, FEW_MATCHES, FEW_INLIERS, DEGENERATE } status = FEW_MATCHES;
    Mat3 R = Mat3::identity();
    Vec3 t{{0, 0, 0}};  // unit length; scaled with the ground-truth step in the composition stage
    int n_matches = 0, n_inliers = 0, iterations = 0;
    Mat3 F = Mat3::zero();           // the fitted model of an OK pair (what the reference's long-lived model would hold)
    std::vector<PointPair> inliers;
};

VisualOdometry::VisualOdometry(const std::string&, std::size_t num_threads, int device)
    : number_of_threads_(num_threads ? num_threads : 1), device_(device) {
    vo_params_preset(&params, VO_PRESET_VO_GPU_REF);
}

VisualOdometry::~VisualOdometry() {
    if (ctx_) vo_destroy(ctx_);
}

void VisualOdometry::adopt_context(const Context& c) {
    if (!c.ctx) return;
    if (ctx_ || c.device != device_ || c.max_keypoints != params.max_keypoints || c.batch != batch || c.cand_cap != cand_cap) {
        vo_destroy(c.ctx);
        return;
    }
    ctx_ = c.ctx; ctx_rows_ = c.rows; ctx_cols_ = c.cols; ctx_batch_ = c.batch; ctx_max_kp_ = c.max_keypoints;
}

VisualOdometry::Context VisualOdometry::release_context() {
    Context c;
    c.ctx = ctx_; c.rows = ctx_rows_; c.cols = ctx_cols_; c.max_keypoints = ctx_max_kp_; c.batch = ctx_batch_; c.cand_cap = cand_cap;
    c.device = device_;
    ctx_ = nullptr; ctx_rows_ = ctx_cols_ = 0;
    return c;
}

void VisualOdometry::ensure_ctx(int rows, int cols) {
    if (ctx_ && rows <= ctx_rows_ && cols <= ctx_cols_ && ctx_batch_ == batch && ctx_max_kp_ == params.max_keypoints) return;
    if (ctx_) { vo_destroy(ctx_); ctx_ = nullptr; }
    const int slots = 2 * (batch + 1) + 2;  // two chunk sets of batch+1 frames, plus two for on-demand pairs
    const int rc = vo_create_ex(&ctx_, device_, rows, cols, params.max_keypoints, slots, cand_cap);
    if (rc != VO_OK) {
        const std::string msg = vo_last_error(nullptr);
        ctx_ = nullptr;
        throw std::runtime_error("vo_create failed: " + msg);  // like the reference's "No OpenCL devices found" (VisualOdometry.cpp:18)
    }
    ctx_rows_ = rows; ctx_cols_ = cols; ctx_batch_ = batch; ctx_max_kp_ = params.max_keypoints;
}

// VisualOdometry.cpp:28-31
std::pair<std::vector<std::vector<uint8_t>>, std::vector<KeyPoint>> VisualOdometry::compute_descriptor_with_key_points(const GrayImage& image) {
    std::pair<std::vector<std::vector<uint8_t>>, std::vector<KeyPoint>> out;
    if (image.empty()) return out;
    ensure_ctx(image.rows, image.cols);
    check(ctx_, vo_extract(ctx_, image.data.data(), image.rows, image.cols, static_cast<size_t>(image.cols), &params, 0), "vo_extract");
    int n = 0;
    std::vector<float> kp(2 * static_cast<size_t>(params.max_keypoints));
    std::vector<uint8_t> desc(512 * static_cast<size_t>(params.max_keypoints));
    check(ctx_, vo_frame_fetch(ctx_, 0, &n, kp.data(), desc.data(), nullptr), "vo_frame_fetch");
    out.first.resize(n);
    out.second.resize(n);
    for (int i = 0; i < n; ++i) {
        out.first[i].assign(desc.begin() + 512 * static_cast<size_t>(i), desc.begin() + 512 * static_cast<size_t>(i + 1));
        out.second[i].pt.x = kp[2 * i]; out.second[i].pt.y = kp[2 * i + 1];
    }
    return out;
}

// VisualOdometry.cpp:34-36
std::vector<std::pair<int, int>> VisualOdometry::match_descriptors(const std::vector<std::vector<uint8_t>>& d1,
                                                                   const std::vector<std::vector<uint8_t>>& d2) {
    std::vector<std::pair<int, int>> out;
    if (d1.empty() || d2.empty()) return out;
    if (!ctx_) throw std::runtime_error("match_descriptors before any extraction: no context");
    auto flatten = [](const std::vector<std::vector<uint8_t>>& d) {
        std::vector<uint8_t> f(d.size() * 512);
        for (size_t i = 0; i < d.size(); ++i) {
            if (d[i].size() != 512) throw std::runtime_error("descriptor rows must hold 512 bytes");
            std::memcpy(&f[i * 512], d[i].data(), 512);
        }
        return f;
    };
    const std::vector<uint8_t> f1 = flatten(d1), f2 = flatten(d2);
    std::vector<int32_t> pairs(2 * d1.size());
    int n = 0;
    check(ctx_, vo_match_host(ctx_, f1.data(), static_cast<int>(d1.size()), f2.data(), static_cast<int>(d2.size()), 512,
                              params.ratio, params.match_bits, pairs.data(), static_cast<int>(d1.size()), &n), "vo_match_host");
    out.resize(n);
    for (int i = 0; i < n; ++i) out[i] = {pairs[2 * i], pairs[2 * i + 1]};
    return out;
}

// VisualOdometry.cpp:108-177 for one frame pair: matched points -> RANSAC -> inliers -> getPose (unit translation)
void VisualOdometry::solve_pair(const float* kp1, const float* kp2, const int32_t* pairs, int n_matches, size_t pair_index,
                                PairResult& out) const {
    out = PairResult();
    out.n_matches = n_matches;
    if (n_matches < 8) { out.status = PairResult::FEW_MATCHES; return; }
    std::vector<PointPair> matched(static_cast<size_t>(n_matches));
    for (int m = 0; m < n_matches; ++m) {
        const int i1 = pairs[2 * m], i2 = pairs[2 * m + 1];
        matched[m] = {Point{static_cast<double>(kp1[2 * i1]), static_cast<double>(kp1[2 * i1 + 1])},
                      Point{static_cast<double>(kp2[2 * i2]), static_cast<double>(kp2[2 * i2 + 1])}};
    }
    FundamentalMatrix model;  // fresh per pair (the reference reuses one object across frames, VisualOdometry.cpp:49)
    Ransac ransac;
    const RansacStats rs = ransac.run(model, matched, ransac_probability, sampson_threshold, static_cast<int>(number_of_threads_),
                                      seed ? seed + pair_index : 0);
    out.iterations = rs.iterations;
    const auto& inl = model.getInliers();
    out.n_inliers = static_cast<int>(inl.size());
    if (inl.size() < 8) { out.status = PairResult::FEW_INLIERS; return; }
    std::vector<Point2f> p1(inl.size()), p2(inl.size());
    for (size_t i = 0; i < inl.size(); ++i) {
        p1[i] = Point2f{static_cast<float>(inl[i].first.x), static_cast<float>(inl[i].first.y)};
        p2[i] = Point2f{static_cast<float>(inl[i].second.x), static_cast<float>(inl[i].second.y)};
    }
    try {
        PoseUpdate estimator;
        const PoseResult pr = estimator.getPoseEx(model.getMatrix(), p1, p2, 1.0);
        out.R = pr.R; out.t = pr.t;
        out.status = PairResult::OK;
        if (reuse_model) { out.F = model.getMatrix(); out.inliers = inl; }
    } catch (const std::runtime_error&) {
        out.status = PairResult::DEGENERATE;  // the reference would terminate on the uncaught exception
    }
}

namespace {

struct Chunk {
    size_t first = 0, last = 0;   // frame indices [first, last) this chunk is responsible for
    std::vector<size_t> ids;      // frames held in `pixels`, in order; ids[0] is the previous readable frame
    std::vector<char> missing;    // per frame of [first, last)
    uint8_t* pixels = nullptr;    // pinned, (batch + 1) frames
    // results of vo_sequence_run, pinned as well: its device-to-host copies run asynchronously on their own stream and
    // overlap the next chunk's kernels only into page-locked memory
    int32_t* nkp = nullptr; int32_t* nmatch = nullptr; int32_t* pairs = nullptr;
    float* kp = nullptr;
    size_t cap[5] = {0, 0, 0, 0, 0};  // pool capacities of pixels, nkp, nmatch, kp, pairs
    double decode_s = 0;
};

}  // namespace

// VisualOdometry.cpp:38-193
void VisualOdometry::run(const std::string image_dir, const size_t num_images, const std::string pose_file,
                         const std::string output_csv) {
    const double t_run = now_s();
    stats = RunStats();
    estimated_poses_.clear();
    std::vector<Mat4> gt_poses;
    if (!readGTFile(pose_file, gt_poses)) {
        std::fprintf(stderr, "Failed to open pose file.\n");
        return;
    }
    Mat4 T_curr = Mat4::identity();
    estimated_poses_.push_back(T_curr);
    std::string err;
    const GrayImage first = readGray(frameFileName(image_dir, 0), &err);
    if (first.empty()) {
        std::fprintf(stderr, "Failed to load image: %s (%s)\n", frameFileName(image_dir, 0).c_str(), err.c_str());
        return;
    }
    const int rows = first.rows, cols = first.cols;
    ensure_ctx(rows, cols);
    const size_t fbytes = static_cast<size_t>(rows) * cols;
    const int maxkp = params.max_keypoints;
    const size_t nchunks = num_images > 1 ? (num_images - 1 + batch - 1) / batch : 0;

    Chunk chunks[3];
    auto free_chunks = [&] {
        for (Chunk& c : chunks) {
            pinned_pool().release(c.pixels, c.cap[0]); pinned_pool().release(c.nkp, c.cap[1]); pinned_pool().release(c.nmatch, c.cap[2]);
            pinned_pool().release(c.kp, c.cap[3]); pinned_pool().release(c.pairs, c.cap[4]);
            c.pixels = nullptr; c.nkp = c.nmatch = c.pairs = nullptr; c.kp = nullptr;
        }
    };
    for (Chunk& c : chunks) {
        const size_t nf = static_cast<size_t>(batch) + 1;
        c.pixels = static_cast<uint8_t*>(pinned_pool().acquire(fbytes * nf, &c.cap[0]));
        c.nkp = static_cast<int32_t*>(pinned_pool().acquire(sizeof(int32_t) * nf, &c.cap[1]));
        c.nmatch = static_cast<int32_t*>(pinned_pool().acquire(sizeof(int32_t) * nf, &c.cap[2]));
        c.kp = static_cast<float*>(pinned_pool().acquire(sizeof(float) * nf * maxkp * 2, &c.cap[3]));
        c.pairs = static_cast<int32_t*>(pinned_pool().acquire(sizeof(int32_t) * nf * maxkp * 2, &c.cap[4]));
        if (!c.pixels || !c.nkp || !c.nmatch || !c.kp || !c.pairs) { free_chunks(); throw std::runtime_error("vo_host_alloc failed"); }
    }
    std::vector<uint8_t> anchor(fbytes);  // pixels of the query frame while a failure streak lasts
    size_t anchor_id = static_cast<size_t>(-1);

    // ---- stage 1: decode chunk k (frames 1 + k*batch ...), prefixed with the last readable frame before it
    auto decode = [&](size_t k) {
        const double t0 = now_s();
        Chunk& c = chunks[k % 3];
        c.first = 1 + k * batch;
        c.last = std::min(num_images, c.first + batch);
        const size_t cnt = c.last - c.first;
        c.ids.clear();
        c.missing.assign(cnt, 0);
        if (k == 0) {
            std::memcpy(c.pixels, first.data.data(), fbytes);
            c.ids.push_back(0);
        } else {
            const Chunk& p = chunks[(k - 1) % 3];
            std::memcpy(c.pixels, p.pixels + (p.ids.size() - 1) * fbytes, fbytes);
            c.ids.push_back(p.ids.back());
        }
        // decode into per-frame scratch positions, then compact over missing frames
        std::vector<char> ok(cnt, 0);
        parallel_for(cnt, number_of_threads_, [&](size_t j) {
            const GrayImage g = readGray(frameFileName(image_dir, c.first + j));
            if (!g.empty() && g.rows == rows && g.cols == cols) {
                std::memcpy(c.pixels + (1 + j) * fbytes, g.data.data(), fbytes);
                ok[j] = 1;
            }
        });
        size_t w = 1;
        for (size_t j = 0; j < cnt; ++j) {
            if (!ok[j]) { c.missing[j] = 1; continue; }
            if (w != 1 + j) std::memmove(c.pixels + w * fbytes, c.pixels + (1 + j) * fbytes, fbytes);
            c.ids.push_back(c.first + j);
            ++w;
        }
        c.decode_s = now_s() - t0;
    };

    // ---- stage 2: the GPU front end on every frame of the chunk, consecutive pairs matched
    auto frontend = [&](Chunk& c) {
        const int n = static_cast<int>(c.ids.size());
        std::fill(c.nkp, c.nkp + n, 0);
        std::fill(c.nmatch, c.nmatch + (n > 1 ? n - 1 : 0), 0);
        if (n < 2) return;
        const double t0 = now_s();
        vo_seq_result out{c.nkp, c.kp, c.nmatch, c.pairs};
        const int rc = vo_sequence_run(ctx_, c.pixels, n, rows, cols, static_cast<size_t>(cols), fbytes, &params, batch + 1, &out);
        if (rc == VO_ERR_CAPACITY) {
            // a frame had more NMS candidates than the context holds (vo_create_ex sizes the list): its keypoints would come
            // from an arbitrary subset, so the frame is treated like one without features - its pairs carry no matches and
            // take the reference's "too few matches" branch (VisualOdometry.cpp:108-115) - instead of ending the run
            std::vector<int32_t> flags(n);
            check(ctx_, vo_sequence_flags(ctx_, flags.data(), n), "vo_sequence_flags");
            for (int k = 0; k < n; ++k) {
                if (!(flags[k] & 1)) continue;
                if (k > 0 || c.first == 1) ++stats.capacity_frames;  // ids[0] of a later chunk was counted as the last frame of the one before
                c.nkp[k] = 0;
                if (k > 0) c.nmatch[k - 1] = 0;
                if (k + 1 < n) c.nmatch[k] = 0;
            }
        } else {
            check(ctx_, rc, "vo_sequence_run");
        }
        stats.frontend_s += now_s() - t0;
    };

    // ---- stage 3: RANSAC + pose for every speculative pair (ids[j-1], ids[j]) of the chunk
    std::vector<PairResult> results[3];
    auto host_stage = [&](size_t k) {
        const double t0 = now_s();
        Chunk& c = chunks[k % 3];
        std::vector<PairResult>& res = results[k % 3];
        const size_t np = c.ids.size() > 1 ? c.ids.size() - 1 : 0;
        res.assign(np, PairResult());
        parallel_for(np, number_of_threads_, [&](size_t j) {
            solve_pair(&c.kp[j * maxkp * 2], &c.kp[(j + 1) * maxkp * 2], &c.pairs[j * maxkp * 2], c.nmatch[j], c.ids[j + 1], res[j]);
        });
        return now_s() - t0;
    };

    // ---- stage 4: sequential composition with the reference's skip rules
    size_t last_valid = 0;
    Mat3 model_F = Mat3::zero();               // reuse_model: the state of the reference's long-lived model
    std::vector<PointPair> model_inliers;
    double sum_kp = 0, sum_m = 0, sum_inl = 0;
    size_t n_kp = 0, n_m = 0;
    bool warned_gt = false;
    std::vector<float> kpa(static_cast<size_t>(maxkp) * 2), kpb(static_cast<size_t>(maxkp) * 2);
    std::vector<int32_t> prs(static_cast<size_t>(maxkp) * 2);
    auto compose = [&](size_t k) {
        const double t0 = now_s();
        Chunk& c = chunks[k % 3];
        const std::vector<PairResult>& res = results[k % 3];
        size_t j = 0;  // position in ids of the last readable frame seen
        for (size_t i = c.first; i < c.last; ++i) {
            if (c.missing[i - c.first]) {  // VisualOdometry.cpp:78-83: pose repeated without the flip
                std::fprintf(stderr, "Failed to load image: %s\n", frameFileName(image_dir, i).c_str());
                estimated_poses_.push_back(T_curr);
                ++stats.missing_frames;
                continue;
            }
            ++j;
            PairResult r;
            if (last_valid == c.ids[j - 1]) {
                r = res[j - 1];
                sum_kp += c.nkp[j]; ++n_kp;
            } else {  // the speculative pair started from a frame that failed: redo (last valid frame, i)
                ++stats.respeculated;
                const int sa = 2 * (batch + 1), sb = sa + 1;
                int na = 0, nb = 0, nm = 0;
                // A frame that overflows its candidate list does so here as well as on the chunk path (same pixels, same
                // parameters), where it was flagged and counted: VO_ERR_CAPACITY from any of these calls makes the pair one
                // without matches (r stays FEW_MATCHES, VisualOdometry.cpp:108-115), every other error still ends the run.
                bool overflow = false;
                auto soft = [&](int rc, const char* what) {
                    if (rc == VO_ERR_CAPACITY) overflow = true;
                    else check(ctx_, rc, what);
                    return !overflow;
                };
                if (soft(vo_extract(ctx_, anchor.data(), rows, cols, static_cast<size_t>(cols), &params, sa), "vo_extract") &&
                    soft(vo_extract(ctx_, c.pixels + j * fbytes, rows, cols, static_cast<size_t>(cols), &params, sb), "vo_extract") &&
                    soft(vo_frame_fetch(ctx_, sa, &na, kpa.data(), nullptr, nullptr), "vo_frame_fetch") &&
                    soft(vo_frame_fetch(ctx_, sb, &nb, kpb.data(), nullptr, nullptr), "vo_frame_fetch") &&
                    soft(vo_match(ctx_, sa, sb, params.ratio, params.match_bits, prs.data(), maxkp, &nm), "vo_match"))
                    solve_pair(kpa.data(), kpb.data(), prs.data(), nm, i, r);
            }
            stats.ransac_iterations += r.iterations;
            sum_m += r.n_matches; ++n_m;
            if (reuse_model) {
                if (r.status == PairResult::OK) {
                    model_F = r.F; model_inliers = r.inliers;
                } else if (r.status == PairResult::FEW_INLIERS && model_inliers.size() >= 8) {
                    // the reference's model object still holds the previous pair's fit: pose from that (see vo_pipeline.hpp)
                    std::vector<Point2f> p1(model_inliers.size()), p2(model_inliers.size());
                    for (size_t q = 0; q < model_inliers.size(); ++q) {
                        p1[q] = Point2f{static_cast<float>(model_inliers[q].first.x), static_cast<float>(model_inliers[q].first.y)};
                        p2[q] = Point2f{static_cast<float>(model_inliers[q].second.x), static_cast<float>(model_inliers[q].second.y)};
                    }
                    PoseUpdate estimator;
                    const PoseResult pr = estimator.getPoseEx(model_F, p1, p2, 1.0);
                    r.R = pr.R; r.t = pr.t; r.n_inliers = static_cast<int>(model_inliers.size());
                    r.status = PairResult::OK;
                    ++stats.stale_model_pairs;
                }
            }
            if (r.status != PairResult::OK) {
                if (r.status == PairResult::FEW_MATCHES && verbose) std::fprintf(stderr, "Too few matches at frame %zu\n", i);
                if (anchor_id != last_valid) {  // keep the query frame's pixels for the pairs to come
                    if (last_valid == c.ids[j - 1]) std::memcpy(anchor.data(), c.pixels + (j - 1) * fbytes, fbytes);
                    anchor_id = last_valid;
                }
                estimated_poses_.push_back(mul(flipZ, T_curr));
                ++stats.skipped_frames;
                continue;
            }
            sum_inl += r.n_inliers;
            double gt_scale = 1.0;
            if (i < gt_poses.size() && last_valid < gt_poses.size()) {
                const Mat4 rel = mul(inverse_affine(gt_poses[i]), gt_poses[last_valid]);
                gt_scale = std::sqrt(rel(0, 3) * rel(0, 3) + rel(1, 3) * rel(1, 3) + rel(2, 3) * rel(2, 3));
            } else if (!warned_gt) {
                std::fprintf(stderr, "Pose file has fewer lines than frames: unit step length used from frame %zu\n", i);
                warned_gt = true;
            }
            last_valid = i;
            Mat4 T_rel = Mat4::identity();
            for (int a = 0; a < 3; ++a) {
                for (int b = 0; b < 3; ++b) T_rel(a, b) = r.R(a, b);
                T_rel(a, 3) = r.t(a) * gt_scale;
            }
            T_curr = mul(T_curr, T_rel);
            estimated_poses_.push_back(mul(flipZ, T_curr));
            ++stats.pairs_ok;
        }
        stats.compose_s += now_s() - t0;
    };

    try {
        std::future<void> fut_dec;
        std::future<double> fut_host;
        if (nchunks) fut_dec = std::async(std::launch::async, decode, size_t(0));
        for (size_t k = 0; k < nchunks; ++k) {
            fut_dec.get();
            stats.decode_s += chunks[k % 3].decode_s;
            if (k + 1 < nchunks) fut_dec = std::async(std::launch::async, decode, k + 1);
            frontend(chunks[k % 3]);
            if (k > 0) { stats.host_s += fut_host.get(); }
            fut_host = std::async(std::launch::async, host_stage, k);
            if (k > 0) compose(k - 1);
        }
        if (nchunks) { stats.host_s += fut_host.get(); compose(nchunks - 1); }
    } catch (...) {
        free_chunks();
        throw;
    }
    free_chunks();

    if (!writePoseCSV(output_csv, estimated_poses_)) std::fprintf(stderr, "Failed to open output CSV file.\n");
    else if (verbose) std::printf("Wrote estimated poses to: %s\n", output_csv.c_str());
    stats.frames = num_images;
    stats.mean_keypoints = n_kp ? sum_kp / n_kp : 0;
    stats.mean_matches = n_m ? sum_m / n_m : 0;
    stats.mean_inliers = stats.pairs_ok ? sum_inl / stats.pairs_ok : 0;
    stats.wall_s = now_s() - t_run;
}

}  // namespace vo
