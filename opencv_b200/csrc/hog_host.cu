// hog_host.cu -- host side of the B200-native HOG detector: the C ABI of include/hogb200.h.
//
// Mirrors HOG_Impl (modules/cudaobjdetect/src/hog.cpp:139-689): same parameters, defaults, argument checks and
// call sequence, but
//   * the whole pyramid of a frame is three launches (resize, fused gradient+histogram, SVM+compaction) instead
//     of ~5 launches + 11 constant copies + 2 device syncs + 1 blocking download PER LEVEL (hog.cpp:449-482);
//   * all scratch lives in a per-handle workspace sized once per frame shape (no BufferPool traffic);
//   * hits come back as one compact candidate list per frame (or per chunk of frames), sorted on the host into
//     the reference's (level, row-major) order before rect scaling and groupRectangles.
// No CPU fallback: every work call needs the CUDA device the handle was created on.
#include "../../include/hogb200.h"
#include "hog_internal.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cfloat>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

using namespace hogk;

namespace {

thread_local std::string g_err;

int fail(int code, const std::string& msg) { g_err = msg; return code; }

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess) {                                                                  \
            cudaGetLastError();                                                                    \
            return fail(HOGB200_CUDA_ERROR, std::string(#call) + ": " + cudaGetErrorString(e__)); \
        }                                                                                          \
    } while (0)

inline int cvRound(double v) { return (int)lrint(v); }          // round-half-even, as the reference's cvRound
inline int cvFloorF(float v) { int i = (int)v; return i - (v < (float)i); }

template <typename T>
struct DevBuf {
    T* p = nullptr; size_t cap = 0;
    cudaError_t reserve(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = n + n / 8 + 64;
        cudaError_t e = cudaMalloc((void**)&p, want * sizeof(T));
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

template <typename T>
struct PinBuf {
    T* p = nullptr; size_t cap = 0;
    cudaError_t reserve(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFreeHost(p);
        p = nullptr; cap = 0;
        size_t want = n + n / 8 + 64;
        cudaError_t e = cudaMallocHost((void**)&p, want * sizeof(T));
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

struct LevelHost {
    double scale; int w, h; int nbx, nby; int nwx, nwy;
    long long hist_off, score_off; size_t img_off; int pitch; int mode;
};

// Everything that depends only on (frame shape, channels, parameters): level table, resize coefficient
// tables, kernel work lists.  Built once, reused for every frame of that shape.
struct Plan {
    int W = 0, H = 0, cn = 0;
    int nlevels_param = 0; double scale0 = 0; int win_w = 0, win_h = 0, sx = 0, sy = 0;
    bool single_level = false;
    int group = 1;                       // frames per launch the work lists were built for
    int lane_group = 1;                  // frames between the two workspace lanes (the chunk's full group size, >= group)
    long long pyr_stride = 0, hist_stride = 0;   // per-frame strides inside d_pyr (bytes) / d_hist (floats)
    int strip = 0;                       // block rows per histogram work item (chosen by balance_strip)
    std::vector<LevelHost> lv;
    long long hist_floats = 0, score_floats = 0, pyr_pixels = 0, nblocks = 0, nwindows = 0;
    size_t pyr_bytes = 0;
    int n_resize = 0, n_hist = 0, n_svm = 0, n_svm2 = 0;
    DevBuf<LevelDesc> d_levels, d_levels_l[3];   // level tables of workspace lane 0 / lanes 1..3 (same levels, pyramids lane_group frames apart)
    DevBuf<ResizeCoef> d_xtab, d_ytab;
    DevBuf<ResizeItem> d_ritems; DevBuf<HistItem> d_hitems; DevBuf<SvmItem> d_sitems, d_sitems2;   // tiles of one-detector / two-detector scoring passes
    const uint8_t* pyr_base = nullptr;   // pyramid buffer the level pointers were built for
    bool valid = false;
};

// One chunk (consecutive frames of one shape, at most 64) of an asynchronous batch: where its candidates go and what the
// tail needs once the kernels are done.
struct ChunkJob {
    int f0 = 0, nf = 0;
    size_t cand_off = 0, cap = 0;           // region of the job's candidate buffer: [16-byte header][Candidate * cap]
    std::vector<LevelHost> lv;              // level scales of the plan the chunk ran on
    long long nwindows = 0;
    cudaEvent_t done = nullptr;             // kernels finished and the candidate count is in h_counts[4 * chunk]
};
// A batch in flight (hogb200_submit_batch .. hogb200_collect_batch).
struct BatchJob {
    bool active = false;
    std::vector<hogb200_image> frames;      // the caller's descriptors (the pixel buffers stay the caller's)
    int nslots = 0;
    double hit_threshold = 0; int group_threshold = 0, stride_x = 0, stride_y = 0;     // parameters at submit time
    std::vector<ChunkJob> chunks;
    DevBuf<unsigned char> d_cand;
    PinBuf<unsigned int> h_counts;
    std::vector<cudaEvent_t> events;        // pool, reused from batch to batch
    int launches = 0;
    bool finished = false;                  // the tail has run; results wait for a collect call with enough capacity
    std::vector<hogb200_rect> result; std::vector<size_t> result_off;
};

}  // namespace

namespace hogk { int storage_fail(int code, const std::string& msg) { return fail(code, msg); } }

struct hogb200 {
    int device = 0;
    hogb200_geom g;
    // parameters (defaults hog.cpp:241-249)
    BlockWeights weights;            // separable block window for win_sigma (recomputed by the setter)
    double win_sigma = -1.0, l2hys = 0.2; int gamma = 1; int nlevels = 64; double hit_threshold = 0.0;
    int win_stride_x = 8, win_stride_y = 8; double scale0 = 1.05; int group_threshold = 2;
    int descr_format = HOGB200_DESCR_FORMAT_COL_BY_COL;
    // detectors: reordered to row-major blocks (hog.cpp:303-317) on the device, bias on the host
    DevBuf<float> d_det[kMaxDet]; float bias[kMaxDet] = {0, 0, 0, 0}; bool have_det[kMaxDet] = {false, false, false, false};
    std::vector<float> det_host[kMaxDet];           // as given (column-major blocks + optional bias), for hogb200_save
    // streams
    cudaStream_t own_stream = nullptr, stream = nullptr, copy_stream = nullptr;
    // The batch path deals its frame groups round-robin to LANES: lane 0 = `stream`, lanes 1.. = `lane_stream[]`, each with
    // its own pyramid / histogram workspace, level table and histogram work queue, so that the ragged end of one group's
    // persistent kernels (and the launch gaps between its three kernels) are filled by the next group's kernels
    // (HOGB200_LANES=n, 1 .. kMaxLanes; 1 switches it off).
    static constexpr int kMaxLanes = 4;
    cudaStream_t lane_stream[kMaxLanes - 1] = {nullptr, nullptr, nullptr};
    cudaEvent_t lane_fork = nullptr, lane_join[kMaxLanes - 1] = {nullptr, nullptr, nullptr};
    // staging slots of the batch path rotate ACROSS calls (a batch never starts on the slot the previous one ended on)
    static constexpr int kStage = 4;
    cudaEvent_t up_done[kStage] = {nullptr, nullptr, nullptr, nullptr}, fr_done[kStage] = {nullptr, nullptr, nullptr, nullptr};
    unsigned stage_seq = 0;
    // workspace
    DevBuf<uint8_t> d_frame[kStage * kMaxGroup]; DevBuf<uint8_t> d_raw[kStage * kMaxGroup]; DevBuf<uint8_t> d_pyr; DevBuf<float> d_hist; DevBuf<float> d_scores;
    DevBuf<unsigned int> d_queue;   // work-queue counters of the fused histogram kernel (self re-arming)
    int num_sms = 148;
    DevBuf<unsigned char> d_cand;   // [16-byte header: counter][Candidate * cap]
    size_t cand_cap = 0;
    PinBuf<unsigned char> h_cand;
    DevBuf<float> d_desc;
    DevBuf<unsigned char> d_group;  // workspace of the device groupRectangles (group.cu)
    PinBuf<int> h_post;             // batch path, device post-processing: [total, per-segment counts][rects]
    DevBuf<uint8_t> d_roi_level, d_roi_pad;    // ROI path: the resized level and its reflect-101 padded copy
    Plan plan_roi;                             // single-level plan of the padded ROI level
    BatchJob jobs[2];                          // asynchronous batches in flight (ticket = index)
    size_t job_cap = (size_t)1 << 18;          // candidates per chunk region of a job; grows when a chunk overflows
    cudaStream_t post_stream = nullptr;        // device tails of asynchronous batches (runs beside the next batch's kernels)
    cudaEvent_t post_done = nullptr;
    Plan plan, plan1, plan_grp;     // multi-scale plan, single-scale plan, multi-scale plan of the batch path (frame groups)
    Plan plan_part[kMaxGroup];      // batch path: plans sized for the smaller groups (lead-in, ragged last group)
    // timing
    int timing = 0; cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    hogb200_timings last = {};
};

namespace {

int win_bx(const hogb200* h) { return (h->g.win_w - kBlock) / kCell + 1; }
int win_by(const hogb200* h) { return (h->g.win_h - kBlock) / kCell + 1; }

int use_device(const hogb200* h)
{
    CK(cudaSetDevice(h->device));
    return HOGB200_OK;
}

// Separable block window (HOGCache::init, objdetect/src/hog.cpp:653-728 Gaussian, :763-835 cell interpolation):
// weight(i,j) = gauss(i)gauss(j) * wcell_cy(i) * wcell_cx(j); a0/a1 = the 1-D factor for cell 0 / cell 1.
BlockWeights make_weights(double win_sigma)
{
    BlockWeights W;
    const float sigma = (float)(win_sigma >= 0 ? win_sigma : (kBlock + kBlock) / 8.0);   // hog.cpp:288-291
    const float scale = 1.f / (sigma * sigma * 2);
    for (int i = 0; i < 16; i++) {
        float d = i - kBlock * 0.5f;
        const float gss = std::exp(-(d * d) * scale);
        float cell = (i + 0.5f) / kCell - 0.5f;
        const int c0 = cvFloorF(cell);
        cell -= c0;
        float w0 = 0.f, w1 = 0.f;                  // contributions to cell 0 and cell 1; cells -1 and 2 are dropped
        if (c0 == 0) w0 = 1.f - cell; else if (c0 + 1 == 0) w0 = cell;
        if (c0 == 1) w1 = 1.f - cell; else if (c0 + 1 == 1) w1 = cell;
        W.a0[i] = gss * w0;
        W.a1[i] = gss * w1;
    }
    return W;
}

}  // namespace

namespace {

// ---- level enumeration: HOG_Impl::detectMultiScale, cudaobjdetect/src/hog.cpp:424-441,453 -----------------
void enumerate_levels(int W, int H, int win_w, int win_h, double scale0, int nlevels, std::vector<double>& scales)
{
    scales.clear();
    double scale = 1.0;
    int levels = 0;
    for (levels = 0; levels < nlevels; levels++) {
        scales.push_back(scale);
        if (cvRound(W / scale) < win_w || cvRound(H / scale) < win_h || scale0 <= 1) break;
        scale *= scale0;
    }
    levels = std::max(levels, 1);
    scales.resize(levels);
}

// ---- cv::resize INTER_LINEAR coefficient tables, imgproc/src/resize.cpp:3609-3696 (double -> float -> Q11)
void resize_table(int src, int dst, bool is_x, std::vector<ResizeCoef>& out)
{
    const double inv_scale = (double)dst / src;     // resize.cpp:3730-3731
    const double scale = 1. / inv_scale;            // :3481
    auto clip = [](int x, int a, int b) { return x >= a ? (x < b ? x : b - 1) : a; };   // resize.cpp:1773-1776
    for (int d = 0; d < dst; d++) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = cvFloorF(f);
        f -= s;
        if (is_x) {
            if (s < 0) { f = 0; s = 0; }
            if (s >= src - 1) { f = 0; s = src - 1; }   // :3631-3636 (xmax handling: second tap has weight 0)
        }
        const unsigned c0 = (unsigned)(unsigned short)(short)lrintf((1.f - f) * 2048);
        const unsigned c1 = (unsigned)(unsigned short)(short)lrintf(f * 2048);
        ResizeCoef c;
        if (is_x) {
            c.s0 = s; c.s1 = std::min(s + 1, src - 1); c.c0 = c0 | (c1 << 16); c.c1 = 0;
            // the last column has weight (2048, 0): express it as (0, 2048) on columns src-2, src-1, so that the
            // second tap is ALWAYS the pixel right of the first one and the kernels need a single column pointer
            if (s >= src - 1 && src >= 2) { c.s0 = src - 2; c.s1 = src - 1; c.c0 = (c0 << 16) | c1; }
        }
        else      { c.s0 = clip(s, 0, src); c.s1 = clip(s + 1, 0, src); c.c0 = c0; c.c1 = c1; }   // rows: :1821-1825
        out.push_back(c);
    }
}

bool is_exact_half(int src_w, int src_h, int dst_w, int dst_h)
{
    // resize.cpp:3481-3489,3523-3524: scale == 2 in both directions (tested on the doubles the reference forms)
    const double sx = 1. / ((double)dst_w / src_w), sy = 1. / ((double)dst_h / src_h);
    return std::abs(sx - 2) < DBL_EPSILON && std::abs(sy - 2) < DBL_EPSILON && lrint(sx) == 2 && lrint(sy) == 2;
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// workspace lanes of the batch path (HOGB200_LANES=n overrides, 1 = everything on one stream)
int batch_lanes()
{
    static const int v = [] {
        const char* e = getenv("HOGB200_LANES");
        const int n = e ? atoi(e) : 2;
        return n >= 1 && n <= hogb200::kMaxLanes ? n : 2;
    }();
    return v;
}

// Block rows per histogram work item.  The kernel's workers (warps) pull items longest-first from a queue, so the
// finishing time is that of greedy list scheduling: simulate it for every candidate height and keep the best
// (a taller strip recomputes less -- one cell row per strip is shared with the strip above -- but balances worse).
int balance_strip(const std::vector<LevelHost>& lv, int group, int workers)
{
    if (hist_strip_blocks_y() > 0) return hist_strip_blocks_y();
    const int tbx = hist_tile_blocks_x();
    int best = 8; double best_t = 1e300;
    for (int S = 4; S <= 16; S++) {
        std::vector<int> cost;
        for (const LevelHost& L : lv)
            for (int by = 0; by < L.nby; by += S)
                for (int bx = 0; bx < L.nbx; bx += tbx) cost.push_back(std::min(S, L.nby - by) + 2);   // + shared cell row + set-up
        std::sort(cost.begin(), cost.end(), std::greater<int>());
        std::vector<long long> heap(workers, 0);          // min-heap of worker finishing times
        auto cmp = std::greater<long long>();
        for (int g = 0; g < group; g++)
            for (int c : cost) {
                std::pop_heap(heap.begin(), heap.end(), cmp);
                heap.back() += c;
                std::push_heap(heap.begin(), heap.end(), cmp);
            }
        // the model assumes a worker's speed does not depend on how many others are running; that only holds with a
        // few items per worker, so strips that leave fewer than ~2 items per worker are not considered
        if ((long long)cost.size() * group < (long long)workers * 9 / 5 && S > 4) continue;
        const double t = (double)*std::max_element(heap.begin(), heap.end());
        if (t <= best_t) { best_t = t; best = S; }
    }
    return best;
}

int build_plan(hogb200* h, Plan& P, int W, int H, int cn, bool single, int group = 1, int lane_group = 0)
{
    if (lane_group < group) lane_group = group;
    if (P.valid && P.W == W && P.H == H && P.cn == cn && P.single_level == single && P.group == group && P.lane_group == lane_group &&
        P.nlevels_param == h->nlevels && P.scale0 == h->scale0 && P.win_w == h->g.win_w && P.win_h == h->g.win_h &&
        P.sx == h->win_stride_x && P.sy == h->win_stride_y && P.pyr_base == h->d_pyr.p)
        return HOGB200_OK;
    P.valid = false;
    P.W = W; P.H = H; P.cn = cn; P.single_level = single; P.group = group; P.lane_group = lane_group;
    P.nlevels_param = h->nlevels; P.scale0 = h->scale0; P.win_w = h->g.win_w; P.win_h = h->g.win_h;
    P.sx = h->win_stride_x; P.sy = h->win_stride_y;

    std::vector<double> scales;
    if (single) scales.assign(1, 1.0);
    else enumerate_levels(W, H, h->g.win_w, h->g.win_h, h->scale0, h->nlevels, scales);
    if ((int)scales.size() > kMaxLevels) return fail(HOGB200_BAD_ARG, "too many pyramid levels (max 256)");

    const int wbx = win_bx(h), wby = win_by(h);
    const bool svm_ok = svm_supports(wbx, wby);          // else the plan serves compute / blockHistograms only
    const int tw1 = svm_ok ? svm_tile_wx(wbx, wby, 1) : 16, tw2 = 16;
    const bool pass2 = svm_ok && svm_two_detector_pass(wbx, wby);
    P.lv.clear();
    P.hist_floats = P.score_floats = P.pyr_pixels = P.nblocks = P.nwindows = 0;
    size_t pyr = 0;
    std::vector<ResizeCoef> xtab, ytab;
    std::vector<ResizeItem> ritems; std::vector<HistItem> hitems; std::vector<SvmItem> sitems, sitems2;
    std::vector<int> scost, scost2;
    std::vector<LevelDesc> descs;
    for (size_t i = 0; i < scales.size(); i++) {
        LevelHost L;
        L.scale = scales[i];
        L.w = cvRound(W / L.scale); L.h = cvRound(H / L.scale);                 // hog.cpp:453
        L.nbx = L.w >= kBlock ? (L.w - kBlock) / kCell + 1 : 0;
        L.nby = L.h >= kBlock ? (L.h - kBlock) / kCell + 1 : 0;
        if (L.nbx <= 0 || L.nby <= 0) L.nbx = L.nby = 0;
        // numPartsWithin, hog.cpp:267-276; an image smaller than the window has no windows
        L.nwx = L.w >= h->g.win_w ? (L.w - h->g.win_w + h->win_stride_x) / h->win_stride_x : 0;
        L.nwy = L.h >= h->g.win_h ? (L.h - h->g.win_h + h->win_stride_y) / h->win_stride_y : 0;
        if (L.nwx <= 0 || L.nwy <= 0) L.nwx = L.nwy = 0;
        L.hist_off = P.hist_floats; P.hist_floats += (long long)L.nbx * L.nby * kHist;
        L.score_off = P.score_floats; P.score_floats += (long long)L.nwx * L.nwy;
        P.pyr_pixels += (long long)L.w * L.h; P.nblocks += (long long)L.nbx * L.nby; P.nwindows += (long long)L.nwx * L.nwy;
        LevelDesc D; memset(&D, 0, sizeof D);
        if (L.w == W && L.h == H) {                 // hog.cpp:456-459: the frame itself
            L.mode = 0; L.img_off = 0; L.pitch = 0;
            D.img = nullptr; D.pitch = 0;
        } else {
            L.mode = is_exact_half(W, H, L.w, L.h) ? 2 : 1;
            L.pitch = (int)align_up((size_t)L.w * cn, 64);
            L.img_off = pyr; pyr += align_up((size_t)L.pitch * L.h, 256);
            D.xtab_off = (int)xtab.size(); D.ytab_off = (int)ytab.size();
            if (L.mode == 1) { resize_table(W, L.w, true, xtab); resize_table(H, L.h, false, ytab); }
            if (L.mode == 1) D.rs_rows = resize_plan_rows(&xtab[D.xtab_off], L.w, &ytab[D.ytab_off], L.h, W, cn);
            const int tw = resize_tile_w(cn), th = resize_tile_h(cn, D.rs_rows);
            for (int y = 0; y < L.h; y += th)
                for (int x = 0; x < L.w; x += tw) ritems.push_back(ResizeItem{(short)i, (short)x, (short)y, 0});
        }
        D.w = L.w; D.h = L.h; D.nbx = L.nbx; D.nby = L.nby; D.nwx = L.nwx; D.nwy = L.nwy;
        D.hist_off = L.hist_off; D.score_off = L.score_off; D.mode = L.mode; D.pitch = L.pitch;
        const int vwx = L.nbx - wbx + 1, vwy = L.nby - wby + 1;
        if (svm_ok && L.nwx > 0 && vwx > 0 && vwy > 0)
            for (int wy = 0; wy < vwy; wy += svm_tile_wy()) {
                for (int wx = 0; wx < vwx; wx += tw1) {
                    sitems.push_back(SvmItem{(short)i, (short)wx, (short)wy, 0});
                    scost.push_back(svm_tile_cost(vwx - wx, vwy - wy, tw1 > 16 && vwx - wx > 16 ? 2 : 1));
                }
                for (int wx = 0; pass2 && wx < vwx; wx += tw2) {
                    sitems2.push_back(SvmItem{(short)i, (short)wx, (short)wy, 0});
                    scost2.push_back(svm_tile_cost(vwx - wx, vwy - wy, 2));
                }
            }
        P.lv.push_back(L); descs.push_back(D);
    }
    // histogram work items: strips of P.strip block rows x tiles of 14 blocks, for every frame of the group;
    // the kernel's workers pull them from a queue, longest first, so that the tail is made of short ones
    P.strip = balance_strip(P.lv, group, h->num_sms * hist_workers_per_sm());
    for (size_t i = 0; i < P.lv.size(); i++) {
        const LevelHost& L = P.lv[i];
        const int tbx = hist_tile_blocks_x();
        for (int by = 0; by < L.nby; by += P.strip)
            for (int bx = 0; bx < L.nbx; bx += tbx)
                hitems.push_back(HistItem{(short)i, (short)bx, (short)by, (short)std::min(P.strip, L.nby - by)});
    }
    std::stable_sort(hitems.begin(), hitems.end(), [](const HistItem& a, const HistItem& b) {
        return a.nby > b.nby;       // a partial tile costs a warp as much as a full one
    });
    // SVM tiles are dealt round-robin to the persistent CTAs: most expensive first, so that the partly filled edge
    // tiles even out the tail
    auto by_cost = [](std::vector<SvmItem>& items, const std::vector<int>& cost) {
        std::vector<int> order(items.size());
        for (size_t i = 0; i < order.size(); i++) order[i] = (int)i;
        std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return cost[a] > cost[b]; });
        std::vector<SvmItem> sorted; sorted.reserve(items.size());
        for (int i : order) sorted.push_back(items[i]);
        items.swap(sorted);
    };
    by_cost(sitems, scost); by_cost(sitems2, scost2);
    // replicate the lists for frames 1 .. group-1 (frame-minor, so that equal-sized items of all frames are adjacent)
    if (group > 1) {
        std::vector<ResizeItem> r2; std::vector<HistItem> h2; std::vector<SvmItem> s2;
        for (const ResizeItem& it : ritems) for (int f = 0; f < group; f++) { ResizeItem t = it; t.frame = (short)f; r2.push_back(t); }
        for (const HistItem& it : hitems) for (int f = 0; f < group; f++) { HistItem t = it; t.level = (short)(it.level | (f << 8)); h2.push_back(t); }
        for (const SvmItem& it : sitems) for (int f = 0; f < group; f++) { SvmItem t = it; t.frame = (short)f; s2.push_back(t); }
        ritems.swap(r2); hitems.swap(h2); sitems.swap(s2);
        s2.clear();
        for (const SvmItem& it : sitems2) for (int f = 0; f < group; f++) { SvmItem t = it; t.frame = (short)f; s2.push_back(t); }
        sitems2.swap(s2);
    }
    P.pyr_stride = (long long)align_up(pyr, 256);
    P.hist_stride = P.hist_floats;
    P.pyr_bytes = pyr;
    const int nlanes = single ? 1 : batch_lanes();
    CK(h->d_pyr.reserve((size_t)P.pyr_stride * lane_group * nlanes + 256));   // one workspace per lane (see hogb200::lane_stream)
    for (size_t i = 0; i < descs.size(); i++)
        if (P.lv[i].mode != 0) descs[i].img = h->d_pyr.p + P.lv[i].img_off;
    P.pyr_base = h->d_pyr.p;
    CK(h->d_hist.reserve((size_t)P.hist_stride * lane_group * nlanes + 64));
    P.n_resize = (int)ritems.size(); P.n_hist = (int)hitems.size(); P.n_svm = (int)sitems.size(); P.n_svm2 = (int)sitems2.size();
    cudaStream_t st = h->stream;
    CK(cudaStreamSynchronize(st));      // the previous plan's tables may still be in use,
    for (cudaStream_t ls : h->lane_stream) CK(cudaStreamSynchronize(ls));      // on any lane
    auto up = [&](auto& buf, const auto& v) -> cudaError_t {
        cudaError_t e = buf.reserve(v.size() + 1);
        if (e != cudaSuccess || v.empty()) return e;
        return cudaMemcpy(buf.p, v.data(), v.size() * sizeof(v[0]), cudaMemcpyHostToDevice);
    };
    CK(up(P.d_levels, descs));
    for (int l = 1; l < nlanes; l++) {
        std::vector<LevelDesc> dl = descs;
        for (size_t i = 0; i < dl.size(); i++)
            if (P.lv[i].mode != 0) dl[i].img = descs[i].img + (size_t)P.pyr_stride * lane_group * l;
        CK(up(P.d_levels_l[l - 1], dl));
    }
    CK(up(P.d_xtab, xtab)); CK(up(P.d_ytab, ytab));
    CK(up(P.d_ritems, ritems)); CK(up(P.d_hitems, hitems)); CK(up(P.d_sitems, sitems)); CK(up(P.d_sitems2, sitems2));
    P.valid = true;
    return HOGB200_OK;
}

}  // namespace

namespace {

// dev / pitch: what the kernels read.  raw != nullptr: the frame was uploaded flat (rows `raw_pitch` bytes apart, not a
// multiple of 16) and still has to be re-pitched into dev on the compute stream (queue_group does it for the whole group)
struct FrameRef { const uint8_t* dev; int pitch; const uint8_t* raw = nullptr; int raw_pitch = 0; int row_bytes = 0, rows = 0; };

int check_image(const hogb200_image* img)
{
    if (!img || !img->data) return fail(HOGB200_BAD_ARG, "image is null");
    // 8UC1 / 8UC4 as cv::cuda::HOG (cudaobjdetect/src/hog.cpp:340); 8UC3 (B,G,R) is widened to 8UC4 on the device,
    // which is the cvtColor example_gpu_tbd runs on the host before its upload (samples/gpu/tbd.cpp:573-599)
    if (img->channels != 1 && img->channels != 3 && img->channels != 4)
        return fail(HOGB200_BAD_ARG, "image type must be CV_8UC1, CV_8UC3 or CV_8UC4");
    if (img->width <= 0 || img->height <= 0 || img->pitch < (size_t)img->width * img->channels)
        return fail(HOGB200_BAD_ARG, "bad image size / pitch");
    if (img->width > 32000 || img->height > 32000) return fail(HOGB200_BAD_ARG, "image too large");
    return HOGB200_OK;
}

// HOGB200_FLAT_UPLOAD=0: host frames with unaligned rows take the 2-D copy again
bool flat_upload()
{
    static const bool v = [] { const char* e = getenv("HOGB200_FLAT_UPLOAD"); return !e || atoi(e) != 0; }();
    return v;
}

// Make the frame device-resident and 4-byte addressable.  buf = 0/1 selects the staging buffer (double
// buffering in the batch path).  Copies are issued on `st`.
int stage_frame(hogb200* h, const hogb200_image* img, int buf, cudaStream_t st, FrameRef& out, bool allow_raw = false)
{
    out.raw = nullptr;
    const int cn = img->channels;
    const size_t row = (size_t)img->width * cn;
    if (cn == 3) {                   // B,G,R -> B,G,R,255 on the device; host frames travel as 3 bytes per pixel
        const uint8_t* src = (const uint8_t*)img->data;
        size_t spitch = img->pitch;
        if (!img->on_device) {
            spitch = align_up(row, 64);
            CK(h->d_raw[buf].reserve(spitch * img->height + 64));
            CK(cudaMemcpy2DAsync(h->d_raw[buf].p, spitch, img->data, img->pitch, row, img->height, cudaMemcpyHostToDevice, st));
            src = h->d_raw[buf].p;
        }
        const size_t pitch = align_up((size_t)img->width * 4, 64);
        CK(h->d_frame[buf].reserve(pitch * img->height + 64));
        launch_bgr_to_bgra(src, spitch, h->d_frame[buf].p, pitch, img->width, img->height, st);
        CK(cudaGetLastError());
        out.dev = h->d_frame[buf].p; out.pitch = (int)pitch;
        return HOGB200_OK;
    }
    // the kernels read rows in 4-pixel vectors: 16-byte aligned rows for 8UC4, 4-byte for 8UC1, and (so that
    // the last vector of the last row stays inside the caller's buffer) a width that is a multiple of 4
    // unless the pitch leaves room; anything else goes through the staging copy below
    const size_t al = cn == 4 ? 16 : 4;
    if (img->on_device && ((uintptr_t)img->data % al == 0) && (img->pitch % al == 0) && img->pitch < (1u << 30) &&
        (img->width % 4 == 0 || img->pitch >= align_up(row, al))) {
        out.dev = (const uint8_t*)img->data; out.pitch = (int)img->pitch;
        return HOGB200_OK;
    }
    const size_t pitch = align_up(row, 64);
    CK(h->d_frame[buf].reserve(pitch * img->height + 64));
    if (allow_raw && img->on_device && row % 4 == 0 && img->pitch % 4 == 0 && (uintptr_t)img->data % 4 == 0 && flat_upload()) {
        // a device frame whose rows are not 16-byte aligned: the re-pitch kernel reads it where it lies
        out.dev = h->d_frame[buf].p; out.pitch = (int)pitch;
        out.raw = (const uint8_t*)img->data; out.raw_pitch = (int)img->pitch; out.row_bytes = (int)row; out.rows = img->height;
        return HOGB200_OK;
    }
    if (allow_raw && !img->on_device && img->pitch == row && pitch != row && row % 4 == 0 && flat_upload()) {
        // a contiguous host frame whose rows are not 64-byte multiples (1242 x 4 = 4968 bytes): a 2-D host->device copy of
        // such rows reaches 31 GB/s, a flat one 54 GB/s; the rows are moved to their aligned places by a kernel on the
        // compute stream instead (a device-to-device 2-D copy on the copy stream was measured too: slower than either)
        CK(h->d_raw[buf].reserve(row * img->height + 64));
        CK(cudaMemcpyAsync(h->d_raw[buf].p, img->data, row * img->height, cudaMemcpyHostToDevice, st));
        out.dev = h->d_frame[buf].p; out.pitch = (int)pitch;
        out.raw = h->d_raw[buf].p; out.raw_pitch = (int)row; out.row_bytes = (int)row; out.rows = img->height;
        return HOGB200_OK;
    }
    CK(cudaMemcpy2DAsync(h->d_frame[buf].p, pitch, img->data, img->pitch, row, img->height,
                         img->on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
    out.dev = h->d_frame[buf].p; out.pitch = (int)pitch;
    return HOGB200_OK;
}

int ensure_cand(hogb200* h, size_t cap)
{
    if (cap < 4096) cap = 4096;
    if (cap > h->cand_cap || !h->d_cand.p) {
        CK(cudaStreamSynchronize(h->stream));
        CK(h->d_cand.reserve(16 + cap * sizeof(Candidate)));
        h->cand_cap = (h->d_cand.cap - 16) / sizeof(Candidate);
    }
    return HOGB200_OK;
}

// Queue the three kernels for a group of frames (1 .. P.group) on the handle's stream.
// Debugging aid (HOGB200_WATCHDOG=<seconds>): an event after every kernel of the batch path; if the stream has not
// drained after that many seconds, the first unfinished event names the kernel that hangs, and the process aborts.
struct Watchdog {
    struct Mark { cudaEvent_t ev; const char* what; int group_frames; };
    std::vector<Mark> marks; size_t used = 0;
    double seconds = 0;
    Watchdog() { const char* e = getenv("HOGB200_WATCHDOG"); seconds = e ? atof(e) : 0; }
    void mark(cudaStream_t st, const char* what, int nf)
    {
        if (seconds <= 0) return;
        if (used == marks.size()) { Mark m; cudaEventCreateWithFlags(&m.ev, cudaEventDisableTiming); marks.push_back(m); }
        marks[used].what = what; marks[used].group_frames = nf;
        cudaEventRecord(marks[used].ev, st);
        used++;
    }
    void drain(cudaStream_t st)
    {
        if (seconds <= 0) return;
        const auto t0 = std::chrono::steady_clock::now();
        while (cudaStreamQuery(st) == cudaErrorNotReady) {
            if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > seconds) {
                for (size_t i = 0; i < used; i++)
                    if (cudaEventQuery(marks[i].ev) == cudaErrorNotReady) {
                        fprintf(stderr, "hogb200 watchdog: kernel %zu of %zu since the last drain (%s, %d frames) has not finished after %.0f s\n",
                                i, used, marks[i].what, marks[i].group_frames, seconds);
                        break;
                    }
                abort();
            }
        }
        used = 0;
    }
};
Watchdog g_watchdog;

// Debugging aid (HOGB200_TRACE=1): the upload and kernel intervals of every frame group of a batch call, in ms from
// the first group's upload, printed to stderr when the call has drained.  Off: no events, no synchronisation.
struct GroupTrace {
    struct Row { cudaEvent_t ev[4]; };       // upload begin / end (copy stream), kernels begin / end (compute stream)
    std::vector<Row> rows; size_t used = 0;
    const bool on = getenv("HOGB200_TRACE") != nullptr;
    void begin() { used = 0; }
    void mark(int which, cudaStream_t st)
    {
        if (!on) return;
        if (which == 0) {
            if (used == rows.size()) { Row r; for (cudaEvent_t& e : r.ev) cudaEventCreate(&e); rows.push_back(r); }
            used++;
        }
        cudaEventRecord(rows[used - 1].ev[which], st);
    }
    void print()
    {
        if (!on || !used) return;
        cudaDeviceSynchronize();
        for (size_t g = 0; g < used; g++) {
            float t[4] = {-1, -1, -1, -1};
            for (int k = 0; k < 4; k++) cudaEventElapsedTime(&t[k], rows[0].ev[0], rows[g].ev[k]);
            fprintf(stderr, "group %2zu: upload %.3f .. %.3f   kernels %.3f .. %.3f\n", g, t[0], t[1], t[2], t[3]);
        }
    }
};
GroupTrace g_trace;

// where a pass appends its candidates: [16-byte header: counter][Candidate * cap] in device memory
struct CandTarget { unsigned char* p; size_t cap; };

int queue_group(hogb200* h, Plan& P, const FrameRef* f, int nframes, int ndet, const int* slots, float thresh,
                float* d_scores, int tag, int* launches, const CandTarget* target = nullptr, int lane = 0)
{
    const CandTarget own{h->d_cand.p, h->cand_cap};
    const CandTarget& ct = target ? *target : own;
    cudaStream_t st = lane ? h->lane_stream[lane - 1] : h->stream;
    const LevelDesc* d_levels = lane ? P.d_levels_l[lane - 1].p : P.d_levels.p;
    float* d_hist = h->d_hist.p + (size_t)P.hist_stride * P.lane_group * lane;
    unsigned int* d_queue = h->d_queue.p + 2 * lane;
    FrameSet fs; memset(&fs, 0, sizeof fs);
    for (int i = 0; i < nframes; i++) { fs.base[i] = f[i].dev; fs.pitch[i] = f[i].pitch; }
    fs.nframes = nframes; fs.pyr_stride = P.pyr_stride; fs.hist_stride = P.hist_stride;
    {   // frames that were uploaded flat: rows to their aligned places (stage_frame)
        RepitchSet rp; memset(&rp, 0, sizeof rp);
        for (int i = 0; i < nframes; i++)
            if (f[i].raw) {
                rp.src[rp.n] = f[i].raw; rp.dst[rp.n] = const_cast<uint8_t*>(f[i].dev);
                rp.spitch = f[i].raw_pitch; rp.dpitch = f[i].pitch; rp.row_bytes = f[i].row_bytes; rp.rows = f[i].rows;
                rp.n++;
            }
        if (rp.n) { launch_repitch(rp, st); (*launches)++; }
    }
    if (h->timing) cudaEventRecord(h->ev[1], st);
    if (P.n_resize) {
        launch_resize(fs, P.W, P.H, P.cn, d_levels, P.d_xtab.p, P.d_ytab.p, P.d_ritems.p, P.n_resize, st);
        (*launches)++;
        g_watchdog.mark(st, "resize", nframes);
    }
    if (h->timing) cudaEventRecord(h->ev[2], st);
    if (P.n_hist) {
        launch_hist(fs, d_levels, P.d_hitems.p, P.n_hist, d_queue, h->num_sms, P.cn, h->gamma,
                    h->weights, (float)h->l2hys, d_hist, st);
        (*launches)++;
        g_watchdog.mark(st, "hist", nframes);
    }
    if (h->timing) cudaEventRecord(h->ev[3], st);
    if (ndet > 0 && P.n_svm) {
        SvmParams sp; memset(&sp, 0, sizeof sp);
        sp.win_bx = win_bx(h); sp.win_by = win_by(h);
        sp.stride_bx = h->win_stride_x / kCell; sp.stride_by = h->win_stride_y / kCell;
        sp.ndet = ndet;
        for (int d = 0; d < ndet; d++) { sp.det[d] = h->d_det[slots[d]].p; sp.bias[d] = h->bias[slots[d]]; }
        sp.thresh = thresh; sp.scores = d_scores; sp.score_plane = P.score_floats;
        sp.cand = reinterpret_cast<Candidate*>(ct.p + 16);
        sp.cand_count = reinterpret_cast<unsigned int*>(ct.p);
        sp.cand_cap = (unsigned int)ct.cap; sp.tag = tag;
        launch_svm(fs, d_levels, P.d_sitems.p, P.n_svm, P.d_sitems2.p, P.n_svm2, d_hist, sp, h->num_sms, st);
        (*launches) += ndet >= 2 && P.n_svm2 ? (ndet + 1) / 2 : ndet;
        g_watchdog.mark(st, "svm", nframes);
    }
    if (h->timing) cudaEventRecord(h->ev[4], st);
    CK(cudaGetLastError());
    return HOGB200_OK;
}

// Bring the candidate list of the finished chunk to the host.  Returns HOGB200_CAPACITY (with *count set) when
// the device list overflowed and the chunk must be re-run with a larger list.
int fetch_candidates(hogb200* h, std::vector<Candidate>& out, size_t* count)
{
    cudaStream_t st = h->stream;
    const size_t first = std::min<size_t>(h->cand_cap, 4096);
    CK(h->h_cand.reserve(16 + h->cand_cap * sizeof(Candidate)));
    CK(cudaMemcpyAsync(h->h_cand.p, h->d_cand.p, 16 + first * sizeof(Candidate), cudaMemcpyDeviceToHost, st));
    g_watchdog.drain(st);
    CK(cudaStreamSynchronize(st));
    const size_t n = *reinterpret_cast<unsigned int*>(h->h_cand.p);
    *count = n;
    if (n > h->cand_cap) return HOGB200_CAPACITY;
    if (n > first) {
        CK(cudaMemcpyAsync(h->h_cand.p + 16 + first * sizeof(Candidate), h->d_cand.p + 16 + first * sizeof(Candidate),
                           (n - first) * sizeof(Candidate), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
    }
    const Candidate* c = reinterpret_cast<const Candidate*>(h->h_cand.p + 16);
    out.assign(c, c + n);
    return HOGB200_OK;
}

inline bool cand_less(const Candidate& a, const Candidate& b)
{
    if (a.level != b.level) return (unsigned)a.level < (unsigned)b.level;   // frame, detector, level
    if (a.y != b.y) return a.y < b.y;
    return a.x < b.x;
}

// ---- cv::groupRectangles(rectList, groupThreshold, eps), objdetect/src/cascadedetect.cpp:64-174 ----------
// Same result as the reference: classes are the connected components of the SimilarRects relation
// (objdetect.hpp:121-134), numbered by first appearance (cv::partition, core/operations.hpp:481-569), but the
// O(N^2) predicate scan is cut down by sorting on x: similar rects differ in x by at most
// eps*(w_i+h_i)/2, so only a window of the sorted order is examined.
struct Uf {
    std::vector<int> p;
    explicit Uf(int n) : p(n) { for (int i = 0; i < n; i++) p[i] = i; }
    int find(int a) { while (p[a] != a) { p[a] = p[p[a]]; a = p[a]; } return a; }
    void unite(int a, int b) { a = find(a); b = find(b); if (a != b) p[std::max(a, b)] = std::min(a, b); }
};

inline bool similar(const hogb200_rect& r1, const hogb200_rect& r2, double eps)
{
    const double delta = eps * (std::min(r1.width, r2.width) + std::min(r1.height, r2.height)) * 0.5;
    return std::abs(r1.x - r2.x) <= delta && std::abs(r1.y - r2.y) <= delta &&
           std::abs(r1.x + r1.width - r2.x - r2.width) <= delta &&
           std::abs(r1.y + r1.height - r2.y - r2.height) <= delta;
}

void group_rectangles(std::vector<hogb200_rect>& rects, int groupThreshold, double eps)
{
    if (groupThreshold <= 0 || rects.empty()) return;
    const int N = (int)rects.size();
    std::vector<int> order(N);
    for (int i = 0; i < N; i++) order[i] = i;
    std::sort(order.begin(), order.end(), [&](int a, int b) { return rects[a].x < rects[b].x; });
    Uf uf(N);
    for (int a = 0; a < N; a++) {
        const hogb200_rect& ra = rects[order[a]];
        const double reach = eps * (ra.width + ra.height) * 0.5;
        for (int b = a + 1; b < N; b++) {
            const hogb200_rect& rb = rects[order[b]];
            if (rb.x - ra.x > reach) break;
            if (similar(ra, rb, eps)) uf.unite(order[a], order[b]);
        }
    }
    // class labels in order of first appearance (root = smallest member index => first appearance of the class)
    std::vector<int> label(N, -1);
    int nclasses = 0;
    std::vector<int> cls(N);
    for (int i = 0; i < N; i++) {
        const int r = uf.find(i);
        if (label[r] < 0) label[r] = nclasses++;
        cls[i] = label[r];
    }
    std::vector<hogb200_rect> rr(nclasses, hogb200_rect{0, 0, 0, 0});
    std::vector<int> rw(nclasses, 0);
    for (int i = 0; i < N; i++) {
        hogb200_rect& r = rr[cls[i]];
        r.x += rects[i].x; r.y += rects[i].y; r.width += rects[i].width; r.height += rects[i].height;
        rw[cls[i]]++;
    }
    for (int i = 0; i < nclasses; i++) {
        const float s = 1.f / rw[i];
        hogb200_rect r = rr[i];
        rr[i] = hogb200_rect{(int)lrintf(r.x * s), (int)lrintf(r.y * s), (int)lrintf(r.width * s), (int)lrintf(r.height * s)};
    }
    rects.clear();
    for (int i = 0; i < nclasses; i++) {
        const hogb200_rect r1 = rr[i];
        const int n1 = rw[i];
        if (n1 <= groupThreshold) continue;
        int j;
        for (j = 0; j < nclasses; j++) {
            const int n2 = rw[j];
            if (j == i || n2 <= groupThreshold) continue;
            const hogb200_rect r2 = rr[j];
            const int dx = cvRound(r2.width * eps), dy = cvRound(r2.height * eps);
            if (r1.x >= r2.x - dx && r1.y >= r2.y - dy && r1.x + r1.width <= r2.x + r2.width + dx &&
                r1.y + r1.height <= r2.y + r2.height + dy && (n2 > std::max(3, n1) || n1 < 3))
                break;
        }
        if (j == nclasses) rects.push_back(r1);
    }
}

// cv::groupRectangles on the device (group.cu): same classes, same order, same rects as group_rectangles above
int group_rectangles_device(hogb200* h, std::vector<hogb200_rect>& rects, int groupThreshold, double eps)
{
    if (groupThreshold <= 0 || rects.empty()) return HOGB200_OK;
    const int n = (int)rects.size();
    CK(cudaSetDevice(h->device));
    CK(h->d_group.reserve(hogk::group_workspace_bytes(n)));
    std::vector<hogb200_rect> out(n);
    const int m = hogk::group_rectangles_device(&rects[0].x, n, groupThreshold, eps, &out[0].x, h->d_group.p, h->stream);
    if (m < 0) return fail(HOGB200_CUDA_ERROR, std::string("groupRectangles on the device: ") + cudaGetErrorString(cudaGetLastError()));
    out.resize(m);
    rects.swap(out);
    return HOGB200_OK;
}

// frames with this many candidates or more are grouped on the device (HOGB200_GPU_GROUP_MIN overrides; 0 = never)
size_t gpu_group_min()
{
    static const size_t v = [] {
        const char* e = getenv("HOGB200_GPU_GROUP_MIN");
        return e ? (size_t)atoll(e) : (size_t)1024;
    }();
    return v;
}

// a CUDA failure of the device path is reported, never papered over by regrouping on the host
int group_rectangles_auto(hogb200* h, std::vector<hogb200_rect>& rects, int groupThreshold, double eps)
{
    if (gpu_group_min() > 0 && rects.size() >= gpu_group_min()) return group_rectangles_device(h, rects, groupThreshold, eps);
    group_rectangles(rects, groupThreshold, eps);
    return HOGB200_OK;
}

// candidates of one (frame, detector), already in (level, y, x) order -> rects (hog.cpp:471-481) [+ grouping :484-487]
int candidates_to_rects(hogb200* h, const Plan& P, const Candidate* c, size_t n,
                        std::vector<hogb200_rect>& rects, std::vector<double>* conf)
{
    rects.clear();
    if (conf) conf->clear();
    for (size_t i = 0; i < n; i++) {
        const int lvl = c[i].level & 0xffff;
        const double scale = P.lv[lvl].scale;
        hogb200_rect r;
        r.x = cvRound((double)(c[i].x * h->win_stride_x) * scale);
        r.y = cvRound((double)(c[i].y * h->win_stride_y) * scale);
        r.width = cvRound(h->g.win_w * scale);
        r.height = cvRound(h->g.win_h * scale);
        rects.push_back(r);
        if (conf) conf->push_back((double)c[i].score);
    }
    if (h->group_threshold > 0) return group_rectangles_auto(h, rects, h->group_threshold, 0.2);
    return HOGB200_OK;
}

void finish_timing(hogb200* h, const Plan& P, int launches, size_t ncand)
{
    h->last = hogb200_timings{};
    h->last.kernel_launches = launches;
    h->last.pyramid_pixels = P.pyr_pixels; h->last.blocks = P.nblocks; h->last.windows = P.nwindows;
    h->last.candidates = (long long)ncand;
    if (!h->timing) return;
    cudaEventSynchronize(h->ev[5]);
    cudaEventElapsedTime(&h->last.upload_ms, h->ev[0], h->ev[1]);
    cudaEventElapsedTime(&h->last.resize_ms, h->ev[1], h->ev[2]);
    cudaEventElapsedTime(&h->last.hist_ms, h->ev[2], h->ev[3]);
    cudaEventElapsedTime(&h->last.svm_ms, h->ev[3], h->ev[4]);
    cudaEventElapsedTime(&h->last.download_ms, h->ev[4], h->ev[5]);
    cudaEventElapsedTime(&h->last.total_ms, h->ev[0], h->ev[5]);
}

}  // namespace

// =============================================================================== C ABI
extern "C" {

const char* hogb200_last_error(void) { return g_err.c_str(); }
const char* hogb200_version(void) { return "hogb200 0.1 (sm_100a)"; }

int hogb200_create(const hogb200_geom* geom, int device, hogb200_t** out)
{
    if (!out) return fail(HOGB200_BAD_ARG, "out is null");
    *out = nullptr;
    hogb200_geom g = {64, 128, 16, 16, 8, 8, 8, 8, 9};
    if (geom) g = *geom;
    // ctor asserts, hog.cpp:252-264
    if (g.block_stride_w <= 0 || g.block_stride_h <= 0 || g.cell_w <= 0 || g.cell_h <= 0 || g.block_w <= 0 || g.block_h <= 0)
        return fail(HOGB200_BAD_ARG, "sizes must be positive");
    if ((g.win_w - g.block_w) % g.block_stride_w != 0 || (g.win_h - g.block_h) % g.block_stride_h != 0)
        return fail(HOGB200_BAD_ARG, "(win_size - block_size) % block_stride != 0");
    if (g.block_w % g.cell_w != 0 || g.block_h % g.cell_h != 0) return fail(HOGB200_BAD_ARG, "block_size % cell_size != 0");
    if (g.block_stride_w * 2 != g.block_w || g.block_stride_h * 2 != g.block_h)
        return fail(HOGB200_BAD_ARG, "block_stride must equal block_size / 2");
    if (g.cell_w != g.cell_h) return fail(HOGB200_BAD_ARG, "cells must be square");
    if (g.block_w != kBlock || g.block_h != kBlock || g.cell_w != kCell || g.nbins != kBins)
        return fail(HOGB200_UNSUPPORTED, "only block 16x16 / cell 8x8 / 9 bins is supported (cudaobjdetect.hpp:89-93)");
    if (g.win_w < kBlock || g.win_h < kBlock) return fail(HOGB200_BAD_ARG, "window smaller than a block");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
        cudaGetLastError();
        return fail(HOGB200_CUDA_ERROR, "no CUDA device: hogb200 has no CPU fallback");
    }
    if (device < 0 || device >= ndev) return fail(HOGB200_BAD_ARG, "bad device index");
    hogb200* h = new hogb200();
    h->device = device; h->g = g;
    h->weights = make_weights(h->win_sigma);
    h->win_stride_x = g.block_stride_w; h->win_stride_y = g.block_stride_h;   // hog.cpp:246
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking);
    for (int l = 0; l < hogb200::kMaxLanes - 1 && e == cudaSuccess; l++) {
        e = cudaStreamCreateWithFlags(&h->lane_stream[l], cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->lane_join[l], cudaEventDisableTiming);
    }
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->lane_fork, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->post_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->post_done, cudaEventDisableTiming);
    for (int i = 0; i < hogb200::kStage && e == cudaSuccess; i++) {
        e = cudaEventCreateWithFlags(&h->up_done[i], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->fr_done[i], cudaEventDisableTiming);
    }
    for (int i = 0; i < 6 && e == cudaSuccess; i++) e = cudaEventCreate(&h->ev[i]);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&h->num_sms, cudaDevAttrMultiProcessorCount, device);
    if (e == cudaSuccess) e = h->d_queue.reserve(2 * hogb200::kMaxLanes);
    if (e == cudaSuccess) e = cudaMemset(h->d_queue.p, 0, 2 * hogb200::kMaxLanes * sizeof(unsigned int));     // [head, finished] per lane
    if (e != cudaSuccess) {
        cudaGetLastError();
        hogb200_destroy(h);
        return fail(HOGB200_CUDA_ERROR, std::string("create: ") + cudaGetErrorString(e));
    }
    h->stream = h->own_stream;
    *out = h;
    return HOGB200_OK;
}

void hogb200_destroy(hogb200_t* h)
{
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    for (int i = 0; i < kMaxDet; i++) h->d_det[i].release();
    for (auto& b : h->d_frame) b.release();
    for (auto& b : h->d_raw) b.release();
    h->d_pyr.release(); h->d_hist.release();
    h->d_scores.release(); h->d_queue.release(); h->d_cand.release(); h->h_cand.release(); h->d_desc.release(); h->d_group.release(); h->h_post.release();
    std::vector<Plan*> plans = {&h->plan, &h->plan1, &h->plan_grp, &h->plan_roi};
    h->d_roi_level.release(); h->d_roi_pad.release();
    for (Plan& P : h->plan_part) plans.push_back(&P);
    for (Plan* P : plans) {
        P->d_levels.release(); for (auto& b : P->d_levels_l) b.release(); P->d_xtab.release(); P->d_ytab.release();
        P->d_ritems.release(); P->d_hitems.release(); P->d_sitems.release(); P->d_sitems2.release();
    }
    for (int i = 0; i < hogb200::kStage; i++) { if (h->up_done[i]) cudaEventDestroy(h->up_done[i]); if (h->fr_done[i]) cudaEventDestroy(h->fr_done[i]); }
    for (int i = 0; i < 6; i++) if (h->ev[i]) cudaEventDestroy(h->ev[i]);
    for (BatchJob& J : h->jobs) { J.d_cand.release(); J.h_counts.release(); for (cudaEvent_t e : J.events) cudaEventDestroy(e); }
    if (h->post_stream) cudaStreamDestroy(h->post_stream);
    if (h->post_done) cudaEventDestroy(h->post_done);
    for (cudaStream_t ls : h->lane_stream) if (ls) { cudaStreamSynchronize(ls); cudaStreamDestroy(ls); }
    if (h->lane_fork) cudaEventDestroy(h->lane_fork);
    for (cudaEvent_t ev : h->lane_join) if (ev) cudaEventDestroy(ev);
    if (h->own_stream) cudaStreamDestroy(h->own_stream);
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    delete h;
}

int hogb200_set_stream(hogb200_t* h, void* s)
{
    if (!h) return fail(HOGB200_BAD_ARG, "null handle");
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->stream);
    h->stream = s ? (cudaStream_t)s : h->own_stream;
    return HOGB200_OK;
}

#define H_OR_FAIL if (!h) return fail(HOGB200_BAD_ARG, "null handle")
// getters on a null handle return the reference's defaults (hog.cpp:241-249) instead of dereferencing it
int hogb200_set_win_sigma(hogb200_t* h, double v) { H_OR_FAIL; h->win_sigma = v; h->weights = make_weights(v); return HOGB200_OK; }
double hogb200_get_win_sigma(const hogb200_t* h) { return !h ? 4.0 : h->win_sigma >= 0 ? h->win_sigma : (h->g.block_w + h->g.block_h) / 8.0; }
int hogb200_set_l2hys_threshold(hogb200_t* h, double v) { H_OR_FAIL; h->l2hys = v; return HOGB200_OK; }
double hogb200_get_l2hys_threshold(const hogb200_t* h) { return h ? h->l2hys : 0.2; }
int hogb200_set_gamma_correction(hogb200_t* h, int v) { H_OR_FAIL; h->gamma = v != 0; return HOGB200_OK; }
int hogb200_get_gamma_correction(const hogb200_t* h) { return h ? h->gamma : 1; }
int hogb200_set_num_levels(hogb200_t* h, int v)
{
    H_OR_FAIL;
    // the reference takes any int and then forms scale 0 -> cvRound(W / 0.0) for nlevels <= 0 (hog.cpp:424-441): refused here
    if (v < 1) return fail(HOGB200_BAD_ARG, "nlevels must be >= 1");
    h->nlevels = v; return HOGB200_OK;
}
int hogb200_get_num_levels(const hogb200_t* h) { return h ? h->nlevels : 64; }
int hogb200_set_hit_threshold(hogb200_t* h, double v) { H_OR_FAIL; h->hit_threshold = v; return HOGB200_OK; }
double hogb200_get_hit_threshold(const hogb200_t* h) { return h ? h->hit_threshold : 0.0; }
int hogb200_set_win_stride(hogb200_t* h, int sx, int sy)
{
    H_OR_FAIL;
    if (sx <= 0 || sy <= 0) return fail(HOGB200_BAD_ARG, "win_stride must be positive");
    h->win_stride_x = sx; h->win_stride_y = sy;     // divisibility is checked by the work calls, as hog.cpp:341,497 do
    return HOGB200_OK;
}
void hogb200_get_win_stride(const hogb200_t* h, int* sx, int* sy) { if (sx) *sx = h ? h->win_stride_x : 8; if (sy) *sy = h ? h->win_stride_y : 8; }
int hogb200_set_scale_factor(hogb200_t* h, double v)
{
    H_OR_FAIL;
    // scale0 <= 1 is legal (one level, hog.cpp:433); zero, negative or non-finite factors only produce garbage sizes
    if (!(v > 0) || !std::isfinite(v)) return fail(HOGB200_BAD_ARG, "scale factor must be positive and finite");
    h->scale0 = v; return HOGB200_OK;
}
double hogb200_get_scale_factor(const hogb200_t* h) { return h ? h->scale0 : 1.05; }
int hogb200_set_group_threshold(hogb200_t* h, int v) { H_OR_FAIL; h->group_threshold = v; return HOGB200_OK; }
int hogb200_get_group_threshold(const hogb200_t* h) { return h ? h->group_threshold : 2; }
int hogb200_set_descriptor_format(hogb200_t* h, int v)
{
    H_OR_FAIL;
    if (v != HOGB200_DESCR_FORMAT_COL_BY_COL && v != HOGB200_DESCR_FORMAT_ROW_BY_ROW)
        return fail(HOGB200_BAD_ARG, "Unknown descriptor format");
    h->descr_format = v; return HOGB200_OK;
}
int hogb200_get_descriptor_format(const hogb200_t* h) { return h ? h->descr_format : HOGB200_DESCR_FORMAT_COL_BY_COL; }
size_t hogb200_get_descriptor_size(const hogb200_t* h) { return h ? (size_t)win_bx(h) * win_by(h) * kHist : 0; }
size_t hogb200_get_block_histogram_size(const hogb200_t*) { return kHist; }
int hogb200_enable_timing(hogb200_t* h, int on) { H_OR_FAIL; h->timing = on != 0; return HOGB200_OK; }
int hogb200_last_timings(const hogb200_t* h, hogb200_timings* out) { if (!h || !out) return fail(HOGB200_BAD_ARG, "null"); *out = h->last; return HOGB200_OK; }

int hogb200_set_svm_detector(hogb200_t* h, int slot, const float* coeffs, size_t n)
{
    H_OR_FAIL;
    if (slot < 0 || slot >= kMaxDet) return fail(HOGB200_BAD_ARG, "bad detector slot");
    if (use_device(h)) return HOGB200_CUDA_ERROR;
    if (n == 0) { CK(cudaStreamSynchronize(h->stream)); h->have_det[slot] = false; h->det_host[slot].clear(); return HOGB200_OK; }
    const size_t ds = hogb200_get_descriptor_size(h);
    if (!coeffs || (n != ds && n != ds + 1))
        return fail(HOGB200_BAD_ARG, "detector.cols must be descriptor_size or descriptor_size + 1 (hog.cpp:301)");
    // reorder blocks from column-major (public / CPU order) to row-major (hog.cpp:303-317)
    const int bw = win_bx(h), bh = win_by(h);
    std::vector<float> re(ds);
    for (int i = 0; i < bh; i++)
        for (int j = 0; j < bw; j++)
            memcpy(&re[(size_t)(i * bw + j) * kHist], coeffs + (size_t)(j * bh + i) * kHist, kHist * sizeof(float));
    CK(cudaStreamSynchronize(h->stream));
    CK(h->d_det[slot].reserve(ds + 4));
    CK(cudaMemcpy(h->d_det[slot].p, re.data(), ds * sizeof(float), cudaMemcpyHostToDevice));
    h->bias[slot] = n > ds ? coeffs[ds] : 0.f;      // hog.cpp:320
    h->have_det[slot] = true;
    h->det_host[slot].assign(coeffs, coeffs + n);
    return HOGB200_OK;
}

int hogb200_save(const hogb200_t* h, const char* filename, const char* objname, int slot)
{
    H_OR_FAIL;
    if (slot < 0 || slot >= kMaxDet) return fail(HOGB200_BAD_ARG, "bad detector slot");
    hogb200_model m;
    m.geom = h->g; m.win_sigma = h->win_sigma; m.l2hys_threshold = h->l2hys; m.gamma_correction = h->gamma; m.nlevels = h->nlevels;
    const std::vector<float>& d = h->det_host[slot];
    return hogb200_model_write(filename, objname, &m, d.empty() ? nullptr : d.data(), d.size());
}

int hogb200_create_from_file(const char* filename, const char* objname, int device, hogb200_t** out)
{
    if (!out) return fail(HOGB200_BAD_ARG, "out is null");
    *out = nullptr;
    hogb200_model m;
    std::vector<float> det(1 << 16);
    size_t n = 0;
    int rc = hogb200_model_read(filename, objname, &m, det.data(), det.size(), &n);
    if (rc) return rc;
    hogb200_t* h = nullptr;
    if ((rc = hogb200_create(&m.geom, device, &h))) return rc;
    // HOG_Impl::read (cudaobjdetect/src/hog.cpp:599-641): the stored parameters, defaults for the rest
    h->win_sigma = m.win_sigma; h->weights = make_weights(m.win_sigma); h->l2hys = m.l2hys_threshold; h->gamma = m.gamma_correction; h->nlevels = m.nlevels;
    if (n && (rc = hogb200_set_svm_detector(h, 0, det.data(), n))) { hogb200_destroy(h); return rc; }
    *out = h;
    return HOGB200_OK;
}

}  // extern "C"

namespace {

// chunks with at least this many candidates keep them on the device: ordering, rect scaling and groupRectangles of all
// (frame, detector) segments run there (group.cu cand_group_device); smaller lists take the host path.  HOGB200_GPU_POST_MIN
// overrides (0 = never on the device, 1 = always).
size_t gpu_post_min()
{
    const char* e = getenv("HOGB200_GPU_POST_MIN");       // read per call: the tests switch paths inside one process
    return e ? (size_t)atoll(e) : (size_t)1024;
}

// header of the candidate buffer only: the number of candidates the chunk produced
int fetch_candidate_count(hogb200* h, size_t* count)
{
    cudaStream_t st = h->stream;
    CK(h->h_cand.reserve(16 + 4096 * sizeof(Candidate)));
    CK(cudaMemcpyAsync(h->h_cand.p, h->d_cand.p, 16, cudaMemcpyDeviceToHost, st));
    g_watchdog.drain(st);
    CK(cudaStreamSynchronize(st));
    *count = *reinterpret_cast<unsigned int*>(h->h_cand.p);
    return HOGB200_OK;
}

// The rest of a chunk's candidate list (the header is already on the host).
int fetch_candidate_list(hogb200* h, size_t n, std::vector<Candidate>& out)
{
    cudaStream_t st = h->stream;
    CK(h->h_cand.reserve(16 + n * sizeof(Candidate)));
    if (n) {
        CK(cudaMemcpyAsync(h->h_cand.p + 16, h->d_cand.p + 16, n * sizeof(Candidate), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
    }
    const Candidate* c = reinterpret_cast<const Candidate*>(h->h_cand.p + 16);
    out.assign(c, c + n);
    return HOGB200_OK;
}

// Device tail of a chunk: n candidates in d_cand -> grouped rects of every (frame, detector) segment, appended to `all`
// with their offsets.  One synchronisation for the counts, one for the rects.
// host tail of a chunk: candidates -> reference order -> rects (+ grouping) per (frame, detector) segment
int host_tail(hogb200* h, const Plan& P, std::vector<Candidate>& c, int nf, int nslots, std::vector<hogb200_rect>& all, size_t* offsets,
              size_t* out_idx)
{
    std::sort(c.begin(), c.end(), cand_less);
    std::vector<hogb200_rect> rects;
    size_t pos = 0;
    for (int i = 0; i < nf; i++)
        for (int d = 0; d < nslots; d++) {
            const int key = (i << 20) | (d << 16);
            size_t e = pos;
            while (e < c.size() && (c[e].level & ~0xffff) == key) e++;
            const int rc = candidates_to_rects(h, P, c.data() + pos, e - pos, rects, nullptr);
            if (rc) return rc;
            pos = e;
            all.insert(all.end(), rects.begin(), rects.end());
            offsets[++*out_idx] = all.size();
        }
    return HOGB200_OK;
}

CandGeom cand_geom(const hogb200* h, const Plan& P)
{
    CandGeom G; memset(&G, 0, sizeof G);
    for (size_t i = 0; i < P.lv.size(); i++) G.scale[i] = P.lv[i].scale;
    G.win_w = h->g.win_w; G.win_h = h->g.win_h; G.stride_x = h->win_stride_x; G.stride_y = h->win_stride_y;
    // widths of the sort-key fields (group.cu): the radix sort makes one pass per 8 key bits
    auto bits = [](long long v) { int b = 1; while ((1LL << b) <= v) b++; return b; };
    int mx = 1, my = 1; long long rx = 1;
    for (const LevelHost& L : P.lv) {
        mx = std::max(mx, L.nwx); my = std::max(my, L.nwy);
        rx = std::max(rx, (long long)std::ceil((double)L.nwx * h->win_stride_x * L.scale) + 1);    // > cvRound(x * stride * scale)
    }
    G.bits_x = bits(mx); G.bits_y = bits(my); G.bits_level = bits((long long)P.lv.size());
    G.bits_rx = bits(rx);                          // (from the levels alone: collect() rebuilds the geometry from a job's level list)
    return G;
}

int post_process_device(hogb200* h, const CandGeom& G, const unsigned char* d_cand, cudaStream_t st, size_t n, int nf, int nslots,
                        int group_threshold, std::vector<hogb200_rect>& all, size_t* offsets, size_t* out_idx)
{
    const int nseg = nf * nslots;
    CK(h->d_group.reserve(cand_group_workspace_bytes((int)n, nseg)));
    CK(h->h_post.reserve((size_t)nseg + 1 + 4 * n));
    const int* d_out = nullptr; const int* d_hdr = nullptr;
    if (cand_group_device(d_cand + 16, (int)n, nseg, nslots, G, group_threshold, 0.2, h->d_group.p, st, &d_out, &d_hdr))
        return fail(HOGB200_CUDA_ERROR, std::string("candidate post-processing on the device: ") + cudaGetErrorString(cudaGetLastError()));
    CK(cudaMemcpyAsync(h->h_post.p, d_hdr, sizeof(int) * (nseg + 1), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    const size_t total = (size_t)h->h_post.p[0];
    if (total > n) return fail(HOGB200_CUDA_ERROR, "candidate post-processing on the device: inconsistent counts");
    int* rects = h->h_post.p + nseg + 1;
    if (total) {
        CK(cudaMemcpyAsync(rects, d_out, total * sizeof(hogb200_rect), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
    }
    const hogb200_rect* r = reinterpret_cast<const hogb200_rect*>(rects);
    all.insert(all.end(), r, r + total);
    size_t acc = offsets[*out_idx];
    for (int sgi = 0; sgi < nseg; sgi++) { acc += (size_t)h->h_post.p[1 + sgi]; offsets[++*out_idx] = acc; }
    return HOGB200_OK;
}

// what the SVM scoring kernel can take (svm.cu); checked before anything is queued, by every entry point that scores
int check_svm_geometry(const hogb200* h)
{
    // any block-aligned window (hog.cpp:252-264) whose coefficient table leaves room for the histogram ring in shared
    // memory: up to 21 blocks (176 px) wide and roughly 300 blocks in all (svm.cu, plan_pass)
    if (!svm_supports(win_bx(h), win_by(h)))
        return fail(HOGB200_UNSUPPORTED, "SVM scoring: window too large (more than 176 pixels wide or more than ~300 blocks)");
    return HOGB200_OK;
}

int check_work_params(const hogb200* h)
{
    if (h->win_stride_x % h->g.block_stride_w != 0 || h->win_stride_y % h->g.block_stride_h != 0)
        return fail(HOGB200_BAD_ARG, "win_stride must be a multiple of block_stride (hog.cpp:341)");
    return HOGB200_OK;
}

// first group of a batch of host frames = one frame (HOGB200_LEAD_IN=0 disables)
bool batch_lead_in()
{
    static const bool v = [] { const char* e = getenv("HOGB200_LEAD_IN"); return !e || atoi(e) != 0; }();
    return v;
}


// Frames per launch in the batch path (HOGB200_GROUP overrides, 1 .. kMaxGroup).  Measured on B200 (bench.py
// --workload ..., profiles/): 4 suits 1080p and KITTI-sized frames; frames whose whole pyramid is a few megapixels
// (VGA) need more per launch to fill the persistent kernels' queues; at 4K two base frames (66 MB) still sit in L2
// for the resize kernel, four do not.
int batch_group(const hogb200* h, int W, int H, int cn)
{
    static const int forced = [] {
        const char* e = getenv("HOGB200_GROUP");
        const int n = e ? atoi(e) : 0;
        return n >= 1 && n <= kMaxGroup ? n : 0;
    }();
    if (forced) return forced;
    std::vector<double> scales;
    enumerate_levels(W, H, h->g.win_w, h->g.win_h, h->scale0, h->nlevels, scales);
    double px = 0;
    for (double s : scales) px += (double)cvRound(W / s) * cvRound(H / s);
    // re-measured with two lanes (profiles/r02h_group_lanes.txt): 4K 1 225 / 1 202 / 1 143 fps at 1 / 2 / 4 frames per launch,
    // 1080p 4 850 / 4 886 / 4 769 at 2 / 4 / 8, KITTI 16 762 / 17 652 at 4 / 8, VGA 33 372 / 33 726 at 8 / 16
    if ((double)W * H * cn > 16e6) return 1;
    if (px < 4e6) return 16;
    if (px < 8e6) return 8;
    return 4;
}

// One frame through the pipeline; candidates come back sorted by (detector, level, y, x).
int run_single(hogb200* h, Plan& P, const hogb200_image* img, bool single, int ndet, const int* slots,
               float thresh, float* d_scores, std::vector<Candidate>& cands)
{
    int rc;
    if ((rc = check_image(img))) return rc;
    if ((rc = check_work_params(h))) return rc;
    if ((rc = use_device(h))) return rc;
    if ((rc = build_plan(h, P, img->width, img->height, img->channels == 3 ? 4 : img->channels, single))) return rc;
    if (ndet > 0 && (rc = check_svm_geometry(h))) return rc;
    size_t cap = std::max<size_t>(h->cand_cap, 1 << 16);
    for (int attempt = 0; attempt < 2; attempt++) {
        if ((rc = ensure_cand(h, cap))) return rc;
        cudaStream_t st = h->stream;
        int launches = 0;
        if (h->timing) cudaEventRecord(h->ev[0], st);
        CK(cudaMemsetAsync(h->d_cand.p, 0, 16, st));
        FrameRef f;
        if ((rc = stage_frame(h, img, 0, st, f))) return rc;
        if ((rc = queue_group(h, P, &f, 1, ndet, slots, thresh, d_scores, 0, &launches))) return rc;
        size_t n = 0;
        rc = fetch_candidates(h, cands, &n);
        if (h->timing) cudaEventRecord(h->ev[5], st);
        if (rc == HOGB200_CAPACITY && attempt == 0) { cap = (size_t)P.nwindows * std::max(ndet, 1) + 16; continue; }
        if (rc) return rc;
        finish_timing(h, P, launches, n);
        break;
    }
    std::sort(cands.begin(), cands.end(), cand_less);
    return HOGB200_OK;
}

template <typename T>
int emit(const std::vector<T>& v, T* out, size_t capacity, size_t* count)
{
    if (count) *count = v.size();
    if (v.size() > capacity) return fail(HOGB200_CAPACITY, "output capacity too small");
    if (!v.empty() && !out) return fail(HOGB200_BAD_ARG, "output pointer is null");
    if (!v.empty()) memcpy(out, v.data(), v.size() * sizeof(T));
    return HOGB200_OK;
}

}  // namespace

namespace {

// Queue every frame group of a chunk (consecutive frames of one shape, plan P built for groups of `grp` frames) on the
// handle's streams: uploads on the copy stream, double-buffered per group, kernels on the compute stream.
int queue_chunk(hogb200* h, Plan& P, const hogb200_image* fr, int nf, int grp, int nslots, const int* slots, const CandTarget* target,
                int* launches, bool allow_lead_in = true)
{
    int rc;
    cudaStream_t st = h->stream, cs = h->copy_stream;
    int& total_launches = *launches;
    g_trace.begin();
    // Nothing hides the upload of the first group, so when frames come from host memory the batch opens with
    // groups of 1 and grp - 1 frames: the kernels start after one frame's copy instead of a whole group's,
    // and the copies (faster than the kernels per frame) stay ahead from then on.
    bool lead_in = false;
    for (int j = 0; j < nf; j++) lead_in |= !fr[j].on_device;
    // (not when a batch is already in flight: its kernels hide this batch's first uploads)
    lead_in = lead_in && allow_lead_in && nf > grp && batch_lead_in() && !h->jobs[0].active && !h->jobs[1].active;
    // groups are dealt round-robin to the lanes (lane 0 = the handle's stream); the other lanes start behind everything
    // queued on the handle's stream so far and are joined back before the chunk's count is read
    const int nlanes = (h->timing || nf <= grp) ? 1 : batch_lanes();
    if (nlanes > 1) {
        CK(cudaEventRecord(h->lane_fork, st));
        for (int l = 1; l < nlanes; l++) CK(cudaStreamWaitEvent(h->lane_stream[l - 1], h->lane_fork, 0));
    }
    for (int g0 = 0, gi = 0, ng = 0; g0 < nf; g0 += ng, gi++) {
        ng = (lead_in && gi == 0) ? 1 : (lead_in && gi == 1 && grp > 2) ? grp - 1 : std::min(grp, nf - g0);
        const int b = (int)(h->stage_seq++ % hogb200::kStage);     // staging slots: uploads run up to two groups ahead of the kernels
        const int lane = gi % nlanes;
        cudaStream_t ks = lane ? h->lane_stream[lane - 1] : st;    // the stream of this group's kernels
        FrameRef f[kMaxGroup];
        bool any_host = false;
        for (int j = 0; j < ng; j++) any_host |= !fr[g0 + j].on_device;
        if (any_host) CK(cudaStreamWaitEvent(cs, h->fr_done[b], 0));
        else CK(cudaStreamWaitEvent(ks, h->fr_done[b], 0));
        g_trace.mark(0, any_host ? cs : ks);
        for (int j = 0; j < ng; j++) {
            const hogb200_image* im = fr + g0 + j;
            // host frames: upload on the copy stream, overlapping the previous group's kernels
            if ((rc = stage_frame(h, im, b * kMaxGroup + j, im->on_device ? ks : cs, f[j], true))) return rc;
        }
        if (any_host) {
            CK(cudaEventRecord(h->up_done[b], cs));
            CK(cudaStreamWaitEvent(ks, h->up_done[b], 0));
        }
        g_trace.mark(1, any_host ? cs : ks);
        g_trace.mark(2, ks);
        // a group smaller than grp runs on a plan of its own size: work lists dealt for grp frames leave the
        // persistent kernels unbalanced when most of their items belong to absent frames
        Plan* Pg = &P;
        if (ng != grp) {
            Pg = &h->plan_part[ng];
            if ((rc = build_plan(h, *Pg, fr->width, fr->height, fr->channels == 3 ? 4 : fr->channels, false, ng, grp))) return rc;
        }
        if ((rc = queue_group(h, *Pg, f, ng, nslots, slots, (float)h->hit_threshold, nullptr, g0 << 20, &total_launches, target, lane))) return rc;
        CK(cudaEventRecord(h->fr_done[b], ks));
        g_trace.mark(3, ks);
    }
    for (int l = 1; l < nlanes; l++) { CK(cudaEventRecord(h->lane_join[l - 1], h->lane_stream[l - 1])); CK(cudaStreamWaitEvent(st, h->lane_join[l - 1], 0)); }
    return HOGB200_OK;
}

}  // namespace

extern "C" {

int hogb200_detect(hogb200_t* h, const hogb200_image* img, hogb200_point* hits, double* conf, size_t capacity, size_t* count)
{
    H_OR_FAIL;
    if (count) *count = 0;
    if (!h->have_det[0]) { int rc = check_image(img); return rc ? rc : check_work_params(h); }   // hog.cpp:343-345
    std::vector<Candidate> c;
    const int slot0 = 0;
    int rc = run_single(h, h->plan1, img, true, 1, &slot0, (float)h->hit_threshold, nullptr, c);
    if (rc) return rc;
    std::vector<hogb200_point> pts; std::vector<double> cf;
    for (const Candidate& k : c) {
        if (conf && !((double)k.score >= h->hit_threshold)) continue;        // hog.cpp:405: float score vs double threshold
        pts.push_back(hogb200_point{k.x * h->win_stride_x, k.y * h->win_stride_y});
        cf.push_back((double)k.score);
    }
    rc = emit(pts, hits, capacity, count);
    if (rc == HOGB200_OK && conf && !cf.empty()) memcpy(conf, cf.data(), cf.size() * sizeof(double));
    return rc;
}

int hogb200_detect_multiscale_slot(hogb200_t* h, int slot, const hogb200_image* img, hogb200_rect* found, double* conf,
                                   size_t capacity, size_t* count)
{
    H_OR_FAIL;
    if (count) *count = 0;
    if (slot < 0 || slot >= kMaxDet) return fail(HOGB200_BAD_ARG, "bad detector slot");
    if (conf && h->group_threshold != 0)
        return fail(HOGB200_BAD_ARG, "confidences require group_threshold == 0 (hog.cpp:422)");
    if (!h->have_det[slot]) { int rc = check_image(img); return rc ? rc : check_work_params(h); }
    std::vector<Candidate> c;
    int rc = run_single(h, h->plan, img, false, 1, &slot, (float)h->hit_threshold, nullptr, c);
    if (rc) return rc;
    if (conf) {   // confidence path compares the float score with the double threshold (hog.cpp:405)
        c.erase(std::remove_if(c.begin(), c.end(), [&](const Candidate& k) { return !((double)k.score >= h->hit_threshold); }), c.end());
    }
    std::vector<hogb200_rect> rects; std::vector<double> cf;
    if ((rc = candidates_to_rects(h, h->plan, c.data(), c.size(), rects, conf ? &cf : nullptr))) return rc;
    rc = emit(rects, found, capacity, count);
    if (rc == HOGB200_OK && conf && !cf.empty()) memcpy(conf, cf.data(), cf.size() * sizeof(double));
    return rc;
}

int hogb200_detect_multiscale(hogb200_t* h, const hogb200_image* img, hogb200_rect* found, double* conf, size_t capacity, size_t* count)
{
    return hogb200_detect_multiscale_slot(h, 0, img, found, conf, capacity, count);
}

int hogb200_detect_multiscale_batch(hogb200_t* h, const hogb200_image* frames, int nframes, int nslots,
                                    hogb200_rect* found, size_t capacity, size_t* offsets)
{
    H_OR_FAIL;
    if (nframes < 0 || !offsets || (nframes > 0 && !frames)) return fail(HOGB200_BAD_ARG, "bad batch arguments");
    if (nslots < 1 || nslots > kMaxDet) return fail(HOGB200_BAD_ARG, "bad nslots");
    for (int d = 0; d < nslots; d++) if (!h->have_det[d]) return fail(HOGB200_BAD_ARG, "detector slot not set");
    int rc;
    if ((rc = check_work_params(h))) return rc;
    if ((rc = check_svm_geometry(h))) return rc;
    if ((rc = use_device(h))) return rc;
    int slots[kMaxDet] = {0, 1, 2, 3};
    std::vector<hogb200_rect> all; all.reserve(1024);
    offsets[0] = 0;
    size_t out_idx = 0;
    int total_launches = 0;
    size_t total_cands = 0;
    const int kChunk = 64;
    std::vector<Candidate> c;
    cudaStream_t st = h->stream;
    if (h->timing) cudaEventRecord(h->ev[0], st);
    for (int i = 0; i < nframes; i++) if ((rc = check_image(&frames[i]))) return rc;
    int f0 = 0;
    while (f0 < nframes) {
        // a chunk = consecutive frames of one shape (they share a plan), at most kChunk
        int nf = 1;
        while (f0 + nf < nframes && nf < kChunk && frames[f0 + nf].width == frames[f0].width &&
               frames[f0 + nf].height == frames[f0].height && frames[f0 + nf].channels == frames[f0].channels) nf++;
        const hogb200_image* fr = &frames[f0];
        // frames go through the kernels in groups: one launch of each kernel covers `grp` frames, which evens out the
        // tail of the persistent histogram kernel and the last wave of the other two
        const int grp = std::min(batch_group(h, fr->width, fr->height, fr->channels == 3 ? 4 : fr->channels), nf);
        if ((rc = build_plan(h, h->plan_grp, fr->width, fr->height, fr->channels == 3 ? 4 : fr->channels, false, grp))) return rc;
        Plan& P = h->plan_grp;
        size_t cap = std::max<size_t>(h->cand_cap, 1 << 18);
        size_t ncand = 0;
        for (int attempt = 0; attempt < 2; attempt++) {
            if ((rc = ensure_cand(h, cap))) return rc;
            CK(cudaMemsetAsync(h->d_cand.p, 0, 16, st));
            if ((rc = queue_chunk(h, P, fr, nf, grp, nslots, slots, nullptr, &total_launches))) return rc;
            if ((rc = fetch_candidate_count(h, &ncand))) return rc;
            g_trace.print();
            if (ncand > h->cand_cap && attempt == 0) { cap = (size_t)P.nwindows * nslots * nf + 16; continue; }
            if (ncand > h->cand_cap) return fail(HOGB200_CAPACITY, "candidate list overflow");
            break;
        }
        total_cands += ncand;
        if (ncand == 0) {
            for (int i = 0; i < nf * nslots; i++) { offsets[out_idx + 1] = offsets[out_idx]; out_idx++; }
        } else if (gpu_post_min() > 0 && ncand >= gpu_post_min()) {
            if ((rc = post_process_device(h, cand_geom(h, P), h->d_cand.p, st, ncand, nf, nslots, h->group_threshold, all, offsets, &out_idx))) return rc;
        } else {
            if ((rc = fetch_candidate_list(h, ncand, c))) return rc;
            if ((rc = host_tail(h, P, c, nf, nslots, all, offsets, &out_idx))) return rc;
        }
        f0 += nf;
    }
    if (h->timing) cudaEventRecord(h->ev[5], st);
    finish_timing(h, h->plan_grp, total_launches, total_cands);
    size_t cnt = 0;
    return emit(all, found, capacity, &cnt);
}

// ---- asynchronous batches: submit queues everything and returns; collect waits for that batch only and runs its tail.
// With two batches in flight the uploads and kernels of batch k + 1 overlap the host-side tail of batch k, and the
// device tail (group.cu) runs on its own stream beside the next batch's kernels.
int hogb200_submit_batch(hogb200_t* h, const hogb200_image* frames, int nframes, int nslots, int* ticket)
{
    H_OR_FAIL;
    if (!ticket || nframes < 0 || (nframes > 0 && !frames)) return fail(HOGB200_BAD_ARG, "bad batch arguments");
    *ticket = -1;
    if (nslots < 1 || nslots > kMaxDet) return fail(HOGB200_BAD_ARG, "bad nslots");
    for (int d = 0; d < nslots; d++) if (!h->have_det[d]) return fail(HOGB200_BAD_ARG, "detector slot not set");
    int rc;
    if ((rc = check_work_params(h))) return rc;
    if ((rc = check_svm_geometry(h))) return rc;
    for (int i = 0; i < nframes; i++) if ((rc = check_image(&frames[i]))) return rc;
    if ((rc = use_device(h))) return rc;
    int slot = -1;
    for (int i = 0; i < 2 && slot < 0; i++) if (!h->jobs[i].active) slot = i;
    if (slot < 0) return fail(HOGB200_BAD_ARG, "two batches are already in flight: collect one first");
    BatchJob& J = h->jobs[slot];
    J.frames.assign(frames, frames + nframes);
    J.nslots = nslots; J.launches = 0; J.chunks.clear(); J.finished = false;
    J.hit_threshold = h->hit_threshold; J.group_threshold = h->group_threshold; J.stride_x = h->win_stride_x; J.stride_y = h->win_stride_y;
    const int kChunk = 64;
    for (int f0 = 0; f0 < nframes;) {
        int nf = 1;
        while (f0 + nf < nframes && nf < kChunk && frames[f0 + nf].width == frames[f0].width &&
               frames[f0 + nf].height == frames[f0].height && frames[f0 + nf].channels == frames[f0].channels) nf++;
        ChunkJob c; c.f0 = f0; c.nf = nf;
        J.chunks.push_back(c);
        f0 += nf;
    }
    const size_t nchunks = J.chunks.size();
    // one candidate region per chunk (a chunk that overflows its region is re-run by collect)
    const size_t cap = h->job_cap, region = align_up(16 + cap * sizeof(Candidate), 256);
    cudaStream_t st = h->stream;
    if (J.d_cand.cap < region * nchunks) { CK(cudaStreamSynchronize(st)); CK(cudaStreamSynchronize(h->post_stream)); }
    CK(J.d_cand.reserve(region * nchunks));
    CK(J.h_counts.reserve(4 * nchunks + 4));
    while (J.events.size() < nchunks) { cudaEvent_t e; CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); J.events.push_back(e); }
    int slots[kMaxDet] = {0, 1, 2, 3};
    for (size_t k = 0; k < nchunks; k++) {
        ChunkJob& c = J.chunks[k];
        const hogb200_image* fr = &J.frames[c.f0];
        const int cn = fr->channels == 3 ? 4 : fr->channels;
        const int grp = std::min(batch_group(h, fr->width, fr->height, cn), c.nf);
        if ((rc = build_plan(h, h->plan_grp, fr->width, fr->height, cn, false, grp))) return rc;
        c.cand_off = k * region; c.cap = cap; c.lv = h->plan_grp.lv; c.nwindows = h->plan_grp.nwindows; c.done = J.events[k];
        const CandTarget target{J.d_cand.p + c.cand_off, cap};
        CK(cudaMemsetAsync(target.p, 0, 16, st));
        // (only the first chunk of a batch can find the stream idle)
        if ((rc = queue_chunk(h, h->plan_grp, fr, c.nf, grp, nslots, slots, &target, &J.launches, k == 0))) return rc;
        CK(cudaMemcpyAsync(J.h_counts.p + 4 * k, target.p, 16, cudaMemcpyDeviceToHost, st));
        CK(cudaEventRecord(c.done, st));
    }
    J.active = true;
    *ticket = slot;
    return HOGB200_OK;
}

int hogb200_collect_batch(hogb200_t* h, int ticket, hogb200_rect* found, size_t capacity, size_t* offsets)
{
    H_OR_FAIL;
    if (ticket < 0 || ticket >= 2 || !h->jobs[ticket].active) return fail(HOGB200_BAD_ARG, "no batch in flight under this ticket");
    if (!offsets) return fail(HOGB200_BAD_ARG, "offsets is null");
    BatchJob& J = h->jobs[ticket];
    int rc;
    if ((rc = use_device(h))) return rc;
    if (J.finished) {                                      // a previous collect found `capacity` too small
        memcpy(offsets, J.result_off.data(), J.result_off.size() * sizeof(size_t));
        size_t cnt = 0;
        rc = emit(J.result, found, capacity, &cnt);
        if (rc != HOGB200_CAPACITY) J.active = false;
        return rc;
    }
    std::vector<hogb200_rect> all; all.reserve(1024);
    std::vector<Candidate> cands;
    offsets[0] = 0;
    size_t out_idx = 0, total_cands = 0;
    const int nslots = J.nslots;
    cudaStream_t ps = h->post_stream;
    // the tail reads the parameters the batch was submitted with
    const int g_now = h->group_threshold, sx_now = h->win_stride_x, sy_now = h->win_stride_y;
    h->group_threshold = J.group_threshold; h->win_stride_x = J.stride_x; h->win_stride_y = J.stride_y;
    auto restore = [&](int code) { h->group_threshold = g_now; h->win_stride_x = sx_now; h->win_stride_y = sy_now; J.active = false; return code; };
    for (size_t k = 0; k < J.chunks.size(); k++) {
        ChunkJob& c = J.chunks[k];
        if (cudaEventSynchronize(c.done) != cudaSuccess)
            return restore(fail(HOGB200_CUDA_ERROR, std::string("collect: ") + cudaGetErrorString(cudaGetLastError())));
        const size_t n = J.h_counts.p[4 * k];
        const unsigned char* d_cand = J.d_cand.p + c.cand_off;
        Plan P; P.lv = c.lv;                               // the level scales are all the tail needs of the plan
        if (n > c.cap) {
            // the chunk produced more candidates than its region holds: run it again through the blocking entry
            // point, which sizes its list to the worst case; later batches get regions twice the size of this demand
            {
                size_t want = h->job_cap;
                while (want < 2 * n) want *= 2;
                h->job_cap = std::min<size_t>(want, (size_t)std::max<long long>(c.nwindows, 1) * nslots * c.nf + 16);
            }
            if ((rc = ensure_cand(h, 2 * n))) return restore(rc);     // its first attempt then holds the list
            std::vector<hogb200_rect> part((size_t)std::max<long long>(c.nwindows, 1) * nslots * c.nf);
            std::vector<size_t> off((size_t)c.nf * nslots + 1);
            const double hit_now = h->hit_threshold; h->hit_threshold = J.hit_threshold;
            rc = hogb200_detect_multiscale_batch(h, &J.frames[c.f0], c.nf, nslots, part.data(), part.size(), off.data());
            h->hit_threshold = hit_now;
            if (rc) return restore(rc);
            const size_t base = all.size();
            all.insert(all.end(), part.begin(), part.begin() + off.back());
            for (size_t i = 1; i < off.size(); i++) offsets[++out_idx] = base + off[i];
            total_cands += (size_t)h->last.candidates;
            continue;
        }
        total_cands += n;
        if (n == 0) {
            for (int i = 0; i < c.nf * nslots; i++) { offsets[out_idx + 1] = offsets[out_idx]; out_idx++; }
        } else if (gpu_post_min() > 0 && n >= gpu_post_min()) {
            if ((rc = post_process_device(h, cand_geom(h, P), d_cand, ps, n, c.nf, nslots, J.group_threshold, all, offsets, &out_idx)))
                return restore(rc);
        } else {
            cudaError_t e = h->h_cand.reserve(16 + n * sizeof(Candidate));
            if (e == cudaSuccess) e = cudaMemcpyAsync(h->h_cand.p + 16, d_cand + 16, n * sizeof(Candidate), cudaMemcpyDeviceToHost, ps);
            if (e == cudaSuccess) e = cudaStreamSynchronize(ps);
            if (e != cudaSuccess) { cudaGetLastError(); return restore(fail(HOGB200_CUDA_ERROR, std::string("collect: ") + cudaGetErrorString(e))); }
            const Candidate* cp = reinterpret_cast<const Candidate*>(h->h_cand.p + 16);
            cands.assign(cp, cp + n);
            if ((rc = host_tail(h, P, cands, c.nf, nslots, all, offsets, &out_idx))) return restore(rc);
        }
    }
    h->last = hogb200_timings{};
    h->last.kernel_launches = J.launches; h->last.candidates = (long long)total_cands;
    size_t cnt = 0;
    rc = restore(emit(all, found, capacity, &cnt));
    if (rc == HOGB200_CAPACITY) {                          // keep the results: collect again with all.size() entries
        J.result.swap(all); J.result_off.assign(offsets, offsets + out_idx + 1);
        J.finished = true; J.active = true;
    }
    return rc;
}

}  // extern "C"

#include "people_detector.inc"

extern "C" {

int hogb200_default_people_detector(int win_w, int win_h, const float** coeffs, size_t* n)
{
    if (!coeffs || !n) return fail(HOGB200_BAD_ARG, "null output");
    if (win_w == 64 && win_h == 128) { *coeffs = reinterpret_cast<const float*>(kPeople64x128); *n = 3781; return HOGB200_OK; }
    if (win_w == 48 && win_h == 96) { *coeffs = reinterpret_cast<const float*>(kPeople48x96); *n = 1981; return HOGB200_OK; }
    return fail(HOGB200_BAD_ARG, "getDefaultPeopleDetector: win_size must be 64x128 or 48x96 (hog.cpp:328)");
}

int hogb200_levels(const hogb200_t* h, int width, int height, double* scales, int* widths, int* heights,
                   size_t capacity, size_t* count)
{
    H_OR_FAIL;
    std::vector<double> sc;
    enumerate_levels(width, height, h->g.win_w, h->g.win_h, h->scale0, h->nlevels, sc);
    if (count) *count = sc.size();
    if (sc.size() > capacity) return fail(HOGB200_CAPACITY, "output capacity too small");
    for (size_t i = 0; i < sc.size(); i++) {
        if (scales) scales[i] = sc[i];
        if (widths) widths[i] = cvRound(width / sc[i]);
        if (heights) heights[i] = cvRound(height / sc[i]);
    }
    return HOGB200_OK;
}

int hogb200_enumerate_levels(int width, int height, int win_w, int win_h, double scale0, int nlevels,
                             double* scales, int* widths, int* heights, size_t capacity, size_t* count)
{
    std::vector<double> sc;
    enumerate_levels(width, height, win_w, win_h, scale0, nlevels, sc);
    if (count) *count = sc.size();
    if (sc.size() > capacity) return fail(HOGB200_CAPACITY, "output capacity too small");
    for (size_t i = 0; i < sc.size(); i++) {
        if (scales) scales[i] = sc[i];
        if (widths) widths[i] = cvRound(width / sc[i]);
        if (heights) heights[i] = cvRound(height / sc[i]);
    }
    return HOGB200_OK;
}

int hogb200_group_rectangles(hogb200_rect* rects, size_t n, int group_threshold, double eps, size_t* count)
{
    if (n && !rects) return fail(HOGB200_BAD_ARG, "null rects");
    std::vector<hogb200_rect> v(rects, rects + n);
    group_rectangles(v, group_threshold, eps);
    if (!v.empty()) memcpy(rects, v.data(), v.size() * sizeof(hogb200_rect));
    if (count) *count = v.size();
    return HOGB200_OK;
}

int hogb200_batch_group(const hogb200_t* h, int width, int height, int channels)
{
    if (!h || width <= 0 || height <= 0) return 0;
    return batch_group(h, width, height, channels == 3 ? 4 : channels);
}

int hogb200_group_rectangles_device(hogb200_t* h, hogb200_rect* rects, size_t n, int group_threshold, double eps, size_t* count)
{
    H_OR_FAIL;
    if (n && !rects) return fail(HOGB200_BAD_ARG, "null rects");
    if (n > (size_t)1 << 26) return fail(HOGB200_BAD_ARG, "too many rects");
    std::vector<hogb200_rect> v(rects, rects + n);
    const int rc = group_rectangles_device(h, v, group_threshold, eps);
    if (rc) return rc;
    if (!v.empty()) memcpy(rects, v.data(), v.size() * sizeof(hogb200_rect));
    if (count) *count = v.size();
    return HOGB200_OK;
}

int hogb200_resize(hogb200_t* h, const hogb200_image* img, int dst_w, int dst_h, uint8_t* out)
{
    H_OR_FAIL;
    int rc;
    if ((rc = check_image(img))) return rc;
    if (img->channels == 3) return fail(HOGB200_BAD_ARG, "resize: 8UC1 or 8UC4 (the pyramid is built from the widened frame)");
    if (dst_w <= 0 || dst_h <= 0 || !out) return fail(HOGB200_BAD_ARG, "bad destination");
    if ((rc = use_device(h))) return rc;
    cudaStream_t st = h->stream;
    const int cn = img->channels, W = img->width, H = img->height;
    FrameRef f;
    if ((rc = stage_frame(h, img, 0, st, f))) return rc;
    const size_t row = (size_t)dst_w * cn;
    if (dst_w == W && dst_h == H) {            // resize.cpp:3745-3750: plain copy
        CK(cudaMemcpy2DAsync(out, row, f.dev, f.pitch, row, H, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        return HOGB200_OK;
    }
    LevelDesc D; memset(&D, 0, sizeof D);
    D.w = dst_w; D.h = dst_h; D.pitch = (int)align_up(row, 64);
    D.mode = is_exact_half(W, H, dst_w, dst_h) ? 2 : 1;
    std::vector<ResizeCoef> xt, yt; std::vector<ResizeItem> items;
    if (D.mode == 1) { resize_table(W, dst_w, true, xt); resize_table(H, dst_h, false, yt); }
    if (D.mode == 1) D.rs_rows = resize_plan_rows(xt.data(), dst_w, yt.data(), dst_h, W, cn);
    const int tw = resize_tile_w(cn), th = resize_tile_h(cn, D.rs_rows);
    for (int y = 0; y < dst_h; y += th) for (int x = 0; x < dst_w; x += tw) items.push_back(ResizeItem{0, (short)x, (short)y, 0});
    DevBuf<uint8_t> d_out; DevBuf<LevelDesc> d_lv; DevBuf<ResizeCoef> d_xt, d_yt; DevBuf<ResizeItem> d_it;
    cudaError_t e = d_out.reserve((size_t)D.pitch * dst_h + 64);
    D.img = d_out.p;
    if (e == cudaSuccess) e = d_lv.reserve(1);
    if (e == cudaSuccess) e = d_xt.reserve(xt.size() + 1);
    if (e == cudaSuccess) e = d_yt.reserve(yt.size() + 1);
    if (e == cudaSuccess) e = d_it.reserve(items.size() + 1);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_lv.p, &D, sizeof D, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess && !xt.empty()) e = cudaMemcpyAsync(d_xt.p, xt.data(), xt.size() * sizeof(ResizeCoef), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess && !yt.empty()) e = cudaMemcpyAsync(d_yt.p, yt.data(), yt.size() * sizeof(ResizeCoef), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_it.p, items.data(), items.size() * sizeof(ResizeItem), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) {
        FrameSet fs; memset(&fs, 0, sizeof fs);
        fs.base[0] = f.dev; fs.pitch[0] = f.pitch; fs.nframes = 1;
        launch_resize(fs, W, H, cn, d_lv.p, d_xt.p, d_yt.p, d_it.p, (int)items.size(), st);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpy2DAsync(out, row, d_out.p, D.pitch, row, dst_h, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st); else cudaStreamSynchronize(st);
    d_out.release(); d_lv.release(); d_xt.release(); d_yt.release(); d_it.release();
    if (e != cudaSuccess) { cudaGetLastError(); return fail(HOGB200_CUDA_ERROR, std::string("resize: ") + cudaGetErrorString(e)); }
    return HOGB200_OK;
}

int hogb200_block_histograms(hogb200_t* h, const hogb200_image* img, float* out, size_t capacity_floats,
                             int* blocks_x, int* blocks_y)
{
    H_OR_FAIL;
    std::vector<Candidate> c;
    int rc = run_single(h, h->plan1, img, true, 0, nullptr, 0.f, nullptr, c);
    if (rc) return rc;
    const LevelHost& L = h->plan1.lv[0];
    if (blocks_x) *blocks_x = L.nbx;
    if (blocks_y) *blocks_y = L.nby;
    const size_t n = (size_t)L.nbx * L.nby * kHist;
    if (n > capacity_floats) return fail(HOGB200_CAPACITY, "output capacity too small");
    if (n) {
        if (!out) return fail(HOGB200_BAD_ARG, "null output");
        CK(cudaMemcpyAsync(out, h->d_hist.p, n * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
    }
    return HOGB200_OK;
}

}  // extern "C"

namespace {

// One level of the frame, device resident: the frame itself when the size matches (hog.cpp:3522-3525), else
// cv::resize INTER_LINEAR semantics into h->d_roi_level (the mandated pyramid resize; the CPU class uses
// INTER_LINEAR_EXACT here).
int roi_level_image(hogb200* h, const FrameRef& f, int W, int H, int cn, int dw, int dh, FrameRef& out)
{
    if (dw == W && dh == H) { out = f; return HOGB200_OK; }
    cudaStream_t st = h->stream;
    LevelDesc D; memset(&D, 0, sizeof D);
    D.w = dw; D.h = dh; D.pitch = (int)align_up((size_t)dw * cn, 64);
    D.mode = is_exact_half(W, H, dw, dh) ? 2 : 1;
    std::vector<ResizeCoef> xt, yt; std::vector<ResizeItem> items;
    if (D.mode == 1) { resize_table(W, dw, true, xt); resize_table(H, dh, false, yt); }
    if (D.mode == 1) D.rs_rows = resize_plan_rows(xt.data(), dw, yt.data(), dh, W, cn);
    const int tw = resize_tile_w(cn), th = resize_tile_h(cn, D.rs_rows);
    for (int y = 0; y < dh; y += th) for (int x = 0; x < dw; x += tw) items.push_back(ResizeItem{0, (short)x, (short)y, 0});
    CK(cudaStreamSynchronize(st));          // the tables below are freed at the end of this call: nothing may still read older ones
    CK(h->d_roi_level.reserve((size_t)D.pitch * dh + 64));
    D.img = h->d_roi_level.p;
    DevBuf<LevelDesc> d_lv; DevBuf<ResizeCoef> d_xt, d_yt; DevBuf<ResizeItem> d_it;
    cudaError_t e = d_lv.reserve(1);
    if (e == cudaSuccess) e = d_xt.reserve(xt.size() + 1);
    if (e == cudaSuccess) e = d_yt.reserve(yt.size() + 1);
    if (e == cudaSuccess) e = d_it.reserve(items.size() + 1);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_lv.p, &D, sizeof D, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess && !xt.empty()) e = cudaMemcpyAsync(d_xt.p, xt.data(), xt.size() * sizeof(ResizeCoef), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess && !yt.empty()) e = cudaMemcpyAsync(d_yt.p, yt.data(), yt.size() * sizeof(ResizeCoef), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_it.p, items.data(), items.size() * sizeof(ResizeItem), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) {
        FrameSet fs; memset(&fs, 0, sizeof fs);
        fs.base[0] = f.dev; fs.pitch[0] = f.pitch; fs.nframes = 1;
        launch_resize(fs, W, H, cn, d_lv.p, d_xt.p, d_yt.p, d_it.p, (int)items.size(), st);
        e = cudaGetLastError();
    }
    cudaError_t e2 = cudaStreamSynchronize(st);
    d_lv.release(); d_xt.release(); d_yt.release(); d_it.release();
    if (e == cudaSuccess) e = e2;
    if (e != cudaSuccess) { cudaGetLastError(); return fail(HOGB200_CUDA_ERROR, std::string("roi level resize: ") + cudaGetErrorString(e)); }
    out.dev = h->d_roi_level.p; out.pitch = D.pitch;
    return HOGB200_OK;
}

// HOGDescriptor::detectROI (objdetect/src/hog.cpp:3549-3660) on a device-resident level image of lw x lh pixels.
// The CPU class scores windows on the image PADDED by (pad_w, pad_h) with BORDER_REFLECT_101 gradients; here the level is
// materialised with a frame of pad + 8 pixels (one block more than the padding, so that every gradient the padded
// grid needs has both neighbours inside the materialised image: its values equal the reference's xmap/ymap look-ups),
// histograms and dense stride-8 scores are computed by the usual two kernels, and the requested locations are read
// out of the score plane.  conf gets one entry per location (-10.0 = outside the padded image, hog.cpp:3596-3602).
int roi_score_level(hogb200* h, int slot, const FrameRef& lv, int lw, int lh, int cn, const hogb200_point* loc, size_t n,
                    int pad_w, int pad_h, std::vector<double>& conf, std::vector<float>& plane)
{
    const int cs = kCell;                               // cacheStride = gcd(winStride = cellSize, blockStride) = 8
    pad_w = (int)align_up((size_t)std::max(pad_w, 0), cs); pad_h = (int)align_up((size_t)std::max(pad_h, 0), cs);
    conf.assign(n, -10.0);
    const int win_w = h->g.win_w, win_h = h->g.win_h;
    bool any = false;
    for (size_t i = 0; i < n; i++) {
        const hogb200_point p = loc[i];
        if (p.x < -pad_w || p.x > lw + pad_w - win_w || p.y < -pad_h || p.y > lh + pad_h - win_h) continue;
        if (p.x % cs != 0 || p.y % cs != 0)
            return fail(HOGB200_BAD_ARG, "detectROI: locations must be multiples of the cache stride 8 (HOGCache::getBlock asserts, hog.cpp:918)");
        any = true;
    }
    if (!any) return HOGB200_OK;
    const int px = pad_w + kCell, py = pad_h + kCell;
    const int mw = lw + 2 * px, mh = lh + 2 * py;
    const size_t mpitch = align_up((size_t)mw * cn, 64);
    int rc;
    CK(cudaStreamSynchronize(h->stream));
    CK(h->d_roi_pad.reserve(mpitch * mh + 64));
    launch_pad_reflect101(lv.dev, (size_t)lv.pitch, lw, lh, cn, h->d_roi_pad.p, mpitch, px, py, h->stream);
    CK(cudaGetLastError());
    hogb200_image im; im.data = h->d_roi_pad.p; im.pitch = mpitch; im.width = mw; im.height = mh; im.channels = cn; im.on_device = 1;
    // dense scores of the materialised level at window stride 8 x 8, whatever the handle's stride is
    const int sx = h->win_stride_x, sy = h->win_stride_y;
    h->win_stride_x = h->win_stride_y = kCell;
    std::vector<Candidate> c;
    rc = build_plan(h, h->plan_roi, mw, mh, cn, true);
    size_t np = 0;
    if (!rc) {
        np = (size_t)h->plan_roi.lv[0].nwx * h->plan_roi.lv[0].nwy;
        cudaError_t e = cudaStreamSynchronize(h->stream);
        if (e == cudaSuccess) e = h->d_scores.reserve(np + 16);
        if (e != cudaSuccess) rc = fail(HOGB200_CUDA_ERROR, cudaGetErrorString(e));
    }
    if (!rc && np) rc = run_single(h, h->plan_roi, &im, true, 1, &slot, INFINITY, h->d_scores.p, c);
    h->win_stride_x = sx; h->win_stride_y = sy;
    if (rc) return rc;
    if (!np) return HOGB200_OK;
    plane.resize(np);
    CK(cudaMemcpyAsync(plane.data(), h->d_scores.p, np * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    const int nwx = h->plan_roi.lv[0].nwx;
    for (size_t i = 0; i < n; i++) {
        const hogb200_point p = loc[i];
        if (p.x < -pad_w || p.x > lw + pad_w - win_w || p.y < -pad_h || p.y > lh + pad_h - win_h) continue;
        conf[i] = (double)plane[(size_t)((p.y + py) / kCell) * nwx + (p.x + px) / kCell];
    }
    return HOGB200_OK;
}

int roi_prepare(hogb200* h, const hogb200_image* img, FrameRef& f, int* cn)
{
    int rc;
    if ((rc = check_image(img))) return rc;
    if ((rc = check_svm_geometry(h))) return rc;
    if ((rc = use_device(h))) return rc;
    if ((rc = stage_frame(h, img, 0, h->stream, f))) return rc;
    *cn = img->channels == 3 ? 4 : img->channels;
    return HOGB200_OK;
}

}  // namespace

extern "C" {

int hogb200_detect_roi(hogb200_t* h, const hogb200_image* img, const hogb200_point* locations, size_t n, int pad_w, int pad_h,
                       hogb200_point* found, double* confidences, size_t capacity, size_t* count)
{
    H_OR_FAIL;
    if (count) *count = 0;
    if (n && !locations) return fail(HOGB200_BAD_ARG, "null locations");
    if (!h->have_det[0] || n == 0) { int rc = check_image(img); return rc; }          // hog.cpp:3558-3559
    FrameRef f; int cn = 0, rc;
    if ((rc = roi_prepare(h, img, f, &cn))) return rc;
    std::vector<double> conf; std::vector<float> plane;
    if ((rc = roi_score_level(h, 0, f, img->width, img->height, cn, locations, n, pad_w, pad_h, conf, plane))) return rc;
    if (confidences) memcpy(confidences, conf.data(), n * sizeof(double));
    std::vector<hogb200_point> hits;
    for (size_t i = 0; i < n; i++) if (conf[i] != -10.0 && conf[i] >= h->hit_threshold) hits.push_back(locations[i]);
    return emit(hits, found, capacity, count);
}

int hogb200_detect_multiscale_roi(hogb200_t* h, const hogb200_image* img, hogb200_roi* rois, size_t nrois, int group_threshold,
                                  hogb200_rect* found, size_t capacity, size_t* count)
{
    H_OR_FAIL;
    if (count) *count = 0;
    if (nrois && !rois) return fail(HOGB200_BAD_ARG, "null rois");
    for (size_t i = 0; i < nrois; i++)
        if (!(rois[i].scale > 0) || !std::isfinite(rois[i].scale) || (rois[i].nlocations && !rois[i].locations))
            return fail(HOGB200_BAD_ARG, "bad roi (scale must be positive, locations non-null)");
    if (!h->have_det[0] || nrois == 0) return check_image(img);
    FrameRef f; int cn = 0, rc;
    if ((rc = roi_prepare(h, img, f, &cn))) return rc;
    const int W = img->width, H = img->height;
    std::vector<hogb200_rect> all; std::vector<double> conf; std::vector<float> plane;
    for (size_t i = 0; i < nrois; i++) {
        hogb200_roi& R = rois[i];
        if (R.nlocations == 0) continue;                                         // detectROI returns at once (hog.cpp:3558)
        const double scale = R.scale;
        const int lw = cvRound(W / scale), lh = cvRound(H / scale);             // HOGConfInvoker, hog.cpp:3518-3520
        if (lw <= 0 || lh <= 0) return fail(HOGB200_BAD_ARG, "roi scale leaves an empty image");
        FrameRef lv;
        if ((rc = roi_level_image(h, f, W, H, cn, lw, lh, lv))) return rc;
        if ((rc = roi_score_level(h, 0, lv, lw, lh, cn, R.locations, R.nlocations, 8, 8, conf, plane))) return rc;   // padding 8 x 8, hog.cpp:3672
        if (R.confidences) memcpy(R.confidences, conf.data(), R.nlocations * sizeof(double));
        const int sw = cvRound(h->g.win_w * scale), sh = cvRound(h->g.win_h * scale);
        for (size_t j = 0; j < R.nlocations; j++)
            if (conf[j] != -10.0 && conf[j] >= h->hit_threshold)
                all.push_back(hogb200_rect{cvRound(R.locations[j].x * scale), cvRound(R.locations[j].y * scale), sw, sh});
    }
    if (group_threshold > 0 && (rc = group_rectangles_auto(h, all, group_threshold, 0.2))) return rc;
    return emit(all, found, capacity, count);
}

size_t hogb200_noalloc_scratch_bytes(void) { return hist_noalloc_scratch_bytes(); }

int hogb200_block_histograms_noalloc(hogb200_t* h, const hogb200_image* img, float* d_out, size_t capacity_floats,
                                     int* blocks_x, int* blocks_y)
{
    H_OR_FAIL;
    int rc;
    if ((rc = check_image(img))) return rc;
    if (!img->on_device) return fail(HOGB200_BAD_ARG, "block_histograms_noalloc: the image must be in device memory (no staging buffer is allocated)");
    const int cn = img->channels;
    if (cn != 1 && cn != 4) return fail(HOGB200_BAD_ARG, "block_histograms_noalloc: CV_8UC1 or CV_8UC4 (widening 8UC3 needs a buffer)");
    // the kernel reads rows in 4-pixel vectors straight from the caller's buffer: nothing can be staged here
    const size_t al = cn == 4 ? 16 : 4, row = (size_t)img->width * cn;
    if ((uintptr_t)img->data % al != 0 || img->pitch % al != 0 || img->pitch >= (1u << 30) ||
        !(img->width % 4 == 0 || img->pitch >= align_up(row, al)))
        return fail(HOGB200_BAD_ARG, "block_histograms_noalloc: rows must be 16-byte (8UC4) / 4-byte (8UC1) aligned, and the pitch must "
                                     "cover the width rounded up to 4 pixels");
    const int nbx = img->width >= kBlock ? (img->width - kBlock) / kCell + 1 : 0;
    const int nby = img->height >= kBlock ? (img->height - kBlock) / kCell + 1 : 0;
    if (blocks_x) *blocks_x = nbx;
    if (blocks_y) *blocks_y = nby;
    const size_t n = (size_t)nbx * nby * kHist;
    if (n > capacity_floats) return fail(HOGB200_CAPACITY, "output capacity too small");
    if (!n) return HOGB200_OK;
    if (!d_out || (uintptr_t)d_out % 16 != 0) return fail(HOGB200_BAD_ARG, "output must be a 16-byte aligned device pointer");
    if ((rc = use_device(h))) return rc;
    long long need = 0;
    const int st = launch_hist_noalloc((const uint8_t*)img->data, (int)img->pitch, img->width, img->height, cn, h->gamma, h->weights,
                                       (float)h->l2hys, d_out, h->num_sms, h->stream, &need);
    if (st == 1)
        return fail(HOGB200_CAPACITY, "block_histograms_noalloc: image needs " + std::to_string(need) + " work items, the static scratch (" +
                                      std::to_string(hist_noalloc_scratch_bytes()) + " bytes) holds 32704: use hogb200_block_histograms");
    if (st) return fail(HOGB200_CUDA_ERROR, std::string("block_histograms_noalloc: ") + cudaGetErrorString(cudaGetLastError()));
    return HOGB200_OK;
}

int hogb200_window_scores(hogb200_t* h, int slot, const hogb200_image* img, float* out, size_t capacity_floats,
                          int* wins_x, int* wins_y)
{
    H_OR_FAIL;
    if (slot < 0 || slot >= kMaxDet || !h->have_det[slot]) return fail(HOGB200_BAD_ARG, "detector slot not set");
    int rc;
    if ((rc = check_image(img))) return rc;
    if ((rc = check_work_params(h))) return rc;
    if ((rc = use_device(h))) return rc;
    if ((rc = build_plan(h, h->plan1, img->width, img->height, img->channels == 3 ? 4 : img->channels, true))) return rc;
    const LevelHost& L = h->plan1.lv[0];
    if (wins_x) *wins_x = L.nwx;
    if (wins_y) *wins_y = L.nwy;
    const size_t n = (size_t)L.nwx * L.nwy;
    if (n > capacity_floats) return fail(HOGB200_CAPACITY, "output capacity too small");
    if (!n) return HOGB200_OK;
    if (!out) return fail(HOGB200_BAD_ARG, "null output");
    CK(cudaStreamSynchronize(h->stream));
    CK(h->d_scores.reserve(n + 16));
    std::vector<Candidate> c;
    rc = run_single(h, h->plan1, img, true, 1, &slot, INFINITY, h->d_scores.p, c);
    if (rc) return rc;
    CK(cudaMemcpyAsync(out, h->d_scores.p, n * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return HOGB200_OK;
}

int hogb200_compute(hogb200_t* h, const hogb200_image* img, float* out, size_t capacity_floats, int out_on_device,
                    size_t* nwindows)
{
    H_OR_FAIL;
    if (nwindows) *nwindows = 0;
    std::vector<Candidate> c;
    int rc = run_single(h, h->plan1, img, true, 0, nullptr, 0.f, nullptr, c);
    if (rc) return rc;
    const LevelHost& L = h->plan1.lv[0];
    const size_t nw = (size_t)L.nwx * L.nwy, ds = hogb200_get_descriptor_size(h);
    if (nwindows) *nwindows = nw;
    if (nw * ds > capacity_floats) return fail(HOGB200_CAPACITY, "output capacity too small");
    if (!nw) return HOGB200_OK;
    if (!out) return fail(HOGB200_BAD_ARG, "null output");
    float* d_out = out;
    if (!out_on_device) { CK(h->d_desc.reserve(nw * ds + 16)); d_out = h->d_desc.p; }
    launch_extract(h->plan1.d_levels.p, L.nwx, L.nwy, h->d_hist.p, win_bx(h), win_by(h), h->win_stride_x / kCell,
                   h->win_stride_y / kCell, h->descr_format == HOGB200_DESCR_FORMAT_COL_BY_COL, d_out, h->stream);
    CK(cudaGetLastError());
    if (!out_on_device) CK(cudaMemcpyAsync(out, d_out, nw * ds * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return HOGB200_OK;
}

}  // extern "C"
