// tbd_host.cpp -- tracking-by-detection step behind include/tbdb200.h.
//
// Host C++ (a frame carries a few dozen boxes; nothing here is worth a kernel).  Restates cv::tbd::Tracker
// (modules/trackingbydetection/src/tbd.cpp:177-1106) so that tracks come out IDENTICAL to the reference's:
//   * prediction: constant velocity between the last two boxes, scaled by the frame gap (tbd.cpp:1046-1081), the
//     predicted corner rounded like cv::Point2d -> cv::Point (cvRound, ties to even)            (tbd.cpp:300-317)
//   * cost: 1 - intersection/union in double precision                                          (tbd.cpp:346-366, 1083-1106)
//   * assignment: the reference's zero-cover procedure -- NOT a generic optimal solver: its greedy zero picking can
//     stop with rows unassigned, and which of several equal-cost matchings comes out depends on its scan order, so
//     the same scans are performed here (on a flat matrix with byte masks)                      (tbd.cpp:402-880)
//   * update: box size stabilised over the last <= 4 boxes in unsigned integer arithmetic, confidence window,
//     deletion rule, creation from unassigned detections                                        (tbd.cpp:897-1044)
//   * trajectories: per ground-truth id and frame, whether the detection was tracked, by which track, the track's
//     newest and predicted boxes and the IoU with the registered ground-truth box                (tbd.cpp:144-175, 236-272)
// What is not carried over: the reference's progress prints to stdout, the random display colour of a track, and
// world coordinates (never read by the tracker).
#include "../../include/tbdb200.h"

#include <algorithm>
#include <cmath>
#include <map>
#include <string>
#include <utility>
#include <vector>

namespace {

thread_local std::string g_err;
int fail(const char* msg) { g_err = msg; return 1; }

struct Box { int x, y, w, h; };

// cv::saturate_cast<int>(double) = cvRound: round to nearest even under the default rounding mode (core/fast_math.hpp:200-214)
inline int round_half_even(double v) { return (int)lrint(v); }

// computeBoundingBoxOverlap (tbd.cpp:1083-1106)
double iou(const Box& a, const Box& b)
{
    const double left = std::max(a.x, b.x), right = std::min(a.x + a.w, b.x + b.w);
    const double top = std::max(a.y, b.y), bottom = std::min(a.y + a.h, b.y + b.h);
    if (right < left || bottom < top) return 0.0;
    const double inter = (right - left) * (bottom - top);
    const double uni = a.w * a.h + b.w * b.h - inter;
    return inter / uni;
}

struct TrackState {
    unsigned id = 0, age = 0, visible = 0;
    double max_conf = 0, avg_conf = 0, overlap = 0;
    Box pred = {0, 0, 0, 0};
    std::vector<Box> boxes;
    std::vector<double> scores;
    std::vector<int> frames;
};

// constantVelocityMotionModel (tbd.cpp:1046-1081): the predicted CENTRE of the track at `frame`
void predict_centre(const TrackState& t, int frame, double& cx, double& cy)
{
    const Box& b1 = t.boxes.back();
    if (t.age == 1) {
        cx = b1.x + b1.w / 2;           // integer halves, as in the reference
        cy = b1.y + b1.h / 2;
        return;
    }
    const Box& b2 = t.boxes[t.boxes.size() - 2];
    const int f1 = t.frames[t.frames.size() - 1], f2 = t.frames[t.frames.size() - 2];
    const double ratio = ((double)(frame - f1)) / (f1 - f2);
    const double dx = ratio * (b1.x - b2.x), dy = ratio * (b1.y - b2.y);
    const double w = (b1.w + b2.w) / 2.0, h = (b1.h + b2.h) / 2.0;
    cx = b1.x + w / 2 + dx;
    cy = b1.y + h / 2 + dy;
}

// The reference's assignment procedure on an n x n matrix (rows = tracks, columns = detections, both padded).
// Returns, per row, the assigned column or n.
class ZeroCoverAssigner {
public:
    ZeroCoverAssigner(std::vector<double>& m, unsigned n) : a_(m), n_(n) {}

    std::vector<unsigned> solve()
    {
        const unsigned n = n_;
        std::vector<unsigned> col_of_row(n, n);
        if (n == 0) return col_of_row;
        reduce_rows();
        reduce_cols();
        std::vector<unsigned char> row_used(n), col_used(n), row_mark(n), col_mark(n);
        // Every pass either completes the assignment, proves (by the reference's own test) that it cannot be
        // extended, or creates at least one new zero.  The cap only guards against a cycle the reference would
        // spin in forever.
        const unsigned long long max_passes = 16ull * n * n + 64;
        for (unsigned long long pass = 0; pass < max_passes; pass++) {
            const unsigned assigned = pick_zeros(col_of_row, row_used, col_used);
            if (assigned == n) break;
            // minimum line cover of the zeros: marked columns + unmarked rows (tbd.cpp:744-812)
            for (unsigned r = 0; r < n; r++) row_mark[r] = !row_used[r];
            std::fill(col_mark.begin(), col_mark.end(), 0);
            for (;;) {
                unsigned fresh = 0;
                for (unsigned r = 0; r < n; r++) {
                    if (!row_mark[r]) continue;
                    for (unsigned c = 0; c < n; c++)
                        if (!col_mark[c] && is_zero(at(r, c))) { col_mark[c] = 1; fresh++; }
                }
                for (unsigned c = 0; c < n; c++) {
                    if (!col_mark[c]) continue;
                    for (unsigned r = 0; r < n; r++)
                        if (col_of_row[r] == c) row_mark[r] = 1;
                }
                if (!fresh) break;
            }
            if (std::find(row_mark.begin(), row_mark.end(), 0) == row_mark.end()) break;   // "all marked": give up
            // smallest element no line covers (marked row, unmarked column), then shift (tbd.cpp:842-873)
            double delta = kHuge;
            for (unsigned r = 0; r < n; r++) {
                if (!row_mark[r]) continue;
                for (unsigned c = 0; c < n; c++)
                    if (!col_mark[c] && at(r, c) < delta) delta = at(r, c);
            }
            for (unsigned r = 0; r < n; r++)
                for (unsigned c = 0; c < n; c++) {
                    if (row_mark[r] && !col_mark[c]) at(r, c) -= delta;
                    else if (!row_mark[r] && col_mark[c]) at(r, c) += delta;
                }
        }
        return col_of_row;
    }

private:
    static constexpr double kHuge = 10000000.0;
    std::vector<double>& a_;
    unsigned n_;

    double& at(unsigned r, unsigned c) { return a_[(size_t)r * n_ + c]; }
    // Tracker::equalsZero (tbd.hpp:179-182)
    static bool is_zero(double v) { return v < 0.0 ? v > -0.00000001 : v < 0.00000001; }

    void reduce_rows()
    {
        for (unsigned r = 0; r < n_; r++) {
            double m = kHuge;
            for (unsigned c = 0; c < n_; c++) m = at(r, c) < m ? at(r, c) : m;
            for (unsigned c = 0; c < n_; c++) at(r, c) -= m;
        }
    }
    void reduce_cols()
    {
        for (unsigned c = 0; c < n_; c++) {
            double m = kHuge;
            for (unsigned r = 0; r < n_; r++) m = at(r, c) < m ? at(r, c) : m;
            for (unsigned r = 0; r < n_; r++) at(r, c) -= m;
        }
    }

    // Greedy zero picking from scratch (tbd.cpp:586-722): rows, then columns, that hold exactly one zero (counted
    // over ALL columns / rows, assigned or not) take it if its partner is free; a sweep that assigns nothing is
    // followed by "every free row takes its first zero in a free column"; repeat until a sweep assigns nothing.
    unsigned pick_zeros(std::vector<unsigned>& col_of_row, std::vector<unsigned char>& row_used,
                        std::vector<unsigned char>& col_used)
    {
        const unsigned n = n_;
        std::fill(row_used.begin(), row_used.end(), 0);
        std::fill(col_used.begin(), col_used.end(), 0);
        std::fill(col_of_row.begin(), col_of_row.end(), n);
        unsigned assigned = 0;
        auto take = [&](unsigned r, unsigned c) { row_used[r] = 1; col_used[c] = 1; col_of_row[r] = c; assigned++; };
        for (bool progress = true; progress;) {
            progress = false;
            for (unsigned r = 0; r < n; r++) {
                if (row_used[r]) continue;
                unsigned zeros = 0, where = 0;
                for (unsigned c = 0; c < n; c++)
                    if (is_zero(at(r, c))) { if (!zeros) where = c; zeros++; }
                if (zeros == 1 && !col_used[where]) { take(r, where); progress = true; }
            }
            for (unsigned c = 0; c < n; c++) {
                if (col_used[c]) continue;
                unsigned zeros = 0, where = 0;
                for (unsigned r = 0; r < n; r++)
                    if (is_zero(at(r, c))) { if (!zeros) where = r; zeros++; }
                if (zeros == 1 && !row_used[where]) { take(where, c); progress = true; }
            }
            if (!progress) {
                for (unsigned r = 0; r < n; r++) {
                    if (row_used[r]) continue;
                    for (unsigned c = 0; c < n; c++)
                        if (!col_used[c] && is_zero(at(r, c))) { take(r, c); progress = true; break; }
                }
            }
        }
        return assigned;
    }
};

}  // namespace

struct tbdb200_trackset { std::vector<TrackState> tracks; };

struct TrajectoryInfo { bool tracked; unsigned track_id; Box track_pos, pred_pos; double overlap; };

struct tbdb200 {
    std::map<std::pair<int, int>, Box> gt_position;            // (id, frame) -> Trajectory::positionPerFrame
    std::map<std::pair<int, int>, TrajectoryInfo> trajectory;  // (id, frame) -> Trajectory::*PerFrame
    tbdb200_args args;
    int xmin = 0, xmax = 1280, ymin = 0, ymax = 720;      // tbd.cpp:222
    unsigned next_id = 0;
    std::vector<TrackState> tracks;
    std::vector<tbdb200_frame_metrics> metrics;
};

namespace {

// Tracker::updateTrackConfidence (tbd.cpp:882-895)
void update_confidence(const tbdb200* t, TrackState& tr)
{
    const size_t use = std::min<size_t>(tr.scores.size(), t->args.time_window_size);
    double best = 0.0, sum = 0.0;
    for (size_t i = tr.scores.size() - use; i < tr.scores.size(); i++) {
        sum += tr.scores[i];
        if (tr.scores[i] > best) best = tr.scores[i];
    }
    tr.max_conf = best;
    tr.avg_conf = sum / use;
}

}  // namespace

extern "C" {

const char* tbdb200_last_error(void) { return g_err.c_str(); }

int tbdb200_create(const tbdb200_args* args, tbdb200_t** out)
{
    if (!args || !out) return fail("tbdb200_create: null argument");
    tbdb200* t = new tbdb200();
    t->args = *args;
    *out = t;
    return 0;
}

void tbdb200_destroy(tbdb200_t* t) { delete t; }

int tbdb200_reset(tbdb200_t* t)
{
    if (!t) return fail("tbdb200_reset: null handle");
    t->next_id = 0;
    t->tracks.clear();
    t->metrics.clear();
    return 0;
}

int tbdb200_set_bounds(tbdb200_t* t, int xmin, int xmax, int ymin, int ymax)
{
    if (!t) return fail("tbdb200_set_bounds: null handle");
    t->xmin = xmin; t->xmax = xmax; t->ymin = ymin; t->ymax = ymax;
    return 0;
}

int tbdb200_step(tbdb200_t* t, const tbdb200_detection* det, int n, int frame_id, int* track_of_detection)
{
    if (!t || n < 0 || (n > 0 && !det)) return fail("tbdb200_step: bad argument");
    std::vector<TrackState>& tracks = t->tracks;

    // predict (tbd.cpp:300-317) and drop tracks predicted outside the window (tbd.cpp:319-344)
    for (TrackState& tr : tracks) {
        const Box& b = tr.boxes.back();
        double cx, cy;
        predict_centre(tr, frame_id, cx, cy);
        tr.pred = Box{round_half_even(cx - b.w / 2), round_half_even(cy - b.h / 2), b.w, b.h};
    }
    tracks.erase(std::remove_if(tracks.begin(), tracks.end(), [&](const TrackState& tr) {
        const Box& p = tr.pred;
        return p.x + p.w < t->xmin || p.x >= t->xmax || p.y + p.h < t->ymin || p.y >= t->ymax;
    }), tracks.end());

    // cost matrix, padded to square with 2 x costOfNonAssignment (tbd.cpp:346-366, 412-452)
    const unsigned nt = (unsigned)tracks.size(), nd = (unsigned)n, dim = std::max(nt, nd);
    std::vector<double> cost((size_t)dim * dim, t->args.cost_of_non_assignment * 2);
    for (unsigned i = 0; i < nt; i++)
        for (unsigned j = 0; j < nd; j++)
            cost[(size_t)i * dim + j] = 1.0 - iou(tracks[i].pred, Box{det[j].x, det[j].y, det[j].width, det[j].height});
    const std::vector<unsigned> col_of_row = ZeroCoverAssigner(cost, dim).solve();

    // classifyAssignments (tbd.cpp:368-400): only real track rows count; a detection taken by a padding row is free
    std::vector<int> det_of_track(nt, -1);
    std::vector<unsigned char> det_taken(nd, 0);
    std::vector<unsigned> free_tracks, free_dets;
    for (unsigned i = 0; i < nt; i++) {
        if (col_of_row[i] < nd) { det_of_track[i] = (int)col_of_row[i]; det_taken[col_of_row[i]] = 1; }
        else free_tracks.push_back(i);
    }
    for (unsigned j = 0; j < nd; j++)
        if (!det_taken[j]) free_dets.push_back(j);
    if (track_of_detection) std::fill(track_of_detection, track_of_detection + n, -1);

    // updateAssignedTracks (tbd.cpp:897-951)
    unsigned num_assigned = 0;
    for (unsigned i = 0; i < nt; i++) {
        if (det_of_track[i] < 0) continue;
        num_assigned++;
        TrackState& tr = tracks[i];
        const tbdb200_detection& d = det[det_of_track[i]];
        const unsigned prior = (unsigned)std::min<size_t>(tr.boxes.size(), 4);
        unsigned wsum = 0, hsum = 0;
        for (size_t k = tr.boxes.size() - prior; k < tr.boxes.size(); k++) { wsum += tr.boxes[k].w; hsum += tr.boxes[k].h; }
        const int w = (wsum + d.width) / (prior + 1), h = (hsum + d.height) / (prior + 1);
        double px = d.x, py = d.y;
        px += (d.width / 2) - (w / 2);
        py += (d.height / 2) - (h / 2);
        tr.boxes.push_back(Box{round_half_even(px), round_half_even(py), w, h});
        tr.frames.push_back(d.frame_id);
        tr.overlap = iou(Box{d.x, d.y, d.width, d.height}, tr.pred);
        tr.age++;
        tr.visible++;
        tr.scores.push_back(d.confidence);
        update_confidence(t, tr);
        if (track_of_detection) track_of_detection[det_of_track[i]] = (int)tr.id;
    }
    // updateUnassignedTracks (tbd.cpp:956-982): coast on the prediction with score 0
    for (unsigned i : free_tracks) {
        TrackState& tr = tracks[i];
        tr.age++;
        tr.boxes.push_back(tr.pred);
        tr.scores.push_back(0.0);
        tr.frames.push_back(frame_id);
        tr.overlap = 0.0;
        update_confidence(t, tr);
    }
    // trajectories of detections that carry a ground-truth id (tbd.cpp:236-272; Trajectory::addTrackingInfo :151-175):
    // assigned ones in track order, then the unassigned ones; before any track is deleted or created
    for (unsigned i = 0; i < nt; i++) {
        if (det_of_track[i] < 0 || det[det_of_track[i]].id < 0) continue;
        const TrackState& tr = tracks[i];
        const std::pair<int, int> key(det[det_of_track[i]].id, frame_id);
        const Box gt = t->gt_position[key];            // a missing position reads as Rect(), as map::operator[] gives it
        t->trajectory[key] = TrajectoryInfo{true, tr.id, tr.boxes.back(), tr.pred, iou(tr.boxes.back(), gt)};
    }
    for (unsigned j : free_dets) {
        if (det[j].id < 0) continue;
        TrajectoryInfo& ti = t->trajectory[std::make_pair(det[j].id, frame_id)];
        ti.tracked = false;                            // the other fields keep what an earlier report left (reference: untouched maps)
    }
    // deleteLostTracks (tbd.cpp:984-1012)
    tracks.erase(std::remove_if(tracks.begin(), tracks.end(), [&](const TrackState& tr) {
        const double visibility = ((double)tr.visible) / tr.age;
        return (tr.age <= t->args.track_age_threshold && visibility <= t->args.track_visibility_threshold) ||
               (tr.max_conf >= 0.0 && tr.max_conf <= t->args.track_confidence_threshold);
    }), tracks.end());
    // createNewTracks (tbd.cpp:1018-1029, Track::Track tbd.cpp:68-93)
    for (unsigned j : free_dets) {
        TrackState tr;
        tr.id = t->next_id++;
        tr.boxes.push_back(Box{det[j].x, det[j].y, det[j].width, det[j].height});
        tr.scores.push_back(det[j].confidence);
        tr.frames.push_back(det[j].frame_id);
        tr.age = 1; tr.visible = 1;
        tr.max_conf = tr.avg_conf = det[j].confidence;
        tr.pred = tr.boxes.back();
        tr.overlap = 1.0;
        tracks.push_back(std::move(tr));
    }
    if (t->args.should_store_metrics) {       // tbd.cpp:283-297
        tbdb200_frame_metrics m;
        m.true_positives = (int)num_assigned; m.false_negatives = (int)free_dets.size();
        m.false_positives = (int)free_tracks.size(); m.ground_truths = n; m.num_matches = (int)num_assigned;
        m.bbox_overlap = 0.0;
        for (const TrackState& tr : tracks) m.bbox_overlap += tr.overlap;
        t->metrics.push_back(m);
    }
    return 0;
}

int tbdb200_num_tracks(const tbdb200_t* t) { return t ? (int)t->tracks.size() : 0; }

int tbdb200_get_tracks(const tbdb200_t* t, tbdb200_track* out, int capacity, int* count)
{
    if (!t || !count || (capacity > 0 && !out)) return fail("tbdb200_get_tracks: bad argument");
    *count = (int)t->tracks.size();
    for (int i = 0; i < std::min(capacity, *count); i++) {
        const TrackState& tr = t->tracks[i];
        tbdb200_track& o = out[i];
        o.id = tr.id; o.age = tr.age; o.total_visible_count = tr.visible; o.history_len = (int)tr.boxes.size();
        o.max_confidence = tr.max_conf; o.avg_confidence = tr.avg_conf; o.bbox_overlap = tr.overlap;
        o.pred_x = tr.pred.x; o.pred_y = tr.pred.y; o.pred_width = tr.pred.w; o.pred_height = tr.pred.h;
        const Box& b = tr.boxes.back();
        o.x = b.x; o.y = b.y; o.width = b.w; o.height = b.h;
        o.last_frame = tr.frames.back();
    }
    return 0;
}

int tbdb200_track_history(const tbdb200_t* t, int index, int* frames, int* boxes, double* scores, int capacity, int* count)
{
    if (!t || !count || index < 0 || index >= (int)t->tracks.size()) return fail("tbdb200_track_history: bad argument");
    const TrackState& tr = t->tracks[index];
    *count = (int)tr.boxes.size();
    for (int i = 0; i < std::min(capacity, *count); i++) {
        if (frames) frames[i] = tr.frames[i];
        if (scores) scores[i] = tr.scores[i];
        if (boxes) { boxes[4 * i] = tr.boxes[i].x; boxes[4 * i + 1] = tr.boxes[i].y; boxes[4 * i + 2] = tr.boxes[i].w; boxes[4 * i + 3] = tr.boxes[i].h; }
    }
    return 0;
}

int tbdb200_copy_tracks(const tbdb200_t* t, tbdb200_trackset_t** out)
{
    if (!t || !out) return fail("tbdb200_copy_tracks: null argument");
    *out = new tbdb200_trackset{t->tracks};
    return 0;
}

int tbdb200_set_tracks(tbdb200_t* t, const tbdb200_trackset_t* s)
{
    if (!t) return fail("tbdb200_set_tracks: null handle");
    if (s) t->tracks = s->tracks; else t->tracks.clear();
    return 0;
}

void tbdb200_trackset_destroy(tbdb200_trackset_t* s) { delete s; }

int tbdb200_trajectory_add_position(tbdb200_t* t, int id, int frame, const int* box)
{
    if (!t || !box) return fail("tbdb200_trajectory_add_position: null argument");
    t->gt_position[std::make_pair(id, frame)] = Box{box[0], box[1], box[2], box[3]};
    return 0;
}

int tbdb200_trajectory_get(const tbdb200_t* t, int id, int frame, tbdb200_trajectory_frame* out)
{
    if (!t || !out) return fail("tbdb200_trajectory_get: null argument");
    *out = tbdb200_trajectory_frame();
    const auto it = t->trajectory.find(std::make_pair(id, frame));
    if (it == t->trajectory.end()) return 0;
    const TrajectoryInfo& ti = it->second;
    out->has_info = 1; out->is_tracked = ti.tracked; out->track_id = (int)ti.track_id;
    const Box* b[2] = {&ti.track_pos, &ti.pred_pos};
    int* o[2] = {out->track_pos, out->pred_pos};
    for (int k = 0; k < 2; k++) { o[k][0] = b[k]->x; o[k][1] = b[k]->y; o[k][2] = b[k]->w; o[k][3] = b[k]->h; }
    out->bbox_overlap = ti.overlap;
    return 0;
}

int tbdb200_trajectory_clear(tbdb200_t* t)
{
    if (!t) return fail("tbdb200_trajectory_clear: null handle");
    t->gt_position.clear(); t->trajectory.clear();
    return 0;
}

int tbdb200_metrics(const tbdb200_t* t, tbdb200_frame_metrics* out, int capacity, int* count)
{
    if (!t || !count || (capacity > 0 && !out)) return fail("tbdb200_metrics: bad argument");
    *count = (int)t->metrics.size();
    for (int i = 0; i < std::min(capacity, *count); i++) out[i] = t->metrics[i];
    return 0;
}

double tbdb200_bbox_overlap(const int* a, const int* b)
{
    return iou(Box{a[0], a[1], a[2], a[3]}, Box{b[0], b[1], b[2], b[3]});
}

}  // extern "C"
