// ref_tbd_shim.cpp -- TEST INFRASTRUCTURE, never the product path.
//
// extern "C" wrapper over the UNMODIFIED reference tracker cv::tbd::Tracker
// (modules/trackingbydetection/src/tbd.cpp, compiled from where it lies under /root/reference by oracle/Makefile
// into oracle/_ref/libcvref.so).  tests/test_tbd.py drives it and opencv_b200's tbdb200_* with the same detections
// and compares every track after every step; tests/golden/make_tbd_golden.py records its outputs as fixtures.
#include <opencv2/tbd.hpp>

#include <iostream>
#include <map>
#include <vector>

using namespace cv;
using namespace cv::tbd;

struct RefTracker {
    TbdArgs args;
    Tracker tracker;
    std::map<int, Trajectory> trajectories;
    std::vector<std::vector<Track> > snapshots;
    RefTracker(double c, unsigned tw, unsigned age, double vis, double conf, bool metrics)
        : args(c, tw, age, vis, conf, metrics), tracker(&args) {}
};

// The reference narrates to std::cout ("Filtering out track ...").  Inside a Python process that has numpy loaded,
// formatting a number through this library's std::cout crashes in the locale facets, so the stream is put in the
// failed state (every insertion is then skipped) for the duration of a step.  The reference code is untouched.
struct MuteCout {
    std::ios_base::iostate saved;
    MuteCout() : saved(std::cout.rdstate()) { std::cout.setstate(std::ios_base::failbit); }
    ~MuteCout() { std::cout.clear(saved); }
};

extern "C" {

void* cvref_tbd_create(double cost_non_assign, unsigned time_window, unsigned age_thr, double vis_thr, double conf_thr, int metrics)
{
    return new RefTracker(cost_non_assign, time_window, age_thr, vis_thr, conf_thr, metrics != 0);
}
void cvref_tbd_destroy(void* h) { delete (RefTracker*)h; }
void cvref_tbd_reset(void* h) { ((RefTracker*)h)->tracker.reset(); }

// dets: n x (id, frame_id, x, y, w, h) ints + conf[n]
void cvref_tbd_step(void* h, const int* dets, const double* conf, int n, int frame_id)
{
    RefTracker* r = (RefTracker*)h;
    std::vector<Detection> v;
    for (int i = 0; i < n; i++) {
        Rect bb(dets[6 * i + 2], dets[6 * i + 3], dets[6 * i + 4], dets[6 * i + 5]);
        v.push_back(Detection(dets[6 * i], dets[6 * i + 1], bb, conf[i]));
    }
    MuteCout mute;
    r->tracker.performTrackingStep(v, r->trajectories, frame_id);
}

int cvref_tbd_num_tracks(void* h) { return (int)((RefTracker*)h)->tracker.getTracks().size(); }

// per track: ints (id, age, visible, history_len, pred x,y,w,h, box x,y,w,h, last_frame) = 13, doubles (max, avg, overlap) = 3
void cvref_tbd_get_tracks(void* h, int* ints, double* dbls)
{
    std::vector<Track>& t = ((RefTracker*)h)->tracker.getTracks();
    for (size_t i = 0; i < t.size(); i++) {
        int* o = ints + 13 * i;
        o[0] = (int)t[i].id; o[1] = (int)t[i].age; o[2] = (int)t[i].totalVisibleCount; o[3] = (int)t[i].bboxes.size();
        o[4] = t[i].predPosition.x; o[5] = t[i].predPosition.y; o[6] = t[i].predPosition.width; o[7] = t[i].predPosition.height;
        const Rect& b = t[i].bboxes.back();
        o[8] = b.x; o[9] = b.y; o[10] = b.width; o[11] = b.height; o[12] = t[i].frames.back();
        dbls[3 * i] = t[i].maxConfidence; dbls[3 * i + 1] = t[i].avgConfidence; dbls[3 * i + 2] = t[i].bboxOverlap;
    }
}

int cvref_tbd_track_history(void* h, int index, int* frames, int* boxes, double* scores, int cap)
{
    std::vector<Track>& t = ((RefTracker*)h)->tracker.getTracks();
    const Track& tr = t[index];
    const int n = (int)tr.bboxes.size();
    for (int i = 0; i < n && i < cap; i++) {
        frames[i] = tr.frames[i]; scores[i] = tr.scores[i];
        boxes[4 * i] = tr.bboxes[i].x; boxes[4 * i + 1] = tr.bboxes[i].y; boxes[4 * i + 2] = tr.bboxes[i].width; boxes[4 * i + 3] = tr.bboxes[i].height;
    }
    return n;
}

// getTracks() copied into a numbered snapshot / setTracks() from one (the sample's history-age buffers)
int cvref_tbd_snapshot(void* h)
{
    RefTracker* r = (RefTracker*)h;
    std::vector<Track> copy;
    std::vector<Track>& t = r->tracker.getTracks();
    for (size_t i = 0; i < t.size(); i++) copy.push_back(Track(t[i]));
    r->snapshots.push_back(copy);
    return (int)r->snapshots.size() - 1;
}
void cvref_tbd_restore(void* h, int snapshot)
{
    RefTracker* r = (RefTracker*)h;
    std::vector<Track> copy;
    if (snapshot >= 0) copy = r->snapshots[snapshot];
    r->tracker.setTracks(copy);
}

// Trajectory::addPosition for ground-truth object `id` (creates the trajectory on first use, like the sample's parser)
void cvref_tbd_add_position(void* h, int id, int frame, const int* box)
{
    RefTracker* r = (RefTracker*)h;
    if (!r->trajectories.count(id)) r->trajectories[id] = Trajectory(id);
    Rect bb(box[0], box[1], box[2], box[3]);
    Vec3d w;
    r->trajectories[id].addPosition(frame, bb, w);
}
// ints: is_tracked, track_id, track_pos[4], pred_pos[4]; returns 0 when the (id, frame) pair was never reported
int cvref_tbd_trajectory_get(void* h, int id, int frame, int* ints, double* overlap)
{
    RefTracker* r = (RefTracker*)h;
    if (!r->trajectories.count(id)) return 0;
    Trajectory& t = r->trajectories[id];
    if (!t.isTrackedPerFrame.count(frame)) return 0;
    ints[0] = t.isTrackedPerFrame[frame] ? 1 : 0;
    if (t.trackIdPerFrame.count(frame)) {
        ints[1] = t.trackIdPerFrame[frame];
        const Rect a = t.trackPosPerFrame[frame], b = t.predPosPerFrame[frame];
        ints[2] = a.x; ints[3] = a.y; ints[4] = a.width; ints[5] = a.height;
        ints[6] = b.x; ints[7] = b.y; ints[8] = b.width; ints[9] = b.height;
        *overlap = t.bboxOverlapPerFrame[frame];
    }
    return 1;
}

// metrics: count, then per frame ints (tp, fn, fp, gt, matches) and double overlap
int cvref_tbd_metrics(void* h, int* ints, double* overlap, int cap)
{
    Tracker& t = ((RefTracker*)h)->tracker;
    const int n = (int)t.truePositives.size();
    for (int i = 0; i < n && i < cap; i++) {
        ints[5 * i] = t.truePositives[i]; ints[5 * i + 1] = t.falseNegatives[i]; ints[5 * i + 2] = t.falsePositives[i];
        ints[5 * i + 3] = t.groundTruths[i]; ints[5 * i + 4] = t.numMatches[i];
        overlap[i] = t.bboxOverlap[i];
    }
    return n;
}

}  // extern "C"
