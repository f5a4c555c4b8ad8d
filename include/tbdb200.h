/* tbdb200.h -- C ABI of the tracking-by-detection step that consumes the detector's boxes.
 *
 * Replaces, for example_gpu_tbd (samples/gpu/tbd.cpp:556-700), the host class cv::tbd::Tracker
 * (modules/trackingbydetection/include/opencv2/tbd.hpp:131-184, src/tbd.cpp:177-1106): constant-velocity prediction,
 * 1 - IoU cost matrix, the reference's own zero-cover assignment procedure, track update / deletion / creation and the
 * per-frame CLEAR-MOT counters.  The arithmetic (integer box stabilisation, double-precision IoU, cvRound of predicted
 * corners) follows the reference operation by operation, so tracks are identical, not merely similar
 * (tests/test_tbd.py pins them against the unmodified reference compiled into oracle/_ref).
 *
 * Host code only (the step is a few dozen boxes per frame); lives in the same library as the detector
 * (opencv_b200/libhogb200.so).  Plain C types, no exceptions across the boundary: every call returns a status
 * (hogb200.h codes: 0 ok, 1 bad argument) and tbdb200_last_error() describes the last failure of the calling thread.
 * One handle per host thread; a handle is not thread-safe, like the reference object.
 */
#ifndef TBDB200_H
#define TBDB200_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct tbdb200 tbdb200_t;              /* cv::tbd::Tracker           tbd.hpp:131 */
typedef struct tbdb200_trackset tbdb200_trackset_t; /* a copy of vector<Track>  (Tracker::getTracks / setTracks) */

/* cv::tbd::TbdArgs (tbd.hpp:25-41, tbd.cpp:53-66) */
typedef struct tbdb200_args {
    double cost_of_non_assignment;
    unsigned int time_window_size;
    unsigned int track_age_threshold;
    double track_visibility_threshold;
    double track_confidence_threshold;
    int should_store_metrics;
} tbdb200_args;

/* cv::tbd::Detection (tbd.hpp:79-93): id = ground-truth id or -1, confidence in [0,1] or -1.0 when unknown.
 * The sample builds one per grouped rect: Detection(-1, frame_id, rect, -1.0) (samples/gpu/tbd.cpp:614-620). */
typedef struct tbdb200_detection {
    int id;
    int frame_id;
    int x, y, width, height;
    double confidence;
} tbdb200_detection;

/* The scalar state of a cv::tbd::Track (tbd.hpp:99-126); its per-frame history is read with tbdb200_track_history. */
typedef struct tbdb200_track {
    unsigned int id;
    unsigned int age;
    unsigned int total_visible_count;
    int history_len;                  /* bboxes.size() == scores.size() == frames.size() */
    double max_confidence;
    double avg_confidence;
    double bbox_overlap;
    int pred_x, pred_y, pred_width, pred_height;     /* predPosition */
    int x, y, width, height;                          /* bboxes.back() */
    int last_frame;                                   /* frames.back() */
} tbdb200_track;

/* per-frame counters kept when should_store_metrics (tbd.hpp:150-156, tbd.cpp:283-297) */
typedef struct tbdb200_frame_metrics {
    int true_positives, false_negatives, false_positives, ground_truths, num_matches;
    double bbox_overlap;
} tbdb200_frame_metrics;

const char* tbdb200_last_error(void);

/* Tracker::Tracker(const TbdArgs*)  tbd.cpp:177-181.  The arguments are copied. */
int tbdb200_create(const tbdb200_args* args, tbdb200_t** out);
void tbdb200_destroy(tbdb200_t* t);

/* Tracker::reset()  tbd.cpp:198-209 */
int tbdb200_reset(tbdb200_t* t);

/* The window outside which predicted tracks are dropped.  The reference hard-codes (0, 1280, 0, 720)
 * (tbd.cpp:222); that is the default here. */
int tbdb200_set_bounds(tbdb200_t* t, int xmin, int xmax, int ymin, int ymax);

/* Tracker::performTrackingStep(detections, trajectoryMap, frame_id)  tbd.cpp:211-298.
 * track_of_detection (optional, n ints): the id of the track each detection was assigned to, -1 if it was left
 * unassigned (and therefore started a new track).  Detections with id >= 0 are also entered into the handle's
 * trajectory records (below), as the reference enters them into the caller's trajectoryMap. */
int tbdb200_step(tbdb200_t* t, const tbdb200_detection* detections, int n, int frame_id, int* track_of_detection);

/* Tracker::getTracks()  tbd.cpp:193-196: number of live tracks / their scalar state, in the reference's order. */
int tbdb200_num_tracks(const tbdb200_t* t);
int tbdb200_get_tracks(const tbdb200_t* t, tbdb200_track* out, int capacity, int* count);
/* history of track `index`: frames[cap], boxes[4*cap] (x, y, width, height), scores[cap]; any may be NULL */
int tbdb200_track_history(const tbdb200_t* t, int index, int* frames, int* boxes, double* scores, int capacity, int* count);

/* Tracker::getTracks() copied / Tracker::setTracks(): the sample restarts every step from the track set of an
 * earlier frame (its "history age" buffers, samples/gpu/tbd.cpp:660-680). */
int tbdb200_copy_tracks(const tbdb200_t* t, tbdb200_trackset_t** out);
int tbdb200_set_tracks(tbdb200_t* t, const tbdb200_trackset_t* tracks);   /* NULL = no tracks */
void tbdb200_trackset_destroy(tbdb200_trackset_t* s);

/* ---- ground-truth trajectories (cv::tbd::Trajectory, tbd.hpp:47-73): the sample registers the ground-truth box of
 * object `id` in frame `frame` (Trajectory::addPosition, tbd.cpp:144-149; samples/gpu/tbd.cpp parseDetections), and a
 * step records, for every detection with id >= 0, whether it was tracked and by which track
 * (Trajectory::addTrackingInfo, tbd.cpp:151-175).  The maps live with the handle; tbdb200_reset leaves them alone, as
 * Tracker::reset leaves the caller's trajectoryMap alone. */
typedef struct tbdb200_trajectory_frame {
    int has_info;                 /* 0: this (id, frame) was never reported by a step */
    int is_tracked;               /* isTrackedPerFrame */
    int track_id;                 /* trackIdPerFrame            (valid when is_tracked) */
    int track_pos[4];             /* trackPosPerFrame: the track's newest box, x y width height */
    int pred_pos[4];              /* predPosPerFrame */
    double bbox_overlap;          /* bboxOverlapPerFrame: IoU of track_pos with the registered ground-truth box */
} tbdb200_trajectory_frame;
int tbdb200_trajectory_add_position(tbdb200_t* t, int id, int frame, const int* box);
int tbdb200_trajectory_get(const tbdb200_t* t, int id, int frame, tbdb200_trajectory_frame* out);
int tbdb200_trajectory_clear(tbdb200_t* t);

/* the per-frame metric vectors (one entry per step since the last reset) */
int tbdb200_metrics(const tbdb200_t* t, tbdb200_frame_metrics* out, int capacity, int* count);

/* computeBoundingBoxOverlap  tbd.cpp:1083-1106 (intersection over union of two x,y,width,height boxes) */
double tbdb200_bbox_overlap(const int* a, const int* b);

#ifdef __cplusplus
}
#endif
#endif /* TBDB200_H */
