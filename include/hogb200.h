/* hogb200.h -- C ABI of the B200-native HOG multi-scale detector.
 *
 * Drop-in boundary for the hot path behind tkortz/opencv's example_gpu_tbd: everything a binding of
 * cv::cuda::HOG (modules/cudaobjdetect/include/opencv2/cudaobjdetect.hpp:78-212, implemented by HOG_Impl in
 * modules/cudaobjdetect/src/hog.cpp:139-689) needs.  Plain pointers and sizes only; no exceptions cross
 * this boundary: every call returns a hogb200_status and hogb200_last_error() holds the text.
 *
 * Threading: one handle per host thread per device (the reference object is not thread-safe either);
 * handles never share global/constant state, so several handles with different geometry may coexist.
 * All work of a handle is issued on its stream (hogb200_set_stream, default: a private non-blocking
 * stream); the blocking calls synchronise that stream before returning.
 *
 * There is NO CPU fallback: without a CUDA device hogb200_create fails with HOGB200_CUDA_ERROR.
 */
#ifndef HOGB200_H
#define HOGB200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef enum hogb200_status {
    HOGB200_OK = 0,
    HOGB200_BAD_ARG = 1,      /* the reference would CV_Assert / CV_Error here */
    HOGB200_CUDA_ERROR = 2,   /* the reference would throw from cudaSafeCall (core/cuda/common.hpp:66-75) */
    HOGB200_CAPACITY = 3,     /* caller-provided output array too small; *count holds the needed size */
    HOGB200_UNSUPPORTED = 4   /* geometry outside block 16x16 / cell 8x8 / stride 8x8 / 9 bins */
} hogb200_status;

/* cv::cuda::HOG::DESCR_FORMAT_* (cudaobjdetect.hpp:82-86) */
enum { HOGB200_DESCR_FORMAT_COL_BY_COL = 0, HOGB200_DESCR_FORMAT_ROW_BY_ROW = 1 };

typedef struct hogb200 hogb200_t;

/* HOG::create(win_size, block_size, block_stride, cell_size, nbins), cudaobjdetect.hpp:94-98 */
typedef struct hogb200_geom {
    int win_w, win_h;                    /* 64 x 128 (or 48 x 96)            */
    int block_w, block_h;                /* 16 x 16                           */
    int block_stride_w, block_stride_h;  /* 8 x 8                             */
    int cell_w, cell_h;                  /* 8 x 8                             */
    int nbins;                           /* 9                                 */
} hogb200_geom;

/* One frame: 8-bit, 1 or 4 interleaved channels (CV_8UC1 / CV_8UC4, hog.cpp:340,421,496), pitched rows.
 * 3 channels (CV_8UC3, B,G,R) are accepted too and widened to B,G,R,255 on the device: the cvtColor that
 * example_gpu_tbd runs on the host before its upload (samples/gpu/tbd.cpp:573-599); results equal those of the
 * 8UC4 frame.  hogb200_resize takes 1 or 4 channels only.
 * `data` is a device pointer when on_device != 0 (a GpuMat), else a host pointer (pinned memory makes the
 * upload asynchronous).  A ROI is expressed by data/pitch as with GpuMat. */
typedef struct hogb200_image {
    const void* data;
    size_t pitch;      /* bytes between rows */
    int width, height;
    int channels;      /* 1, 3 or 4 */
    int on_device;
} hogb200_image;

typedef struct hogb200_rect { int x, y, width, height; } hogb200_rect;
typedef struct hogb200_point { int x, y; } hogb200_point;

const char* hogb200_last_error(void);
const char* hogb200_version(void);

/* ---- lifetime: HOG::create, hog.cpp:682-689 + ctor asserts :252-264.  geom == NULL -> defaults. */
/* ---- detector files (widening row f3) ------------------------------------------------------------------
 * The "opencv-object-detector-hog" FileStorage node, XML or YAML by file extension, that
 * cv::HOGDescriptor::save/load (modules/objdetect/src/hog.cpp:139-209) and cv::cuda::HOG::save/load
 * (modules/cudaobjdetect/src/hog.cpp:599-679) exchange -- including SVMDetector, which the fork's GPU writer
 * omits.  Files written here load in the reference and vice versa.  Host-only: no CUDA device is needed. */
typedef struct hogb200_model {
    hogb200_geom geom;
    double win_sigma;          /* < 0: (block_w + block_h) / 8 (getWinSigma, hog.cpp:104-107) */
    double l2hys_threshold;
    int gamma_correction;
    int nlevels;
} hogb200_model;
/* objname NULL: the first top-level node (load) / FileStorage::getDefaultObjectName(filename) (save) */
int hogb200_model_read(const char* filename, const char* objname, hogb200_model* m, float* detector, size_t capacity,
                       size_t* count);
int hogb200_model_write(const char* filename, const char* objname, const hogb200_model* m, const float* detector, size_t n);
/* cv::cuda::HOG::save with the detector of `slot` included; HOG::load as a factory (the geometry is fixed at creation) */
int hogb200_save(const hogb200_t* h, const char* filename, const char* objname, int slot);
int hogb200_create_from_file(const char* filename, const char* objname, int device, hogb200_t** out);

int  hogb200_create(const hogb200_geom* geom, int device, hogb200_t** out);
void hogb200_destroy(hogb200_t* h);
/* cudaStream_t as void*; NULL restores the handle's private stream.  (HOG::compute's Stream& argument.) */
int  hogb200_set_stream(hogb200_t* h, void* cuda_stream);

/* ---- parameters: virtual setters/getters of cv::cuda::HOG, cudaobjdetect.hpp:101-142; defaults hog.cpp:241-249 */
int    hogb200_set_win_sigma(hogb200_t* h, double v);         double hogb200_get_win_sigma(const hogb200_t* h);
int    hogb200_set_l2hys_threshold(hogb200_t* h, double v);   double hogb200_get_l2hys_threshold(const hogb200_t* h);
int    hogb200_set_gamma_correction(hogb200_t* h, int v);     int    hogb200_get_gamma_correction(const hogb200_t* h);
int    hogb200_set_num_levels(hogb200_t* h, int v);           int    hogb200_get_num_levels(const hogb200_t* h);
int    hogb200_set_hit_threshold(hogb200_t* h, double v);     double hogb200_get_hit_threshold(const hogb200_t* h);
int    hogb200_set_win_stride(hogb200_t* h, int sx, int sy);  void   hogb200_get_win_stride(const hogb200_t* h, int* sx, int* sy);
int    hogb200_set_scale_factor(hogb200_t* h, double v);      double hogb200_get_scale_factor(const hogb200_t* h);
int    hogb200_set_group_threshold(hogb200_t* h, int v);      int    hogb200_get_group_threshold(const hogb200_t* h);
int    hogb200_set_descriptor_format(hogb200_t* h, int v);    int    hogb200_get_descriptor_format(const hogb200_t* h);
size_t hogb200_get_descriptor_size(const hogb200_t* h);       /* hog.cpp:278-281 */
size_t hogb200_get_block_histogram_size(const hogb200_t* h);  /* hog.cpp:283-286 */

/* ---- detectors.  setSVMDetector (hog.cpp:293-321): n == descriptor_size or descriptor_size + 1 (bias last),
 * PUBLIC order (blocks column-major inside the window, as HOGDescriptor::getDefaultPeopleDetector returns it).
 * slot 0 is the detector the cv::cuda::HOG-shaped calls use; slots 1..HOGB200_MAX_DETECTORS-1 hold further
 * detectors (e.g. a vehicle SVM) scored in the same pass over the block histograms.  n == 0 clears the slot. */
#define HOGB200_MAX_DETECTORS 4
int hogb200_set_svm_detector(hogb200_t* h, int slot, const float* coeffs, size_t n);
/* getDefaultPeopleDetector (hog.cpp:326-334): 64x128 -> 3781 floats, 48x96 -> 1981 floats, else BAD_ARG.
 * The pointer refers to static storage. */
int hogb200_default_people_detector(int win_w, int win_h, const float** coeffs, size_t* n);

/* ---- work calls (blocking, like the reference's).  Output arrays are caller-owned host memory with
 * `capacity` entries; *count receives the number produced (or needed, with HOGB200_CAPACITY).
 *
 * detect (hog.cpp:336-413): single scale; hits are window origins in row-major window order;
 * conf (nullable) receives the SVM scores of the hits as doubles. */
int hogb200_detect(hogb200_t* h, const hogb200_image* img,
                   hogb200_point* hits, double* conf, size_t capacity, size_t* count);
/* detectMultiScale (hog.cpp:415-488): levels, INTER_LINEAR pyramid, per-level detect, rect scaling,
 * cv::groupRectangles(found, group_threshold, 0.2) when group_threshold > 0.
 * conf != NULL requires group_threshold == 0 (hog.cpp:422). */
int hogb200_detect_multiscale(hogb200_t* h, const hogb200_image* img,
                              hogb200_rect* found, double* conf, size_t capacity, size_t* count);
/* The same for detector `slot` (e.g. the vehicle SVM), reusing nothing from previous calls. */
int hogb200_detect_multiscale_slot(hogb200_t* h, int slot, const hogb200_image* img,
                                   hogb200_rect* found, double* conf, size_t capacity, size_t* count);
/* Frame batch (example_gpu_tbd's frame loop, samples/gpu/tbd.cpp:568-621, with the per-frame host sync
 * removed): frames are queued back to back on the stream; rects of frame i are
 * found[offsets[i] .. offsets[i+1]) ; offsets has nframes+1 entries.  With nslots > 1 every frame is scored
 * against detector slots 0..nslots-1 in one pass and the output is frame-major, slot-minor
 * (offsets has nframes*nslots+1 entries). */
int hogb200_detect_multiscale_batch(hogb200_t* h, const hogb200_image* frames, int nframes, int nslots,
                                    hogb200_rect* found, size_t capacity, size_t* offsets);
/* The same batch, asynchronously -- the counterpart of the Stream& argument of HOG::compute (hog.cpp:490-534) for the
 * detection path, whose reference form blocks on the null stream (hog.cpp:336-413): submit queues the uploads and kernels
 * of the whole batch on the handle's streams and returns a ticket; collect waits for THAT batch, runs its tail (candidate
 * ordering, rect scaling, groupRectangles -- on the device from HOGB200_GPU_POST_MIN = 1024 candidates per chunk) and
 * fills found / offsets exactly as hogb200_detect_multiscale_batch does.  Two batches may be in flight per handle, so the
 * uploads and kernels of batch k + 1 overlap the tail of batch k.  The frames' pixel buffers must stay valid, and the
 * handle's parameters and detectors unchanged, until the batch is collected; no other work call of the handle may run
 * in between.  Batches are collected in any order. */
int hogb200_submit_batch(hogb200_t* h, const hogb200_image* frames, int nframes, int nslots, int* ticket);
int hogb200_collect_batch(hogb200_t* h, int ticket, hogb200_rect* found, size_t capacity, size_t* offsets);
/* compute (hog.cpp:490-534): descriptors of every window, wins x descriptor_size floats in
 * descriptor_format order.  out is a host pointer (out_on_device == 0) or a device pointer. */
int hogb200_compute(hogb200_t* h, const hogb200_image* img, float* out, size_t capacity_floats,
                    int out_on_device, size_t* nwindows);

/* ---- stage entry points: what cv::cuda::device::hog::* exposes to HOG_Impl (hog.cpp:63-134).
 * Used by the parity tests to check each stage against the oracle; all outputs are host arrays. */
/* level enumeration (hog.cpp:424-441,453): scales/widths/heights with `capacity` entries */
int hogb200_levels(const hogb200_t* h, int width, int height,
                   double* scales, int* widths, int* heights, size_t capacity, size_t* count);
/* the same rule without a handle (host-only) */
int hogb200_enumerate_levels(int width, int height, int win_w, int win_h, double scale0, int nlevels,
                             double* scales, int* widths, int* heights, size_t capacity, size_t* count);
/* hog::resize_8UC1 / resize_8UC4 (hog.cpp:130-131) with cv::resize INTER_LINEAR semantics; out is tightly packed */
int hogb200_resize(hogb200_t* h, const hogb200_image* img, int dst_w, int dst_h, uint8_t* out);
/* compute_gradients + compute_hists + normalize_hists (hog.cpp:543-595): blocks_y x blocks_x x 36 floats,
 * cell order (cellX*2 + cellY) inside a block as in hog.cu:202 */
int hogb200_block_histograms(hogb200_t* h, const hogb200_image* img, float* out, size_t capacity_floats,
                             int* blocks_x, int* blocks_y);
/* hogb200_block_histograms without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same arithmetic and output layout as hogb200_block_histograms, but
 *   - the kernel's scratch (work-queue counters, level descriptor, work-item list) lives in a file-scope
 *     `static __device__` array of hogb200_noalloc_scratch_bytes() = 256 KiB that every GPU loading the library holds
 *     until exit; the call makes no cudaMalloc / cudaFree / allocating library call;
 *   - img must be a DEVICE image, CV_8UC1 or CV_8UC4, rows 4-byte / 16-byte aligned, pitch covering the width rounded
 *     up to 4 pixels (nothing can be staged); d_out is a 16-byte aligned DEVICE buffer of capacity_floats floats;
 *   - limit: at most 32 704 work items of 14 block columns x 4..16 block rows (any image up to about 8192 x 8192);
 *     larger images are refused with HOGB200_CAPACITY and a message naming the limit;
 *   - the work is QUEUED on the handle's stream and the call returns without synchronising;
 *   - every call (of every handle on that GPU) shares the one scratch array, which a set-up kernel rewrites: calls
 *     that overlap on the GPU (different streams / handles) race.  Serialise them on one stream. */
size_t hogb200_noalloc_scratch_bytes(void);
int hogb200_block_histograms_noalloc(hogb200_t* h, const hogb200_image* img, float* d_out, size_t capacity_floats,
                                     int* blocks_x, int* blocks_y);
/* compute_confidence_hists (hog.cpp:383-397): SVM score of every window of one image for detector `slot`,
 * wins_y x wins_x floats row-major */
int hogb200_window_scores(hogb200_t* h, int slot, const hogb200_image* img, float* out, size_t capacity_floats,
                          int* wins_x, int* wins_y);
/* cv::groupRectangles(rects, group_threshold, eps) (objdetect/src/cascadedetect.cpp:64-174, 366-371), in place */
int hogb200_group_rectangles(hogb200_rect* rects, size_t n, int group_threshold, double eps, size_t* count);
/* The same function computed on the handle's device (csrc/group.cu: sort by x, warp-per-rect SimilarRects sweep,
 * lock-free union-find, class sums, filter, ordered compaction): identical output, for candidate lists of tens of
 * thousands of rects.  detectMultiScale switches to it by itself from 1024 candidates
 * per frame (environment HOGB200_GPU_GROUP_MIN overrides, 0 = never). */
int hogb200_group_rectangles_device(hogb200_t* h, hogb200_rect* rects, size_t n, int group_threshold, double eps,
                                    size_t* count);

/* cv::groupRectangles_meanshift(rectList, foundWeights, foundScales, detectThreshold, winDetSize)
 * (objdetect/src/cascadedetect.cpp:176-395): the grouping cv::HOGDescriptor::detectMultiScale uses with
 * useMeanshiftGrouping (objdetect/src/hog.cpp:2100-2101).  rects / weights (SVM scores) / scales (pyramid scale of each
 * hit) have n entries; in place, *count modes survive (weights then hold the mode densities).  Host-only. */
int hogb200_group_rectangles_meanshift(hogb200_rect* rects, double* weights, const double* scales, size_t n,
                                       double detect_threshold, int win_w, int win_h, size_t* count);

/* ---- ROI variants: score only requested window origins, e.g. around tracker-predicted boxes.
 * cv::HOGDescriptor::detectROI (objdetect/src/hog.cpp:3549-3660): `locations` are window origins in pixels of `img`
 * (multiples of 8: the cache stride; the reference asserts the same in HOGCache::getBlock); windows may reach pad_w / pad_h
 * pixels (rounded up to 8) outside the image, where gradients follow BORDER_REFLECT_101 as in the CPU class.
 * confidences (nullable) receives one score per location, -10.0 for a location outside the padded image (:3596-3602);
 * found / *count the locations with score >= hit threshold, in input order.  Detector slot 0. */
int hogb200_detect_roi(hogb200_t* h, const hogb200_image* img, const hogb200_point* locations, size_t n, int pad_w, int pad_h,
                       hogb200_point* found, double* confidences, size_t capacity, size_t* count);
/* cv::DetectionROI (objdetect.hpp:345-353): one pyramid scale, the locations to evaluate IN THAT LEVEL's pixels, and (out,
 * nullable) their confidences. */
typedef struct hogb200_roi {
    double scale;
    const hogb200_point* locations;
    size_t nlocations;
    double* confidences;
} hogb200_roi;
/* cv::HOGDescriptor::detectMultiScaleROI (hog.cpp:3662-3677, worker :3489-3537): per roi the level cvRound(size / scale)
 * (resized with the INTER_LINEAR pyramid semantics of this library), detectROI with padding 8 x 8, rects
 * (cvRound(x*scale), cvRound(y*scale), cvRound(win*scale)) in roi order, then cv::groupRectangles(found, group_threshold, 0.2). */
int hogb200_detect_multiscale_roi(hogb200_t* h, const hogb200_image* img, hogb200_roi* rois, size_t nrois, int group_threshold,
                                  hogb200_rect* found, size_t capacity, size_t* count);

/* frames the batch entry point hands to one launch of each kernel for frames of this shape under the handle's current
 * scale factor / nlevels (instrumentation: bench.py quotes its roofline on launches of this size) */
int hogb200_batch_group(const hogb200_t* h, int width, int height, int channels);

/* ---- instrumentation for bench.py: device time (ms, CUDA events on the handle's stream) of the stages of
 * the last work call, and the number of kernels it launched. */
typedef struct hogb200_timings {
    float upload_ms, resize_ms, hist_ms, svm_ms, download_ms, total_ms;
    int   kernel_launches;
    long long pyramid_pixels, blocks, windows, candidates;
} hogb200_timings;
int hogb200_enable_timing(hogb200_t* h, int on);
int hogb200_last_timings(const hogb200_t* h, hogb200_timings* out);

#ifdef __cplusplus
}
#endif
#endif /* HOGB200_H */
