// hog_internal.h -- structures shared by the host pipeline (hog_host.cu) and the sm_100a kernels.
//
// Geometry is fixed to what the north-star scope names (and what cudaobjdetect.hpp:89-93 documents as
// supported): block 16x16, cell 8x8, block stride 8x8, 9 bins -> 36 floats per block.  The window size
// (64x128, 48x96, ...) and the window stride (a multiple of 8) are run-time parameters.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace hogk {

constexpr int kCell = 8;
constexpr int kBlock = 16;
constexpr int kBins = 9;
constexpr int kHist = 36;            // floats per block histogram
constexpr int kMaxLevels = 256;
constexpr int kMaxDet = 4;
#ifndef HOGB200_MAX_GROUP
#define HOGB200_MAX_GROUP 16
#endif
constexpr int kMaxGroup = HOGB200_MAX_GROUP;        // frames one launch of each kernel can process (frame groups of the batch path)

// The frames of one group.  Frame f keeps its pyramid levels at LevelDesc::img + f * pyr_stride and its block
// histograms at hist + f * hist_stride; level 0 aliases base[f].  Work items carry the frame index.
struct FrameSet {
    const uint8_t* base[kMaxGroup];
    int pitch[kMaxGroup];
    int nframes;                     // items of frames >= nframes are skipped (partial group)
    int pad;
    long long pyr_stride;            // bytes
    long long hist_stride;           // floats
};

// One pyramid level of one frame.  Lives in device memory (array indexed by level).
struct LevelDesc {
    const uint8_t* img;   // level image (level 0 may alias the caller's frame)
    int pitch;            // bytes
    int w, h;             // pixels
    int nbx, nby;         // block grid: (w-16)/8+1, (h-16)/8+1 (0 if smaller than a block)
    int nwx, nwy;         // window grid for the current win size / stride
    long long hist_off;   // float offset of this level's [nby][nbx][36] plane in the hist buffer
    long long score_off;  // float offset of this level's [nwy][nwx] plane in the score buffer
    int xtab_off, ytab_off;  // offsets into the resize coefficient tables (entries)
    int mode;             // 0 = alias/copy of base, 1 = bilinear, 2 = exact-2x area
    int rs_rows;          // resize: destination rows per warp and staging scheme (resize_plan_rows)
};

// resize coefficient table entry (one per destination column / row of a level), everything the kernel would
// otherwise recompute per pixel is folded in on the host:
//   column entry: s0 = left source column, s1 = min(s0 + 1, W - 1) (its weight is 0 at the right edge),
//                 c0 = a0 | a1 << 16  (11-bit fixed-point weights, resize.cpp:3647-3655), c1 unused
//   row entry:    s0 / s1 = the two source rows AFTER the clip of resize.cpp:1821-1825,
//                 c0 = b0, c1 = b1  (11-bit fixed-point weights; b * (S >> 4) fits 32 bits)
struct __align__(16) ResizeCoef { int s0, s1; unsigned c0, c1; };

// work item of the fused gradient+histogram kernel (one warp): a strip of block rows x a tile of block columns;
// level = pyramid level | frame-in-group << 8
struct HistItem { short level; short bx0; short by0; short nby; };   // nby = block rows in this strip
// work item of the SVM kernel: a tile of windows of one level of one frame
struct SvmItem { short level; short wx0; short wy0; short frame; };
// work item of the resize kernel: a tile of destination pixels of one level of one frame
struct ResizeItem { short level; short x0; short y0; short frame; };

// separable block window: vertical (or horizontal) weight of pixel offset i in [0,16) for cell 0 / cell 1 of
// the block: gauss(i) * bilinear cell weight (HOGCache::init, objdetect/src/hog.cpp:653-839)
struct BlockWeights { float a0[16]; float a1[16]; };

// x,y = window index (not pixels); level = level | det << 16 | frame-in-chunk << 20
struct Candidate { int level; int x; int y; float score; };

// ---- kernel launchers (defined in the .cu files) ----------------------------------------------------
int resize_tile_w(int cn);   // tile geometry the host uses to build ResizeItems
int resize_tile_h(int cn, int rs_rows);
int resize_plan_rows(const ResizeCoef* xt, int dst_w, const ResizeCoef* yt, int dst_h, int src_w, int cn);
// 8UC3 (B,G,R) -> 8UC4 (B,G,R,255): the colour conversion example_gpu_tbd does on the host before the upload
// (samples/gpu/tbd.cpp:573-599), moved to the device.  dst rows are 16-byte aligned.
void launch_bgr_to_bgra(const uint8_t* src, size_t spitch, uint8_t* dst, size_t dpitch, int w, int h, cudaStream_t st);
// dst = src with a BORDER_REFLECT_101 frame of px / py pixels (cn = 1 or 4; 4-byte aligned rows for cn = 4)
// rows of up to kMaxGroup flat-uploaded frames (same geometry) to their aligned pitch, 4 bytes per thread
struct RepitchSet { const uint8_t* src[kMaxGroup]; uint8_t* dst[kMaxGroup]; int n, spitch, dpitch, row_bytes, rows; };
void launch_repitch(const RepitchSet& rp, cudaStream_t st);
void launch_pad_reflect101(const uint8_t* src, size_t spitch, int w, int h, int cn, uint8_t* dst, size_t dpitch, int px, int py,
                           cudaStream_t st);
void launch_resize(const FrameSet& fs, int bw, int bh, int cn,
                   const LevelDesc* d_levels, const ResizeCoef* d_xtab, const ResizeCoef* d_ytab,
                   const ResizeItem* d_items, int nitems, cudaStream_t st);

// base/base_pitch: the frame itself, used for levels whose LevelDesc::img is null (level 0 aliases the frame)
// d_queue: two zero-initialised counters in device memory (work queue head, finished workers); the kernel
// re-arms them itself, so consecutive launches on one stream need no host reset.  Rows of the level images
// must be 16-byte aligned for cn == 4 (4-byte for cn == 1).
void launch_hist(const FrameSet& fs, const LevelDesc* d_levels, const HistItem* d_items,
                 int nitems, unsigned int* d_queue, int num_sms, int cn, int gamma, const BlockWeights& bw,
                 float l2hys, float* d_hist, cudaStream_t st);
// allocation-free variant: queue counters, level descriptor and work items live in a file-scope __device__ array of
// hist.cu (hist_noalloc_scratch_bytes()); single image, output straight into the caller's device buffer.
// 0 = queued, 1 = image needs more work items than the static scratch holds (*items_needed), 2 = CUDA error
int launch_hist_noalloc(const uint8_t* d_img, int pitch, int w, int h, int cn, int gamma, const BlockWeights& bw, float l2hys,
                        float* d_out, int num_sms, cudaStream_t st, long long* items_needed);
size_t hist_noalloc_scratch_bytes();
int hist_workers_per_sm();  // warps pulling items from the queue, per SM (the host balances the item list for them)

struct SvmParams {
    int win_bx, win_by;        // window size in blocks (7 x 15)
    int stride_bx, stride_by;  // window stride in blocks
    int ndet;                  // detectors scored in this pass
    const float* det[kMaxDet]; // reordered (row-major blocks) coefficients, win_by*win_bx*36 floats each
    float bias[kMaxDet];
    float thresh;              // hit threshold
    float* scores;             // optional dense output (nullptr = none); detector d at + d*score_plane
    long long score_plane;
    Candidate* cand;           // compacted hits of all frames / detectors of a chunk
    unsigned int* cand_count;  // one counter
    unsigned int cand_cap;
    int tag;                   // OR-ed into Candidate::level: frame-in-chunk << 20 (detector goes to bits 16..19)
};
// items1 / items2: tile lists for passes scoring one detector (tiles svm_tile_wx(.., 1) windows wide) and two detectors
// at once (16 wide); detectors are taken two per pass where the window geometry allows it
void launch_svm(const FrameSet& fs, const LevelDesc* d_levels, const SvmItem* d_items1, int nitems1,
                const SvmItem* d_items2, int nitems2, const float* d_hist, const SvmParams& p, int num_sms, cudaStream_t st);
bool svm_supports(int win_bx, int win_by);            // window geometries the scoring kernel takes (any up to 21 x ~14 blocks^2 of shared memory)
bool svm_two_detector_pass(int win_bx, int win_by);
int svm_tile_wx(int win_bx, int win_by, int ndet_pass);   // tile geometry the host uses to build SvmItems
int svm_tile_wy();
int svm_tile_cost(int cols_left, int rows_left, int units);   // relative time of a tile with that many valid windows right / below its origin

void launch_extract(const LevelDesc* d_level0, int nwx, int nwy, const float* d_hist, int win_bx, int win_by,
                    int stride_bx, int stride_by, int col_by_col, float* d_out, cudaStream_t st);

// cv::groupRectangles on the device (group.cu): rects are 4 ints (x, y, width, height) in host memory, in the
// reference's input order; returns the number of grouped rects written to h_out, -1 on a CUDA error
size_t group_workspace_bytes(int n);
int group_rectangles_device(const int* h_rects, int n, int thr, double eps, int* h_out, void* ws, cudaStream_t st);

// The tail of detectMultiScale for all (frame, detector) segments of a chunk on the device (group.cu): candidate list ->
// reference order -> scaled rects -> groupRectangles per segment.  Nothing is synchronised; see group.cu.
struct CandGeom {
    double scale[kMaxLevels]; int win_w, win_h, stride_x, stride_y;
    int bits_x, bits_y, bits_level, bits_rx;     // key widths: window x / y, level, scaled rect x (radix passes follow the key width)
};
size_t cand_group_workspace_bytes(int n, int nseg);
int cand_group_device(const void* d_cand, int n, int nseg, int ndet, const CandGeom& G, int thr, double eps, void* ws,
                      cudaStream_t st, const int** d_out_p, const int** d_hdr_p);

int hist_tile_blocks_x();   // blocks per tile in x (14)
int hist_strip_blocks_y();  // block rows per strip: 0 = let the host choose, else the HOGB200_STRIP override

}  // namespace hogk
