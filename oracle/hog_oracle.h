/* oracle/hog_oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C, single-threaded CPU restatement of the reference's HOG multi-scale detection path
 * (tkortz/opencv == OpenCV 3.4.7).  It is the checker the CUDA product is compared with; nothing
 * in opencv_b200/ may include, link or call it.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs use it.
 *
 * Parity pin: every function here is checked bit-for-bit (integers/bytes) or to float equality
 * against the UNMODIFIED reference compiled into oracle/_ref/libcvref.so (tests/test_oracle_vs_ref.py,
 * run in the build container) and against the committed golden vectors under tests/golden/
 * produced by that same library (tests/golden/make_golden.py).
 *
 * All images are 8-bit, `cn` interleaved channels (1, 3 or 4; channel 3 of a 4-channel image is
 * ignored by the gradient, as cv::cuda::HOG does, modules/cudaobjdetect/src/cuda/hog.cu:640-642).
 */
#ifndef HOG_ORACLE_H
#define HOG_ORACLE_H
#ifdef __cplusplus
extern "C" {
#endif

typedef struct hogo_geom {
    int win_w, win_h;       /* 64 x 128 */
    int block_w, block_h;   /* 16 x 16  */
    int stride_w, stride_h; /* block stride 8 x 8 */
    int cell_w, cell_h;     /* 8 x 8 */
    int nbins;              /* 9 */
    double win_sigma;       /* <0 -> (block_w+block_h)/8 */
    double l2hys;           /* 0.2 */
    int gamma;              /* 1 */
} hogo_geom;

void hogo_default_geom(hogo_geom* g);

/* Level enumeration: modules/cudaobjdetect/src/hog.cpp:424-441,453 == modules/objdetect/src/hog.cpp:2058-2073.
 * Writes up to `cap` scales / sizes; returns the number of levels. */
int hogo_levels(int W, int H, int win_w, int win_h, double scale0, int nlevels,
                double* scales, int* widths, int* heights, int cap);

/* cv::resize(..., INTER_LINEAR) for 8U, modules/imgproc/src/resize.cpp:3713-3753, 3470-3706,
 * 1491-1542, 1577-1603, 1780-1857 (incl. the dsize==ssize copy and the exact-2x INTER_AREA switch). */
void hogo_resize_linear(const unsigned char* src, int sw, int sh, int cn, long spitch,
                        unsigned char* dst, int dw, int dh, long dpitch);

/* HOGDescriptor::computeGradient, padding 0: modules/objdetect/src/hog.cpp:239-550 with
 * cartToPolar (core/src/mathfuncs.cpp:267-313) and fastAtan32f / magnitude32f
 * (core/src/mathfuncs_core.simd.hpp:45-71,123-159,194-225).
 * grad: h*w*2 floats; qangle: h*w*2 bytes. */
void hogo_compute_gradient(const hogo_geom* g, const unsigned char* img, int w, int h, int cn, long pitch,
                           float* grad, unsigned char* qangle);

/* Normalised block histograms for every block position of the image (block stride grid), CPU
 * arithmetic order: HOGCache::init / getBlock / normalizeBlockHistogram, hog.cpp:619-858, 860-1117,
 * 1119-1248.  out: [blocks_y][blocks_x][ncells*nbins] floats, cell order (cellX*ncellsY + cellY).
 * blocks_x = (w - block_w)/stride_w + 1 (and likewise y).  Returns blocks_x*blocks_y (0 if the image is
 * smaller than a block). */
long hogo_block_hists(const hogo_geom* g, const unsigned char* img, int w, int h, int cn, long pitch,
                      float* out, int* blocks_x, int* blocks_y);

/* HOGDescriptor::compute with padding 0: windows row-major, blocks column-major inside a window
 * (DESCR_FORMAT_COL_BY_COL), hog.cpp:1589-1653, 851-857.  Returns the number of windows. */
long hogo_compute(const hogo_geom* g, const unsigned char* img, int w, int h, int cn, long pitch,
                  int win_stride_x, int win_stride_y, float* descriptors);

/* SVM score of every window (hog.cpp:1655-1767 arithmetic: float 4-lane partial sums per block, double
 * accumulation).  svm: nsvm floats in the PUBLIC (column-major block) order, bias = svm[descsize] if
 * present.  scores: wins_y*wins_x doubles, row-major.  Returns the number of windows. */
long hogo_scores(const hogo_geom* g, const unsigned char* img, int w, int h, int cn, long pitch,
                 const float* svm, int nsvm, int win_stride_x, int win_stride_y,
                 double* scores, int* wins_x, int* wins_y);

/* cv::groupRectangles(rects, groupThreshold, eps): objdetect/src/cascadedetect.cpp:64-174,
 * cv::partition core/include/opencv2/core/operations.hpp:481-569, SimilarRects objdetect.hpp:121-134.
 * rects: n x (x,y,w,h) ints, grouped in place; returns the new count. */
long hogo_group_rectangles(int* rects, long n, int group_threshold, double eps);

/* The whole path with cv::cuda::HOG::detectMultiScale's host logic (cudaobjdetect/src/hog.cpp:415-488)
 * on the mandated INTER_LINEAR pyramid: levels -> resize -> scores -> hits (score >= hit_threshold,
 * row-major) -> Rect(cvRound(x*s), cvRound(y*s), cvRound(win*s)) -> groupRectangles if group_threshold>0.
 * rects: capacity x 4 ints; conf (may be NULL): capacity doubles, the hit scores (only meaningful when
 * group_threshold == 0).  Returns the rect count, or -1 if capacity is too small. */
long hogo_detect_multiscale(const hogo_geom* g, const unsigned char* img, int w, int h, int cn, long pitch,
                            const float* svm, int nsvm, double hit_threshold,
                            int win_stride_x, int win_stride_y, double scale0, int nlevels,
                            int group_threshold, int* rects, double* conf, long capacity);

#ifdef __cplusplus
}
#endif
#endif
