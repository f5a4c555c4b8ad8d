// oracle/ref_shim.cpp -- TEST INFRASTRUCTURE, not product code.
//
// A thin extern "C" wrapper over the UNMODIFIED reference (tkortz/opencv, OpenCV 3.4.7)
// CPU implementation of the HOG path.  It is compiled together with the reference's own
// core / imgproc / objdetect sources (from /root/reference, where they lie) into
// oracle/_ref/libcvref.so by oracle/Makefile.  Only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs may load it.
//
// Reference entry points wrapped here:
//   cv::resize(INTER_LINEAR / INTER_LINEAR_EXACT)   modules/imgproc/src/resize.cpp:3713-3753
//   cv::HOGDescriptor::computeGradient              modules/objdetect/src/hog.cpp:239-550
//   cv::HOGDescriptor::compute                      modules/objdetect/src/hog.cpp:1589-1653
//   cv::HOGDescriptor::detect                       modules/objdetect/src/hog.cpp:1655-1767
//   cv::HOGDescriptor::detectMultiScale             modules/objdetect/src/hog.cpp:2051-2105
//   cv::groupRectangles (free function)             modules/objdetect/src/cascadedetect.cpp:366-371
//   HOGDescriptor::getDefaultPeopleDetector / getDaimlerPeopleDetector   hog.cpp:2174-3487
//   cv::HOGDescriptor::save / load (FileStorage XML / YAML)              modules/objdetect/src/hog.cpp:139-209
#include <opencv2/core.hpp>
#include <opencv2/core/utility.hpp>
#include <opencv2/imgproc.hpp>
#include <opencv2/objdetect.hpp>
#include <cstring>
#include <vector>
#include <string>

using namespace cv;

namespace {
thread_local std::string g_err;

Mat wrap(const unsigned char* p, int w, int h, int cn, long pitch)
{
    return Mat(h, w, CV_8UC(cn), const_cast<unsigned char*>(p), (size_t)pitch);
}

// BGRA (or BGR) -> BGR view material for the CPU HOG, which asserts 8UC1/8UC3
// (objdetect/src/hog.cpp:244).  Alpha is dropped, as the GPU kernel ignores .w.
Mat toCpuHogInput(const Mat& m)
{
    if (m.channels() == 4) {
        Mat bgr(m.rows, m.cols, CV_8UC3);
        int fromTo[] = {0,0, 1,1, 2,2};
        mixChannels(&m, 1, &bgr, 1, fromTo, 3);
        return bgr;
    }
    return m;
}

struct HogParams {
    int win_w, win_h, block_w, block_h, stride_w, stride_h, cell_w, cell_h, nbins;
    double win_sigma, l2hys; int gamma;
};

HOGDescriptor makeHog(const int* geom /*9 ints*/, double win_sigma, double l2hys, int gamma, int nlevels)
{
    return HOGDescriptor(Size(geom[0], geom[1]), Size(geom[2], geom[3]), Size(geom[4], geom[5]),
                         Size(geom[6], geom[7]), geom[8], 1, win_sigma, HOGDescriptor::L2Hys,
                         l2hys, gamma != 0, nlevels, false);
}
} // namespace

#define REF_TRY try {
#define REF_CATCH } catch (const cv::Exception& e) { g_err = e.what(); return -1; } \
                    catch (const std::exception& e) { g_err = e.what(); return -1; }

extern "C" {

const char* cvref_last_error() { return g_err.c_str(); }

const char* cvref_version() { return CV_VERSION; }

void cvref_set_num_threads(int n) { cv::setNumThreads(n); }
int  cvref_get_num_threads() { return cv::getNumThreads(); }

// interp: 0 = INTER_LINEAR (the mandated pyramid), 1 = INTER_LINEAR_EXACT (what the CPU HOG uses)
int cvref_resize(const unsigned char* src, int sw, int sh, int cn, long spitch,
                 unsigned char* dst, int dw, int dh, long dpitch, int interp)
{
    REF_TRY
    Mat s = wrap(src, sw, sh, cn, spitch);
    Mat d = wrap(dst, dw, dh, cn, dpitch);
    Mat out = d;
    resize(s, out, Size(dw, dh), 0, 0, interp == 1 ? INTER_LINEAR_EXACT : INTER_LINEAR);
    if (out.data != d.data) out.copyTo(d);
    return 0;
    REF_CATCH
}

// grad: h*w*2 floats, qangle: h*w*2 bytes (no padding)
int cvref_compute_gradient(const int* geom, double win_sigma, double l2hys, int gamma,
                           const unsigned char* img, int w, int h, int cn, long pitch,
                           float* grad, unsigned char* qangle)
{
    REF_TRY
    HOGDescriptor hog = makeHog(geom, win_sigma, l2hys, gamma, 64);
    Mat m = toCpuHogInput(wrap(img, w, h, cn, pitch));
    Mat g, q;
    hog.computeGradient(m, g, q, Size(0, 0), Size(0, 0));
    CV_Assert(g.type() == CV_32FC2 && q.type() == CV_8UC2 && g.size() == m.size());
    for (int y = 0; y < h; y++) {
        memcpy(grad + (size_t)y * w * 2, g.ptr<float>(y), sizeof(float) * w * 2);
        memcpy(qangle + (size_t)y * w * 2, q.ptr<uchar>(y), (size_t)w * 2);
    }
    return 0;
    REF_CATCH
}

// HOGDescriptor::compute with winStride, padding 0.  Returns the number of floats written
// (windows * descriptorSize), or -1.  capacity in floats.
long cvref_compute(const int* geom, double win_sigma, double l2hys, int gamma,
                   const unsigned char* img, int w, int h, int cn, long pitch,
                   int stride_x, int stride_y, float* out, long capacity)
{
    REF_TRY
    HOGDescriptor hog = makeHog(geom, win_sigma, l2hys, gamma, 64);
    Mat m = toCpuHogInput(wrap(img, w, h, cn, pitch));
    std::vector<float> d;
    hog.compute(m, d, Size(stride_x, stride_y), Size(0, 0));
    if ((long)d.size() > capacity) { g_err = "capacity"; return -1; }
    memcpy(out, d.data(), d.size() * sizeof(float));
    return (long)d.size();
    REF_CATCH
}

// Single-scale HOGDescriptor::detect.  Writes (x,y) pairs and weights; returns count.
long cvref_detect(const int* geom, double win_sigma, double l2hys, int gamma,
                  const float* svm, int nsvm,
                  const unsigned char* img, int w, int h, int cn, long pitch,
                  double hit_threshold, int stride_x, int stride_y,
                  int* xy, double* weights, long capacity)
{
    REF_TRY
    HOGDescriptor hog = makeHog(geom, win_sigma, l2hys, gamma, 64);
    hog.setSVMDetector(std::vector<float>(svm, svm + nsvm));
    Mat m = toCpuHogInput(wrap(img, w, h, cn, pitch));
    std::vector<Point> loc; std::vector<double> wts;
    hog.detect(m, loc, wts, hit_threshold, Size(stride_x, stride_y), Size(0, 0));
    if ((long)loc.size() > capacity) { g_err = "capacity"; return -1; }
    for (size_t i = 0; i < loc.size(); i++) { xy[2*i] = loc[i].x; xy[2*i+1] = loc[i].y; weights[i] = wts[i]; }
    return (long)loc.size();
    REF_CATCH
}

// Stock HOGDescriptor::detectMultiScale (INTER_LINEAR_EXACT pyramid, member groupRectangles,
// clipObjects) -- the timed CPU baseline of record.  rects: x,y,w,h ints.
long cvref_detect_multiscale(const int* geom, double win_sigma, double l2hys, int gamma, int nlevels,
                             const float* svm, int nsvm,
                             const unsigned char* img, int w, int h, int cn, long pitch,
                             double hit_threshold, int stride_x, int stride_y,
                             double scale0, double final_threshold,
                             int* rects, double* weights, long capacity)
{
    REF_TRY
    HOGDescriptor hog = makeHog(geom, win_sigma, l2hys, gamma, nlevels);
    hog.setSVMDetector(std::vector<float>(svm, svm + nsvm));
    Mat m = toCpuHogInput(wrap(img, w, h, cn, pitch));
    std::vector<Rect> found; std::vector<double> wts;
    hog.detectMultiScale(m, found, wts, hit_threshold, Size(stride_x, stride_y), Size(0, 0),
                         scale0, final_threshold, false);
    if ((long)found.size() > capacity) { g_err = "capacity"; return -1; }
    for (size_t i = 0; i < found.size(); i++) {
        rects[4*i] = found[i].x; rects[4*i+1] = found[i].y; rects[4*i+2] = found[i].width; rects[4*i+3] = found[i].height;
        if (weights) weights[i] = i < wts.size() ? wts[i] : 0.0;
    }
    return (long)found.size();
    REF_CATCH
}

// Free function cv::groupRectangles(rectList, groupThreshold, eps) -- what cv::cuda::HOG calls
// (modules/cudaobjdetect/src/hog.cpp:486).  In place; returns new count.
long cvref_group_rectangles(int* rects, long n, int group_threshold, double eps)
{
    REF_TRY
    std::vector<Rect> v((size_t)n);
    for (long i = 0; i < n; i++) v[i] = Rect(rects[4*i], rects[4*i+1], rects[4*i+2], rects[4*i+3]);
    groupRectangles(v, group_threshold, eps);
    for (size_t i = 0; i < v.size(); i++) {
        rects[4*i] = v[i].x; rects[4*i+1] = v[i].y; rects[4*i+2] = v[i].width; rects[4*i+3] = v[i].height;
    }
    return (long)v.size();
    REF_CATCH
}

// cv::groupRectangles_meanshift(rectList, foundWeights, foundScales, detectThreshold, winDetSize)
// (modules/objdetect/src/cascadedetect.cpp:387-394) -- the grouping HOGDescriptor::detectMultiScale uses with
// useMeanshiftGrouping (objdetect/src/hog.cpp:2100-2101).  In place; returns the new count.
long cvref_group_rectangles_meanshift(int* rects, double* weights, const double* scales, long n,
                                      double detect_threshold, int win_w, int win_h)
{
    REF_TRY
    std::vector<Rect> v((size_t)n); std::vector<double> w(weights, weights + n), sc(scales, scales + n);
    for (long i = 0; i < n; i++) v[i] = Rect(rects[4*i], rects[4*i+1], rects[4*i+2], rects[4*i+3]);
    groupRectangles_meanshift(v, w, sc, detect_threshold, Size(win_w, win_h));
    for (size_t i = 0; i < v.size(); i++) {
        rects[4*i] = v[i].x; rects[4*i+1] = v[i].y; rects[4*i+2] = v[i].width; rects[4*i+3] = v[i].height;
        weights[i] = w[i];
    }
    return (long)v.size();
    REF_CATCH
}

// HOGDescriptor::detectROI (modules/objdetect/src/hog.cpp:3549-3660): scores the requested window origins of ONE image.
// confidences gets one entry per location (-10 for a location outside the padded image); returns the number of hits.
long cvref_detect_roi(const int* geom, double win_sigma, double l2hys, int gamma, const float* svm, int nsvm,
                      const unsigned char* img, int w, int h, int cn, long pitch,
                      const int* locations, long nloc, double hit_threshold, int pad_w, int pad_h,
                      int* found_xy, double* confidences)
{
    REF_TRY
    HOGDescriptor hog = makeHog(geom, win_sigma, l2hys, gamma, 64);
    hog.setSVMDetector(std::vector<float>(svm, svm + nsvm));
    Mat m = toCpuHogInput(wrap(img, w, h, cn, pitch));
    std::vector<Point> loc((size_t)nloc), found; std::vector<double> conf;
    for (long i = 0; i < nloc; i++) loc[i] = Point(locations[2*i], locations[2*i+1]);
    hog.detectROI(m, loc, found, conf, hit_threshold, Size(), Size(pad_w, pad_h));
    for (size_t i = 0; i < conf.size(); i++) confidences[i] = conf[i];
    for (size_t i = 0; i < found.size(); i++) { found_xy[2*i] = found[i].x; found_xy[2*i+1] = found[i].y; }
    return (long)found.size();
    REF_CATCH
}

// HOGDescriptor::detectMultiScaleROI (hog.cpp:3662-3677), unmodified (INTER_LINEAR_EXACT levels, padding 8x8,
// groupRectangles(found, groupThreshold, 0.2)).  scales[nroi]; the locations of roi i are
// locations[2*off[i] .. 2*off[i+1]); confidences (same indexing) receives every location's score.
long cvref_detect_multiscale_roi(const int* geom, double win_sigma, double l2hys, int gamma, const float* svm, int nsvm,
                                 const unsigned char* img, int w, int h, int cn, long pitch,
                                 const double* scales, const long* off, long nroi, const int* locations,
                                 double hit_threshold, int group_threshold,
                                 int* rects, long capacity, double* confidences)
{
    REF_TRY
    HOGDescriptor hog = makeHog(geom, win_sigma, l2hys, gamma, 64);
    hog.setSVMDetector(std::vector<float>(svm, svm + nsvm));
    Mat m = toCpuHogInput(wrap(img, w, h, cn, pitch));
    std::vector<DetectionROI> rois((size_t)nroi);
    for (long i = 0; i < nroi; i++) {
        rois[i].scale = scales[i];
        for (long j = off[i]; j < off[i + 1]; j++) rois[i].locations.push_back(Point(locations[2*j], locations[2*j+1]));
    }
    std::vector<Rect> found;
    hog.detectMultiScaleROI(m, found, rois, hit_threshold, group_threshold);
    for (long i = 0; i < nroi; i++)
        for (size_t j = 0; j < rois[i].confidences.size(); j++) confidences[off[i] + j] = rois[i].confidences[j];
    if ((long)found.size() > capacity) { g_err = "capacity"; return -1; }
    for (size_t i = 0; i < found.size(); i++) {
        rects[4*i] = found[i].x; rects[4*i+1] = found[i].y; rects[4*i+2] = found[i].width; rects[4*i+3] = found[i].height;
    }
    return (long)found.size();
    REF_CATCH
}

// which: 0 = 64x128 INRIA default people detector (3781), 1 = 48x96 Daimler (1981)
long cvref_people_detector(int which, float* out, long capacity)
{
    REF_TRY
    std::vector<float> d = which == 1 ? HOGDescriptor::getDaimlerPeopleDetector()
                                      : HOGDescriptor::getDefaultPeopleDetector();
    if ((long)d.size() > capacity) { g_err = "capacity"; return -1; }
    memcpy(out, d.data(), d.size() * sizeof(float));
    return (long)d.size();
    REF_CATCH
}

// cv::RNG fill, as the reference's Strict test builds its inputs (test_cascadeandhog.cpp:1335):
// uniform [0,256) per byte.
int cvref_rng_fill(unsigned long long seed, unsigned char* img, int w, int h, int cn, long pitch)
{
    REF_TRY
    RNG rng(seed);
    Mat m = wrap(img, w, h, cn, pitch);
    rng.fill(m, RNG::UNIFORM, 0, 256);
    return 0;
    REF_CATCH
}

// cv::HOGDescriptor::save (objdetect/src/hog.cpp:205-209): "opencv-object-detector-hog" node incl. SVMDetector
int cvref_hog_save(const char* path, const char* objname, const int* geom, double win_sigma, double l2hys, int gamma,
                   int nlevels, const float* det, long ndet)
{
    REF_TRY
    HOGDescriptor hog = makeHog(geom, win_sigma, l2hys, gamma, nlevels);
    if (det && ndet > 0) hog.svmDetector.assign(det, det + ndet);
    hog.save(path, objname ? objname : "");
    return 0;
    REF_CATCH
}

// cv::HOGDescriptor::load (hog.cpp:198-203).  Returns the detector length (0 = none), -1 on error.
long cvref_hog_load(const char* path, const char* objname, int* geom, double* win_sigma, double* l2hys, int* gamma,
                    int* nlevels, float* det, long capacity)
{
    REF_TRY
    HOGDescriptor hog;
    if (!hog.load(path, objname ? objname : "")) { g_err = "load returned false"; return -1; }
    geom[0] = hog.winSize.width; geom[1] = hog.winSize.height; geom[2] = hog.blockSize.width; geom[3] = hog.blockSize.height;
    geom[4] = hog.blockStride.width; geom[5] = hog.blockStride.height; geom[6] = hog.cellSize.width; geom[7] = hog.cellSize.height;
    geom[8] = hog.nbins;
    *win_sigma = hog.winSigma; *l2hys = hog.L2HysThreshold; *gamma = hog.gammaCorrection ? 1 : 0; *nlevels = hog.nlevels;
    if ((long)hog.svmDetector.size() > capacity) { g_err = "capacity"; return -1; }
    if (!hog.svmDetector.empty()) memcpy(det, hog.svmDetector.data(), hog.svmDetector.size() * sizeof(float));
    return (long)hog.svmDetector.size();
    REF_CATCH
}

} // extern "C"
