// hogb200.hpp -- header-only C++ face of the C ABI (hogb200.h, tbdb200.h) with the reference's names, so that C++
// callers of cv::cuda::HOG (modules/cudaobjdetect/include/opencv2/cudaobjdetect.hpp:78-212) and cv::tbd::Tracker
// (modules/trackingbydetection/include/opencv2/tbd.hpp:131-184) port by changing a namespace.  No OpenCV types:
// Size / Rect / Point are plain structs with the reference's member names, Image stands in for cv::cuda::GpuMat /
// cv::Mat (pointer, pitch, size, channels, where it lives).  Errors throw hogb200::Error where the reference throws
// cv::Exception.  examples/tbd_loop.cpp is example_gpu_tbd's frame loop (samples/gpu/tbd.cpp:568-700) on top of it.
#ifndef HOGB200_HPP
#define HOGB200_HPP

#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "hogb200.h"
#include "tbdb200.h"

namespace hogb {

struct Size { int width, height; };
typedef hogb200_rect Rect;        // x, y, width, height
typedef hogb200_point Point;      // x, y

struct Error : std::runtime_error {
    int status;
    Error(int s, const std::string& m) : std::runtime_error(m), status(s) {}
};

// a frame: host memory (cv::Mat) or device memory (cv::cuda::GpuMat), CV_8UC1 / CV_8UC3 (B,G,R) / CV_8UC4
struct Image {
    const void* data; size_t step; int cols, rows, channels; bool on_device;
    hogb200_image c() const { hogb200_image i = {data, step, cols, rows, channels, on_device ? 1 : 0}; return i; }
};

class HOG {
public:
    enum { DESCR_FORMAT_ROW_BY_ROW = HOGB200_DESCR_FORMAT_ROW_BY_ROW, DESCR_FORMAT_COL_BY_COL = HOGB200_DESCR_FORMAT_COL_BY_COL };

    // cv::cuda::HOG::create (cudaobjdetect.hpp:92-96)
    static std::shared_ptr<HOG> create(Size win_size = Size{64, 128}, Size block_size = Size{16, 16},
                                       Size block_stride = Size{8, 8}, Size cell_size = Size{8, 8}, int nbins = 9,
                                       int device = 0)
    {
        hogb200_geom g = {win_size.width, win_size.height, block_size.width, block_size.height, block_stride.width,
                          block_stride.height, cell_size.width, cell_size.height, nbins};
        hogb200_t* h = nullptr;
        check(hogb200_create(&g, device, &h));
        return std::shared_ptr<HOG>(new HOG(h, win_size));
    }
    // Algorithm::load<HOG>(file) (hog.cpp:661-679)
    static std::shared_ptr<HOG> load(const std::string& filename, const std::string& objname = std::string(), int device = 0)
    {
        hogb200_t* h = nullptr;
        check(hogb200_create_from_file(filename.c_str(), objname.empty() ? nullptr : objname.c_str(), device, &h));
        return std::shared_ptr<HOG>(new HOG(h, Size{0, 0}));
    }
    ~HOG() { hogb200_destroy(h_); }

    void setWinSigma(double v) { check(hogb200_set_win_sigma(h_, v)); }
    double getWinSigma() const { return hogb200_get_win_sigma(h_); }
    void setL2HysThreshold(double v) { check(hogb200_set_l2hys_threshold(h_, v)); }
    double getL2HysThreshold() const { return hogb200_get_l2hys_threshold(h_); }
    void setGammaCorrection(bool v) { check(hogb200_set_gamma_correction(h_, v)); }
    bool getGammaCorrection() const { return hogb200_get_gamma_correction(h_) != 0; }
    void setNumLevels(int v) { check(hogb200_set_num_levels(h_, v)); }
    int getNumLevels() const { return hogb200_get_num_levels(h_); }
    void setHitThreshold(double v) { check(hogb200_set_hit_threshold(h_, v)); }
    double getHitThreshold() const { return hogb200_get_hit_threshold(h_); }
    void setWinStride(Size s) { check(hogb200_set_win_stride(h_, s.width, s.height)); }
    Size getWinStride() const { Size s; hogb200_get_win_stride(h_, &s.width, &s.height); return s; }
    void setScaleFactor(double v) { check(hogb200_set_scale_factor(h_, v)); }
    double getScaleFactor() const { return hogb200_get_scale_factor(h_); }
    void setGroupThreshold(int v) { check(hogb200_set_group_threshold(h_, v)); }
    int getGroupThreshold() const { return hogb200_get_group_threshold(h_); }
    void setDescriptorFormat(int v) { check(hogb200_set_descriptor_format(h_, v)); }
    int getDescriptorFormat() const { return hogb200_get_descriptor_format(h_); }
    size_t getDescriptorSize() const { return hogb200_get_descriptor_size(h_); }
    size_t getBlockHistogramSize() const { return hogb200_get_block_histogram_size(h_); }
    void setStream(void* cuda_stream) { check(hogb200_set_stream(h_, cuda_stream)); }

    // setSVMDetector (hog.cpp:293-321); slot > 0: further detectors scored in the same pass
    void setSVMDetector(const std::vector<float>& detector, int slot = 0)
    {
        check(hogb200_set_svm_detector(h_, slot, detector.data(), detector.size()));
    }
    // getDefaultPeopleDetector (hog.cpp:326-334)
    std::vector<float> getDefaultPeopleDetector() const
    {
        const float* p = nullptr; size_t n = 0;
        check(hogb200_default_people_detector(win_.width, win_.height, &p, &n));
        return std::vector<float>(p, p + n);
    }

    // detect (hog.cpp:336-413)
    void detect(const Image& img, std::vector<Point>& found_locations, std::vector<double>* confidences = nullptr)
    {
        const hogb200_image im = img.c();
        grow(found_locations, confidences, [&](Point* p, double* c, size_t cap, size_t* n) {
            return hogb200_detect(h_, &im, p, c, cap, n);
        });
    }
    // detectMultiScale (hog.cpp:415-488)
    void detectMultiScale(const Image& img, std::vector<Rect>& found_locations, std::vector<double>* confidences = nullptr,
                          int slot = 0)
    {
        const hogb200_image im = img.c();
        grow(found_locations, confidences, [&](Rect* p, double* c, size_t cap, size_t* n) {
            return hogb200_detect_multiscale_slot(h_, slot, &im, p, c, cap, n);
        });
    }
    // the sample's frame loop in one call: found[f * nslots + d] = boxes of frame f, detector d
    void detectMultiScaleBatch(const std::vector<Image>& frames, std::vector<std::vector<Rect> >& found, int nslots = 1)
    {
        std::vector<hogb200_image> im;
        for (const Image& f : frames) im.push_back(f.c());
        std::vector<size_t> offs(frames.size() * nslots + 1);
        for (;;) {
            const int rc = hogb200_detect_multiscale_batch(h_, im.data(), (int)im.size(), nslots, rects_.data(), rects_.size(), offs.data());
            if (rc == HOGB200_CAPACITY) { rects_.resize(rects_.size() * 8); continue; }
            check(rc);
            break;
        }
        found.assign(frames.size() * nslots, std::vector<Rect>());
        for (size_t i = 0; i + 1 < offs.size(); i++) found[i].assign(rects_.begin() + offs[i], rects_.begin() + offs[i + 1]);
    }
    // the same batch asynchronously (two may be in flight): submit batch k + 1, then collect batch k, so that the uploads
    // and kernels of one batch overlap the tail of the other.  The frames' pixels must stay valid until collect.
    int submitBatch(const std::vector<Image>& frames, int nslots = 1)
    {
        std::vector<hogb200_image> im;
        for (const Image& f : frames) im.push_back(f.c());
        int ticket = -1;
        check(hogb200_submit_batch(h_, im.data(), (int)im.size(), nslots, &ticket));
        inflight_[ticket] = frames.size() * nslots;
        return ticket;
    }
    void collectBatch(int ticket, std::vector<std::vector<Rect> >& found)
    {
        const size_t nseg = ticket >= 0 && ticket < 2 ? inflight_[ticket] : 0;
        std::vector<size_t> offs(nseg + 1);
        for (;;) {
            const int rc = hogb200_collect_batch(h_, ticket, rects_.data(), rects_.size(), offs.data());
            if (rc == HOGB200_CAPACITY) { rects_.resize(rects_.size() * 8); continue; }     // the library keeps the results
            check(rc);
            break;
        }
        found.assign(nseg, std::vector<Rect>());
        for (size_t i = 0; i + 1 < offs.size(); i++) found[i].assign(rects_.begin() + offs[i], rects_.begin() + offs[i + 1]);
    }
    // compute (hog.cpp:490-534): descriptors of every window, getDescriptorSize() floats each
    void compute(const Image& img, std::vector<float>& descriptors)
    {
        const hogb200_image im = img.c();
        size_t nwin = 0;
        int rc = hogb200_compute(h_, &im, descriptors.data(), descriptors.size(), 0, &nwin);
        if (rc == HOGB200_CAPACITY) {
            descriptors.resize(nwin * getDescriptorSize());
            rc = hogb200_compute(h_, &im, descriptors.data(), descriptors.size(), 0, &nwin);
        }
        check(rc);
        descriptors.resize(nwin * getDescriptorSize());
    }
    void save(const std::string& filename, const std::string& objname = std::string(), int slot = 0) const
    {
        check(hogb200_save(h_, filename.c_str(), objname.empty() ? nullptr : objname.c_str(), slot));
    }

    hogb200_t* handle() { return h_; }

private:
    HOG(hogb200_t* h, Size win) : h_(h), win_(win), rects_(1 << 14) {}
    HOG(const HOG&); HOG& operator=(const HOG&);
    static void check(int rc)
    {
        if (rc != HOGB200_OK) throw Error(rc, hogb200_last_error());
    }
    template <class T, class F>
    static void grow(std::vector<T>& out, std::vector<double>* conf, F call)
    {
        size_t cap = out.capacity() < 256 ? 256 : out.capacity(), n = 0;
        for (;;) {
            out.resize(cap);
            if (conf) conf->resize(cap);
            const int rc = call(out.data(), conf ? conf->data() : nullptr, cap, &n);
            if (rc == HOGB200_CAPACITY) { cap = n > cap ? n : cap * 2; continue; }
            check(rc);
            break;
        }
        out.resize(n);
        if (conf) conf->resize(n);
    }
    hogb200_t* h_; Size win_;
    std::vector<Rect> rects_;
    size_t inflight_[2] = {0, 0};        // output segments of the batches in flight, by ticket
};

// ---- cv::tbd (tbd.hpp:25-184)
namespace tbd {

typedef tbdb200_args TbdArgs;            // costOfNonAssignment ... shouldStoreMetrics, in the reference's order
typedef tbdb200_track Track;
typedef tbdb200_frame_metrics FrameMetrics;

struct Detection : tbdb200_detection {
    Detection(int id_, int frame_id_, const Rect& bbox, double confidence_)
    {
        id = id_; frame_id = frame_id_; x = bbox.x; y = bbox.y; width = bbox.width; height = bbox.height; confidence = confidence_;
    }
};

class Tracker {
public:
    explicit Tracker(const TbdArgs* args) : t_(nullptr)
    {
        if (tbdb200_create(args, &t_)) throw Error(1, tbdb200_last_error());
    }
    ~Tracker() { tbdb200_destroy(t_); }
    void reset() { check(tbdb200_reset(t_)); }
    void setBounds(int xmin, int xmax, int ymin, int ymax) { check(tbdb200_set_bounds(t_, xmin, xmax, ymin, ymax)); }
    // performTrackingStep (tbd.cpp:211-298); returns the track id of every detection (-1: started a new track)
    std::vector<int> performTrackingStep(const std::vector<Detection>& detections, int frame_id)
    {
        std::vector<tbdb200_detection> d(detections.begin(), detections.end());
        std::vector<int> track_of(d.size());
        check(tbdb200_step(t_, d.data(), (int)d.size(), frame_id, track_of.data()));
        return track_of;
    }
    std::vector<Track> getTracks() const
    {
        std::vector<Track> out(tbdb200_num_tracks(t_));
        int n = 0;
        check(tbdb200_get_tracks(t_, out.data(), (int)out.size(), &n));
        return out;
    }
    std::vector<FrameMetrics> metrics() const
    {
        int n = 0;
        check(tbdb200_metrics(t_, nullptr, 0, &n));
        std::vector<FrameMetrics> out(n);
        check(tbdb200_metrics(t_, out.data(), n, &n));
        return out;
    }
    tbdb200_t* handle() { return t_; }

private:
    Tracker(const Tracker&); Tracker& operator=(const Tracker&);
    static void check(int rc) { if (rc) throw Error(rc, tbdb200_last_error()); }
    tbdb200_t* t_;
};

}  // namespace tbd
}  // namespace hogb

#endif  // HOGB200_HPP
