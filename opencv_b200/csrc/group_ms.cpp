// group_ms.cpp -- cv::groupRectangles_meanshift (modules/objdetect/src/cascadedetect.cpp:176-395), the grouping that
// cv::HOGDescriptor::detectMultiScale switches to with useMeanshiftGrouping (objdetect/src/hog.cpp:1908-1911): every hit
// becomes a point (centre x, centre y, log scale), each point climbs to a mode of the weighted kernel density
// (bandwidth 8 x 16 pixels scaled by e^z, log 1.3 in scale), modes closer than 1 (in bandwidth units) are merged in
// first-come order, and a mode survives if its density exceeds the detection threshold.  Host code, double arithmetic in
// the reference's operation order (the outputs are truncated to ints, so the order is part of the contract).
// SURVEY.md section 8 row f4.
#include "../../include/hogb200.h"

#include <cmath>
#include <vector>

namespace {

struct P3 { double x, y, z; };

struct Density {
    const std::vector<P3>& pos; const std::vector<double>& wt; P3 band;

    // bandwidth of sample a: the x/y kernel widths grow with the sample's scale (cascadedetect.cpp:246-247)
    P3 width(const P3& a) const { const double e = std::exp(a.z); return P3{band.x * e, band.y * e, band.z}; }

    // one mean-shift step from q (getNewValue, :239-273)
    P3 step(const P3& q) const
    {
        P3 num{0, 0, 0}, den{0, 0, 0};
        for (size_t i = 0; i < pos.size(); i++) {
            const P3 s = width(pos[i]);
            const P3 a{pos[i].x / s.x, pos[i].y / s.y, pos[i].z / s.z}, b{q.x / s.x, q.y / s.y, q.z / s.z};
            const P3 d{a.x - b.x, a.y - b.y, a.z - b.z};
            const double w = wt[i] * std::exp(-(d.x * d.x + d.y * d.y + d.z * d.z) / 2) / std::sqrt(s.x + s.y + s.z);
            num.x += w * a.x; num.y += w * a.y; num.z += w * a.z;
            den.x += w / s.x; den.y += w / s.y; den.z += w / s.z;
        }
        return P3{num.x / den.x, num.y / den.y, num.z / den.z};
    }

    // kernel density at q (getResultWeight, :275-295)
    double at(const P3& q) const
    {
        double sum = 0;
        for (size_t i = 0; i < pos.size(); i++) {
            const P3 s = width(pos[i]);
            const P3 d{(pos[i].x - q.x) / s.x, (pos[i].y - q.y) / s.y, (pos[i].z - q.z) / s.z};
            sum += wt[i] * std::exp(-(d.x * d.x + d.y * d.y + d.z * d.z) / 2) / std::sqrt(s.x + s.y + s.z);
        }
        return sum;
    }

    // squared distance in units of the bandwidth AT b (getDistance, :312-322)
    double dist2(const P3& a, const P3& b) const
    {
        const P3 s = width(b);
        const P3 d{(b.x - a.x) / s.x, (b.y - a.y) / s.y, (b.z - a.z) / s.z};
        return d.x * d.x + d.y * d.y + d.z * d.z;
    }
};

}  // namespace

extern "C" int hogb200_group_rectangles_meanshift(hogb200_rect* rects, double* weights, const double* scales, size_t n,
                                                  double detect_threshold, int win_w, int win_h, size_t* count)
{
    if (count) *count = 0;
    if (n && (!rects || !weights || !scales)) return HOGB200_BAD_ARG;
    std::vector<P3> pos(n); std::vector<double> wt(weights, weights + n);
    for (size_t i = 0; i < n; i++) {
        // centre = (tl + br) * 0.5 evaluated on INTEGER points: Point_<int> * double rounds each component back to int
        // with cvRound (core/types.hpp:1336-1339) before it is widened to a Point2d (:335)
        const int sx = rects[i].x + (rects[i].x + rects[i].width), sy = rects[i].y + (rects[i].y + rects[i].height);
        pos[i] = P3{(double)lrint(sx * 0.5), (double)lrint(sy * 0.5), std::log(scales[i])};
    }
    const Density D{pos, wt, P3{8, 16, std::log(1.3)}};
    // every sample climbs: one step, then up to 100 more until a step moves less than 1e-5 (:194-199, :297-310)
    std::vector<P3> modes;
    for (size_t i = 0; i < n; i++) {
        P3 a = D.step(pos[i]);
        for (int it = 0; it < 100; it++) {
            const P3 b = a;
            a = D.step(b);
            if (D.dist2(a, b) <= 1e-5) break;
        }
        bool known = false;                       // getModes (:201-227): first-come modes, merge radius 1
        for (size_t j = 0; j < modes.size() && !known; j++) known = D.dist2(a, modes[j]) < 1;
        if (!known) modes.push_back(a);
    }
    size_t m = 0;
    for (const P3& q : modes) {
        const double dens = D.at(q);
        const double scale = std::exp(q.z);
        const int w = (int)(win_w * scale), h = (int)(win_h * scale);
        if (dens > detect_threshold) {            // never more modes than samples: the output fits the input arrays
            rects[m] = hogb200_rect{(int)(q.x - w / 2), (int)(q.y - h / 2), w, h};
            weights[m] = dens;
            m++;
        }
    }
    if (count) *count = m;
    return HOGB200_OK;
}
