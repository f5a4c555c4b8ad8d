// tbd_loop.cpp -- example_gpu_tbd's frame loop (samples/gpu/tbd.cpp:556-700) on hogb200.hpp: synthetic frames with a
// few walking silhouettes instead of a video file, the sample's detector settings, one Detection per grouped rect,
// one tracking step per frame.  Prints the tracks alive at the end.
// `tbd_loop --pipelined` runs the same loop on the asynchronous batch API: batches of 3 frames, batch k + 1 submitted before
// batch k is collected (two in flight), tracking steps in frame order -- identical output.
//
//   g++ -std=c++14 -O2 -Iinclude examples/tbd_loop.cpp -o tbd_loop -Lopencv_b200 -lhogb200 -Wl,-rpath,$PWD/opencv_b200
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

#include "hogb200.hpp"

using namespace hogb;

static std::vector<uint8_t> make_frame(int W, int H, int k)
{
    std::vector<uint8_t> f((size_t)W * H * 4);
    unsigned s = 12345u;
    for (int y = 0; y < H; y++)
        for (int x = 0; x < W; x++) {
            s = s * 1664525u + 1013904223u;                                  // fixed texture, same every frame
            uint8_t v = (uint8_t)(96 + ((s >> 24) & 63));
            const int walkers[3][5] = {{80, 150, 7, 1, 150}, {520, 300, -9, -2, 110}, {300, 240, 4, 0, 190}};
            for (const int* w : walkers) {
                const double cx = w[0] + w[2] * k, cy = w[1] + w[3] * k, hh = w[4], ww = hh / 3.0;
                const double dy = (y - cy) / (hh / 2), dx = (x - cx) / (ww / 2);
                if (dx * dx + dy * dy <= 1) v = (uint8_t)(v * 0.2 + 200);
            }
            uint8_t* p = &f[((size_t)y * W + x) * 4];
            p[0] = p[1] = p[2] = v; p[3] = 255;
        }
    return f;
}

int main(int argc, char** argv)
{
    const bool pipelined = argc > 1 && !std::strcmp(argv[1], "--pipelined");
    try {
        const int W = 640, H = 480, nframes = 12;
        std::shared_ptr<HOG> gpu_hog = HOG::create();                       // tbd.cpp:390
        gpu_hog->setSVMDetector(gpu_hog->getDefaultPeopleDetector());      // tbd.cpp:430-436
        gpu_hog->setNumLevels(15); gpu_hog->setHitThreshold(-1.6); gpu_hog->setWinStride(Size{8, 8});
        gpu_hog->setScaleFactor(1.05); gpu_hog->setGroupThreshold(2);       // tbd.cpp:599-604 (threshold lowered: synthetic people)
        tbd::TbdArgs args = {10.0, 16, 4, 0.3, -1.0, 1};
        tbd::Tracker pedestrianTracker(&args);
        pedestrianTracker.setBounds(0, W, 0, H);
        auto track = [&](int frame_id, const std::vector<Rect>& foundRects) {
            std::vector<tbd::Detection> detections;
            for (const Rect& r : foundRects) detections.push_back(tbd::Detection(-1, frame_id, r, -1.0));   // tbd.cpp:614-620
            pedestrianTracker.performTrackingStep(detections, frame_id);    // tbd.cpp:684-688
            std::printf("frame %2d: %zu detections, %zu tracks\n", frame_id, foundRects.size(), pedestrianTracker.getTracks().size());
        };
        if (!pipelined) {
            for (int frame_id = 0; frame_id < nframes; frame_id++) {
                const std::vector<uint8_t> frame = make_frame(W, H, frame_id);
                const Image img = {frame.data(), (size_t)W * 4, W, H, 4, false};
                std::vector<Rect> foundRects;
                gpu_hog->detectMultiScale(img, foundRects);                 // tbd.cpp:605
                track(frame_id, foundRects);
            }
        } else {
            const int B = 3;                                                // frames per batch
            std::vector<std::vector<uint8_t> > pixels;                      // stay valid until their batch is collected
            for (int k = 0; k < nframes; k++) pixels.push_back(make_frame(W, H, k));
            auto submit = [&](int first) {
                std::vector<Image> batch;
                for (int k = first; k < first + B && k < nframes; k++) batch.push_back(Image{pixels[k].data(), (size_t)W * 4, W, H, 4, false});
                return gpu_hog->submitBatch(batch);
            };
            int pending = submit(0);
            for (int first = 0; first < nframes; first += B) {
                const int next = first + B < nframes ? submit(first + B) : -1;       // batch k + 1 goes in before batch k comes out
                std::vector<std::vector<Rect> > found;
                gpu_hog->collectBatch(pending, found);
                for (size_t j = 0; j < found.size(); j++) track(first + (int)j, found[j]);
                pending = next;
            }
        }
        for (const tbd::Track& t : pedestrianTracker.getTracks())
            std::printf("track %u: age %u, visible %u, box (%d, %d, %d, %d)\n", t.id, t.age, t.total_visible_count, t.x, t.y, t.width, t.height);
    } catch (const Error& e) {
        std::fprintf(stderr, "error %d: %s\n", e.status, e.what());
        return 1;
    }
    return 0;
}
