// ref_cuda_driver.cu -- times the reference's own CUDA HOG kernels (compiled by baseline/build_ref_cuda.sh from
// /root/reference/modules/cudaobjdetect/src/cuda/hog.cu) on this GPU, sequenced as HOG_Impl::computeBlockHistograms +
// HOG_Impl::detect do per pyramid level (modules/cudaobjdetect/src/hog.cpp:336-413, 543-595): set_up_constants,
// compute_gradients_8UC4, compute_hists, normalize_hists, classify_hists.  The resize (texture references, does not
// compile), the label download and the CPU scan are left out, so the figure is a LOWER bound of the reference's time.
// Comparator only: nothing here is part of the product.
#include <cuda_runtime.h>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "opencv2/core/cuda/common.hpp"

namespace cv { namespace cuda { namespace device { namespace hog {
    void set_up_constants(int nbins, int block_stride_x, int block_stride_y, int nblocks_win_x, int nblocks_win_y,
                          int ncells_block_x, int ncells_block_y, const cudaStream_t& stream);
    void compute_hists(int nbins, int block_stride_x, int block_stride_y, int height, int width, const PtrStepSzf& grad,
                       const PtrStepSzb& qangle, float sigma, float* block_hists, int cell_size_x, int cell_size_y,
                       int ncells_block_x, int ncells_block_y, const cudaStream_t& stream);
    void normalize_hists(int nbins, int block_stride_x, int block_stride_y, int height, int width, float* block_hists,
                         float threshold, int cell_size_x, int cell_size_y, int ncells_block_x, int ncells_block_y,
                         const cudaStream_t& stream);
    void classify_hists(int win_height, int win_width, int block_stride_y, int block_stride_x, int win_stride_y, int win_stride_x,
                        int height, int width, float* block_hists, float* coefs, float free_coef, float threshold, int cell_size_x,
                        int ncells_block_x, unsigned char* labels);
    void compute_gradients_8UC4(int nbins, int height, int width, const PtrStepSzb& img, float angle_scale, PtrStepSzf grad,
                                PtrStepSzb qangle, bool correct_gamma, const cudaStream_t& stream);
}}}}
using namespace cv::cuda;
namespace hog = cv::cuda::device::hog;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

struct Level { int w, h; };

int main(int argc, char** argv)
{
    const int W = argc > 2 ? atoi(argv[1]) : 1920, H = argc > 2 ? atoi(argv[2]) : 1080;
    const double scale0 = 1.05; const int nlevels = 64;
    std::vector<Level> lv;                      // HOG_Impl::detectMultiScale level rule (hog.cpp:424-441,453)
    double scale = 1.0;
    for (int i = 0; i < nlevels; i++) {
        lv.push_back(Level{(int)lrint(W / scale), (int)lrint(H / scale)});
        if (lrint(W / scale) < 64 || lrint(H / scale) < 128 || scale0 <= 1) break;
        scale *= scale0;
    }
    if (lv.size() > 1 && (lv.back().w < 64 || lv.back().h < 128)) lv.pop_back();
    unsigned char *img, *qangle, *labels; float *grad, *hists, *coefs;
    size_t ipitch, gpitch, qpitch;
    CK(cudaMallocPitch((void**)&img, &ipitch, (size_t)W * 4, H));
    CK(cudaMallocPitch((void**)&grad, &gpitch, (size_t)W * 8, H));
    CK(cudaMallocPitch((void**)&qangle, &qpitch, (size_t)W * 2, H));
    CK(cudaMalloc((void**)&hists, (size_t)(W / 8) * (H / 8) * 36 * 4));
    CK(cudaMalloc((void**)&labels, (size_t)(W / 8) * (H / 8)));
    CK(cudaMalloc((void**)&coefs, 3780 * 4));
    std::vector<unsigned char> h((size_t)ipitch * H);
    srand(1);
    for (auto& b : h) b = (unsigned char)(rand() >> 7);
    CK(cudaMemcpy(img, h.data(), h.size(), cudaMemcpyHostToDevice));
    std::vector<float> hc(3780);
    for (auto& c : hc) c = (rand() % 2001 - 1000) * 1e-4f;
    CK(cudaMemcpy(coefs, hc.data(), 3780 * 4, cudaMemcpyHostToDevice));
    cudaEvent_t ev[6];
    for (int i = 0; i < 6; i++) CK(cudaEventCreate(&ev[i]));
    const float angle_scale = (float)(9 / M_PI);
    double tot[5] = {0, 0, 0, 0, 0}, lvl0[5] = {0, 0, 0, 0, 0};
    const int reps = 5;
    for (int rep = -1; rep < reps; rep++)       // rep -1 = warm-up
        for (size_t i = 0; i < lv.size(); i++) {
            const int w = lv[i].w, hh = lv[i].h;        // the level image: the top-left w x h of the buffer (content is irrelevant to the time)
            PtrStepSzb I(hh, w * 4, img, ipitch), Q(hh, w * 2, qangle, qpitch);
            PtrStepSzf G(hh, w * 2, grad, gpitch);
            I.cols = w; Q.cols = w; G.cols = w;
            CK(cudaEventRecord(ev[0]));
            hog::set_up_constants(9, 8, 8, 7, 15, 2, 2, 0);
            CK(cudaEventRecord(ev[1]));
            hog::compute_gradients_8UC4(9, hh, w, I, angle_scale, G, Q, true, 0);
            CK(cudaEventRecord(ev[2]));
            hog::compute_hists(9, 8, 8, hh, w, G, Q, 4.f, hists, 8, 8, 2, 2, 0);
            CK(cudaEventRecord(ev[3]));
            hog::normalize_hists(9, 8, 8, hh, w, hists, 0.2f, 8, 8, 2, 2, 0);
            CK(cudaEventRecord(ev[4]));
            hog::classify_hists(128, 64, 8, 8, 8, 8, hh, w, hists, coefs, -6.6657f, 0.f, 8, 2, labels);
            CK(cudaEventRecord(ev[5]));
            CK(cudaEventSynchronize(ev[5]));
            if (rep < 0) continue;
            for (int s = 0; s < 5; s++) {
                float ms; CK(cudaEventElapsedTime(&ms, ev[s], ev[s + 1]));
                tot[s] += ms; if (i == 0) lvl0[s] += ms;
            }
        }
    const char* names[5] = {"set_up_constants (11 cudaMemcpyToSymbolAsync)", "compute_gradients_8UC4", "compute_hists", "normalize_hists", "classify_hists (+ its cudaDeviceSynchronize)"};
    double sum = 0, sum0 = 0;
    printf("reference CUDA HOG kernels (hog.cu, sm_100a), %dx%d 8UC4, %zu levels, mean of %d passes\n", W, H, lv.size(), reps);
    printf("%-48s %12s %12s\n", "stage", "level 0 us", "all levels us");
    for (int s = 0; s < 5; s++) { printf("%-48s %12.1f %12.1f\n", names[s], 1e3 * lvl0[s] / reps, 1e3 * tot[s] / reps); sum += tot[s]; sum0 += lvl0[s]; }
    printf("%-48s %12.1f %12.1f   (no resize, no label download / CPU scan: a lower bound)\n", "sum", 1e3 * sum0 / reps, 1e3 * sum / reps);
    printf("frames/s bound from the kernels alone: %.1f\n", 1e3 / (sum / reps));
    return 0;
}
