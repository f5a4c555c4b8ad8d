// resize.cu -- image pyramid for the HOG detector: every level is resampled FROM THE BASE FRAME
// (cudaobjdetect/src/hog.cpp:456-468) with the bit-exact semantics of cv::resize(..., INTER_LINEAR) on 8U:
// 11-bit fixed-point separable bilinear with two truncating shifts in the vertical pass
// (imgproc/src/resize.cpp:3609-3706 coefficients, :1491-1542 horizontal, :1577-1603 vertical, :1821-1825 row clip),
// and the silent switch to a 2x2 box average when the level is exactly half the frame (:3523-3554, :2542-2578).
//
// All levels of a frame are produced by ONE launch: the grid is a list of (level, x0, y0) tiles.  The base frame
// (8 MB at 1080p BGRA) stays resident in the 126 MB L2, so HBM sees it once; each level is written once.
// Coefficients come from small per-level tables built on the host in double precision exactly as the reference
// builds them, so the kernel is integer-only.
#include "hog_internal.h"

namespace hogk {

namespace {

constexpr int kRsRows = 8;

__device__ __forceinline__ int clipi(int x, int a, int b) { return x >= a ? (x < b ? x : b - 1) : a; }

__device__ __forceinline__ int vlin(int h0, int h1, int b0, int b1)
{
    // resize.cpp:1594-1601: uchar(( ((b0 * (S0 >> 4)) >> 16) + ((b1 * (S1 >> 4)) >> 16) + 2) >> 2)
    return (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;
}

// 8UC4: one thread = one destination pixel (a 4-byte vector store); tile 32 x 8.
__global__ void __launch_bounds__(256)
resize_c4_kernel(const uint8_t* __restrict__ base, int bpitch, int bw, int bh,
                 const LevelDesc* __restrict__ levels, const ResizeCoef* __restrict__ xtab,
                 const ResizeCoef* __restrict__ ytab, const ResizeItem* __restrict__ items)
{
    const ResizeItem it = items[blockIdx.x];
    const LevelDesc& L = levels[it.level];
    const int x = it.x0 + (threadIdx.x & 31), y = it.y0 + (threadIdx.x >> 5);
    if (x >= L.w || y >= L.h) return;
    uchar4 out;
    if (L.mode == 2) {
        const uchar4* r0 = reinterpret_cast<const uchar4*>(base + (size_t)(2 * y) * bpitch) + 2 * x;
        const uchar4* r1 = reinterpret_cast<const uchar4*>(base + (size_t)(2 * y + 1) * bpitch) + 2 * x;
        uchar4 a = __ldg(r0), b = __ldg(r0 + 1), c = __ldg(r1), d = __ldg(r1 + 1);
        out.x = (a.x + b.x + c.x + d.x + 2) >> 2;
        out.y = (a.y + b.y + c.y + d.y + 2) >> 2;
        out.z = (a.z + b.z + c.z + d.z + 2) >> 2;
        out.w = (a.w + b.w + c.w + d.w + 2) >> 2;
    } else {
        const ResizeCoef cx = xtab[L.xtab_off + x], cy = ytab[L.ytab_off + y];
        const int sy0 = clipi(cy.ofs, 0, bh), sy1 = clipi(cy.ofs + 1, 0, bh);
        const int sx0 = cx.ofs, sx1 = min(cx.ofs + 1, bw - 1);     // weight of sx1 is 0 at the right edge
        const uchar4* r0 = reinterpret_cast<const uchar4*>(base + (size_t)sy0 * bpitch);
        const uchar4* r1 = reinterpret_cast<const uchar4*>(base + (size_t)sy1 * bpitch);
        const uchar4 p00 = __ldg(r0 + sx0), p01 = __ldg(r0 + sx1), p10 = __ldg(r1 + sx0), p11 = __ldg(r1 + sx1);
        const int a0 = cx.c0, a1 = cx.c1, b0 = cy.c0, b1 = cy.c1;
        out.x = vlin(p00.x * a0 + p01.x * a1, p10.x * a0 + p11.x * a1, b0, b1);
        out.y = vlin(p00.y * a0 + p01.y * a1, p10.y * a0 + p11.y * a1, b0, b1);
        out.z = vlin(p00.z * a0 + p01.z * a1, p10.z * a0 + p11.z * a1, b0, b1);
        out.w = vlin(p00.w * a0 + p01.w * a1, p10.w * a0 + p11.w * a1, b0, b1);
    }
    *reinterpret_cast<uchar4*>(const_cast<uint8_t*>(L.img) + (size_t)y * L.pitch + 4 * x) = out;
}

// 8UC1: one thread = four consecutive destination pixels (one 4-byte store); tile 128 x 8.
__global__ void __launch_bounds__(256)
resize_c1_kernel(const uint8_t* __restrict__ base, int bpitch, int bw, int bh,
                 const LevelDesc* __restrict__ levels, const ResizeCoef* __restrict__ xtab,
                 const ResizeCoef* __restrict__ ytab, const ResizeItem* __restrict__ items)
{
    const ResizeItem it = items[blockIdx.x];
    const LevelDesc& L = levels[it.level];
    const int x4 = it.x0 + 4 * (threadIdx.x & 31), y = it.y0 + (threadIdx.x >> 5);
    if (x4 >= L.w || y >= L.h) return;
    uint8_t o[4] = {0, 0, 0, 0};
    if (L.mode == 2) {
        const uint8_t* r0 = base + (size_t)(2 * y) * bpitch;
        const uint8_t* r1 = r0 + bpitch;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            int x = x4 + k;
            if (x < L.w) o[k] = (__ldg(r0 + 2 * x) + __ldg(r0 + 2 * x + 1) + __ldg(r1 + 2 * x) + __ldg(r1 + 2 * x + 1) + 2) >> 2;
        }
    } else {
        const ResizeCoef cy = ytab[L.ytab_off + y];
        const int sy0 = clipi(cy.ofs, 0, bh), sy1 = clipi(cy.ofs + 1, 0, bh);
        const uint8_t* r0 = base + (size_t)sy0 * bpitch;
        const uint8_t* r1 = base + (size_t)sy1 * bpitch;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            int x = x4 + k;
            if (x < L.w) {
                const ResizeCoef cx = xtab[L.xtab_off + x];
                const int sx0 = cx.ofs, sx1 = min(cx.ofs + 1, bw - 1);
                o[k] = vlin(__ldg(r0 + sx0) * cx.c0 + __ldg(r0 + sx1) * cx.c1,
                            __ldg(r1 + sx0) * cx.c0 + __ldg(r1 + sx1) * cx.c1, cy.c0, cy.c1);
            }
        }
    }
    uint8_t* dst = const_cast<uint8_t*>(L.img) + (size_t)y * L.pitch + x4;
    if (x4 + 3 < L.w) {
        *reinterpret_cast<uchar4*>(dst) = make_uchar4(o[0], o[1], o[2], o[3]);   // level pitch is a multiple of 16
    } else {
        for (int k = 0; k < 4 && x4 + k < L.w; k++) dst[k] = o[k];
    }
}

}  // namespace

void launch_resize(const uint8_t* base, int base_pitch, int bw, int bh, int cn,
                   const LevelDesc* d_levels, const ResizeCoef* d_xtab, const ResizeCoef* d_ytab,
                   const ResizeItem* d_items, int nitems, cudaStream_t st)
{
    if (nitems <= 0) return;
    if (cn == 4)
        resize_c4_kernel<<<nitems, 256, 0, st>>>(base, base_pitch, bw, bh, d_levels, d_xtab, d_ytab, d_items);
    else
        resize_c1_kernel<<<nitems, 256, 0, st>>>(base, base_pitch, bw, bh, d_levels, d_xtab, d_ytab, d_items);
}

}  // namespace hogk
