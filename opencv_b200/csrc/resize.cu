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

__device__ __forceinline__ int clipi(int x, int a, int b) { return x >= a ? (x < b ? x : b - 1) : a; }

__device__ __forceinline__ int vlin(int h0, int h1, int b0, int b1)
{
    // resize.cpp:1594-1601: uchar(( ((b0 * (S0 >> 4)) >> 16) + ((b1 * (S1 >> 4)) >> 16) + 2) >> 2)
    return (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;
}

// 8UC4: a CTA owns a tile of 32 columns x 64 rows; thread <-> column, each warp walks DOWN 8 rows.
// The horizontal pass of a source row (resize.cpp:1491-1542) is kept in registers and reused by the next
// destination row when it needs the same source row (always the case for scales below ~1.5, i.e. for most of the
// pyramid's pixels); both taps of a channel are one IDP.2A (16-bit weights x 8-bit samples).
constexpr int kC4TileW = 32, kC4RowsPerWarp = 8, kC4TileH = 8 * kC4RowsPerWarp;

struct HRow { int t[4]; };        // horizontal result >> 4 of the four channels (resize.cpp:1596: S0[x] >> 4)

__device__ __forceinline__ HRow hrow_c4(const uint8_t* __restrict__ row, int sx0, int sx1, unsigned a01)
{
    const unsigned p0 = __ldg(reinterpret_cast<const unsigned*>(row) + sx0);
    const unsigned p1 = __ldg(reinterpret_cast<const unsigned*>(row) + sx1);
    const unsigned lo = __byte_perm(p0, p1, 0x5140);     // (p0.b0, p1.b0, p0.b1, p1.b1)
    const unsigned hi = __byte_perm(p0, p1, 0x7362);     // (p0.b2, p1.b2, p0.b3, p1.b3)
    HRow h;
    h.t[0] = (int)(__dp2a_lo(a01, lo, 0u) >> 4);      // S[sx]*a0 + S[sx+cn]*a1
    h.t[1] = (int)(__dp2a_hi(a01, lo, 0u) >> 4);
    h.t[2] = (int)(__dp2a_lo(a01, hi, 0u) >> 4);
    h.t[3] = (int)(__dp2a_hi(a01, hi, 0u) >> 4);
    return h;
}

__global__ void __launch_bounds__(256)
resize_c4_kernel(const __grid_constant__ FrameSet fs, int bw, int bh,
                 const LevelDesc* __restrict__ levels, const ResizeCoef* __restrict__ xtab,
                 const ResizeCoef* __restrict__ ytab, const ResizeItem* __restrict__ items)
{
    const ResizeItem it = items[blockIdx.x];
    if (it.frame >= fs.nframes) return;
    const uint8_t* __restrict__ base = fs.base[it.frame];
    const int bpitch = fs.pitch[it.frame];
    const LevelDesc& L = levels[it.level];
    const int x = it.x0 + (threadIdx.x & 31);
    const int y0 = it.y0 + (threadIdx.x >> 5) * kC4RowsPerWarp;
    const int w = L.w, h = L.h;
    if (x >= w || y0 >= h) return;
    const int y1 = min(y0 + kC4RowsPerWarp, h);
    uint8_t* __restrict__ dst = const_cast<uint8_t*>(L.img) + it.frame * fs.pyr_stride + (size_t)y0 * L.pitch + 4 * x;
    const int dpitch = L.pitch;
    if (L.mode == 2) {          // exact half: 2x2 box average (resize.cpp:2542-2578)
        for (int y = y0; y < y1; y++, dst += dpitch) {
            const uint2 a = __ldg(reinterpret_cast<const uint2*>(base + (size_t)(2 * y) * bpitch) + x);
            const uint2 b = __ldg(reinterpret_cast<const uint2*>(base + (size_t)(2 * y + 1) * bpitch) + x);
            unsigned o = 0;
#pragma unroll
            for (int c = 0; c < 4; c++) {
                const unsigned sum = ((a.x >> (8 * c)) & 255u) + ((a.y >> (8 * c)) & 255u) +
                                     ((b.x >> (8 * c)) & 255u) + ((b.y >> (8 * c)) & 255u) + 2u;
                o |= (sum >> 2) << (8 * c);
            }
            *reinterpret_cast<unsigned*>(dst) = o;
        }
        return;
    }
    const ResizeCoef cx = xtab[L.xtab_off + x];
    const int sx0 = cx.s0, sx1 = cx.s1;
    const unsigned a01 = cx.c0;
    const ResizeCoef* __restrict__ yt = ytab + L.ytab_off;
    // Horizontal results of two source rows live in A and B.  `r1` is the source row of the one that served as the
    // LOWER tap of the previous destination row; when the next row's upper tap is that row, only one new source row
    // is needed and the two buffers swap roles (par) instead of being copied.  The row index is the same for the
    // whole warp, so none of these branches diverges.
    HRow A, B;
#pragma unroll
    for (int c = 0; c < 4; c++) A.t[c] = B.t[c] = 0;
    int r1 = -1;
    bool par = false;           // false: lower tap of the previous row is in B; true: in A
    for (int y = y0; y < y1; y++, dst += dpitch) {
        const ResizeCoef cy = yt[y];
        const int sy0 = cy.s0, sy1 = cy.s1;
        // vertical pass, resize.cpp:1594-1601: ((b0*(S0>>4))>>16) + ((b1*(S1>>4))>>16) + 2 >> 2; both products are
        // non-negative and below 2^31, so (b<<16) * t >> 32 is the same truncating shift
        const unsigned b0 = cy.c0, b1 = cy.c1;
        auto vertical = [&](const HRow& h0, const HRow& h1) {
            unsigned v[4];
#pragma unroll
            for (int c = 0; c < 4; c++)
                v[c] = (__umulhi(b0, (unsigned)h0.t[c]) + __umulhi(b1, (unsigned)h1.t[c]) + 2u) >> 2;
            *reinterpret_cast<unsigned*>(dst) = v[0] + v[1] * 256u + v[2] * 65536u + v[3] * 16777216u;
        };
        const uint8_t* __restrict__ row1 = base + (size_t)sy1 * bpitch;
        if (sy0 == r1) {
            if (par) { B = hrow_c4(row1, sx0, sx1, a01); vertical(A, B); }
            else     { A = hrow_c4(row1, sx0, sx1, a01); vertical(B, A); }
            par = !par;
        } else {
            A = hrow_c4(base + (size_t)sy0 * bpitch, sx0, sx1, a01);
            B = hrow_c4(row1, sx0, sx1, a01);
            vertical(A, B);
            par = false;
        }
        r1 = sy1;
    }
}

// 8UC1: one thread = four consecutive destination pixels (one 4-byte store); tile 128 x 8.
__global__ void __launch_bounds__(256)
resize_c1_kernel(const __grid_constant__ FrameSet fs, int bw, int bh,
                 const LevelDesc* __restrict__ levels, const ResizeCoef* __restrict__ xtab,
                 const ResizeCoef* __restrict__ ytab, const ResizeItem* __restrict__ items)
{
    const ResizeItem it = items[blockIdx.x];
    if (it.frame >= fs.nframes) return;
    const uint8_t* __restrict__ base = fs.base[it.frame];
    const int bpitch = fs.pitch[it.frame];
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
        const int sy0 = cy.s0, sy1 = cy.s1;
        const int b0 = (int)(cy.c0 >> 16), b1 = (int)(cy.c1 >> 16);
        const uint8_t* r0 = base + (size_t)sy0 * bpitch;
        const uint8_t* r1 = base + (size_t)sy1 * bpitch;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            int x = x4 + k;
            if (x < L.w) {
                const ResizeCoef cx = xtab[L.xtab_off + x];
                const int sx0 = cx.s0, sx1 = cx.s1;
                const int a0 = (int)(cx.c0 & 0xffffu), a1 = (int)(cx.c0 >> 16);
                o[k] = vlin(__ldg(r0 + sx0) * a0 + __ldg(r0 + sx1) * a1,
                            __ldg(r1 + sx0) * a0 + __ldg(r1 + sx1) * a1, b0, b1);
            }
        }
    }
    uint8_t* dst = const_cast<uint8_t*>(L.img) + it.frame * fs.pyr_stride + (size_t)y * L.pitch + x4;
    if (x4 + 3 < L.w) {
        *reinterpret_cast<uchar4*>(dst) = make_uchar4(o[0], o[1], o[2], o[3]);   // level pitch is a multiple of 16
    } else {
        for (int k = 0; k < 4 && x4 + k < L.w; k++) dst[k] = o[k];
    }
}

}  // namespace

int resize_tile_w(int cn) { return cn == 4 ? kC4TileW : 128; }
int resize_tile_h(int cn) { return cn == 4 ? kC4TileH : 8; }

void launch_resize(const FrameSet& fs, int bw, int bh, int cn,
                   const LevelDesc* d_levels, const ResizeCoef* d_xtab, const ResizeCoef* d_ytab,
                   const ResizeItem* d_items, int nitems, cudaStream_t st)
{
    if (nitems <= 0) return;
    if (cn == 4)
        resize_c4_kernel<<<nitems, 256, 0, st>>>(fs, bw, bh, d_levels, d_xtab, d_ytab, d_items);
    else
        resize_c1_kernel<<<nitems, 256, 0, st>>>(fs, bw, bh, d_levels, d_xtab, d_ytab, d_items);
}

}  // namespace hogk
