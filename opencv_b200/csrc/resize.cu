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
#include <algorithm>
#include <cstdlib>

namespace hogk {

namespace {


__device__ __forceinline__ int vlin(int h0, int h1, int b0, int b1)
{
    // resize.cpp:1594-1601: uchar(( ((b0 * (S0 >> 4)) >> 16) + ((b1 * (S1 >> 4)) >> 16) + 2) >> 2)
    return (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;
}

// 8UC4.  A warp owns 64 columns x R rows of a level (lane <-> columns x and x + 32) and walks DOWN the rows.
//   * Staging: the source rows the warp needs (columns [c_lo, c_lo + span)) are brought into its private slice of
//     shared memory by bulk async copies (cp.async.bulk, one per source row, completion on the warp's own
//     mbarrier), so the inner loop addresses pixels with 32-bit shared offsets and never waits on L2.  For
//     scales < 2 the rows are one contiguous range; for scales >= 2 (no source row is shared by two destination
//     rows) only the two rows of every destination row are staged.  R is chosen per level on the host so that the
//     slice fits (resize_plan_rows); levels too coarse even for R = 1 read global memory directly.
//   * The horizontal pass of a source row (resize.cpp:1491-1542) stays in registers and is reused by the next
//     destination row when it needs the same source row (always the case below scale ~1.5, i.e. for most of the
//     pyramid's pixels); both taps of a channel are one IDP.2A (16-bit weights x 8-bit samples).
//   * Vertical pass + pack: the two truncated products of channels c and c+2 are packed by IMADs into 16-bit
//     fields, rounded and shifted together, and one PRMT assembles the four bytes.
#ifndef HOGB200_RS_WARPS
#define HOGB200_RS_WARPS 4
#endif
#ifndef HOGB200_RS_BYTES
#define HOGB200_RS_BYTES 6144
#endif
constexpr int kRsWarps = HOGB200_RS_WARPS;
constexpr int kRsTileW = 64;
constexpr int kRsMaxRows = 16;                              // destination rows per warp, at most
constexpr int kRsHeader = 16 + 16 * kRsMaxRows;             // mbarrier + per-row coefficients
constexpr int kRsWarpBytes = HOGB200_RS_BYTES;                        // staged source pixels per warp
constexpr int kRsWarpStride = kRsHeader + kRsWarpBytes;
constexpr int kRsDirectRows = 16;                           // rows per warp of the unstaged paths
#ifndef HOGB200_RS_CHUNKS
#define HOGB200_RS_CHUNKS 4
#endif
constexpr int kRsChunks = HOGB200_RS_CHUNKS;                // staged path: a warp walks down this many chunks of R rows, restaging its slice for each

struct HRow { int t[4]; };        // horizontal result >> 4 of the four channels (resize.cpp:1596: S0[x] >> 4)

__device__ __forceinline__ HRow hrow_px(unsigned p0, unsigned p1, unsigned a01)
{
    const unsigned lo = __byte_perm(p0, p1, 0x5140);     // (p0.b0, p1.b0, p0.b1, p1.b1)
    const unsigned hi = __byte_perm(p0, p1, 0x7362);     // (p0.b2, p1.b2, p0.b3, p1.b3)
    HRow h;
    h.t[0] = (int)(__dp2a_lo(a01, lo, 0u) >> 4);      // S[sx]*a0 + S[sx+cn]*a1
    h.t[1] = (int)(__dp2a_hi(a01, lo, 0u) >> 4);
    h.t[2] = (int)(__dp2a_lo(a01, hi, 0u) >> 4);
    h.t[3] = (int)(__dp2a_hi(a01, hi, 0u) >> 4);
    return h;
}
__device__ __forceinline__ HRow hrow_smem(unsigned addr, unsigned a01)
{
    unsigned p0, p1;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(p0) : "r"(addr));
    asm volatile("ld.shared.u32 %0, [%1+4];" : "=r"(p1) : "r"(addr));       // weight 0 where sx + 1 is outside the row
    return hrow_px(p0, p1, a01);
}
// resize.cpp:1594-1601: uchar(( ((b0*(S0>>4))>>16) + ((b1*(S1>>4))>>16) + 2) >> 2) for four channels.  b0, b1 <= 2048
// and S>>4 < 2^15, so every product fits 32 bits and its truncated term is the product's upper half: ONE byte
// permute takes the upper halves of channels c and c+2 into the two 16-bit fields of a register (all fields stay
// below 2^11), the two taps and the rounding constant are added field-parallel, and a last permute assembles the bytes.
// The only integer multiplies are the eight products (the multiplier pipe is what bounds this kernel).
__device__ __forceinline__ unsigned vertical_pack(const HRow& h0, const HRow& h1, unsigned b0, unsigned b1)
{
    const unsigned x0 = b0 * (unsigned)h0.t[0], x1 = b0 * (unsigned)h0.t[1], x2 = b0 * (unsigned)h0.t[2], x3 = b0 * (unsigned)h0.t[3];
    const unsigned y0 = b1 * (unsigned)h1.t[0], y1 = b1 * (unsigned)h1.t[1], y2 = b1 * (unsigned)h1.t[2], y3 = b1 * (unsigned)h1.t[3];
    const unsigned p02 = __byte_perm(x0, x2, 0x7632) + __byte_perm(y0, y2, 0x7632) + 0x00020002u;
    const unsigned p13 = __byte_perm(x1, x3, 0x7632) + __byte_perm(y1, y3, 0x7632) + 0x00020002u;
    return __byte_perm(p02 >> 2, p13 >> 2, 0x6240);
}

__device__ __forceinline__ void mbar_init(unsigned mbar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned mbar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned mbar, unsigned parity)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(mbar), "r"(parity) : "memory");
}
// global -> shared bulk copy (16-byte aligned on both sides, size a multiple of 16), completion counted in bytes on mbar
__device__ __forceinline__ void bulk_g2s(unsigned dst, const void* src, unsigned bytes, unsigned mbar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(mbar) : "memory");
}

struct Px2 { unsigned p0, p1; };
__device__ __forceinline__ Px2 load_px2(const uint8_t* __restrict__ col, unsigned off)
{
    const unsigned* p = reinterpret_cast<const unsigned*>(col + off);
    Px2 r; r.p0 = __ldg(p); r.p1 = __ldg(p + 1);
    return r;
}

// The rows of one warp without staging (see resize_c4_direct_kernel).
__device__ __forceinline__ void direct_rows(const uint8_t* __restrict__ base, unsigned bpitch, const ResizeCoef& cxa,
                                            const ResizeCoef& cxb, const ResizeCoef* __restrict__ yt, int nrows,
                                            uint8_t* __restrict__ dst, int dpitch, bool va, bool vb)
{
    const uint8_t* __restrict__ cola = base + 4 * (size_t)cxa.s0;
    const uint8_t* __restrict__ colb = base + 4 * (size_t)cxb.s0;
    const unsigned a01a = cxa.c0, a01b = cxb.c0;

    // Two destination rows per iteration, with the roles of the two pixel / horizontal-pass register sets
    // alternating, so that nothing is copied: row k reads set 0 and fills set 1 with the pixels of row k + 1, row
    // k + 1 reads set 1 and refills set 0.
    struct RowRegs { Px2 ua, ub, la, lb; };          // upper source row (when not reused) and lower source row, 2 columns
    auto request = [&](const ResizeCoef& c, bool reuse_row, RowRegs& r) {
        r.la = load_px2(cola, (unsigned)c.s1 * bpitch); r.lb = load_px2(colb, (unsigned)c.s1 * bpitch);
        if (!reuse_row) { r.ua = load_px2(cola, (unsigned)c.s0 * bpitch); r.ub = load_px2(colb, (unsigned)c.s0 * bpitch); }
    };
    // computes one destination row from r; Hin = horizontal pass of the previous row's lower source row, Hout = this row's
    auto emit = [&](const ResizeCoef& c, bool reuse_row, const RowRegs& r, const HRow& Hina, const HRow& Hinb,
                    HRow& Houta, HRow& Houtb, uint8_t* __restrict__ d) {
        Houta = hrow_px(r.la.p0, r.la.p1, a01a); Houtb = hrow_px(r.lb.p0, r.lb.p1, a01b);
        unsigned outa, outb;
        if (reuse_row) {
            outa = vertical_pack(Hina, Houta, c.c0, c.c1); outb = vertical_pack(Hinb, Houtb, c.c0, c.c1);
        } else {
            const HRow Ua = hrow_px(r.ua.p0, r.ua.p1, a01a), Ub = hrow_px(r.ub.p0, r.ub.p1, a01b);
            outa = vertical_pack(Ua, Houta, c.c0, c.c1); outb = vertical_pack(Ub, Houtb, c.c0, c.c1);
        }
        if (va) *reinterpret_cast<unsigned*>(d) = outa;
        if (vb) *reinterpret_cast<unsigned*>(d + 128) = outb;
    };
    ResizeCoef c0 = yt[0], c1 = yt[min(1, nrows - 1)];
    RowRegs r0, r1;
    HRow H0a, H0b, H1a, H1b;
#pragma unroll
    for (int c = 0; c < 4; c++) H1a.t[c] = H1b.t[c] = 0;
    request(c0, false, r0);
    bool reuse0 = false;                                   // row k reuses the previous row's lower source row
    for (int k = 0; k < nrows; k += 2, dst += 2 * (size_t)dpitch) {
        const ResizeCoef c2 = yt[min(k + 2, nrows - 1)], c3 = yt[min(k + 3, nrows - 1)];
        const bool reuse1 = c1.s0 == c0.s1;                // warp-uniform
        request(c1, reuse1, r1);                           // row k + 1 (a repeat of the last row when k + 1 == nrows)
        emit(c0, reuse0, r0, H1a, H1b, H0a, H0b, dst);
        if (k + 1 >= nrows) break;
        const bool reuse2 = c2.s0 == c1.s1;
        request(c2, reuse2, r0);                           // row k + 2
        emit(c1, reuse1, r1, H0a, H0b, H1a, H1b, dst + dpitch);
        reuse0 = reuse2; c0 = c2; c1 = c3;
    }
}

__global__ void __launch_bounds__(kRsWarps * 32)
resize_c4_kernel(const __grid_constant__ FrameSet fs, int bw, int bh,
                 const LevelDesc* __restrict__ levels, const ResizeCoef* __restrict__ xtab,
                 const ResizeCoef* __restrict__ ytab, const ResizeItem* __restrict__ items)
{
    extern __shared__ __align__(16) unsigned char rs_smem[];
    const ResizeItem it = items[blockIdx.x];
    if (it.frame >= fs.nframes) return;
    const uint8_t* __restrict__ base = fs.base[it.frame];
    const int bpitch = fs.pitch[it.frame];
    const LevelDesc& L = levels[it.level];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int w = L.w, h = L.h, rs = L.rs_rows;
    const int R = rs > 0 ? rs : (rs < 0 ? -rs : kRsDirectRows);
    const int RW = rs != 0 ? R * kRsChunks : R;              // rows of this warp
    const int y0 = it.y0 + warp * RW;
    if (y0 >= h) return;                                     // the whole warp
    const int nrows = min(RW, h - y0);
    const int xa = it.x0 + lane, xb = xa + 32;
    const bool va = xa < w, vb = xb < w;
    const int dpitch = L.pitch;
    uint8_t* __restrict__ dst = const_cast<uint8_t*>(L.img) + it.frame * fs.pyr_stride + (size_t)y0 * dpitch + 4 * xa;

    if (L.mode == 2) {          // exact half: 2x2 box average (resize.cpp:2542-2578)
        for (int k = 0; k < nrows; k++, dst += dpitch) {
            const int y = y0 + k;
#pragma unroll
            for (int c = 0; c < 2; c++) {
                const int x = xa + 32 * c;
                if (x >= w) continue;
                const uint2 a = __ldg(reinterpret_cast<const uint2*>(base + (size_t)(2 * y) * bpitch) + x);
                const uint2 b = __ldg(reinterpret_cast<const uint2*>(base + (size_t)(2 * y + 1) * bpitch) + x);
                unsigned o = 0;
#pragma unroll
                for (int ch = 0; ch < 4; ch++) {
                    const unsigned sum = ((a.x >> (8 * ch)) & 255u) + ((a.y >> (8 * ch)) & 255u) +
                                         ((b.x >> (8 * ch)) & 255u) + ((b.y >> (8 * ch)) & 255u) + 2u;
                    o |= (sum >> 2) << (8 * ch);
                }
                *reinterpret_cast<unsigned*>(dst + 128 * c) = o;
            }
        }
        return;
    }
    const ResizeCoef* __restrict__ xt = xtab + L.xtab_off;
    const ResizeCoef* __restrict__ yt = ytab + L.ytab_off + y0;
    const ResizeCoef cxa = xt[min(xa, w - 1)], cxb = xt[min(xb, w - 1)];
    if (rs == 0) {              // unstaged: levels where staging does not pay (few rows per warp would fit the slice)
        direct_rows(base, (unsigned)bpitch, cxa, cxb, yt, nrows, dst, dpitch, va, vb);
        return;
    }

    // ---- stage the source rows of this warp
    const unsigned ws = (unsigned)__cvta_generic_to_shared(rs_smem) + warp * kRsWarpStride;
    const unsigned mbar = ws, ycoef = ws + 16, tile = ws + kRsHeader;
    const int c_lo = xt[it.x0].s0 & ~3;                                   // first staged source column
    int span = (xt[min(it.x0 + kRsTileW, w) - 1].s0 + 2 - c_lo + 3) & ~3; // through the second tap of the last column
    span = min(span, ((bw + 3) & ~3) - c_lo);                              // rows are readable up to the next multiple of 4
    const unsigned rowbytes = 4u * span;
    const uint8_t* __restrict__ src0 = base + 4 * c_lo;
    const unsigned cola = tile + 4 * (cxa.s0 - c_lo), colb = tile + 4 * (cxb.s0 - c_lo);
    const unsigned a01a = cxa.c0, a01b = cxb.c0;
    if (lane == 0) mbar_init(mbar, 1);
    // Horizontal results of two source rows live in A and B (per column).  `r1` is the staged offset of the row that
    // served as the LOWER tap of the previous destination row; when the next row's upper tap is that row, only one
    // new source row is needed and the two buffers swap roles (par) instead of being copied.  The offsets are the
    // same for the whole warp, so none of these branches diverges.
    HRow Aa, Ba, Ab, Bb;
#pragma unroll
    for (int c = 0; c < 4; c++) Aa.t[c] = Ba.t[c] = Ab.t[c] = Bb.t[c] = 0;
    for (int ch = 0, k0 = 0; k0 < nrows; ch++, k0 += R, yt += R) {
        const int nr = min(R, nrows - k0);                  // destination rows of this chunk
        const ResizeCoef cyl = yt[min(lane, nr - 1)];
        const int r_lo = __shfl_sync(0xffffffffu, cyl.s0, 0);
        const int r_hi = __shfl_sync(0xffffffffu, cyl.s1, nr - 1);
        const int nstaged = rs > 0 ? r_hi - r_lo + 1 : 2 * nr;
        const unsigned off0 = rs > 0 ? (cyl.s0 - r_lo) * rowbytes : 2 * lane * rowbytes;
        const unsigned off1 = rs > 0 ? (cyl.s1 - r_lo) * rowbytes : off0 + rowbytes;
        __syncwarp();                                       // every lane is done with the previous chunk's slice and row table
        if (lane < nr)
            asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(ycoef + 16 * lane), "r"(off0), "r"(off1), "r"(cyl.c0), "r"(cyl.c1) : "memory");
        if (lane == 0) mbar_expect_tx(mbar, nstaged * rowbytes);
        __syncwarp();
        if (rs > 0) {
            for (int j = lane; j < nstaged; j += 32) bulk_g2s(tile + j * rowbytes, src0 + (size_t)(r_lo + j) * bpitch, rowbytes, mbar);
        } else if (lane < nr) {
            bulk_g2s(tile + off0, src0 + (size_t)cyl.s0 * bpitch, rowbytes, mbar);
            bulk_g2s(tile + off1, src0 + (size_t)cyl.s1 * bpitch, rowbytes, mbar);
        }
        mbar_wait(mbar, ch & 1);
#ifdef HOGB200_CHECKS       // the staged rows fit the warp's slice (resize_plan_rows promised it)
        if (nstaged * rowbytes + 16 > (unsigned)kRsWarpBytes) __trap();
#endif

        unsigned r1 = 0xffffffffu;
        bool par = false;           // false: lower tap of the previous row is in B; true: in A
        for (int k = 0; k < nr; k++, dst += dpitch) {
            unsigned o0, o1, b0, b1;
            asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(o0), "=r"(o1), "=r"(b0), "=r"(b1) : "r"(ycoef + 16 * k));
            unsigned outa, outb;
            if (o0 == r1) {
                if (par) {
                    Ba = hrow_smem(cola + o1, a01a); Bb = hrow_smem(colb + o1, a01b);
                    outa = vertical_pack(Aa, Ba, b0, b1); outb = vertical_pack(Ab, Bb, b0, b1);
                } else {
                    Aa = hrow_smem(cola + o1, a01a); Ab = hrow_smem(colb + o1, a01b);
                    outa = vertical_pack(Ba, Aa, b0, b1); outb = vertical_pack(Bb, Ab, b0, b1);
                }
                par = !par;
            } else {
                Aa = hrow_smem(cola + o0, a01a); Ab = hrow_smem(colb + o0, a01b);
                Ba = hrow_smem(cola + o1, a01a); Bb = hrow_smem(colb + o1, a01b);
                outa = vertical_pack(Aa, Ba, b0, b1); outb = vertical_pack(Ab, Bb, b0, b1);
                par = false;
            }
            r1 = o1;
            if (va) *reinterpret_cast<unsigned*>(dst) = outa;
            if (vb) *reinterpret_cast<unsigned*>(dst + 128) = outb;
        }
    }
}

// 8UC4 without staging: a warp owns 64 columns x 16 rows (lane <-> columns x, x+32) and walks down; both taps of a
// pixel come from one per-thread column pointer (the host table guarantees s1 == s0 + 1), the source rows of the
// NEXT destination row are requested before the current one is computed (so L2 latency hides behind arithmetic),
// and the horizontal pass of the lower source row is kept for the next destination row, which reuses it whenever
// its upper row is that row (always below scale ~1.5).
constexpr int kRdWarps = 4, kRdRows = 16;

__global__ void __launch_bounds__(kRdWarps * 32)
resize_c4_direct_kernel(const __grid_constant__ FrameSet fs, const LevelDesc* __restrict__ levels,
                        const ResizeCoef* __restrict__ xtab, const ResizeCoef* __restrict__ ytab,
                        const ResizeItem* __restrict__ items)
{
    const ResizeItem it = items[blockIdx.x];
    if (it.frame >= fs.nframes) return;
    const uint8_t* __restrict__ base = fs.base[it.frame];
    const unsigned bpitch = (unsigned)fs.pitch[it.frame];
    const LevelDesc& L = levels[it.level];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int w = L.w, h = L.h;
    const int y0 = it.y0 + warp * kRdRows;
    if (y0 >= h) return;                                     // the whole warp
    const int nrows = min(kRdRows, h - y0);
    const int xa = it.x0 + lane, xb = xa + 32;
    const bool va = xa < w, vb = xb < w;
    const int dpitch = L.pitch;
    uint8_t* __restrict__ dst = const_cast<uint8_t*>(L.img) + it.frame * fs.pyr_stride + (size_t)y0 * dpitch + 4 * xa;

    if (L.mode == 2) {          // exact half: 2x2 box average (resize.cpp:2542-2578)
        for (int k = 0; k < nrows; k++, dst += dpitch) {
            const int y = y0 + k;
#pragma unroll
            for (int c = 0; c < 2; c++) {
                const int x = xa + 32 * c;
                if (x >= w) continue;
                const uint2 a = __ldg(reinterpret_cast<const uint2*>(base + (size_t)(2 * y) * bpitch) + x);
                const uint2 b = __ldg(reinterpret_cast<const uint2*>(base + (size_t)(2 * y + 1) * bpitch) + x);
                unsigned o = 0;
#pragma unroll
                for (int ch = 0; ch < 4; ch++) {
                    const unsigned sum = ((a.x >> (8 * ch)) & 255u) + ((a.y >> (8 * ch)) & 255u) +
                                         ((b.x >> (8 * ch)) & 255u) + ((b.y >> (8 * ch)) & 255u) + 2u;
                    o |= (sum >> 2) << (8 * ch);
                }
                *reinterpret_cast<unsigned*>(dst + 128 * c) = o;
            }
        }
        return;
    }
    const ResizeCoef* __restrict__ xt = xtab + L.xtab_off;
    const ResizeCoef* __restrict__ yt = ytab + L.ytab_off + y0;
    const ResizeCoef cxa = xt[min(xa, w - 1)], cxb = xt[min(xb, w - 1)];
    direct_rows(base, bpitch, cxa, cxb, yt, nrows, dst, dpitch, va, vb);
}

// 8UC1.  A warp owns 128 columns x R rows of a level (lane <-> 4 adjacent destination pixels, one 4-byte store per row)
// and walks DOWN the rows, kRsChunks chunks of R rows.  The source rows of a chunk are staged in the warp's private slice
// of shared memory with 32-bit loads (an 8UC1 frame only promises 4-byte aligned rows, which rules the 16-byte bulk copy
// out); from there the two taps of a pixel are two aligned 32-bit shared loads + one PRMT that picks bytes (sx, sx + 1)
// -- never a byte load -- and one IDP.2A; the horizontal pass of a source row is reused by the next destination row when
// it needs the same row; the vertical pass is two multiplies per pixel whose upper halves are packed pairwise by PRMT,
// rounded and shifted together and assembled into the 4 output bytes by a last PRMT (resize.cpp:1491-1542, 1577-1603).
constexpr int kR1TileW = 128;
#ifndef HOGB200_R1_CHUNKS
#define HOGB200_R1_CHUNKS 1
#endif
constexpr int kR1Chunks = HOGB200_R1_CHUNKS;       // chunks of R rows per warp: 8UC1 pyramids are small (VGA: 3 Mpx), a 128 x 256 tile
                                                    // (4 chunks) leaves the SMs a handful of CTAs each

struct HRow1 { int t[4]; };      // horizontal result >> 4 of the lane's four pixels

__device__ __forceinline__ HRow1 hrow1_smem(unsigned rowaddr, const unsigned (&woff)[4], const unsigned (&sel)[4], const unsigned (&a01)[4])
{
    HRow1 h;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        unsigned w0, w1;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w0) : "r"(rowaddr + woff[k]));
        asm volatile("ld.shared.u32 %0, [%1+4];" : "=r"(w1) : "r"(rowaddr + woff[k]));
        const unsigned pair = __byte_perm(w0, w1, sel[k]);            // (S[sx], S[sx+1], 0, 0)
        h.t[k] = (int)(__dp2a_lo(a01[k], pair, 0u) >> 4);             // (S[sx]*a0 + S[sx+1]*a1) >> 4
    }
    return h;
}
__device__ __forceinline__ unsigned vertical_pack1(const HRow1& h0, const HRow1& h1, unsigned b0, unsigned b1)
{
    const unsigned x0 = b0 * (unsigned)h0.t[0], x1 = b0 * (unsigned)h0.t[1], x2 = b0 * (unsigned)h0.t[2], x3 = b0 * (unsigned)h0.t[3];
    const unsigned y0 = b1 * (unsigned)h1.t[0], y1 = b1 * (unsigned)h1.t[1], y2 = b1 * (unsigned)h1.t[2], y3 = b1 * (unsigned)h1.t[3];
    const unsigned p01 = __byte_perm(x0, x1, 0x7632) + __byte_perm(y0, y1, 0x7632) + 0x00020002u;
    const unsigned p23 = __byte_perm(x2, x3, 0x7632) + __byte_perm(y2, y3, 0x7632) + 0x00020002u;
    return __byte_perm(p01 >> 2, p23 >> 2, 0x6420);
}

__global__ void __launch_bounds__(kRsWarps * 32)
resize_c1_kernel(const __grid_constant__ FrameSet fs, int bw, int bh,
                 const LevelDesc* __restrict__ levels, const ResizeCoef* __restrict__ xtab,
                 const ResizeCoef* __restrict__ ytab, const ResizeItem* __restrict__ items)
{
    extern __shared__ __align__(16) unsigned char rs_smem[];
    const ResizeItem it = items[blockIdx.x];
    if (it.frame >= fs.nframes) return;
    const uint8_t* __restrict__ base = fs.base[it.frame];
    const int bpitch = fs.pitch[it.frame];
    const LevelDesc& L = levels[it.level];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int w = L.w, h = L.h, rs = L.rs_rows;
    const int R = rs > 0 ? rs : (rs < 0 ? -rs : kRsDirectRows);
    const int RW = rs != 0 ? R * kR1Chunks : R;              // rows of this warp
    const int y0 = it.y0 + warp * RW;
    if (y0 >= h) return;                                     // the whole warp
    const int nrows = min(RW, h - y0);
    const int x4 = it.x0 + 4 * lane;
    const int dpitch = L.pitch;
    uint8_t* __restrict__ dst = const_cast<uint8_t*>(L.img) + it.frame * fs.pyr_stride + (size_t)y0 * dpitch + x4;
    const bool full = x4 + 3 < w;                            // level pitch is a multiple of 64: a whole-word store is in bounds

    if (L.mode == 2 || rs == 0) {
        // exact half (2x2 box average, resize.cpp:2542-2578) and levels too coarse to stage (a handful of small ones): per-pixel loads
        const ResizeCoef* __restrict__ xt = xtab + L.xtab_off;
        for (int k = 0; k < nrows; k++, dst += dpitch) {
            const int y = y0 + k;
            uint8_t o[4] = {0, 0, 0, 0};
            if (L.mode == 2) {
                const uint8_t* r0 = base + (size_t)(2 * y) * bpitch;
                const uint8_t* r1 = r0 + bpitch;
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int x = x4 + j;
                    if (x < w) o[j] = (__ldg(r0 + 2 * x) + __ldg(r0 + 2 * x + 1) + __ldg(r1 + 2 * x) + __ldg(r1 + 2 * x + 1) + 2) >> 2;
                }
            } else {
                const ResizeCoef cy = ytab[L.ytab_off + y];
                const int b0 = (int)cy.c0, b1 = (int)cy.c1;
                const uint8_t* r0 = base + (size_t)cy.s0 * bpitch;
                const uint8_t* r1 = base + (size_t)cy.s1 * bpitch;
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int x = x4 + j;
                    if (x < w) {
                        const ResizeCoef cx = xt[x];
                        const int a0 = (int)(cx.c0 & 0xffffu), a1 = (int)(cx.c0 >> 16);
                        o[j] = vlin(__ldg(r0 + cx.s0) * a0 + __ldg(r0 + cx.s1) * a1, __ldg(r1 + cx.s0) * a0 + __ldg(r1 + cx.s1) * a1, b0, b1);
                    }
                }
            }
            if (full) *reinterpret_cast<uchar4*>(dst) = make_uchar4(o[0], o[1], o[2], o[3]);
            else for (int j = 0; j < 4 && x4 + j < w; j++) dst[j] = o[j];
        }
        return;
    }
    const ResizeCoef* __restrict__ xt = xtab + L.xtab_off;
    const ResizeCoef* __restrict__ yt = ytab + L.ytab_off + y0;

    // ---- the warp's slice: [row table 16 B x R][staged source bytes]
    const unsigned ws = (unsigned)__cvta_generic_to_shared(rs_smem) + warp * kRsWarpStride;
    const unsigned ycoef = ws + 16, tile = ws + kRsHeader;
    const int c_lo = xt[it.x0].s0 & ~3;                                    // first staged source column
    int span = (xt[min(it.x0 + kR1TileW, w) - 1].s0 + 2 - c_lo + 3) & ~3; // through the second tap of the last column
    span = min(span, ((bw + 3) & ~3) - c_lo);                              // rows are readable up to the next multiple of 4
    const unsigned rowbytes = (unsigned)span, rowwords = rowbytes >> 2;
    const uint8_t* __restrict__ src0 = base + c_lo;
    unsigned woff[4], sel[4], a01[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const ResizeCoef c = xt[min(x4 + k, w - 1)];
        const unsigned off = (unsigned)(c.s0 - c_lo);
        woff[k] = off & ~3u;
        sel[k] = 0x4400u | (off & 3u) | (((off & 3u) + 1u) << 4);          // bytes off, off + 1 of the word pair; upper half zero
        a01[k] = c.c0;
    }
    HRow1 A, B;
#pragma unroll
    for (int k = 0; k < 4; k++) A.t[k] = B.t[k] = 0;
    for (int k0 = 0; k0 < nrows; k0 += R, yt += R) {
        const int nr = min(R, nrows - k0);                  // destination rows of this chunk
        const ResizeCoef cyl = yt[min(lane, nr - 1)];
        const int r_lo = __shfl_sync(0xffffffffu, cyl.s0, 0);
        const int r_hi = __shfl_sync(0xffffffffu, cyl.s1, nr - 1);
        const int nstaged = rs > 0 ? r_hi - r_lo + 1 : 2 * nr;
        const unsigned off0 = rs > 0 ? (cyl.s0 - r_lo) * rowbytes : 2 * lane * rowbytes;
        const unsigned off1 = rs > 0 ? (cyl.s1 - r_lo) * rowbytes : off0 + rowbytes;
        __syncwarp();                                       // every lane is done with the previous chunk's slice and row table
        if (lane < nr)
            asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(ycoef + 16 * lane), "r"(off0), "r"(off1), "r"(cyl.c0), "r"(cyl.c1) : "memory");
        // stage: word i of staged row j; the source row of staged row j is r_lo + j, or the j-th of the (s0, s1) pairs
        const int total = nstaged * (int)rowwords;
#ifdef HOGB200_CHECKS
        if (nstaged * rowbytes + 16 > (unsigned)kRsWarpBytes) __trap();
#endif
        for (int i0 = 0; i0 < total; i0 += 128) {           // four words in flight per lane
            unsigned v[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int i = i0 + 32 * u + lane;
                if (i < total) {
                    const int j = i / (int)rowwords, wd = i - j * (int)rowwords;
                    int srow = r_lo + j;
                    if (rs < 0) { const ResizeCoef cj = yt[j >> 1]; srow = (j & 1) ? cj.s1 : cj.s0; }
                    v[u] = __ldg(reinterpret_cast<const unsigned*>(src0 + (size_t)srow * bpitch) + wd);
                }
            }
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int i = i0 + 32 * u + lane;
                if (i < total) asm volatile("st.shared.u32 [%0], %1;" ::"r"(tile + 4 * i), "r"(v[u]) : "memory");
            }
        }
        __syncwarp();

        unsigned r1 = 0xffffffffu;
        bool par = false;           // false: lower tap of the previous row is in B; true: in A
        for (int k = 0; k < nr; k++, dst += dpitch) {
            unsigned o0, o1, b0, b1;
            asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(o0), "=r"(o1), "=r"(b0), "=r"(b1) : "r"(ycoef + 16 * k));
            unsigned out;
            if (o0 == r1) {
                if (par) { B = hrow1_smem(tile + o1, woff, sel, a01); out = vertical_pack1(A, B, b0, b1); }
                else     { A = hrow1_smem(tile + o1, woff, sel, a01); out = vertical_pack1(B, A, b0, b1); }
                par = !par;
            } else {
                A = hrow1_smem(tile + o0, woff, sel, a01);
                B = hrow1_smem(tile + o1, woff, sel, a01);
                out = vertical_pack1(A, B, b0, b1);
                par = false;
            }
            r1 = o1;
            if (full) *reinterpret_cast<unsigned*>(dst) = out;
            else for (int j = 0; j < 4 && x4 + j < w; j++) dst[j] = (uint8_t)(out >> (8 * j));
        }
    }
}

// 8UC3 -> 8UC4, one thread = 4 pixels (three 4-byte loads when the row is 4-byte aligned, one 16-byte store)
__global__ void __launch_bounds__(256)
bgr_to_bgra_kernel(const uint8_t* __restrict__ src, size_t spitch, uint8_t* __restrict__ dst, size_t dpitch, int w, int h,
                   int aligned)
{
    const int x4 = 4 * (blockIdx.x * 64 + (threadIdx.x & 63)), y = blockIdx.y * 4 + (threadIdx.x >> 6);
    if (x4 >= w || y >= h) return;
    const uint8_t* __restrict__ s = src + (size_t)y * spitch + 3 * (size_t)x4;
    uint8_t* __restrict__ d = dst + (size_t)y * dpitch + 4 * (size_t)x4;
    if (aligned && x4 + 3 < w) {
        const unsigned a = __ldg(reinterpret_cast<const unsigned*>(s)), b = __ldg(reinterpret_cast<const unsigned*>(s) + 1),
                       c = __ldg(reinterpret_cast<const unsigned*>(s) + 2);
        uint4 o;                                   // bytes: B0 G0 R0 B1 | G1 R1 B2 G2 | R2 B3 G3 R3
        o.x = (a & 0x00ffffffu) | 0xff000000u;
        o.y = __byte_perm(a, b, 0x0543) | 0xff000000u;
        o.z = __byte_perm(b, c, 0x0432) | 0xff000000u;
        o.w = (c >> 8) | 0xff000000u;
        *reinterpret_cast<uint4*>(d) = o;
    } else {
        for (int k = 0; k < 4 && x4 + k < w; k++) {
            const unsigned v = s[3 * k] | (s[3 * k + 1] << 8) | (s[3 * k + 2] << 16) | 0xff000000u;
            *reinterpret_cast<unsigned*>(d + 4 * k) = v;
        }
    }
}

// Frames uploaded flat (host rows that are no multiple of 16 bytes, e.g. 1242 x 4): every row to its 64-byte aligned
// place.  blockIdx.z = frame, one thread = 4 bytes (row_bytes % 4 == 0, both bases 4-byte aligned).
__global__ void __launch_bounds__(256) repitch_kernel(const __grid_constant__ RepitchSet rp)
{
    const int x = 4 * (blockIdx.x * 256 + threadIdx.x), y = blockIdx.y;
    if (x >= rp.row_bytes) return;
    const unsigned v = __ldg(reinterpret_cast<const unsigned*>(rp.src[blockIdx.z] + (size_t)y * rp.spitch + x));
    *reinterpret_cast<unsigned*>(rp.dst[blockIdx.z] + (size_t)y * rp.dpitch + x) = v;
}

// Image + border, BORDER_REFLECT_101 (cv::borderInterpolate, core/src/copy.cpp): dst(y, x) = src(r(y - py), r(x - px)).
// One thread per destination pixel (BPP = bytes per pixel, 1 or 4).  Feeds the ROI path, whose windows may reach into
// the padding the CPU class puts around an image (objdetect/src/hog.cpp:3566-3570).
__device__ __forceinline__ int reflect101(int p, int len)
{
    if (len == 1) return 0;
    while ((unsigned)p >= (unsigned)len) p = p < 0 ? -p : 2 * len - 2 - p;
    return p;
}
template <typename T>
__global__ void __launch_bounds__(256)
pad_reflect101_kernel(const uint8_t* __restrict__ src, size_t spitch, int w, int h, uint8_t* __restrict__ dst, size_t dpitch,
                      int px, int py)
{
    const int x = blockIdx.x * 64 + (threadIdx.x & 63), y = blockIdx.y * 4 + (threadIdx.x >> 6);
    if (x >= w + 2 * px || y >= h + 2 * py) return;
    const T v = reinterpret_cast<const T*>(src + (size_t)reflect101(y - py, h) * spitch)[reflect101(x - px, w)];
    reinterpret_cast<T*>(dst + (size_t)y * dpitch)[x] = v;
}

}  // namespace

void launch_bgr_to_bgra(const uint8_t* src, size_t spitch, uint8_t* dst, size_t dpitch, int w, int h, cudaStream_t st)
{
    if (w <= 0 || h <= 0) return;
    const int aligned = ((uintptr_t)src % 4 == 0) && (spitch % 4 == 0);
    bgr_to_bgra_kernel<<<dim3((w + 255) / 256, (h + 3) / 4), 256, 0, st>>>(src, spitch, dst, dpitch, w, h, aligned);
}

void launch_repitch(const RepitchSet& rp, cudaStream_t st)
{
    if (rp.n <= 0 || rp.rows <= 0 || rp.row_bytes <= 0) return;
    repitch_kernel<<<dim3((rp.row_bytes / 4 + 255) / 256, rp.rows, rp.n), 256, 0, st>>>(rp);
}

void launch_pad_reflect101(const uint8_t* src, size_t spitch, int w, int h, int cn, uint8_t* dst, size_t dpitch, int px, int py,
                           cudaStream_t st)
{
    const dim3 grid((w + 2 * px + 63) / 64, (h + 2 * py + 3) / 4);
    if (cn == 4) pad_reflect101_kernel<unsigned><<<grid, 256, 0, st>>>(src, spitch, w, h, dst, dpitch, px, py);
    else pad_reflect101_kernel<uint8_t><<<grid, 256, 0, st>>>(src, spitch, w, h, dst, dpitch, px, py);
}

// HOGB200_RS_DIRECT=1: every 8UC4 level through the unstaged kernel (64 x 64 tiles)
bool resize_direct()
{
    static const bool on = [] { const char* e = getenv("HOGB200_RS_DIRECT"); return e && atoi(e) != 0; }();
    return on;
}
int resize_tile_w(int cn) { return cn == 4 ? kRsTileW : 128; }
int resize_tile_h(int cn, int rs_rows)
{
    if (cn == 4 && resize_direct()) return kRdWarps * kRdRows;
    const int chunks = cn == 4 ? kRsChunks : kR1Chunks;
    return kRsWarps * (rs_rows > 0 ? rs_rows * chunks : (rs_rows < 0 ? -rs_rows * chunks : kRsDirectRows));
}

// Destination rows per warp for one 8UC4 level, from the level's own coefficient tables: the largest R whose
// staged source slice fits kRsWarpBytes for every warp tile.  > 0: the rows [s0(first), s1(last)] are staged as
// one range; < 0: only the two source rows of every destination row (cheaper from scale 2 up); 0: no staging.
int resize_plan_rows(const ResizeCoef* xt, int dst_w, const ResizeCoef* yt, int dst_h, int src_w, int cn)
{
    const int tile_w = cn == 4 ? kRsTileW : kR1TileW;
    int span = 0;
    for (int x0 = 0; x0 < dst_w; x0 += tile_w) {
        const int c_lo = xt[x0].s0 & ~3;
        int sp = (xt[std::min(x0 + tile_w, dst_w) - 1].s0 + 2 - c_lo + 3) & ~3;
        sp = std::min(sp, ((src_w + 3) & ~3) - c_lo);
        span = std::max(span, sp);
    }
    const int rowbytes = cn * span, slack = 16;     // the second tap of the last column may lie one pixel past the span
    // staging pays only when the per-warp setup (mbarrier, one bulk copy per source row) is spread over enough rows
    static const int min_rows = [] { const char* e = getenv("HOGB200_RS_MINROWS"); return e ? atoi(e) : 8; }();
    // (8UC1 stages with plain loads, no per-warp barrier set-up: any R pays, and the unstaged path costs eight byte loads per pixel)
    for (int R = kRsMaxRows; R >= (cn == 1 ? 1 : std::max(1, min_rows)); R >>= 1) {
        int range = 0;
        for (int y0 = 0; y0 < dst_h; y0 += R)
            range = std::max(range, yt[std::min(y0 + R, dst_h) - 1].s1 - yt[y0].s0 + 1);
        const int pairs = 2 * R;
        if (range <= pairs && range * rowbytes + slack <= kRsWarpBytes) return R;
        if (pairs < range && pairs * rowbytes + slack <= kRsWarpBytes) return -R;
    }
    return 0;
}

void launch_resize(const FrameSet& fs, int bw, int bh, int cn,
                   const LevelDesc* d_levels, const ResizeCoef* d_xtab, const ResizeCoef* d_ytab,
                   const ResizeItem* d_items, int nitems, cudaStream_t st)
{
    if (nitems <= 0) return;
    if (cn == 4 && resize_direct()) {
        resize_c4_direct_kernel<<<nitems, kRdWarps * 32, 0, st>>>(fs, d_levels, d_xtab, d_ytab, d_items);
    } else if (cn == 4) {
        constexpr int smem = kRsWarps * kRsWarpStride + 16;
        if (smem > 48 * 1024) cudaFuncSetAttribute(resize_c4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        resize_c4_kernel<<<nitems, kRsWarps * 32, smem, st>>>(fs, bw, bh, d_levels, d_xtab, d_ytab, d_items);
    }
    else {
        constexpr int smem = kRsWarps * kRsWarpStride + 16;
        resize_c1_kernel<<<nitems, kRsWarps * 32, smem, st>>>(fs, bw, bh, d_levels, d_xtab, d_ytab, d_items);
    }
}

}  // namespace hogk
