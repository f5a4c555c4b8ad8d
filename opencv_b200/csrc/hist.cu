// hist.cu -- fused gradient + orientation voting + block histogram + L2-Hys kernel (sm_100a).
//
// Replaces, in ONE pass over every pyramid level of a frame, the reference's four kernels
//   compute_gradients_8UC1/8UC4   modules/cudaobjdetect/src/cuda/hog.cu:624-821
//   compute_hists                 hog.cu:133-279
//   normalize_hists               hog.cu:287-386
// and their 10 B/px grad/qangle round trip through HBM: a level pixel is read once, and the only output is the
// normalised 36-float histogram of every 16x16 block (2.25 B/px).  Arithmetic follows the CPU twin
// (objdetect/src/hog.cpp:239-550 gradient with the fastAtan32f polynomial, :619-858 block window,
// :1119-1248 L2-Hys) because that is the implementation the oracle can run; see DESIGN.md.
//
// The kernel is bound by instruction issue and shared-memory bandwidth, not by HBM ("hist kernel"),
// so the design minimises instructions per pixel:
//   * persistent grid, one CTA per SM, warps are INDEPENDENT workers pulling (level, tile, strip) items from
//     a global queue (largest first): no block-wide barrier anywhere in the main loop;
//   * a warp owns a tile of 120 pixel columns (15 cells -> 14 blocks); lane l in 1..30 owns 4 adjacent columns
//     (one 16-byte load per row for 8UC4), lanes 0 and 31 carry the halo column on each side.  The warp walks
//     DOWN the strip: vertical neighbours live in registers, lateral ones in the same thread or one shuffle away;
//   * the block window weight is separable, w(i,j) = A_cy[i] * A_cx[j] (Gaussian x bilinear cell weight), so votes
//     are first accumulated VERTICALLY into per-column, per-bin partial sums in shared memory (private per thread:
//     no atomics, fixed order, bit-reproducible); each pixel feeds 3 live accumulator slots, two scalar
//     read-modify-writes each (bin0 += mag*(1-a), bin0+1 += mag*a); a tenth accumulator row takes the second vote
//     of bin 8 and is folded into bin 0 by the horizontal pass;
//   * orientation binning is branch- and conversion-free: angle in bin units with its sign taken from
//     sign(dx)^sign(dy), floor() by the 2^23+16 magic add, whose mantissa is at the same time the shared-memory
//     row offset of the bin (no wrap needed: the 9 reachable rows are a cyclic rotation of the 9 bins);
//   * every 4 rows one accumulator slot is complete; a HORIZONTAL pass (lane <-> (row, cell)) contracts the 120
//     columns with the 24-tap horizontal weights into the four cell sums of 14 blocks;
//   * once both cell rows of a block row are in, 4 lanes per block do L2-Hys with shuffle reductions; the 2016-byte
//     block row leaves as coalesced 16-byte stores.
#include "hog_internal.h"

#include <cstdlib>
#include <cstring>

namespace hogk {

namespace {

#ifndef HOGB200_HIST_WARPS
#define HOGB200_HIST_WARPS 12
#endif
#ifndef HOGB200_LUT_COPIES
#define HOGB200_LUT_COPIES 32
#endif
constexpr int kWarps = HOGB200_HIST_WARPS;         // independent workers per CTA (one CTA per SM)
constexpr int kThreads = kWarps * 32;
constexpr int kTileCells = 15;                     // 30 voting lanes x 4 columns = 120 columns
constexpr int kTileBlocks = kTileCells - 1;        // 14 blocks per tile row
constexpr int kLutCopies = HOGB200_LUT_COPIES;     // gamma LUT replication (bank spreading; 32 = one copy per lane)
constexpr int kPF = 4;                             // rows of load prefetch

constexpr int kVote = 32;                          // elements per accumulator row and column class: lanes 1..30 vote into 0..29, the
                                                   // halo lanes 0 / 31 into 31 / 30, which nobody reads.  The row stride must stay a
                                                   // multiple of 128 B: lanes vote into DIFFERENT rows (bins) at once, and only then is
                                                   // the bank a function of the lane alone (a 30-element row was tried: +47 % time)
constexpr int kRows = kBins + 1;                   // accumulator rows per slot: row 9 aliases row 0 (second vote of row 8)
constexpr int kColBytes = kVote * 4;               // one column class j of an accumulator row: [idx] float = 128 B
constexpr int kBinBytes = 4 * kColBytes;           // one accumulator row: [j][idx] float = 512 B
constexpr int kSlotBytes = kRows * kBinBytes;      // 5120 B
constexpr int kStgBytes = kTileBlocks * kHist * 4; // 2016 B: the normalisation plane, aliased onto a slot that was just consumed
constexpr float kMagic = 8388624.f;                // 2^23 + 16: q + kMagic has ulp 1 for q in [-4.5, 4.5], mantissa = 16 + n
static_assert(kStgBytes <= kSlotBytes && kStgBytes % 16 == 0 && kSlotBytes % 16 == 0, "staging plane must fit a slot");

// HOGB200_HIST_BULK=1 (measurement variant, tools/build_variant.sh): the rows of a tile reach the warp through a ring of
// cp.async.bulk copies into shared memory (the TMA-staged form north_star (2) prescribes) instead of one LDG.128 per
// lane and row.  8UC4 only; needs HOGB200_LUT_COPIES=16 to make room for the rings.  DESIGN.md 4.1 has both timings.
#ifndef HOGB200_HIST_BULK
#define HOGB200_HIST_BULK 0
#endif
struct __align__(16) WarpSmem {
    float acc[3][kRows][4][kVote];                 // 15,360 B  vertical partial sums
#if HOGB200_HIST_BULK
    unsigned char ring[kPF][512];                  //  2,048 B  staged rows: 32 lanes x 16 B
    unsigned long long bar[kPF];                   //     32 B  one mbarrier per ring slot
#endif
};
struct __align__(16) Smem {
    float4 wrow[3][8];                             //    384 B  vertical weights of row o of a cell row, per slot, for the
                                                   //           three role rotations (slot roles rotate, slots stay put)
    float4 wh[6];                                  //     96 B  horizontal weights, paired like the accumulator loads
    float lut[256 * kLutCopies];                   // 32,768 B  gamma LUT, lut[v*copies + lane%copies]: one copy per lane, no conflicts
    WarpSmem w[kWarps];                            // 12 x 15,360 B
};
static_assert(sizeof(Smem) <= 227 * 1024, "shared memory");

// ---- shared-memory access to the accumulators: 32-bit shared addresses, volatile so that the batches below
// keep their order (loads first, then stores) and ptxas is free to overlap the latencies.
__device__ __forceinline__ float2 lds64(unsigned a)
{
    float2 v;
    asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(a));
    return v;
}
// one vote = two adjacent accumulator rows (bin0 and bin0 + 1), kBinBytes apart
// OFF = byte offset of the slot: the three slots sit at fixed immediates from one address register
template <int OFF>
__device__ __forceinline__ float2 lds32x2(unsigned a)
{
    float2 v;
    asm volatile("ld.shared.f32 %0, [%1+%2];" : "=f"(v.x) : "r"(a), "n"(OFF));
    asm volatile("ld.shared.f32 %0, [%1+%2];" : "=f"(v.y) : "r"(a), "n"(OFF + kBinBytes));
    return v;
}
template <int OFF>
__device__ __forceinline__ void sts32x2(unsigned a, float2 v)
{
    asm volatile("st.shared.f32 [%0+%1], %2;" ::"r"(a), "n"(OFF), "f"(v.x));
    asm volatile("st.shared.f32 [%0+%1], %2;" ::"r"(a), "n"(OFF + kBinBytes), "f"(v.y));
}
__device__ __forceinline__ void sts64_zero(unsigned a)
{
    asm volatile("st.shared.v2.f32 [%0], {%1,%1};" ::"r"(a), "f"(0.f));
}
__device__ __forceinline__ void sts128_zero(unsigned a)
{
    asm volatile("st.shared.v4.f32 [%0], {%1,%1,%1,%1};" ::"r"(a), "f"(0.f));
}

// packed FP32: d = a * b + c on both halves (Blackwell FFMA2)
__device__ __forceinline__ float2 fma2(float w, float2 b, float2 c)
{
    unsigned long long rw, rb, rc, rd;
    asm("mov.b64 %0, {%1,%1};" : "=l"(rw) : "f"(w));
    asm("mov.b64 %0, {%1,%2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
    asm("mov.b64 %0, {%1,%2};" : "=l"(rc) : "f"(c.x), "f"(c.y));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(rw), "l"(rb), "l"(rc));
    float2 d;
    asm("mov.b64 {%0,%1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
    return d;
}

__device__ __forceinline__ float2 fma2v(float2 w, float2 b, float2 c)
{
    unsigned long long rw, rb, rc, rd;
    asm("mov.b64 %0, {%1,%2};" : "=l"(rw) : "f"(w.x), "f"(w.y));
    asm("mov.b64 %0, {%1,%2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
    asm("mov.b64 %0, {%1,%2};" : "=l"(rc) : "f"(c.x), "f"(c.y));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(rw), "l"(rb), "l"(rc));
    float2 d;
    asm("mov.b64 {%0,%1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
    return d;
}

// packed FP32 add / sub / mul (round-to-nearest: same results as the scalar operations)
__device__ __forceinline__ float2 pk_op2(int op, float2 a, float2 b)
{
    unsigned long long ra, rb, rd;
    asm("mov.b64 %0, {%1,%2};" : "=l"(ra) : "f"(a.x), "f"(a.y));
    asm("mov.b64 %0, {%1,%2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
    if (op == 0)      asm("add.rn.f32x2 %0, %1, %2;" : "=l"(rd) : "l"(ra), "l"(rb));
    else if (op == 1) asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(rd) : "l"(ra), "l"(rb));
    else              asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(rd) : "l"(ra), "l"(rb));
    float2 d;
    asm("mov.b64 {%0,%1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
    return d;
}
__device__ __forceinline__ float2 sub2(float2 a, float2 b) { return pk_op2(1, a, b); }
__device__ __forceinline__ float2 mul2(float2 a, float2 b) { return pk_op2(2, a, b); }

__device__ __forceinline__ float sqrt_approx(float x)
{
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float rcp_approx(float x)
{
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

#if HOGB200_HIST_BULK
__device__ __forceinline__ void hb_init(unsigned mbar)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mbar));
}
__device__ __forceinline__ void hb_issue(unsigned dst, const void* src, unsigned bytes, unsigned mbar)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(mbar) : "memory");
}
__device__ __forceinline__ void hb_wait(unsigned mbar, unsigned parity)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "HBW_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"
        "@p bra HBD_%=;\n\t"
        "bra HBW_%=;\n\t"
        "HBD_%=:\n\t}" ::"r"(mbar), "r"(parity), "r"(0x989680u) : "memory");
}
#endif

template <int CN> struct RawT { typedef uint4 T; };
template <> struct RawT<1> { typedef unsigned int T; };

template <int CN>
__device__ __forceinline__ typename RawT<CN>::T load_raw(const uint8_t* p)
{
    return __ldg(reinterpret_cast<const typename RawT<CN>::T*>(p));
}

// gamma LUT (hog.cpp:258-300) of the NC channels of the 4 pixels of one thread-row
template <int CN>
__device__ __forceinline__ void unpack(const typename RawT<CN>::T& raw, const unsigned (&sel)[3],
                                       const float* __restrict__ lutl, float (&v)[4][3]);
template <>
__device__ __forceinline__ void unpack<4>(const uint4& raw, const unsigned (&sel)[3], const float* __restrict__ lutl,
                                          float (&v)[4][3])
{
    const unsigned w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
    for (int j = 0; j < 4; j++)
#pragma unroll
        for (int c = 0; c < 3; c++) v[j][c] = lutl[__byte_perm(w[j], 0, sel[c]) * kLutCopies];
}
template <>
__device__ __forceinline__ void unpack<1>(const unsigned int& raw, const unsigned (&)[3], const float* __restrict__ lutl,
                                          float (&v)[4][3])
{
#pragma unroll
    for (int j = 0; j < 4; j++) v[j][0] = lutl[__byte_perm(raw, 0, 0x4440 + j) * kLutCopies];
}

struct Vote { unsigned rbits; float g0, g1; };

// One pixel: centred differences, strongest channel, magnitude, orientation in bin units, interpolation weights.
// NC = channels examined.  xm = 0 on the first / last column (reflect-101 makes dx zero there, hog.cpp:306-313).
template <int NC>
__device__ __forceinline__ Vote pixel(const float (&L)[3], const float (&R)[3], const float (&U)[3], const float (&D)[3],
                                      float xm)
{
    float dx, dy, m;
    if (NC == 3) {
        // channels 0 and 1 as one packed operation each (same roundings as the scalar form), channel 2 scalar
        const float2 xm2 = make_float2(xm, xm);
        const float2 dx01 = mul2(sub2(make_float2(R[0], R[1]), make_float2(L[0], L[1])), xm2);
        const float2 dy01 = sub2(make_float2(D[0], D[1]), make_float2(U[0], U[1]));
        // unfused, like the CPU's SSE2 (dx*dx)+(dy*dy): channel selection is discontinuous, an FMA here would break
        // exact ties such as (a,b) vs (b,a) differently from the reference.  The two sums stay scalar: ptxas
        // contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2 in spite of the .rn.
        const float2 dxx = mul2(dx01, dx01), dyy = mul2(dy01, dy01);
        const float mm0 = __fadd_rn(dxx.x, dyy.x), mm1 = __fadd_rn(dxx.y, dyy.y);
        const float dx2 = (R[2] - L[2]) * xm, dy2 = D[2] - U[2];
        const float mm2 = __fadd_rn(__fmul_rn(dx2, dx2), __fmul_rn(dy2, dy2));
        // strict >: the first channel in processing order wins ties (hog.cu:694-712, CPU SSE2 body hog.cpp:393-403);
        // the caller reverses the processing order in the last w%4 columns, where the CPU's scalar tail lets the
        // highest channel win (hog.cpp:455-477)
        dx = dx01.x; dy = dy01.x; m = mm0;
        const bool p1 = mm1 > m;
        dx = p1 ? dx01.y : dx; dy = p1 ? dy01.y : dy; m = fmaxf(m, mm1);
        const bool p2 = mm2 > m;
        dx = p2 ? dx2 : dx; dy = p2 ? dy2 : dy; m = fmaxf(m, mm2);
    } else {
        dx = (R[0] - L[0]) * xm; dy = D[0] - U[0];
        m = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
    }
    const float mag = sqrt_approx(m);      // = sqrt(dx*dx + dy*dy) of the winning channel (magnitude32f)
    // fastAtan32f (core/src/mathfuncs_core.simd.hpp:45-71), in units of one bin (20 degrees): coefficients * 9/pi
    const float ax = fabsf(dx), ay = fabsf(dy);
    const float mx = fmaxf(ax, ay), mn = fminf(ax, ay);
    const float c = mn * rcp_approx(mx + 2.220446049250313e-16f);
    const float c2 = c * c;
    float a = ((-0.1269862277f * c2 + 0.4457000067f) * c2 - 0.9333723052f) * c2 + 2.8641811857f;
    a *= c;
    if (ay > ax) a = 4.5f - a;
    // orientation modulo 180 degrees: the first-quadrant angle if dx and dy have the same sign, its negative
    // otherwise.  q is in [-4.5, 4.5]; t = q - 0.5 is the reference's `angle*nbins/pi - 0.5` modulo 9.
    const unsigned s = (__float_as_uint(dx) ^ __float_as_uint(dy)) & 0x80000000u;
    const float q = __uint_as_float(__float_as_uint(a) ^ s);
    // n = round-to-nearest-even(q) = floor(t) + 1 (a tie only moves a zero vote between two bins);
    // the mantissa of r holds n, so r << 9 is n * 512 bytes: the accumulator row of bin floor(t), rows being
    // numbered n + 4 = 0..8 <-> bins 4,5,6,7,8,0,1,2,3 (hog.cpp:531-547 without the wrap)
    const float r = q + kMagic;
    const float nf = r - kMagic;
    const float alpha = (q - nf) + 0.5f;
    Vote v;
    v.rbits = __float_as_uint(r);
    v.g1 = mag * alpha;
    v.g0 = mag - v.g1;
    return v;
}

// Horizontal contraction of one finished vertical accumulator (cell row `cy` of a block row); also zeroes the
// accumulator for its next use.  lane <-> (accumulator row, cell): 2 rows x 15 cells per pass, 5 passes.  The results stay
// in registers: ax[it] = (block `cell`, cx 0), bx[it] = (block `cell` - 1, cx 1), bin of accumulator row 2 it + half.
__device__ __forceinline__ void horizontal_pass(unsigned slot_addr, const float4* __restrict__ wh, int lane, float (&ax)[5], float (&bx)[5])
{
    const int cell = lane & 15, half = lane >> 4;
    // the sixteenth lane of each half has no cell: it reads (and clears) the two elements the halo lanes vote into;
    // whatever they hold only reaches sums nobody stores
    const bool real = true;
    const int ce = cell;
    // weights paired like the loads below: (column j, column 4 + j) of the cell
    // (registers, loaded from shared memory: FFMA2 takes no constant-bank operand)
    float2 w00[4], w11[4], wx[4];
#pragma unroll
    for (int j = 0; j < 2; j++) {
        const float4 a = wh[j], b = wh[2 + j], c = wh[4 + j];
        w00[2 * j] = make_float2(a.x, a.y); w00[2 * j + 1] = make_float2(a.z, a.w);
        w11[2 * j] = make_float2(b.x, b.y); w11[2 * j + 1] = make_float2(b.z, b.w);
        wx[2 * j] = make_float2(c.x, c.y);  wx[2 * j + 1] = make_float2(c.z, c.w);
    }
#pragma unroll
    for (int it = 0; it < 5; it++) {
        const int row = it * 2 + half;
        const bool act = row < kBins;        // row 9 only exists as the alias of row 0, folded in below
        const unsigned base = slot_addr + (act ? row : 0) * kBinBytes + ce * 8;
        float2 v[4];
#pragma unroll
        for (int j = 0; j < 4; j++) v[j] = lds64(base + j * kColBytes);  // columns j and 4+j of the cell
        if (it == 0 && half == 0) {          // row 0 += row 9 (second votes of row 8: both are the same bin)
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const float2 t = lds64(base + 9 * kBinBytes + j * kColBytes);
                v[j].x += t.x; v[j].y += t.y;
                if (real) sts64_zero(base + 9 * kBinBytes + j * kColBytes);
            }
        }
        if (act && real) {
#pragma unroll
            for (int j = 0; j < 4; j++) sts64_zero(base + j * kColBytes);
        }
        float2 p00 = make_float2(0.f, 0.f), p11 = p00, px = p00;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            p00 = fma2v(w00[j], v[j], p00);
            p11 = fma2v(w11[j], v[j], p11);
            px = fma2v(wx[j], v[j], px);
        }
        const float s00 = p00.x + p00.y, s11 = p11.x + p11.y, s01 = px.x, s10 = px.y;
        // block k, cx=0 = own s00 + s01 of cell k+1 ; block k-1, cx=1 = own s11 + s10 of cell k-1
        ax[it] = s00 + __shfl_down_sync(0xffffffffu, s01, 1);
        bx[it] = s11 + __shfl_up_sync(0xffffffffu, s10, 1);
    }
}

// results of the two horizontal passes of a block row -> the staging plane [block][cx*2+cy][bin]
__device__ __forceinline__ void stage_results(unsigned stg, int lane, int cy, const float (&ax)[5], const float (&bx)[5])
{
    const int cell = lane & 15, half = lane >> 4;
#pragma unroll
    for (int it = 0; it < 5; it++) {
        const int row = it * 2 + half;
        if (row >= kBins) continue;
        const int tb = row + 4 >= kBins ? row + 4 - kBins : row + 4;     // bin of this accumulator row
        if (cell < kTileBlocks)
            asm volatile("st.shared.f32 [%0], %1;" ::"r"(stg + 4 * (cell * kHist + (0 * 2 + cy) * kBins + tb)), "f"(ax[it]));
        if (cell >= 1 && cell < kTileCells)
            asm volatile("st.shared.f32 [%0], %1;" ::"r"(stg + 4 * ((cell - 1) * kHist + (1 * 2 + cy) * kBins + tb)), "f"(bx[it]));
    }
}

__device__ __forceinline__ void zero_bytes(unsigned addr, int bytes, int lane)      // bytes: a multiple of 16
{
    for (int i = lane * 16; i < bytes; i += 32 * 16) sts128_zero(addr + i);
}
__device__ __forceinline__ void zero_slot(unsigned slot_addr, int lane) { zero_bytes(slot_addr, kSlotBytes, lane); }

__device__ __forceinline__ float lds32(unsigned a)
{
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
    return v;
}

// L2-Hys (hog.cpp:1119-1248 / hog.cu:333-349) of the staged block row and coalesced store; 4 lanes per block,
// one cell (9 bins) per lane.  The plane sits in a slot that was just consumed: it is zeroed again on the way out.
__device__ __forceinline__ void normalize_store(unsigned stg, int lane, int nblk, float thr, float* __restrict__ out_row)
{
#pragma unroll 1
    for (int g = 0; g < 2; g++) {
        // (block, cell) -> 9 consecutive staging floats; the two lanes past block 13 recompute block 13 (no store)
        const bool live = g * 8 + (lane >> 2) < kTileBlocks;
        const int idx = live ? (g * 32 + lane) * kBins : (kTileBlocks * 4 - 4 + (lane & 3)) * kBins;
        float v[kBins];
        float ss = 0.f;
#pragma unroll
        for (int b = 0; b < kBins; b++) {
            v[b] = lds32(stg + 4 * (idx + b));
            ss = fmaf(v[b], v[b], ss);
        }
        ss += __shfl_xor_sync(0xffffffffu, ss, 1);
        ss += __shfl_xor_sync(0xffffffffu, ss, 2);
        const float sc1 = rcp_approx(sqrt_approx(ss) + kHist * 0.1f);
        float s2 = 0.f;
#pragma unroll
        for (int b = 0; b < kBins; b++) {
            v[b] = fminf(v[b] * sc1, thr);
            s2 = fmaf(v[b], v[b], s2);
        }
        s2 += __shfl_xor_sync(0xffffffffu, s2, 1);
        s2 += __shfl_xor_sync(0xffffffffu, s2, 2);
        const float sc2 = rcp_approx(sqrt_approx(s2) + 1e-3f);
        __syncwarp();       // every lane has read its inputs before any lane overwrites the plane
        if (live) {
#pragma unroll
            for (int b = 0; b < kBins; b++) asm volatile("st.shared.f32 [%0], %1;" ::"r"(stg + 4 * (idx + b)), "f"(v[b] * sc2));
        }
    }
    __syncwarp();
    float4* dst = reinterpret_cast<float4*>(out_row);
    for (int i = lane; i < kTileBlocks * (kHist / 4); i += 32) {
        float4 t;
        asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(t.x), "=f"(t.y), "=f"(t.z), "=f"(t.w) : "r"(stg + 16 * i));
        sts128_zero(stg + 16 * i);           // the slot goes back to its accumulator role
        if (i < nblk * (kHist / 4)) dst[i] = t;
    }
}

template <int CN>
__device__ __forceinline__ void process_item(Smem& s, WarpSmem& ws, const HistItem it, const LevelDesc& L,
                                             const FrameSet& fs, const BlockWeights& W,
                                             float l2hys, float* __restrict__ hist_all, int lane, unsigned& bulk_seq)
{
    const int frame = it.level >> 8;
    const uint8_t* __restrict__ base = fs.base[frame];
    const int base_pitch = fs.pitch[frame];
    float* __restrict__ hist = hist_all + frame * fs.hist_stride;
    constexpr int NC = CN == 4 ? 3 : 1;
    constexpr int kChunk = 4 * CN;                          // bytes of 4 pixels
    const int w = L.w, h = L.h, pitch = L.img ? L.pitch : base_pitch;
    const uint8_t* __restrict__ img = L.img ? L.img + frame * fs.pyr_stride : base;  // level 0 aliases the frame
    const int xt = it.bx0 * kCell - 4 + 4 * lane;           // first of this lane's 4 columns (lane 0 / 31: halo)
    const int chunk = min(max(xt >> 2, 0), (w - 1) >> 2);   // clamped: out-of-range columns never reach an output
    const uint8_t* __restrict__ colp = img + (size_t)chunk * kChunk;
#if HOGB200_HIST_BULK
    // the 16-byte chunks the 32 lanes read are one contiguous (clamped) segment of the row: lane 0's chunk .. lane 31's
    const int c_first = min(max((it.bx0 * kCell - 4) >> 2, 0), (w - 1) >> 2), c_last = min(max((it.bx0 * kCell - 4 + 124) >> 2, 0), (w - 1) >> 2);
    const uint8_t* __restrict__ segp = img + (size_t)c_first * kChunk;
    const unsigned seg_bytes = (unsigned)(c_last - c_first + 1) * 16u, lane_seg = (unsigned)(chunk - c_first) * 16u;
    const unsigned ring0 = (unsigned)__cvta_generic_to_shared(&ws.ring[0][0]), bar0 = (unsigned)__cvta_generic_to_shared(&ws.bar[0]);
#endif
    float xm[4];
#pragma unroll
    for (int j = 0; j < 4; j++) xm[j] = (xt + j <= 0 || xt + j >= w - 1) ? 0.f : 1.f;
    // channel processing order: B,G,R -- reversed in the columns of the CPU's scalar tail (x >= w & ~3)
    const bool tail = xt >= (w & ~3);
    const unsigned sel[3] = {0x4440u + (tail ? 2u : 0u), 0x4441u, 0x4440u + (tail ? 0u : 2u)};
    const bool swap_l = tail != (xt - 4 >= (w & ~3)), swap_r = tail != (xt + 4 >= (w & ~3));
    const float* __restrict__ lutl = s.lut + (lane % kLutCopies);
    const int nblk = min(kTileBlocks, L.nbx - it.bx0);      // blocks of this tile that exist

    // accumulator slots: roles P (block row above, cell row 0), Q (above, cell row 1), R (this, cell row 0);
    // lane l votes into element (l-1)&31 of every [bin][j] row, rows of one slot are 1024 B apart
    const unsigned acc0 = (unsigned)__cvta_generic_to_shared(&ws.acc[0][0][0][0]);
    // row index = n + 4; r's bit pattern is 0x4B000000 + 16 + n, and bits * kBinBytes (mod 2^32) is the row offset plus a
    // constant that is folded into the column base: one IMAD per pixel.  The halo lanes 0 / 31 vote into elements 31 / 30.
    const int vidx = (lane - 1) & 31;
    const unsigned lane_off = vidx * 4 + (4 - 16) * kBinBytes - 0x4B000000u * (unsigned)kBinBytes;
    // The slots stay at fixed offsets 0, 1, 2 x kSlotBytes from one address register (immediates in the inner loop);
    // what rotates after every cell row is which role a slot plays, i.e. which weight column it is fed with:
    // rot = 0: (P,Q,R) = slots (0,1,2); 1: (2,0,1); 2: (1,2,0).
    unsigned colbase[4];
#pragma unroll
    for (int j = 0; j < 4; j++) colbase[j] = acc0 + lane_off + j * kColBytes;
    int rot = 0, slot_p = 0;
    zero_slot(acc0 + 2 * kSlotBytes, lane);                 // R; P and Q are cleared during the first cell row
    __syncwarp();

    float pax[5], pbx[5];                                   // horizontal results of cell row 0 of the block row in flight
#pragma unroll
    for (int i = 0; i < 5; i++) pax[i] = pbx[i] = 0.f;
    int y = it.by0 * kCell;
    float U[4][3], C[4][3], D[4][3];
    typename RawT<CN>::T pf[kPF];
    {
        const int yu = y == 0 ? 1 : y - 1;                  // reflect-101: dy = 0 on the first row
        unpack<CN>(load_raw<CN>(colp + (size_t)yu * pitch), sel, lutl, U);
        unpack<CN>(load_raw<CN>(colp + (size_t)y * pitch), sel, lutl, C);
#pragma unroll
        for (int k = 0; k < kPF; k++) {
            const int yy = y + 1 + k;
#if HOGB200_HIST_BULK
            if (CN == 4) { if (lane == 0) hb_issue(ring0 + 512 * ((bulk_seq + k) % kPF), segp + (size_t)(yy >= h ? h - 2 : yy) * pitch, seg_bytes, bar0 + 8 * ((bulk_seq + k) % kPF)); }
            else
#endif
            pf[k] = load_raw<CN>(colp + (size_t)(yy >= h ? h - 2 : yy) * pitch);
        }
    }

#pragma unroll 1
    for (int cr = 0; cr <= it.nby; cr++) {
#pragma unroll 1
        for (int hf = 0; hf < 2; hf++) {
            float w0[4], w1[4], w2[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const float4 t = s.wrow[rot][hf * 4 + k];
                w0[k] = t.x; w1[k] = t.y; w2[k] = t.z;
            }
#pragma unroll
            for (int k = 0; k < 4; k++, y++) {
                // row y+1 (reflect-101 at the bottom: dy = 0 on the last row), and the prefetch kPF rows ahead
#if HOGB200_HIST_BULK
                if (CN == 4) {
                    const unsigned slot = bulk_seq % kPF;
                    hb_wait(bar0 + 8 * slot, (bulk_seq / kPF) & 1);
                    uint4 raw;
                    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(raw.x), "=r"(raw.y), "=r"(raw.z), "=r"(raw.w) : "r"(ring0 + 512 * slot + lane_seg));
                    pf[0] = *reinterpret_cast<typename RawT<CN>::T*>(&raw);
                    __syncwarp();                               // every lane has its chunk: the slot may be refilled
                    const int yy = y + 1 + kPF;
                    if (lane == 0) hb_issue(ring0 + 512 * slot, segp + (size_t)(yy >= h ? h - 2 : yy) * pitch, seg_bytes, bar0 + 8 * slot);
                    bulk_seq++;
                    unpack<CN>(pf[0], sel, lutl, D);
                } else
#endif
                {
                    unpack<CN>(pf[k % kPF], sel, lutl, D);
                    const int yy = y + 1 + kPF;
                    pf[k % kPF] = load_raw<CN>(colp + (size_t)(yy >= h ? h - 2 : yy) * pitch);
                }
                float lf[3], rt[3];
#pragma unroll
                for (int c = 0; c < NC; c++) {
                    lf[c] = __shfl_up_sync(0xffffffffu, C[3][c], 1);
                    rt[c] = __shfl_down_sync(0xffffffffu, C[0][c], 1);
                }
                if (NC == 3) {      // a neighbour on the other side of the tail boundary stores its channels reversed
                    const float l0 = lf[0], r0 = rt[0];
                    lf[0] = swap_l ? lf[2] : l0; lf[2] = swap_l ? l0 : lf[2];
                    rt[0] = swap_r ? rt[2] : r0; rt[2] = swap_r ? r0 : rt[2];
                }
                Vote v[4];
                v[0] = pixel<NC>(lf, C[1], U[0], D[0], xm[0]);
                v[1] = pixel<NC>(C[0], C[2], U[1], D[1], xm[1]);
                v[2] = pixel<NC>(C[1], C[3], U[2], D[2], xm[2]);
                v[3] = pixel<NC>(C[2], rt, U[3], D[3], xm[3]);
                // ---- vertical accumulation: 12 private read-modify-writes, loads first
                float2 a0[4], a1[4], a2[4];
                unsigned ad[4];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    ad[j] = v[j].rbits * (unsigned)kBinBytes + colbase[j];
#ifdef HOGB200_CHECKS       // tools/build_variant.sh checks -DHOGB200_CHECKS: every vote lands in this warp's own slots, rows 0..9
                    if (ad[j] < acc0 || ad[j] + kBinBytes + 4 > acc0 + kSlotBytes || (ad[j] - acc0) % 4 != 0) __trap();
#endif
                    a0[j] = lds32x2<0>(ad[j]); a1[j] = lds32x2<kSlotBytes>(ad[j]); a2[j] = lds32x2<2 * kSlotBytes>(ad[j]);
                }
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const float2 g = make_float2(v[j].g0, v[j].g1);
                    a0[j] = fma2(w0[k], g, a0[j]); a1[j] = fma2(w1[k], g, a1[j]); a2[j] = fma2(w2[k], g, a2[j]);
                }
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    sts32x2<0>(ad[j], a0[j]); sts32x2<kSlotBytes>(ad[j], a1[j]); sts32x2<2 * kSlotBytes>(ad[j], a2[j]);
                }
#pragma unroll
                for (int j = 0; j < 4; j++)
#pragma unroll
                    for (int c = 0; c < NC; c++) { U[j][c] = C[j][c]; C[j][c] = D[j][c]; }
            }
            __syncwarp();
            // P is complete after 4 rows, Q (the slot after P) after 8
            const int ds = slot_p + hf >= 3 ? slot_p + hf - 3 : slot_p + hf;
            const unsigned done = acc0 + ds * kSlotBytes;
            if (cr == 0) zero_slot(done, lane);
            else if (hf == 0) horizontal_pass(done, s.wh, lane, pax, pbx);     // cell row 0 of the block row: kept in registers
            else {
                float qax[5], qbx[5];
                horizontal_pass(done, s.wh, lane, qax, qbx);
                __syncwarp();                                                   // the slot is read out and zeroed: stage in it
                stage_results(done, lane, 0, pax, pbx);
                stage_results(done, lane, 1, qax, qbx);
                __syncwarp();
                const int by = it.by0 + cr - 1;
                normalize_store(done, lane, nblk, l2hys, hist + L.hist_off + ((long long)by * L.nbx + it.bx0) * kHist);
            }
            __syncwarp();
        }
        // rotate roles: R -> P, T (in P's slot) -> Q, Q's slot (just consumed and zeroed) -> R
        rot = rot == 2 ? 0 : rot + 1;
        slot_p = slot_p == 0 ? 2 : slot_p - 1;
    }
#if HOGB200_HIST_BULK
    if (CN == 4) {              // the rows prefetched past the strip: wait them out so that slot phases stay in step
        for (int k = 0; k < kPF; k++, bulk_seq++) hb_wait(bar0 + 8 * (bulk_seq % kPF), (bulk_seq / kPF) & 1);
        __syncwarp();
    }
#endif
}

template <int CN>
__global__ void __launch_bounds__(kThreads, 1)
hist_kernel(const __grid_constant__ FrameSet fs, const LevelDesc* __restrict__ levels,
            const HistItem* __restrict__ items, int nitems, unsigned int* __restrict__ queue, int gamma,
            const BlockWeights W, float l2hys, float* __restrict__ hist)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Smem& s = *reinterpret_cast<Smem*>(smem_raw);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    // gamma LUT (hog.cpp:258-300): sqrt(v) or v, replicated so that lanes spread over the banks
    for (int i = tid; i < 256 * kLutCopies; i += kThreads) {
        const float v = (float)(i / kLutCopies);
        s.lut[i] = gamma ? sqrtf(v) : v;
    }
    if (tid >= 32 && tid < 44) {
        // pair j = (column j, column 4 + j) of a cell: [0..3] cell k = cx0 of block k; [4..7] cell k = cx1 of block
        // k-1; [8..11] columns 0..3 -> cx0 of block k-1, columns 4..7 -> cx1 of block k
        const int j = (tid - 32) & 3, g = (tid - 32) >> 2;
        float2* wh = reinterpret_cast<float2*>(s.wh);
        wh[g * 4 + j] = g == 0 ? make_float2(W.a0[j], W.a0[4 + j])
                      : g == 1 ? make_float2(W.a1[8 + j], W.a1[12 + j])
                               : make_float2(W.a0[8 + j], W.a1[4 + j]);
    }
    if (tid < 24) {
        // row o of a cell row feeds R: block row cr, cell row 0, i = o;  Q: block row cr-1, cell row 1, i = 8+o;
        // P (o<4): block row cr-1, cell row 0, i = 8+o  or  T (o>=4): block row cr, cell row 1, i = o
        const int o = tid & 7, r = tid >> 3;
        const float wR = W.a0[o], wQ = W.a1[8 + o], wP = o < 4 ? W.a0[8 + o] : W.a1[o];
        // rotation r: P sits in slot (3 - r) % 3, Q in the next, R in the one after
        const float x = r == 0 ? wP : r == 1 ? wQ : wR;     // slot 0
        const float y = r == 0 ? wQ : r == 1 ? wR : wP;     // slot 1
        const float z = r == 0 ? wR : r == 1 ? wP : wQ;     // slot 2
        s.wrow[r][o] = make_float4(x, y, z, 0.f);
    }
    __syncthreads();

    WarpSmem& ws = s.w[warp];
    unsigned bulk_seq = 0;          // rows this warp has pushed through its ring (HOGB200_HIST_BULK)
#if HOGB200_HIST_BULK
    if (lane == 0)
        for (int i = 0; i < kPF; i++) hb_init((unsigned)__cvta_generic_to_shared(&ws.bar[i]));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncwarp();
#endif
    // work queue, one item ahead: the next index and descriptor are in flight while the current item is processed
    auto fetch = [&]() {
        int i = 0;
        if (lane == 0) i = (int)atomicAdd(&queue[0], 1u);
        return __shfl_sync(0xffffffffu, i, 0);
    };
    int idx = fetch();
    HistItem it = items[min(idx, nitems - 1)];
    while (idx < nitems) {
        const int nidx = fetch();
        const HistItem nit = items[min(nidx, nitems - 1)];
        if ((it.level >> 8) < fs.nframes)       // items of absent frames (partial group) are skipped
            process_item<CN>(s, ws, it, levels[it.level & 255], fs, W, l2hys, hist, lane, bulk_seq);
        idx = nidx; it = nit;
    }
    // the last worker to run dry re-arms the queue for the next launch on this stream
    if (lane == 0) {
        const unsigned d = atomicAdd(&queue[1], 1u);
        if (d == gridDim.x * kWarps - 1) { queue[0] = 0u; queue[1] = 0u; }
    }
}

}  // namespace

int hist_tile_blocks_x() { return kTileBlocks; }
int hist_strip_blocks_y()
{
    static const int v = [] {
        const char* e = getenv("HOGB200_STRIP");       // tuning knob: block rows per work item
        const int n = e ? atoi(e) : 0;
        return n > 0 && n <= 512 ? n : 0;
    }();
    return v;
}

int hist_workers_per_sm() { return kWarps; }

void launch_hist(const FrameSet& fs, const LevelDesc* d_levels, const HistItem* d_items,
                 int nitems, unsigned int* d_queue, int num_sms, int cn, int gamma, const BlockWeights& bw, float l2hys,
                 float* d_hist, cudaStream_t st)
{
    if (nitems <= 0) return;
    const int grid = max(1, min(num_sms, (nitems + kWarps - 1) / kWarps));
    // the opt-in shared-memory size is a per-device function attribute: set once per device, not per launch
    static bool attr_set[2][64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    bool& done = attr_set[cn == 4][dev & 63];
    if (cn == 4) {
        if (!done) done = cudaFuncSetAttribute(hist_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem)) == cudaSuccess;
        hist_kernel<4><<<grid, kThreads, sizeof(Smem), st>>>(fs, d_levels, d_items, nitems, d_queue, gamma, bw, l2hys, d_hist);
    } else {
        if (!done) done = cudaFuncSetAttribute(hist_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem)) == cudaSuccess;
        hist_kernel<1><<<grid, kThreads, sizeof(Smem), st>>>(fs, d_levels, d_items, nitems, d_queue, gamma, bw, l2hys, d_hist);
    }
}

// ---- allocation-free variant (no counterpart in the reference) -------
// The fused kernel needs scratch in device memory: the two work-queue counters, the level descriptor and the work-item
// list.  Callers that may not call cudaMalloc (CUDA graph capture, tight serving loops) get them from a file-scope
// array instead: every GPU that loads the library holds these kNoAllocScratchBytes until exit.  Every call shares the
// array and a set-up kernel rewrites it, so calls that overlap on the GPU race (see include/hogb200.h).
constexpr size_t kNoAllocScratchBytes = 256 * 1024;
constexpr int kNoAllocLevelOff = 256, kNoAllocItemsOff = 512;
constexpr int kNoAllocMaxItems = (int)((kNoAllocScratchBytes - kNoAllocItemsOff) / sizeof(HistItem));     // 32 704
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

namespace {

// writes the queue counters (zero), the LevelDesc of the single level and the (tile, strip) work items; device code
// may name the array directly
__global__ void noalloc_setup_kernel(int w, int h, int nbx, int nby, int ntx, int strip, int nitems)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) {
        unsigned int* q = reinterpret_cast<unsigned int*>(g_noalloc_scratch);
        q[0] = q[1] = q[2] = q[3] = 0u;
        LevelDesc L;
        L.img = nullptr; L.pitch = 0;                 // level 0 aliases the caller's frame (FrameSet::base)
        L.w = w; L.h = h; L.nbx = nbx; L.nby = nby; L.nwx = 0; L.nwy = 0;
        L.hist_off = 0; L.score_off = 0; L.xtab_off = 0; L.ytab_off = 0; L.mode = 0; L.rs_rows = 0;
        *reinterpret_cast<LevelDesc*>(g_noalloc_scratch + kNoAllocLevelOff) = L;
    }
    if (i < nitems) {
        const int s = i / ntx, t = i - s * ntx;       // full strips first, the partial last strip at the end of the queue
        HistItem it;
        it.level = 0; it.bx0 = (short)(t * kTileBlocks); it.by0 = (short)(s * strip);
        it.nby = (short)min(strip, nby - s * strip);
        reinterpret_cast<HistItem*>(g_noalloc_scratch + kNoAllocItemsOff)[i] = it;
    }
}

}  // namespace

size_t hist_noalloc_scratch_bytes() { return kNoAllocScratchBytes; }

// 0 = queued; 1 = image too large for the static scratch; 2 = CUDA error (cudaGetLastError holds it)
int launch_hist_noalloc(const uint8_t* d_img, int pitch, int w, int h, int cn, int gamma, const BlockWeights& bw, float l2hys,
                        float* d_out, int num_sms, cudaStream_t st, long long* items_needed)
{
    const int nbx = w >= kBlock ? (w - kBlock) / kCell + 1 : 0, nby = h >= kBlock ? (h - kBlock) / kCell + 1 : 0;
    if (nbx <= 0 || nby <= 0) return 0;
    const int ntx = (nbx + kTileBlocks - 1) / kTileBlocks;
    // strips of 4..16 block rows: about two work items per worker (a taller strip recomputes less, balances worse)
    const long long workers = (long long)num_sms * kWarps;
    int strip = (int)(((long long)nby * ntx + 2 * workers - 1) / (2 * workers));
    strip = strip < 4 ? 4 : strip > 16 ? 16 : strip;
    const long long nitems = (long long)ntx * ((nby + strip - 1) / strip);
    if (items_needed) *items_needed = nitems;
    if (nitems > kNoAllocMaxItems) return 1;
    // The array's name is a host-side stand-in in host code: the device address comes from cudaGetSymbolAddress, once
    // per device, with that device current.
    static unsigned char* scratch_of[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 2;
    if (!scratch_of[dev]) {
        void* p = nullptr;
        if (cudaGetSymbolAddress(&p, g_noalloc_scratch) != cudaSuccess || !p) return 2;
        scratch_of[dev] = static_cast<unsigned char*>(p);
    }
    unsigned char* scratch = scratch_of[dev];
    noalloc_setup_kernel<<<(int)((nitems + 255) / 256), 256, 0, st>>>(w, h, nbx, nby, ntx, strip, (int)nitems);
    FrameSet fs; memset(&fs, 0, sizeof fs);
    fs.base[0] = d_img; fs.pitch[0] = pitch; fs.nframes = 1;
    launch_hist(fs, reinterpret_cast<const LevelDesc*>(scratch + kNoAllocLevelOff),
                reinterpret_cast<const HistItem*>(scratch + kNoAllocItemsOff), (int)nitems,
                reinterpret_cast<unsigned int*>(scratch), num_sms, cn, gamma, bw, l2hys, d_out, st);
    return cudaGetLastError() == cudaSuccess ? 0 : 2;
}

}  // namespace hogk
