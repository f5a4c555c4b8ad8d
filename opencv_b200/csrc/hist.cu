// hist.cu -- fused gradient + orientation voting + block histogram + L2-Hys kernel (sm_100a).
//
// Replaces, in ONE pass over each pyramid level, the reference's four kernels
//   compute_gradients_8UC1/8UC4   modules/cudaobjdetect/src/cuda/hog.cu:624-821
//   compute_hists                 hog.cu:133-279
//   normalize_hists               hog.cu:287-386
// and their 10 B/px grad/qangle round trip through HBM: a level pixel is read once, and the only output is the
// normalised 36-float histogram of every 16x16 block (2.25 B/px).  Arithmetic follows the CPU twin
// (objdetect/src/hog.cpp:239-550 gradient with the fastAtan32f polynomial, :619-858 block window,
// :1119-1248 L2-Hys) because that is the implementation the oracle can run; see DESIGN.md.
//
// Work decomposition (one CTA = 8 warps):
//   * a CTA owns a tile of 240 pixel columns (30 cells -> 29 blocks) and a strip of block rows, and walks
//     DOWN the rows: thread <-> column, so the three vertical neighbours live in registers, the two lateral
//     ones come from warp shuffles (each warp covers 30 columns + 1 halo column each side);
//   * the block window weight is separable, w(i,j) = A_cy[i] * A_cx[j] (Gaussian x bilinear cell weight), so
//     votes are first accumulated VERTICALLY into per-column, per-bin partial sums in shared memory
//     (private per thread: no atomics, fixed order, bit-reproducible), 3 live accumulators per column;
//   * each vote pair (bin0: mag*(1-a), bin0+1: mag*a) is one 64-bit shared-memory read-modify-write: the
//     second vote is kept in a parallel "E" plane indexed by bin0 and folded into bin0+1 at the end;
//   * every 4 rows one accumulator is complete; a HORIZONTAL pass (warp <-> bin, lane <-> cell) contracts the
//     240 columns with the 12-tap horizontal weights into the cell-variant sums of 29 blocks;
//   * once both cell rows of a block row are in, 8 lanes per block do L2-Hys with shuffle reductions and write
//     the 144-byte block histogram, coalesced across the block row.
#include "hog_internal.h"

namespace hogk {

namespace {

constexpr int kWarps = 8;
constexpr int kThreads = kWarps * 32;
constexpr int kColsPerWarp = 30;
constexpr int kCols = kWarps * kColsPerWarp;       // 240 useful columns
constexpr int kCells = kCols / kCell;              // 30
constexpr int kTileBlocks = kCells - 1;            // 29 blocks per tile row
constexpr int kStripBlocks = 16;                   // block rows per strip
constexpr int kLutCopies = 32;                     // gamma LUT replicated per bank -> conflict-free

struct __align__(16) Smem {
    float2 acc[3][kBins][kCols];       // 51,840 B  vertical partial sums, (D,E) pairs
    float stgD[kTileBlocks][kHist];    //  4,176 B  horizontal results, first-vote plane
    float stgE[kTileBlocks][kHist];    //  4,176 B  second-vote plane (belongs to bin+1)
    float lut[256 * kLutCopies];       // 32,768 B  gamma LUT, lut[v*32 + lane]
};

__device__ __forceinline__ float sqrt_approx(float x)
{
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

template <int CN>
__device__ __forceinline__ void load_px(const uint8_t* __restrict__ row, int x, const float* __restrict__ lutl, float (&v)[3])
{
    if (CN == 4) {
        const uchar4 p = __ldg(reinterpret_cast<const uchar4*>(row) + x);
        v[0] = lutl[p.x * kLutCopies]; v[1] = lutl[p.y * kLutCopies]; v[2] = lutl[p.z * kLutCopies];
    } else {
        v[0] = lutl[__ldg(row + x) * kLutCopies];
    }
}

// vote of one pixel into one vertical accumulator: one 64-bit RMW
__device__ __forceinline__ void vote(float2* p, float w, float g0, float g1)
{
    float2 v = *p;
    v.x = fmaf(w, g0, v.x);
    v.y = fmaf(w, g1, v.y);
    *p = v;
}

// Horizontal contraction of one finished vertical accumulator (block row `cy` half) into the staging planes.
// warp <-> bin, lane <-> cell k of the tile.  Also zeroes the accumulator for its next use.
__device__ __forceinline__ void horizontal_pass(Smem& s, int slot, int cy, const BlockWeights& W, int warp, int lane)
{
    for (int b = warp; b < kBins; b += kWarps) {      // warp 0 also takes bin 8
        float2 v[8];
        if (lane < kCells) {
            float4* p = reinterpret_cast<float4*>(&s.acc[slot][b][lane * 8]);
#pragma unroll
            for (int q = 0; q < 4; q++) {
                float4 t = p[q];
                v[2 * q] = make_float2(t.x, t.y);
                v[2 * q + 1] = make_float2(t.z, t.w);
                p[q] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
        } else {
#pragma unroll
            for (int q = 0; q < 8; q++) v[q] = make_float2(0.f, 0.f);
        }
        float2 s00 = make_float2(0.f, 0.f), s11 = s00, s01 = s00, s10 = s00;
#pragma unroll
        for (int p = 0; p < 8; p++) {
            s00.x = fmaf(W.a0[p], v[p].x, s00.x);     s00.y = fmaf(W.a0[p], v[p].y, s00.y);      // cell k = cx0 of block k
            s11.x = fmaf(W.a1[8 + p], v[p].x, s11.x); s11.y = fmaf(W.a1[8 + p], v[p].y, s11.y);  // cell k = cx1 of block k-1
            if (p < 4) { s01.x = fmaf(W.a0[8 + p], v[p].x, s01.x); s01.y = fmaf(W.a0[8 + p], v[p].y, s01.y); }  // -> cx0 of block k-1
            else       { s10.x = fmaf(W.a1[p], v[p].x, s10.x);     s10.y = fmaf(W.a1[p], v[p].y, s10.y); }      // -> cx1 of block k
        }
        // block k, cx=0  = own s00 + s01 of cell k+1 ; block k-1, cx=1 = own s11 + s10 of cell k-1
        const float ax = s00.x + __shfl_down_sync(0xffffffffu, s01.x, 1);
        const float ay = s00.y + __shfl_down_sync(0xffffffffu, s01.y, 1);
        const float bx = s11.x + __shfl_up_sync(0xffffffffu, s10.x, 1);
        const float by = s11.y + __shfl_up_sync(0xffffffffu, s10.y, 1);
        if (lane < kTileBlocks) { s.stgD[lane][(0 * 2 + cy) * kBins + b] = ax; s.stgE[lane][(0 * 2 + cy) * kBins + b] = ay; }
        if (lane >= 1 && lane < kCells) { s.stgD[lane - 1][(1 * 2 + cy) * kBins + b] = bx; s.stgE[lane - 1][(1 * 2 + cy) * kBins + b] = by; }
    }
}

// L2-Hys (hog.cpp:1119-1248 / hog.cu:333-349) of the staged block row and store; 8 lanes per block.
__device__ __forceinline__ void normalize_store(const Smem& s, int tid, int nblk, float thr, float* __restrict__ out_row)
{
    const int blk = tid >> 3, l = tid & 7;
    float v[5];
    float ss = 0.f;
#pragma unroll
    for (int j = 0; j < 5; j++) {
        const int idx = l + 8 * j;
        v[j] = 0.f;
        if (idx < kHist && blk < nblk) {
            const int b = idx % kBins;
            const int eidx = b ? idx - 1 : idx + kBins - 1;     // second votes of bin b-1 land in bin b
            v[j] = s.stgD[blk][idx] + s.stgE[blk][eidx];
        }
        ss = fmaf(v[j], v[j], ss);
    }
    ss += __shfl_xor_sync(0xffffffffu, ss, 1);
    ss += __shfl_xor_sync(0xffffffffu, ss, 2);
    ss += __shfl_xor_sync(0xffffffffu, ss, 4);
    const float sc1 = 1.f / (sqrtf(ss) + kHist * 0.1f);
    float s2 = 0.f;
#pragma unroll
    for (int j = 0; j < 5; j++) {
        v[j] = fminf(v[j] * sc1, thr);
        s2 = fmaf(v[j], v[j], s2);
    }
    s2 += __shfl_xor_sync(0xffffffffu, s2, 1);
    s2 += __shfl_xor_sync(0xffffffffu, s2, 2);
    s2 += __shfl_xor_sync(0xffffffffu, s2, 4);
    const float sc2 = 1.f / (sqrtf(s2) + 1e-3f);
    if (blk < nblk) {
#pragma unroll
        for (int j = 0; j < 5; j++) {
            const int idx = l + 8 * j;
            if (idx < kHist) out_row[blk * kHist + idx] = v[j] * sc2;
        }
    }
}

template <int CN>
__global__ void __launch_bounds__(kThreads, 2)
hist_kernel(const uint8_t* __restrict__ base, int base_pitch, const LevelDesc* __restrict__ levels,
            const HistItem* __restrict__ items, int gamma, const BlockWeights W, float l2hys,
            float* __restrict__ hist)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Smem& s = *reinterpret_cast<Smem*>(smem_raw);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    const HistItem it = items[blockIdx.x];
    const LevelDesc& L = levels[it.level];
    const int w = L.w, h = L.h, pitch = L.img ? L.pitch : base_pitch;
    const uint8_t* __restrict__ img = L.img ? L.img : base;   // level 0 aliases the frame

    // gamma LUT (hog.cpp:258-300): sqrt(v) or v, replicated so that lane l only touches bank l
    for (int i = tid; i < 256 * kLutCopies; i += kThreads) {
        const float v = (float)(i / kLutCopies);
        s.lut[i] = gamma ? sqrtf(v) : v;
    }
    {   // zero the three vertical accumulators once; afterwards the horizontal pass re-zeroes what it consumes
        float4* p = reinterpret_cast<float4*>(&s.acc[0][0][0]);
        for (int i = tid; i < (int)(sizeof(s.acc) / 16); i += kThreads) p[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    __syncthreads();

    const float* lutl = s.lut + lane;
    const int c = warp * kColsPerWarp + lane - 1;          // tile column of this thread, -1 and 240 are halo
    const int x = it.bx0 * kCell + c;                      // level column
    const int xl = min(max(x, 0), w - 1);                  // clamped for loads
    const bool voter = lane >= 1 && lane <= kColsPerWarp;  // owns accumulators
    const bool tail = x >= (w & ~3);
    const bool xedge = (x <= 0) || (x >= w - 1);           // reflect-101 => dx = 0 (hog.cpp:306-313, hog.cu:646-654)
    const int nblk = min(kTileBlocks, L.nbx - it.bx0);     // blocks of this tile that exist
    const int ccl = voter ? c : 0;

    int y = it.by0 * kCell;
    float U[3], C[3], D[3];
    load_px<CN>(img + (size_t)max(y - 1, 0) * pitch, xl, lutl, U);
    load_px<CN>(img + (size_t)min(y, h - 1) * pitch, xl, lutl, C);

    int sp = 0, sq = 1, sr = 2;     // accumulator slots of roles P (block row above, cell row 0), Q (above, cell row 1), R (this, cell row 0)

    for (int cr = 0; cr <= it.nby; cr++) {
        const bool first = cr == 0, last = cr == it.nby;
#pragma unroll
        for (int o = 0; o < 8; o++, y++) {
            load_px<CN>(img + (size_t)min(y + 1, h - 1) * pitch, xl, lutl, D);
            // ---- gradient: centred differences on gamma-corrected values, strongest channel wins
            float dx, dy;
            {
                const bool yedge = (y == 0) || (y >= h - 1);
                float bdx = 0.f, bdy = 0.f, bm = -1.f;
#pragma unroll
                for (int ch = 0; ch < (CN == 4 ? 3 : 1); ch++) {
                    const float lft = __shfl_up_sync(0xffffffffu, C[ch], 1);
                    const float rgt = __shfl_down_sync(0xffffffffu, C[ch], 1);
                    const float ddx = xedge ? 0.f : rgt - lft;
                    const float ddy = yedge ? 0.f : D[ch] - U[ch];
                    // unfused, like the CPU's SSE2 (dx*dx)+(dy*dy): channel selection is discontinuous, an FMA here
                    // would break exact ties such as (a,b) vs (b,a) differently from the reference
                    const float m = __fadd_rn(__fmul_rn(ddx, ddx), __fmul_rn(ddy, ddy));
                    // strict >: the lowest channel wins ties (hog.cu:694-712, CPU SSE2 body hog.cpp:393-403); in the
                    // last w%4 columns the CPU's scalar tail lets the highest channel win them (hog.cpp:455-477)
                    if (m > bm || (tail && m == bm)) { bm = m; bdx = ddx; bdy = ddy; }
                }
                dx = bdx; dy = bdy;
#pragma unroll
                for (int ch = 0; ch < (CN == 4 ? 3 : 1); ch++) { U[ch] = C[ch]; C[ch] = D[ch]; }
            }
            // ---- magnitude and orientation (cartToPolar: mathfuncs_core.simd.hpp:45-71, 194-225)
            const float mag = sqrt_approx(fmaf(dx, dx, dy * dy));
            const float ax = fabsf(dx), ay = fabsf(dy);
            const float mx = fmaxf(ax, ay), mn = fminf(ax, ay);
            const float cc = __fdividef(mn, mx + 2.220446049250313e-16f);
            const float c2 = cc * cc;
            float a = (((-0.04432655554792128f * 57.29577951308232f * c2 + 0.1555786518463281f * 57.29577951308232f) * c2
                        - 0.3258083974640975f * 57.29577951308232f) * c2 + 0.9997878412794807f * 57.29577951308232f) * cc;
            if (ay > ax) a = 90.f - a;
            if (dx < 0.f) a = 180.f - a;
            if (dy < 0.f) a = 360.f - a;
            // ---- bin interpolation (hog.cpp:531-547): 9 bins over 180 degrees
            float t = (a * 0.017453292519943295f) * 2.864788975654116f - 0.5f;
            const float fl = floorf(t);
            const float alpha = t - fl;
            int bin0 = (int)fl;
            bin0 = bin0 < 0 ? bin0 + kBins : (bin0 >= kBins ? bin0 - kBins : bin0);
            const float g1 = mag * alpha, g0 = mag * (1.f - alpha);
            // ---- vertical accumulation: this row feeds 3 of the live accumulators
            if (voter) {
                float2* base = &s.acc[0][bin0][ccl];
                constexpr int kSlot = kBins * kCols;
                if (!last) vote(base + sr * kSlot, W.a0[o], g0, g1);                    // R: block row cr,   cell row 0, i = o
                if (!first) vote(base + sq * kSlot, W.a1[8 + o], g0, g1);               // Q: block row cr-1, cell row 1, i = 8+o
                if (o < 4) { if (!first) vote(base + sp * kSlot, W.a0[8 + o], g0, g1); }  // P: block row cr-1, cell row 0, i = 8+o
                else       { if (!last) vote(base + sp * kSlot, W.a1[o], g0, g1); }       // T: block row cr,   cell row 1, i = o
            }
            if (o == 3 && !first) {
                __syncthreads();
                horizontal_pass(s, sp, 0, W, warp, lane);
                __syncthreads();
            }
        }
        if (!first) {
            __syncthreads();
            horizontal_pass(s, sq, 1, W, warp, lane);
            __syncthreads();
            const int by = it.by0 + cr - 1;
            normalize_store(s, tid, nblk, l2hys, hist + L.hist_off + ((long long)by * L.nbx + it.bx0) * kHist);
        }
        // rotate roles: R -> P, T (in P's slot) -> Q, Q's slot (just consumed and zeroed) -> R
        const int t = sp; sp = sr; sr = sq; sq = t;
    }
}

}  // namespace

int hist_tile_blocks_x() { return kTileBlocks; }
int hist_strip_blocks_y() { return kStripBlocks; }

void launch_hist(const uint8_t* base, int base_pitch, const LevelDesc* d_levels, const HistItem* d_items,
                 int nitems, int cn, int gamma, const BlockWeights& bw, float l2hys, float* d_hist,
                 cudaStream_t st)
{
    if (nitems <= 0) return;
    // per-device attribute (cheap host call); the handle may live on any device
    cudaFuncSetAttribute(hist_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
    cudaFuncSetAttribute(hist_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
    if (cn == 4)
        hist_kernel<4><<<nitems, kThreads, sizeof(Smem), st>>>(base, base_pitch, d_levels, d_items, gamma, bw, l2hys, d_hist);
    else
        hist_kernel<1><<<nitems, kThreads, sizeof(Smem), st>>>(base, base_pitch, d_levels, d_items, gamma, bw, l2hys, d_hist);
}

}  // namespace hogk
