// svm.cu -- linear-SVM window scoring, thresholding and on-GPU candidate compaction; descriptor extraction.
//
// Replaces classify_hists / compute_confidence_hists (modules/cudaobjdetect/src/cuda/hog.cu:394-514), which
// spend one 256-thread CTA per window and re-read every block histogram up to 105 times through L1/L2, the
// blocking per-level label download + CPU scan (cudaobjdetect/src/hog.cpp:369-411), and
// extract_descrs_by_rows/cols (hog.cu:520-618).
//
// score(wy,wx) = bias + sum_{by<WBY} < hist[wy+by][wx .. wx+WBX) , coef[by] >      (WBX*36 contiguous floats)
// is a strided 1-D correlation along x summed over WBY block rows.  It is a thin GEMV per window, so it runs on
// the FP32 pipes (packed FFMA2), and the design goal is to keep that pipe issuing:
//   * persistent kernel, one CTA per SM: 8 consumer warps (two per SM sub-partition -- FFMA2 throughput collapses
//     when the sub-partitions hold unequal numbers of busy warps, tools/micro/probe2.cu) + 1 producer warp;
//   * a tile is 32 window rows of one level, 32 windows wide for one detector or 16 wide for two detectors scored in
//     the same pass.  The producer streams the tile's block rows (and then the next tile's) through a shared-memory
//     ring, one bulk async copy per row, signalling per-slot mbarriers;
//   * a consumer warp owns 8 window rows x 16 window columns of the tile for one detector, for ALL of k: the warps
//     4..7 take the second column group (one detector) or the second detector (two): both detectors read the same
//     ring rows, so the histograms cross shared memory once per pass whatever the number of detectors.
//     lane = (q, p), q = float4 number 0..7 of the 36-float block histogram (k-split across the lanes of a
//     quarter-warp, so its LDS.128 of a block is 128 contiguous bytes: conflict-free), p = pair of window rows.
//     The ninth float4 is shared out: lane q also handles it for windows q and q+8 of the 16.  A loaded histogram
//     float4 feeds the C windows that contain the block in x times the 2 window rows that contain it in y, into
//     32 + 4 register accumulators (float2: even/odd k);
//   * ANY window size: the window's WBX block columns are processed in `nchunks` chunks of C columns (C is the template
//     parameter, 1..8; the coefficient table in shared memory is zero-padded to nchunks * C columns), WBY is a run-time
//     loop bound.  64x128 and 48x96 are C = 7 / 5 with one chunk and no padding;
//   * warps are coupled only through the ring: a warp waits for the one row each step adds and releases the row
//     it read last, so the bands of a tile run staggered, the next tile's rows land while the current one is
//     still being scored, and warps whose 8 x 16 windows lie outside the level only pass the tile's rows on;
//   * the 8 k-partials of a window sit in the 8 lanes of a quarter-warp: a 3-stage shuffle reduce-scatter leaves
//     every lane with 4 finished windows; the bias is added, the score optionally written densely, and windows
//     with score >= hit_threshold are appended to a global candidate list (ballot + one atomic per warp).  The
//     host sorts the (tiny) list into the reference's row-major order.  No shared-memory reduction, no barrier.
#include "hog_internal.h"
#include <algorithm>

namespace hogk {

namespace {

constexpr int kCW = 8;                       // consumer warps
constexpr int kConsumerThreads = kCW * 32;   // 256
constexpr int kSvmThreads = kConsumerThreads + 32;           // + the producer warp
constexpr int kT = 16;                       // adjacent windows per thread (x)
constexpr int kBands = 4;                    // warp = (band of 8 window rows, unit); unit = column group or detector
constexpr int kBandRows = 8;
constexpr int kTileWY = kBands * kBandRows;  // 32
constexpr int kRowBlocks = 38;                             // blocks a ring slot can hold
constexpr int kRowFloats = 1376;                           // 43 x 128 B >= 38 blocks
static_assert(kRowBlocks * kHist <= kRowFloats, "ring slot too small");
constexpr int kMaxRing = 64;                 // slots; a tile reads rows t .. t+31 at step t, the rest is prefetch depth
constexpr int kMinRing = 34;
constexpr int kBarBytes = 2 * kMaxRing * 8;  // full[] + empty[] mbarriers at the start of dynamic shared memory
constexpr int kSmemBytes = 227 * 1024;
constexpr int kMaxChunkC = 8;

#ifdef HOGB200_SVM_TIMING      // experiments only (tools/prof_svm_timing.py): where a consumer warp's cycles go
__device__ unsigned long long g_svm_timing[160][8][8];
#define SVM_T(var) const unsigned var = (unsigned)clock64()
#define SVM_ACC(i, a, b) tacc[i] += (b) - (a)
#else
#define SVM_T(var)
#define SVM_ACC(i, a, b)
#endif

// how the kernel is laid out for one pass (filled by launch_svm)
struct SvmPass {
    int nchunks;        // column chunks of C blocks per window row (WBXp = nchunks * C >= win_bx)
    int wbxp;           // padded window width in blocks
    int tile_wx;        // 32 (one detector, two column groups) or 16
    int two_det;        // warps 4..7 score the second detector of the pass instead of a second column group
    int det_base;       // first detector slot of the pass
    int ring;           // slots
    int ncoefp;         // padded coefficients per detector: win_by * wbxp * 36
    int zero_fill;      // padded columns multiply stale ring contents by zero: make sure they are finite
};

// packed FP32 (Blackwell FFMA2): d = a * b + c on both halves
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c)
{
    unsigned long long ra, rb, rc, rd;
    asm("mov.b64 %0, {%1,%2};" : "=l"(ra) : "f"(a.x), "f"(a.y));
    asm("mov.b64 %0, {%1,%2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
    asm("mov.b64 %0, {%1,%2};" : "=l"(rc) : "f"(c.x), "f"(c.y));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
    float2 d;
    asm("mov.b64 {%0,%1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
    return d;
}
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned mbar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(unsigned mbar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(mbar) : "memory");
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
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(mbar), "r"(parity), "r"(0x989680u) : "memory");   // suspend-time hint: a waiting warp sleeps in
                                                                                  // hardware instead of re-issuing the test (the
                                                                                  // retry loop was 20 % of the kernel's instructions)
}
__device__ __forceinline__ bool mbar_test(unsigned mbar, unsigned parity)      // non-blocking
{
    unsigned ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(mbar), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void bulk_g2s(unsigned dst, const void* src, unsigned bytes, unsigned mbar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(mbar) : "memory");
}
// One step of a tile for one thread: block row `rowp` (the thread's first block) against coefficient row ca for
// the upper window row of the pair (accA) and cb for the lower one (accB); A / B say which of the two window rows
// contain this block row (the first step has no B, the last no A).  The window's columns come in `nchunks` chunks
// of C blocks: chunk c of the coefficients meets the histograms C * c blocks further right.
template <int C, bool MULTI, bool A, bool B>
__device__ __forceinline__ void svm_step(float2 (&accA)[kT], float2 (&accB)[kT], float2 (&exA)[2], float2 (&exB)[2],
                                         const float* __restrict__ ca, const float* __restrict__ cb,
                                         const float* __restrict__ rowp, int q, int nchunks)
{
    // MULTI = false: windows of at most 8 block columns (64x128, 48x96, ...) are one chunk -- straight-line code
#pragma unroll 1
    for (int ch = 0; ch < (MULTI ? nchunks : 1); ch++, ca += C * kHist, cb += C * kHist, rowp += C * kHist) {
        {   // float4 number q of every block, all 16 windows
            float4 wa[C], wb[C];
#pragma unroll
            for (int bx = 0; bx < C; bx++) {
                if (A) wa[bx] = *reinterpret_cast<const float4*>(ca + bx * kHist + 4 * q);
                if (B) wb[bx] = *reinterpret_cast<const float4*>(cb + bx * kHist + 4 * q);
            }
#pragma unroll
            for (int X = 0; X < kT + C - 1; X++) {
                const float4 hv = *reinterpret_cast<const float4*>(rowp + X * kHist + 4 * q);
                // the two halves of the float4 go into the same accumulator: all first halves, then all second halves,
                // so that the two dependent FFMA2 of an accumulator are up to 14 instructions apart
#pragma unroll
                for (int bx = 0; bx < C; bx++) {
                    const int j = X - bx;
                    if (j >= 0 && j < kT) {
                        if (A) accA[j] = fma2(make_float2(hv.x, hv.y), make_float2(wa[bx].x, wa[bx].y), accA[j]);
                        if (B) accB[j] = fma2(make_float2(hv.x, hv.y), make_float2(wb[bx].x, wb[bx].y), accB[j]);
                    }
                }
#pragma unroll
                for (int bx = 0; bx < C; bx++) {
                    const int j = X - bx;
                    if (j >= 0 && j < kT) {
                        if (A) accA[j] = fma2(make_float2(hv.z, hv.w), make_float2(wa[bx].z, wa[bx].w), accA[j]);
                        if (B) accB[j] = fma2(make_float2(hv.z, hv.w), make_float2(wb[bx].z, wb[bx].w), accB[j]);
                    }
                }
            }
        }
        {   // the ninth float4 (floats 32..35): windows q and q + 8 only (lane q reads block q + bx: conflict-free)
            float4 wa[C], wb[C];
#pragma unroll
            for (int bx = 0; bx < C; bx++) {
                if (A) wa[bx] = *reinterpret_cast<const float4*>(ca + bx * kHist + 32);
                if (B) wb[bx] = *reinterpret_cast<const float4*>(cb + bx * kHist + 32);
            }
            const float* r8 = rowp + q * kHist + 32;
#pragma unroll
            for (int e = 0; e < 2; e++) {
#pragma unroll
                for (int bx = 0; bx < C; bx++) {
                    const float4 hv = *reinterpret_cast<const float4*>(r8 + (8 * e + bx) * kHist);
                    if (A) {
                        exA[e] = fma2(make_float2(hv.x, hv.y), make_float2(wa[bx].x, wa[bx].y), exA[e]);
                        exA[e] = fma2(make_float2(hv.z, hv.w), make_float2(wa[bx].z, wa[bx].w), exA[e]);
                    }
                    if (B) {
                        exB[e] = fma2(make_float2(hv.x, hv.y), make_float2(wb[bx].x, wb[bx].y), exB[e]);
                        exB[e] = fma2(make_float2(hv.z, hv.w), make_float2(wb[bx].z, wb[bx].w), exB[e]);
                    }
                }
            }
        }
    }
}

// `n` numbers the rows pushed through the ring since the kernel started; slot = n % ring and the mbarrier phase
// parity is (n / ring) & 1 on both the producer and the consumer side.
//   full[slot]   producer -> consumers: the row's bytes have landed (expect_tx / complete_tx)
//   empty[slot]  consumers -> producer: all 8 warps are done with the row (a warp also waits for and releases the
//                rows it does not read)
template <int C, bool MULTI>
// (9 warps: registers are granted to a CTA in units of 4 warps, so the budget is that of 12 warps -- 168 per thread;
// __maxnreg__(224) compiles to 215 registers without spills but cannot be launched)
__global__ void __launch_bounds__(kSvmThreads, 1)
svm_kernel(const __grid_constant__ FrameSet fs, const LevelDesc* __restrict__ levels, const SvmItem* __restrict__ items,
           int nitems, const float* __restrict__ hist, const __grid_constant__ SvmParams P, const __grid_constant__ SvmPass S)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int tid = threadIdx.x, lane = tid & 31;
    const bool producer = tid >= kConsumerThreads;
    const int ring = S.ring;
    const unsigned full0 = smem_u32(smem_raw), empty0 = full0 + 8 * kMaxRing;
    float* const coef = reinterpret_cast<float*>(smem_raw + kBarBytes);
    const int ndet_pass = S.two_det ? 2 : 1;
    const unsigned coef_bytes = ((unsigned)(ndet_pass * S.ncoefp) * 4u + 127u) & ~127u;
    float* const ringp = reinterpret_cast<float*>(smem_raw + kBarBytes + coef_bytes);
    if (tid == 0) {
        for (int i = 0; i < ring; i++) { mbar_init(full0 + 8 * i, 1); mbar_init(empty0 + 8 * i, kCW); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    const int win_by = P.win_by, win_bx = P.win_bx, wbxp = S.wbxp;
    const int tile_rows = kTileWY + win_by - 1;
    // the detectors of this pass, [detector][block row][padded block column][36], columns >= win_bx are zero
    for (int d = 0; d < ndet_pass; d++) {
        const float4* __restrict__ src = reinterpret_cast<const float4*>(P.det[S.det_base + d]);
        float4* dst = reinterpret_cast<float4*>(coef + d * S.ncoefp);
        for (int i = tid; i < S.ncoefp / 4; i += kSvmThreads) {
            const int blk = i / 9, k4 = i - blk * 9;
            const int by = blk / wbxp, bx = blk - by * wbxp;
            dst[i] = bx < win_bx ? __ldg(src + (by * win_bx + bx) * 9 + k4) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
    if (S.zero_fill)
        for (int i = tid; i < ring * (kRowFloats / 4); i += kSvmThreads) reinterpret_cast<float4*>(ringp)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncthreads();     // barriers initialised, coefficients in place
    unsigned n = 0;

    if (producer) {
        // every lane streams its own rows (lane, lane + 32 of each tile): the waits for free slots overlap, so the
        // 30-odd rows that become loadable at once when a tile ends are issued back to back
        if (S.zero_fill) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // generic-proxy zeros before async-proxy writes
        const unsigned ring0 = smem_u32(ringp);
        const int row_blocks = min(kRowBlocks, S.tile_wx + win_bx - 1);
        for (int ii = blockIdx.x; ii < nitems; ii += gridDim.x) {
            const SvmItem it = items[ii];
            if (it.frame >= fs.nframes) continue;
            const LevelDesc& L = levels[it.level];
            const int nbx = L.nbx, nby = L.nby;
            const float* plane = hist + it.frame * fs.hist_stride + L.hist_off;
            const unsigned bytes = (unsigned)max(0, min(row_blocks, nbx - it.wx0)) * (kHist * 4);
            for (int r0 = 0; r0 < tile_rows; r0 += 32) {
                const int r = r0 + lane;
                if (r < tile_rows) {
                    const unsigned m = n + r, slot = m % ring, use = m / ring;
                    if (use > 0) mbar_wait(empty0 + 8 * slot, (use - 1) & 1);
                    const int by_abs = it.wy0 + r;
                    if (by_abs < nby && bytes > 0) {         // rows below the level feed only discarded windows
                        mbar_expect_tx(full0 + 8 * slot, bytes);
                        bulk_g2s(ring0 + slot * (kRowFloats * 4),
                                 plane + ((long long)by_abs * nbx + it.wx0) * kHist, bytes, full0 + 8 * slot);
                    } else {
                        mbar_arrive(full0 + 8 * slot);
                    }
                }
                // lanes stay within 32 < ring positions of each other: a parity wait on a slot is only sound
                // once the slot's previous use has been issued
                __syncwarp();
            }
            n += tile_rows;
        }
        return;
    }

    // ---- consumers: warp w sits on sub-partition w % 4; (band + 2 * unit) % 4 == w % 4 keeps the
    // sub-partitions evenly loaded when the lower bands or the second unit of a tile have nothing to score
    const int w = tid >> 5;
    const int unit = w >> 2, band = unit ? ((w - 2) & 3) : w;
    const int cg = S.two_det ? 0 : unit;                     // column group of the tile
    const int det = S.two_det ? unit : 0;                    // detector of the pass
    const bool unit_live = S.two_det || cg * kT < S.tile_wx;
    const int dslot = S.det_base + det;
    const float* const mycoef = coef + det * S.ncoefp;
    const int crow = wbxp * kHist;                           // floats per coefficient row
    const int q = lane & 7, pr = lane >> 3;
#ifdef HOGB200_SVM_TIMING
    unsigned tacc[8] = {0, 0, 0, 0, 0, 0, 0, 0};    // dead tiles, first wait, step waits, steps, epilogue, rest wait, live tiles, total
    const unsigned t_begin = (unsigned)clock64();
#endif
    for (int ii = blockIdx.x; ii < nitems; ii += gridDim.x) {
        const SvmItem it = items[ii];
        if (it.frame >= fs.nframes) continue;
        const LevelDesc& L = levels[it.level];
        SVM_T(t_tile);
        const int vwx = L.nbx - win_bx + 1, vwy = L.nby - win_by + 1;   // dense (stride-1) window grid of the level
        const int wx0 = it.wx0 + cg * kT, wy0 = it.wy0 + band * kBandRows;
        // Every warp waits for and releases EVERY row of every tile, in ring order, also the rows (and the whole
        // tiles) it does not read: a parity wait on a slot is only sound for a warp that released the slot's
        // previous use, and the slot cannot be reused twice behind the back of a warp that still owes a release.
        const unsigned tile0 = n, pos0 = n + band * kBandRows;      // ring position of the tile's / the band's first row
        n += tile_rows;
        const int rest = tile_rows - (band * kBandRows + 7 + win_by);   // rows below the band's last one
        if (!unit_live || wx0 >= vwx || wy0 >= vwy) {               // nothing to score here
            // Rows `ring`.. of a tile reuse the slots of its rows 0..: a lane may only test a slot's NEXT use once the
            // slot's current use is known to have landed (a parity test cannot tell "not yet" from "one phase
            // ago"), and rows land out of order.  Hence the warp-wide join between the rounds.
            for (int r0 = 0; r0 < tile_rows; r0 += 32) {
                const unsigned m = tile0 + r0 + lane;
                if (r0 + lane < tile_rows) {
                    mbar_wait(full0 + 8 * (m % ring), (m / ring) & 1);
                    mbar_arrive(empty0 + 8 * (m % ring));
                }
                __syncwarp();
            }
            { SVM_T(t_e); SVM_ACC(0, t_tile, t_e); }
            continue;
        }
        if (lane < band * kBandRows + 7) {                          // the rows above the band + the 7 rows steps 0 and 1 read
            const unsigned m = tile0 + lane;
            mbar_wait(full0 + 8 * (m % ring), (m / ring) & 1);
            if (lane < band * kBandRows) mbar_arrive(empty0 + 8 * (m % ring));
        }
        __syncwarp();
        SVM_T(t_first);
        SVM_ACC(1, t_tile, t_first);
#ifdef HOGB200_SVM_TIMING
        tacc[6] += 1;
#endif

        float2 accA[kT], accB[kT], exA[2], exB[2];    // (even k, odd k) partial sums of window rows 2pr / 2pr+1
#pragma unroll
        for (int j = 0; j < kT; j++) accA[j] = accB[j] = make_float2(0.f, 0.f);
        exA[0] = exA[1] = exB[0] = exB[1] = make_float2(0.f, 0.f);
        unsigned slot_l = (pos0 + 2 * pr) % ring;
        unsigned oslot = pos0 % ring;                                // slot of the band's oldest live row
        unsigned nslot = (pos0 + kBandRows - 1) % ring;              // slot / parity of the row the NEXT step adds
        unsigned npar = ((pos0 + kBandRows - 1) / ring) & 1;
        bool ready = true;
        for (int t = 0; t <= win_by; t++) {
            SVM_T(t_s0);
            if (!ready) {                                            // the row this step adds: one position before (nslot, npar)
                const unsigned ps = nslot == 0 ? ring - 1 : nslot - 1, pp = nslot == 0 ? npar ^ 1 : npar;
                mbar_wait(full0 + 8 * ps, pp);
            }
            // poll the next step's row now, so that the answer is there when this step's arithmetic is done
            ready = t == win_by || mbar_test(full0 + 8 * nslot, npar);
            SVM_T(t_s1);
            SVM_ACC(2, t_s0, t_s1);
            const float* rowp = ringp + slot_l * kRowFloats + cg * kT * kHist;
#ifdef HOGB200_CHECKS       // the row this lane reads is inside the ring, and the producer has declared it full
            if (slot_l >= (unsigned)ring || (cg * kT + kT + wbxp - 1) * kHist > kRowFloats ||
                !mbar_test(full0 + 8 * slot_l, ((pos0 + 2 * pr + t) / ring) & 1)) __trap();
#endif
            const float* ca = mycoef + t * crow;
            const float* cb = ca - crow;
            if (t == 0) svm_step<C, MULTI, true, false>(accA, accB, exA, exB, ca, cb, rowp, q, S.nchunks);
            else if (t == win_by) svm_step<C, MULTI, false, true>(accA, accB, exA, exB, ca, cb, rowp, q, S.nchunks);
            else svm_step<C, MULTI, true, true>(accA, accB, exA, exB, ca, cb, rowp, q, S.nchunks);
            __syncwarp();                                           // every lane has read row pos0 + t for the last time
            if (lane == 0) mbar_arrive(empty0 + 8 * oslot);
            slot_l = slot_l + 1 == ring ? 0 : slot_l + 1;
            oslot = oslot + 1 == ring ? 0 : oslot + 1;
            if (++nslot == ring) { nslot = 0; npar ^= 1; }
            { SVM_T(t_s2); SVM_ACC(3, t_s1, t_s2); }
        }
        SVM_T(t_epi0);
        if (lane < kBandRows - 2) mbar_arrive(empty0 + 8 * ((pos0 + win_by + 1 + lane) % ring));   // the band's last 6 rows

        // ---- the 8 lanes of a quarter-warp hold the k-partials of the same 2 x 16 windows: reduce-scatter
        float v[2 * kT];
#pragma unroll
        for (int j = 0; j < kT; j++) {
            v[j] = accA[j].x + accA[j].y;
            v[kT + j] = accB[j].x + accB[j].y;
            if ((j & 7) == q) {
                v[j] += exA[j >> 3].x + exA[j >> 3].y;
                v[kT + j] += exB[j >> 3].x + exB[j >> 3].y;
            }
        }
        float v4[4];
        {
            const bool h4 = q & 4, h2 = q & 2, h1 = q & 1;
            float v16[16], v8[8];
#pragma unroll
            for (int i = 0; i < 16; i++) {
                const float snd = h4 ? v[i] : v[16 + i], keep = h4 ? v[16 + i] : v[i];
                v16[i] = keep + __shfl_xor_sync(0xffffffffu, snd, 4);
            }
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const float snd = h2 ? v16[i] : v16[8 + i], keep = h2 ? v16[8 + i] : v16[i];
                v8[i] = keep + __shfl_xor_sync(0xffffffffu, snd, 2);
            }
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const float snd = h1 ? v8[i] : v8[4 + i], keep = h1 ? v8[4 + i] : v8[i];
                v4[i] = keep + __shfl_xor_sync(0xffffffffu, snd, 1);
            }
        }
        // lane (q, pr) now owns window row 2pr + (q >> 2), windows 8 * ((q >> 1) & 1) + 4 * (q & 1) .. + 3
        const int vy = wy0 + 2 * pr + (q >> 2);
        const int vxb = wx0 + 8 * ((q >> 1) & 1) + 4 * (q & 1);
        const bool row_ok = vy < vwy && (vy % P.stride_by == 0);
        const int oy = vy / P.stride_by;
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const float sc = v4[i] + P.bias[dslot];
            const int vx = vxb + i;
            const bool valid = row_ok && vx < vwx && (vx % P.stride_bx == 0);
            const int ox = vx / P.stride_bx;
            if (valid && P.scores) P.scores[dslot * P.score_plane + L.score_off + (long long)oy * L.nwx + ox] = sc;
            const bool hit = valid && sc >= P.thresh;
            const unsigned m = __ballot_sync(0xffffffffu, hit);
            if (m) {
                unsigned base = 0;
                if (lane == 0) base = atomicAdd(P.cand_count, __popc(m));
                base = __shfl_sync(0xffffffffu, base, 0);
                if (hit) {
                    const unsigned pos = base + __popc(m & ((1u << lane) - 1u));
                    if (pos < P.cand_cap) {
                        Candidate c2; c2.level = it.level | (dslot << 16) | (P.tag + (it.frame << 20)); c2.x = ox; c2.y = oy; c2.score = sc;
                        P.cand[pos] = c2;
                    }
                }
            }
        }
        SVM_T(t_epi1);
        SVM_ACC(4, t_epi0, t_epi1);
        if (lane < rest) {                                          // rows of the tile below the band
            const unsigned m = pos0 + 7 + win_by + lane;
            mbar_wait(full0 + 8 * (m % ring), (m / ring) & 1);
            mbar_arrive(empty0 + 8 * (m % ring));
        }
        __syncwarp();
        { SVM_T(t_e); SVM_ACC(5, t_epi1, t_e); }
    }
#ifdef HOGB200_SVM_TIMING
    if (lane == 0) {
        tacc[7] = (unsigned)clock64() - t_begin;
        for (int i = 0; i < 8; i++) g_svm_timing[blockIdx.x][w][i] += tacc[i];
    }
#endif
}

// One thread block per window: copies the window's WBY x WBX block histograms into the descriptor matrix, in
// row-by-row or column-by-column block order (extract_descrs_by_rows/cols, hog.cu:520-618).
__global__ void extract_kernel(const LevelDesc* __restrict__ level, const float* __restrict__ hist,
                               int win_bx, int win_by, int stride_bx, int stride_by, int col_by_col,
                               float* __restrict__ out)
{
    const LevelDesc& L = *level;
    const int wx = blockIdx.x, wy = blockIdx.y;
    const float* plane = hist + L.hist_off + ((long long)(wy * stride_by) * L.nbx + wx * stride_bx) * kHist;
    const int dsize = win_bx * win_by * kHist;
    float* dst = out + ((size_t)wy * gridDim.x + wx) * dsize;
    for (int i = threadIdx.x; i < dsize; i += blockDim.x) {
        const int blk = i / kHist, k = i - blk * kHist;
        int by, bx;
        if (col_by_col) { bx = blk / win_by; by = blk - bx * win_by; }
        else            { by = blk / win_bx; bx = blk - by * win_bx; }
        dst[i] = plane[((long long)by * L.nbx + bx) * kHist + k];
    }
}

}  // namespace

// ---- pass geometry (host) ---------------------------------------------------------------------------
// column chunking of a window WBX blocks wide: (C, nchunks) with nchunks * C >= WBX, C <= 8, least padding
static void chunking(int wbx, int* C, int* nchunks)
{
    *nchunks = (wbx + kMaxChunkC - 1) / kMaxChunkC;
    *C = (wbx + *nchunks - 1) / *nchunks;
}

static bool plan_pass(int win_bx, int win_by, int ndet_pass, SvmPass* S, int* C)
{
    int c, nch;
    chunking(win_bx, &c, &nch);
    const int wbxp = c * nch;
    if (kT + wbxp - 1 > kRowBlocks) return false;                     // a 16-window column group must fit a ring slot
    const long long ncoefp = (long long)win_by * wbxp * kHist;
    const long long coef_bytes = ((ndet_pass * ncoefp * 4) + 127) / 128 * 128;
    const long long ring = (kSmemBytes - kBarBytes - coef_bytes) / (kRowFloats * 4);
    if (ring < kMinRing) return false;
    S->nchunks = nch; S->wbxp = wbxp;
    S->two_det = ndet_pass == 2;
    S->tile_wx = (ndet_pass == 1 && 2 * kT + wbxp - 1 <= kRowBlocks) ? 2 * kT : kT;
    S->det_base = 0;
    S->ring = (int)std::min<long long>(ring, kMaxRing);
    S->ncoefp = (int)ncoefp;
    S->zero_fill = wbxp != win_bx;
    *C = c;
    return true;
}

bool svm_supports(int win_bx, int win_by)
{
    SvmPass S; int C;
    return win_bx >= 1 && win_by >= 1 && plan_pass(win_bx, win_by, 1, &S, &C);
}
bool svm_two_detector_pass(int win_bx, int win_by)
{
    SvmPass S; int C;
    return plan_pass(win_bx, win_by, 2, &S, &C);
}
int svm_tile_wx(int win_bx, int win_by, int ndet_pass)
{
    SvmPass S; int C;
    return plan_pass(win_bx, win_by, ndet_pass, &S, &C) ? S.tile_wx : kT;
}
int svm_tile_wy() { return kTileWY; }
int svm_tile_cost(int cols_left, int rows_left, int units)
{
    // a warp = (band, unit) works only if its 8 x 16 windows intersect the level (a second DETECTOR always has the
    // same windows as the first); the tile takes as long as the busiest sub-partition, which hosts (band s, unit 0)
    // and (band (s + 2) % 4, unit 1)
    const int nb = std::min(kBands, (rows_left + kBandRows - 1) / kBandRows);
    int busiest = 0, total = 0;
    for (int sp = 0; sp < 4; sp++) {
        const int a = (sp < nb) + (units > 1 && ((sp + 2) & 3) < nb);
        busiest = std::max(busiest, a); total += a;
    }
    (void)cols_left;
    return busiest * 16 + total;
}

template <int C, bool MULTI>
static void launch_pass(const FrameSet& fs, const LevelDesc* d_levels, const SvmItem* d_items, int nitems, const float* d_hist,
                        const SvmParams& p, const SvmPass& S, int grid, cudaStream_t st)
{
    // the opt-in shared-memory size is a per-device function attribute: set once per device, not per launch
    static bool attr_set[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!attr_set[dev & 63])
        attr_set[dev & 63] = cudaFuncSetAttribute(svm_kernel<C, MULTI>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes) == cudaSuccess;
    svm_kernel<C, MULTI><<<grid, kSvmThreads, kSmemBytes, st>>>(fs, d_levels, d_items, nitems, d_hist, p, S);
}

// items1 / items2: the tile lists for one-detector passes (svm_tile_wx(.., 1) wide) and two-detector passes (16 wide)
void launch_svm(const FrameSet& fs, const LevelDesc* d_levels, const SvmItem* d_items1, int nitems1,
                const SvmItem* d_items2, int nitems2, const float* d_hist, const SvmParams& p, int num_sms, cudaStream_t st)
{
    if (p.ndet <= 0) return;
    const bool can2 = svm_two_detector_pass(p.win_bx, p.win_by) && nitems2 > 0;
    for (int d = 0; d < p.ndet;) {
        const int nd = (p.ndet - d >= 2 && can2) ? 2 : 1;
        SvmPass S; int C = 0;
        if (!plan_pass(p.win_bx, p.win_by, nd, &S, &C)) return;        // refused earlier by the host (svm_supports)
        S.det_base = d;
        const SvmItem* items = nd == 2 ? d_items2 : d_items1;
        const int nitems = nd == 2 ? nitems2 : nitems1;
        if (nitems > 0) {
            const int grid = std::max(1, std::min(num_sms, nitems));
            // windows up to 8 blocks wide are one chunk of C = win_bx columns; wider ones are 2 or 3 chunks of C = 5 .. 8
#define HOGB200_SVM_CASE(c) case c: if (S.nchunks == 1) launch_pass<c, false>(fs, d_levels, items, nitems, d_hist, p, S, grid, st); \
                                    else launch_pass<(c < 5 ? 5 : c), true>(fs, d_levels, items, nitems, d_hist, p, S, grid, st); break;
            switch (C) {
                HOGB200_SVM_CASE(1) HOGB200_SVM_CASE(2) HOGB200_SVM_CASE(3) HOGB200_SVM_CASE(4)
                HOGB200_SVM_CASE(5) HOGB200_SVM_CASE(6) HOGB200_SVM_CASE(7)
                default: if (S.nchunks == 1) launch_pass<8, false>(fs, d_levels, items, nitems, d_hist, p, S, grid, st);
                         else launch_pass<8, true>(fs, d_levels, items, nitems, d_hist, p, S, grid, st);
            }
#undef HOGB200_SVM_CASE
        }
        d += nd;
    }
}

void launch_extract(const LevelDesc* d_level0, int nwx, int nwy, const float* d_hist, int win_bx, int win_by,
                    int stride_bx, int stride_by, int col_by_col, float* d_out, cudaStream_t st)
{
    if (nwx <= 0 || nwy <= 0) return;
    extract_kernel<<<dim3(nwx, nwy), 128, 0, st>>>(d_level0, d_hist, win_bx, win_by, stride_bx, stride_by,
                                                         col_by_col, d_out);
}

}  // namespace hogk

#ifdef HOGB200_SVM_TIMING
extern "C" int hogb200_debug_svm_timing(unsigned long long* out, int reset)
{
    if (out && cudaMemcpyFromSymbol(out, hogk::g_svm_timing, sizeof(unsigned long long) * 160 * 8 * 8) != cudaSuccess) return 1;
    if (reset) {
        void* p = nullptr;
        if (cudaGetSymbolAddress(&p, hogk::g_svm_timing) != cudaSuccess) return 1;
        if (cudaMemset(p, 0, sizeof(unsigned long long) * 160 * 8 * 8) != cudaSuccess) return 1;
    }
    return 0;
}
#endif
