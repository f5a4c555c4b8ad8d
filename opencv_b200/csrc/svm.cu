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
//   * a tile is 32 x 32 windows of one level.  The producer streams the tile's 46 block rows (and then the next
//     tile's) through a 39-slot shared-memory ring, one bulk async copy per row, signalling per-slot mbarriers;
//   * a consumer warp owns 8 window rows x 16 window columns of the tile, for ALL of k: lane = (q, p), q = float4
//     number 0..7 of the 36-float block histogram (k-split across the lanes of a quarter-warp, so its LDS.128 of
//     a block is 128 contiguous bytes: conflict-free), p = pair of window rows.  The ninth float4 is shared out:
//     lane q also handles it for windows q and q+8 of the 16.  A loaded histogram float4 feeds the 7 windows that
//     contain the block in x times the 2 window rows that contain it in y, i.e. 28 FFMA2 per LDS.128, into
//     32 + 4 register accumulators (float2: even/odd k);
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
constexpr int kBands = 4, kColGroups = 2;    // warp = (band of 8 window rows, group of 16 window columns)
constexpr int kBandRows = 8;
constexpr int kTileWX = kColGroups * kT;     // 32
constexpr int kTileWY = kBands * kBandRows;  // 32
constexpr int kRing = 39;                    // a tile reads rows t .. t+31 at step t; the rest is prefetch depth
constexpr int kMaxWBX = 7;
constexpr int kRowBlocks = kTileWX + kMaxWBX - 1;          // 38
constexpr int kRowFloats = 1376;                           // 43 x 128 B >= 38 blocks
static_assert(kRowBlocks * kHist <= kRowFloats, "ring slot too small");
constexpr int kMaxCoef = 7 * 15 * kHist;                   // 3780

struct SvmSmem {
    float ring[kRing][kRowFloats];                         // 214,656 B
    float coef[kMaxCoef + 28];                             //  15,232 B
    unsigned long long full[kRing], empty[kRing];          // mbarriers: row landed / row no longer read
};
static_assert(sizeof(SvmSmem) <= 227 * 1024, "shared memory");

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
__device__ __forceinline__ void mbar_arrive_n(unsigned mbar, unsigned count)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(count) : "memory");
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
__device__ __forceinline__ bool mbar_test(unsigned mbar, unsigned parity)      // non-blocking
{
    unsigned ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(mbar), "r"(parity) : "memory");
    return ok != 0;
}
// slot / parity one position before (nslot, npar)
__device__ __forceinline__ unsigned nslot_prev(unsigned nslot) { return nslot == 0 ? kRing - 1 : nslot - 1; }
__device__ __forceinline__ unsigned npar_prev(unsigned nslot, unsigned npar) { return nslot == 0 ? npar ^ 1 : npar; }
__device__ __forceinline__ void bulk_g2s(unsigned dst, const void* src, unsigned bytes, unsigned mbar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(mbar) : "memory");
}
// One step of a tile for one thread: block row `rowp` (the thread's first block) against coefficient row ca for
// the upper window row of the pair (accA) and cb for the lower one (accB); A / B say which of the two window rows
// contain this block row (the first step has no B, the last no A).
template <int WBX, bool A, bool B>
__device__ __forceinline__ void svm_step(float2 (&accA)[kT], float2 (&accB)[kT], float2 (&exA)[2], float2 (&exB)[2],
                                         const float* __restrict__ ca, const float* __restrict__ cb,
                                         const float* __restrict__ rowp, int q)
{
    {   // float4 number q of every block, all 16 windows
        float4 wa[WBX], wb[WBX];
#pragma unroll
        for (int bx = 0; bx < WBX; bx++) {
            if (A) wa[bx] = *reinterpret_cast<const float4*>(ca + bx * kHist + 4 * q);
            if (B) wb[bx] = *reinterpret_cast<const float4*>(cb + bx * kHist + 4 * q);
        }
#pragma unroll
        for (int X = 0; X < kT + WBX - 1; X++) {
            const float4 hv = *reinterpret_cast<const float4*>(rowp + X * kHist + 4 * q);
#pragma unroll
            for (int bx = 0; bx < WBX; bx++) {
                const int j = X - bx;
                if (j >= 0 && j < kT) {
                    if (A) {
                        accA[j] = fma2(make_float2(hv.x, hv.y), make_float2(wa[bx].x, wa[bx].y), accA[j]);
                        accA[j] = fma2(make_float2(hv.z, hv.w), make_float2(wa[bx].z, wa[bx].w), accA[j]);
                    }
                    if (B) {
                        accB[j] = fma2(make_float2(hv.x, hv.y), make_float2(wb[bx].x, wb[bx].y), accB[j]);
                        accB[j] = fma2(make_float2(hv.z, hv.w), make_float2(wb[bx].z, wb[bx].w), accB[j]);
                    }
                }
            }
        }
    }
    {   // the ninth float4 (floats 32..35): windows q and q + 8 only (lane q reads block q + bx: conflict-free)
        float4 wa[WBX], wb[WBX];
#pragma unroll
        for (int bx = 0; bx < WBX; bx++) {
            if (A) wa[bx] = *reinterpret_cast<const float4*>(ca + bx * kHist + 32);
            if (B) wb[bx] = *reinterpret_cast<const float4*>(cb + bx * kHist + 32);
        }
        const float* r8 = rowp + q * kHist + 32;
#pragma unroll
        for (int e = 0; e < 2; e++) {
#pragma unroll
            for (int bx = 0; bx < WBX; bx++) {
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

// `n` numbers the rows pushed through the ring since the kernel started; slot = n % kRing and the mbarrier phase
// parity is (n / kRing) & 1 on both the producer and the consumer side.
//   full[slot]   producer -> consumers: the row's bytes have landed (expect_tx / complete_tx)
//   empty[slot]  consumers -> producer: all 8 warps are done with the row (a warp also waits for and releases the
//                rows it does not read)
template <int WBX>
__global__ void __launch_bounds__(kSvmThreads, 1)
svm_kernel(const __grid_constant__ FrameSet fs, const LevelDesc* __restrict__ levels, const SvmItem* __restrict__ items,
           int nitems, const float* __restrict__ hist, const __grid_constant__ SvmParams P)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SvmSmem& s = *reinterpret_cast<SvmSmem*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31;
    const bool producer = tid >= kConsumerThreads;
    const unsigned full0 = smem_u32(&s.full[0]), empty0 = smem_u32(&s.empty[0]);
    if (tid == 0) {
        for (int i = 0; i < kRing; i++) { mbar_init(full0 + 8 * i, 1); mbar_init(empty0 + 8 * i, kCW); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    const int win_by = P.win_by;
    const int ncoef = win_by * WBX * kHist;
    const int tile_rows = kTileWY + win_by - 1;
    unsigned n = 0;

    for (int d = 0; d < P.ndet; d++) {
        __syncthreads();     // barriers initialised / everyone is done with the previous detector
        for (int i = tid; i < ncoef / 4; i += kSvmThreads)
            reinterpret_cast<float4*>(s.coef)[i] = __ldg(reinterpret_cast<const float4*>(P.det[d]) + i);
        __syncthreads();

        if (producer) {
            // every lane streams its own rows (lane, lane + 32 of each tile): the waits for free slots overlap, so the
            // 30-odd rows that become loadable at once when a tile ends are issued back to back
            const unsigned ring0 = smem_u32(&s.ring[0][0]);
            for (int ii = blockIdx.x; ii < nitems; ii += gridDim.x) {
                const SvmItem it = items[ii];
                if (it.frame >= fs.nframes) continue;
                const LevelDesc& L = levels[it.level];
                const int nbx = L.nbx, nby = L.nby;
                const float* plane = hist + it.frame * fs.hist_stride + L.hist_off;
                const unsigned bytes = (unsigned)max(0, min(kRowBlocks, nbx - it.wx0)) * (kHist * 4);
                for (int r0 = 0; r0 < tile_rows; r0 += 32) {
                    const int r = r0 + lane;
                    if (r < tile_rows) {
                        const unsigned m = n + r, slot = m % kRing, use = m / kRing;
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
                    // lanes stay within 32 < kRing positions of each other: a parity wait on a slot is only sound
                    // once the slot's previous use has been issued
                    __syncwarp();
                }
                n += tile_rows;
            }
            __syncwarp();
            continue;
        }

        // ---- consumers: warp w sits on sub-partition w % 4; (band + 2 * colgroup) % 4 == w % 4 keeps the
        // sub-partitions evenly loaded when the lower bands or the right column group of a tile are outside the level
        const int w = tid >> 5;
        const int cg = w >> 2, band = cg ? ((w - 2) & 3) : w;
        const int q = lane & 7, pr = lane >> 3;
        for (int ii = blockIdx.x; ii < nitems; ii += gridDim.x) {
            const SvmItem it = items[ii];
            if (it.frame >= fs.nframes) continue;
            const LevelDesc& L = levels[it.level];
            const int vwx = L.nbx - WBX + 1, vwy = L.nby - win_by + 1;   // dense (stride-1) window grid of the level
            const int wx0 = it.wx0 + cg * kT, wy0 = it.wy0 + band * kBandRows;
            // Every warp waits for and releases EVERY row of every tile, in ring order, also the rows (and the whole
            // tiles) it does not read: a parity wait on a slot is only sound for a warp that released the slot's
            // previous use, and the slot cannot be reused twice behind the back of a warp that still owes a release.
            const unsigned tile0 = n, pos0 = n + band * kBandRows;      // ring position of the tile's / the band's first row
            n += tile_rows;
            const int rest = tile_rows - (band * kBandRows + 7 + win_by);   // rows below the band's last one
            if (wx0 >= vwx || wy0 >= vwy) {                             // nothing to score here
                // Rows 39.. of a tile reuse the slots of its rows 0..: a lane may only test a slot's NEXT use once the
                // slot's current use is known to have landed (a parity test cannot tell "not yet" from "one phase
                // ago"), and rows land out of order.  Hence the warp-wide join between the two rounds.
                for (int r0 = 0; r0 < tile_rows; r0 += 32) {
                    const unsigned m = tile0 + r0 + lane;
                    if (r0 + lane < tile_rows) {
                        mbar_wait(full0 + 8 * (m % kRing), (m / kRing) & 1);
                        mbar_arrive(empty0 + 8 * (m % kRing));
                    }
                    __syncwarp();
                }
                continue;
            }
            if (lane < band * kBandRows + 7) {                          // the rows above the band + the 7 rows steps 0 and 1 read
                const unsigned m = tile0 + lane;
                mbar_wait(full0 + 8 * (m % kRing), (m / kRing) & 1);
                if (lane < band * kBandRows) mbar_arrive(empty0 + 8 * (m % kRing));
            }
            __syncwarp();

            float2 accA[kT], accB[kT], exA[2], exB[2];    // (even k, odd k) partial sums of window rows 2pr / 2pr+1
#pragma unroll
            for (int j = 0; j < kT; j++) accA[j] = accB[j] = make_float2(0.f, 0.f);
            exA[0] = exA[1] = exB[0] = exB[1] = make_float2(0.f, 0.f);
            unsigned slot_l = (pos0 + 2 * pr) % kRing;
            unsigned oslot = pos0 % kRing;                               // slot of the band's oldest live row
            unsigned nslot = (pos0 + kBandRows - 1) % kRing;             // slot / parity of the row the NEXT step adds
            unsigned npar = ((pos0 + kBandRows - 1) / kRing) & 1;
            bool ready = true;
            for (int t = 0; t <= win_by; t++) {
                if (!ready) mbar_wait(full0 + 8 * nslot_prev(nslot), npar_prev(nslot, npar));
                // poll the next step's row now, so that the answer is there when this step's arithmetic is done
                ready = t == win_by || mbar_test(full0 + 8 * nslot, npar);
                const float* rowp = &s.ring[slot_l][cg * kT * kHist];
                const float* ca = s.coef + t * (WBX * kHist);
                const float* cb = ca - WBX * kHist;
                if (t == 0) svm_step<WBX, true, false>(accA, accB, exA, exB, ca, cb, rowp, q);
                else if (t == win_by) svm_step<WBX, false, true>(accA, accB, exA, exB, ca, cb, rowp, q);
                else svm_step<WBX, true, true>(accA, accB, exA, exB, ca, cb, rowp, q);
                __syncwarp();                                           // every lane has read row pos0 + t for the last time
                if (lane == 0) mbar_arrive(empty0 + 8 * oslot);
                slot_l = slot_l + 1 == kRing ? 0 : slot_l + 1;
                oslot = oslot + 1 == kRing ? 0 : oslot + 1;
                if (++nslot == kRing) { nslot = 0; npar ^= 1; }
            }
            if (lane < kBandRows - 2) mbar_arrive(empty0 + 8 * ((pos0 + win_by + 1 + lane) % kRing));   // the band's last 6 rows

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
                const float sc = v4[i] + P.bias[d];
                const int vx = vxb + i;
                const bool valid = row_ok && vx < vwx && (vx % P.stride_bx == 0);
                const int ox = vx / P.stride_bx;
                if (valid && P.scores) P.scores[d * P.score_plane + L.score_off + (long long)oy * L.nwx + ox] = sc;
                const bool hit = valid && sc >= P.thresh;
                const unsigned m = __ballot_sync(0xffffffffu, hit);
                if (m) {
                    unsigned base = 0;
                    if (lane == 0) base = atomicAdd(P.cand_count, __popc(m));
                    base = __shfl_sync(0xffffffffu, base, 0);
                    if (hit) {
                        const unsigned pos = base + __popc(m & ((1u << lane) - 1u));
                        if (pos < P.cand_cap) {
                            Candidate c2; c2.level = it.level | (d << 16) | (P.tag + (it.frame << 20)); c2.x = ox; c2.y = oy; c2.score = sc;
                            P.cand[pos] = c2;
                        }
                    }
                }
            }
            if (lane < rest) {                                          // rows of the tile below the band
                const unsigned m = pos0 + 7 + win_by + lane;
                mbar_wait(full0 + 8 * (m % kRing), (m / kRing) & 1);
                mbar_arrive(empty0 + 8 * (m % kRing));
            }
            __syncwarp();
        }
    }
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

int svm_tile_wx() { return kTileWX; }
int svm_tile_wy() { return kTileWY; }
int svm_tile_cost(int cols_left, int rows_left)
{
    // a warp = (band, column group) works only if its 8 x 16 windows intersect the level; the tile takes as long as
    // the busiest sub-partition, which hosts (band s, group 0) and (band (s + 2) % 4, group 1)
    const int nb = std::min(kBands, (rows_left + kBandRows - 1) / kBandRows), nc = cols_left > kT ? 2 : 1;
    int busiest = 0, total = 0;
    for (int sp = 0; sp < 4; sp++) {
        const int a = (sp < nb) + (nc > 1 && ((sp + 2) & 3) < nb);
        busiest = std::max(busiest, a); total += a;
    }
    return busiest * 16 + total;
}

void launch_svm(const FrameSet& fs, const LevelDesc* d_levels, const SvmItem* d_items, int nitems, const float* d_hist,
                const SvmParams& p, int num_sms, cudaStream_t st)
{
    if (nitems <= 0 || p.ndet <= 0) return;
    const int grid = std::max(1, std::min(num_sms, nitems));
    if (p.win_bx == 7) {
        cudaFuncSetAttribute(svm_kernel<7>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SvmSmem));
        svm_kernel<7><<<grid, kSvmThreads, sizeof(SvmSmem), st>>>(fs, d_levels, d_items, nitems, d_hist, p);
    } else if (p.win_bx == 5) {
        cudaFuncSetAttribute(svm_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SvmSmem));
        svm_kernel<5><<<grid, kSvmThreads, sizeof(SvmSmem), st>>>(fs, d_levels, d_items, nitems, d_hist, p);
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
