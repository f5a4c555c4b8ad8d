// svm.cu -- linear-SVM window scoring, thresholding and on-GPU candidate compaction; descriptor extraction.
//
// Replaces classify_hists / compute_confidence_hists (modules/cudaobjdetect/src/cuda/hog.cu:394-514), which
// spend one 256-thread CTA per window and re-read every block histogram up to 105 times through L1/L2, the
// blocking per-level label download + CPU scan (cudaobjdetect/src/hog.cpp:369-411), and
// extract_descrs_by_rows/cols (hog.cu:520-618).
//
// score(wy,wx) = bias + sum_{by<WBY} < hist[wy+by][wx .. wx+WBX) , coef[by] >      (WBX*36 contiguous floats)
// is a strided 1-D correlation along x summed over WBY block rows.  It is a thin GEMV per window, so it runs on
// the FP32 pipes; the design goal is FMA issue rate, i.e. many FMAs per shared-memory load:
//   * a CTA (9 warps) owns a tile of 8 x 32 windows of one level; the block-histogram rows it needs stream
//     through a 9-row shared-memory ring (cp.async prefetch of the next row while the current step computes);
//   * warp q owns floats 4q..4q+3 of every 36-float block histogram (k-split), lane = (window row, group of 8
//     adjacent windows): per block row a thread loads 7 coefficient float4 (broadcast) + 14 histogram float4
//     and issues 224 FMAs into 8 register accumulators;
//   * the 9 k-partials are summed through shared memory, the bias added, the score optionally written densely,
//     and windows with score >= hit_threshold are appended to a global candidate list (ballot + one atomic per
//     warp).  The host sorts the (tiny) list into the reference's row-major order.
#include "hog_internal.h"
#include <algorithm>

namespace hogk {

namespace {

constexpr int kSvmWarps = 9;                 // consumer warps of a group: one per float4 of a block histogram
constexpr int kGroups = 2;                   // independent consumer groups per CTA (each with its own ring and producer)
constexpr int kGroupThreads = kSvmWarps * 32;
constexpr int kConsumerThreads = kGroups * kGroupThreads;                 // 576
constexpr int kSvmThreads = kConsumerThreads + kGroups * 32;              // + one producer warp per group = 640
constexpr int kT = 16;                       // adjacent windows per thread
constexpr int kXG = 2;                       // groups of kT windows across a warp
constexpr int kLY = 32 / kXG;                // window rows across a warp
constexpr int kTileWX = kXG * kT;            // 32
constexpr int kTileWY = kLY;                 // 16
constexpr int kRing = kTileWY + 1;           // rows resident: 16 in use + 1 in flight
constexpr int kMaxWBX = 7;
constexpr int kRowBlocks = kTileWX + kMaxWBX - 1;          // 38
// a ring slot is 11 x 128 B; block row `by_abs` is stored at float offset 4 * (by_abs & 7) inside its slot, so the
// 8 consecutive block rows that the 8 window rows of a quarter-warp read always sit in 8 different 16-byte bank groups
constexpr int kRowFloats = 1408;
static_assert(kRowBlocks * kHist + 28 <= kRowFloats, "ring slot too small");
constexpr int kMaxCoef = 7 * 15 * kHist;                   // 3780
constexpr int kRedCols = kXG * (kT / 2);                   // windows per tile row in one reduction round (16)
constexpr int kRedStride = kRedCols + 1;

struct GroupSmem {
    float ring[kRing][kRowFloats];                         // 95,744 B
    float red[kSvmWarps][kTileWY][kRedStride];             //  9,792 B: k-partials of half a tile
    unsigned long long full[kRing], empty[kRing];          // mbarriers: row landed / row no longer read
};
struct SvmSmem {
    float coef[kMaxCoef];                                  // 15,120 B, shared by both groups
    float pad[4];
    GroupSmem g[kGroups];
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
__device__ __forceinline__ void bulk_g2s(unsigned dst, const void* src, unsigned bytes, unsigned mbar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(mbar) : "memory");
}
__device__ __forceinline__ void group_sync(int g)      // the 9 consumer warps of group g
{
    asm volatile("bar.sync %0, %1;" ::"r"(g + 1), "n"(kGroupThreads) : "memory");
}

// Persistent kernel, one CTA per SM.  Each of the two consumer groups walks its own list of window tiles; its
// producer warp (one elected lane) streams the block rows of those tiles, tile after tile, through the group's ring
// with one bulk async copy per row, so the first rows of the next tile land while the current one is still being
// scored and no consumer warp ever waits for a CTA-wide barrier inside a tile:
//   full[slot]   producer -> consumers: the row's bytes have landed (expect_tx / complete_tx)
//   empty[slot]  consumers -> producer: all 9 warps are past the last step that reads the row
// `n` numbers the rows a group has pushed through its ring since the kernel started; slot = n % kRing and the
// mbarrier phase parity is (n / kRing) & 1 on both sides.
template <int WBX>
__global__ void __launch_bounds__(kSvmThreads, 1)
svm_kernel(const __grid_constant__ FrameSet fs, const LevelDesc* __restrict__ levels, const SvmItem* __restrict__ items,
           int nitems, const float* __restrict__ hist, const __grid_constant__ SvmParams P)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SvmSmem& s = *reinterpret_cast<SvmSmem*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31;
    const bool producer = tid >= kConsumerThreads;
    const int g = producer ? (tid - kConsumerThreads) >> 5 : tid / kGroupThreads;
    GroupSmem& G = s.g[g];
    const unsigned full0 = smem_u32(&G.full[0]), empty0 = smem_u32(&G.empty[0]);
    if (tid == 0) {
        for (int gg = 0; gg < kGroups; gg++)
            for (int i = 0; i < kRing; i++) {
                mbar_init(smem_u32(&s.g[gg].full[i]), 1);
                mbar_init(smem_u32(&s.g[gg].empty[i]), kSvmWarps);
            }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    const int ncoef = P.win_by * WBX * kHist;
    const int tile_rows = kTileWY + P.win_by - 1;
    const int gid = blockIdx.x * kGroups + g, gstride = gridDim.x * kGroups;
    unsigned n = 0;

    for (int d = 0; d < P.ndet; d++) {
        __syncthreads();     // barriers initialised / every group is done with the previous detector
        for (int i = tid; i < ncoef / 4; i += kSvmThreads)
            reinterpret_cast<float4*>(s.coef)[i] = __ldg(reinterpret_cast<const float4*>(P.det[d]) + i);
        __syncthreads();

        if (producer) {
            if (lane == 0) {
                for (int ii = gid; ii < nitems; ii += gstride) {
                    const SvmItem it = items[ii];
                    if (it.frame >= fs.nframes) continue;
                    const LevelDesc& L = levels[it.level];
                    const int nbx = L.nbx, nby = L.nby;
                    const float* plane = hist + it.frame * fs.hist_stride + L.hist_off;
                    const unsigned bytes = (unsigned)max(0, min(kRowBlocks, nbx - it.wx0)) * (kHist * 4);
                    const unsigned ring0 = smem_u32(&G.ring[0][0]);
                    for (int r = 0; r < tile_rows; r++, n++) {
                        const unsigned slot = n % kRing, use = n / kRing;
                        if (use > 0) mbar_wait(empty0 + 8 * slot, (use - 1) & 1);
                        const int by_abs = it.wy0 + r;
                        if (by_abs < nby && bytes > 0) {         // rows below the level feed only windows that are discarded
                            mbar_expect_tx(full0 + 8 * slot, bytes);
                            bulk_g2s(ring0 + slot * (kRowFloats * 4) + 16 * (by_abs & 7),
                                     plane + ((long long)by_abs * nbx + it.wx0) * kHist, bytes, full0 + 8 * slot);
                        } else {
                            mbar_arrive(full0 + 8 * slot);
                        }
                    }
                }
            }
            __syncwarp();
            continue;
        }

        // ---- consumers
        const int gt = tid - g * kGroupThreads, q = gt >> 5;
        const int ly = lane % kLY, xg = lane / kLY;
        for (int ii = gid; ii < nitems; ii += gstride) {
            const SvmItem it = items[ii];
            if (it.frame >= fs.nframes) continue;
            const LevelDesc& L = levels[it.level];
            const int vwx = L.nbx - WBX + 1, vwy = L.nby - P.win_by + 1;   // dense (stride-1) window grid of the level
            const int wx0 = it.wx0, wy0 = it.wy0;

            float2 acc[kT];         // (even k, odd k) partial sums: two packed FFMA2 per float4 pair
#pragma unroll
            for (int j = 0; j < kT; j++) acc[j] = make_float2(0.f, 0.f);
            unsigned slot_l = (n + ly) % kRing;
            if (lane < kTileWY) {                                   // the tile's first 16 rows, one lane each
                const unsigned m = n + lane;
                mbar_wait(full0 + 8 * (m % kRing), (m / kRing) & 1);
            }
            __syncwarp();
            for (int by = 0; by < P.win_by; by++) {
                if (by > 0) {                                       // the row this step adds
                    const unsigned m = n + by + kTileWY - 1;
                    mbar_wait(full0 + 8 * (m % kRing), (m / kRing) & 1);
                }
                float4 wv[WBX];
#pragma unroll
                for (int bx = 0; bx < WBX; bx++)
                    wv[bx] = *reinterpret_cast<const float4*>(&s.coef[(by * WBX + bx) * kHist + 4 * q]);
                const float* rowp = &G.ring[slot_l][4 * ((wy0 + by + ly) & 7) + xg * kT * kHist + 4 * q];
#pragma unroll
                for (int X = 0; X < kT + WBX - 1; X++) {
#if defined(SVM_EXP_NOLDS)
                    const float4 hv = *reinterpret_cast<const float4*>(rowp + (X & 1) * kHist);
#else
                    const float4 hv = *reinterpret_cast<const float4*>(rowp + X * kHist);
#endif
#pragma unroll
                    for (int bx = 0; bx < WBX; bx++) {
                        const int j = X - bx;
#if defined(SVM_EXP_NOFMA)
                        if (j >= 0 && j < kT && bx == 0) {
#else
                        if (j >= 0 && j < kT) {
#endif
                            acc[j] = fma2(make_float2(hv.x, hv.y), make_float2(wv[bx].x, wv[bx].y), acc[j]);
                            acc[j] = fma2(make_float2(hv.z, hv.w), make_float2(wv[bx].z, wv[bx].w), acc[j]);
                        }
                    }
                }
                __syncwarp();                                       // lanes 0 and 16 have read row n + by for the last time
                if (lane == 0) mbar_arrive(empty0 + 8 * ((n + by) % kRing));
                slot_l = slot_l + 1 == kRing ? 0 : slot_l + 1;
            }
            if (lane < kTileWY - 1) mbar_arrive(empty0 + 8 * ((n + P.win_by + lane) % kRing));   // the tile's last 15 rows
            n += tile_rows;

            // ---- sum the 9 k-partials through shared memory, half a tile (8 of a thread's 16 windows) per round
#pragma unroll
            for (int r = 0; r < 2; r++) {
#pragma unroll
                for (int jj = 0; jj < kT / 2; jj++)
                    G.red[q][ly][xg * (kT / 2) + jj] = acc[r * (kT / 2) + jj].x + acc[r * (kT / 2) + jj].y;
                group_sync(g);
                if (gt < kTileWY * kRedCols) {                      // 256 threads = 8 whole warps
                    const int wy_l = gt / kRedCols, c = gt % kRedCols;
                    const int wx_l = (c / (kT / 2)) * kT + r * (kT / 2) + c % (kT / 2);
                    float sc = 0.f;
#pragma unroll
                    for (int k = 0; k < kSvmWarps; k++) sc += G.red[k][wy_l][c];
                    sc += P.bias[d];
                    const int vx = wx0 + wx_l, vy = wy0 + wy_l;
                    const bool valid = vx < vwx && vy < vwy && (vx % P.stride_bx == 0) && (vy % P.stride_by == 0);
                    const int ox = vx / P.stride_bx, oy = vy / P.stride_by;
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
                group_sync(g);
            }
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

void launch_svm(const FrameSet& fs, const LevelDesc* d_levels, const SvmItem* d_items, int nitems, const float* d_hist,
                const SvmParams& p, int num_sms, cudaStream_t st)
{
    if (nitems <= 0 || p.ndet <= 0) return;
    const int grid = std::max(1, std::min(num_sms, (nitems + kGroups - 1) / kGroups));
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
