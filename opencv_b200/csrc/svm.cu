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

namespace hogk {

namespace {

constexpr int kSvmWarps = 9;                 // one per float4 of a block histogram
constexpr int kSvmThreads = kSvmWarps * 32;
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
constexpr int kRedStride = kTileWX + 1;

struct SvmSmem {
    float ring[kRing][kRowFloats];           // 95,744 B
    float coef[kMaxCoef];                    // 15,120 B
};

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem)
{
    unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gmem));
}
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
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// Load block row `by_abs` (blocks wx0 .. wx0+kRowBlocks) of the level's histogram plane into a ring slot;
// out-of-range blocks / rows are zero-filled.
__device__ __forceinline__ void load_row(float* dst, const float* __restrict__ plane, int nbx, int nby,
                                         int by_abs, int bx0, int tid)
{
    const int nvalid = (by_abs < nby) ? max(0, min(kRowBlocks, nbx - bx0)) * (kHist / 4) : 0;   // float4 chunks
    const float4* src = reinterpret_cast<const float4*>(plane + ((long long)by_abs * nbx + bx0) * kHist);
    dst += 4 * (by_abs & 7);
    for (int i = tid; i < kRowBlocks * (kHist / 4); i += kSvmThreads) {
        if (i < nvalid) cp_async16(dst + 4 * i, src + i);
        else *reinterpret_cast<float4*>(dst + 4 * i) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
}

template <int WBX>
__global__ void __launch_bounds__(kSvmThreads, 2)
svm_kernel(const __grid_constant__ FrameSet fs, const LevelDesc* __restrict__ levels, const SvmItem* __restrict__ items,
           const float* __restrict__ hist, const __grid_constant__ SvmParams P)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SvmSmem& s = *reinterpret_cast<SvmSmem*>(smem_raw);
    const int tid = threadIdx.x, q = tid >> 5, lane = tid & 31;
    const int ly = lane % kLY, xg = lane / kLY;

    const SvmItem it = items[blockIdx.x];
    if (it.frame >= fs.nframes) return;
    const LevelDesc& L = levels[it.level];
    const float* plane = hist + it.frame * fs.hist_stride + L.hist_off;
    const int nbx = L.nbx, nby = L.nby;
    // dense (stride-1) window grid of the level; the requested stride subsamples it at output time
    const int vwx = nbx - WBX + 1, vwy = nby - P.win_by + 1;
    const int wx0 = it.wx0, wy0 = it.wy0;
    const int ncoef = P.win_by * WBX * kHist;

    for (int d = 0; d < P.ndet; d++) {
        __syncthreads();     // previous detector's reduction reads are done
        for (int i = tid; i < ncoef / 4; i += kSvmThreads)
            reinterpret_cast<float4*>(s.coef)[i] = __ldg(reinterpret_cast<const float4*>(P.det[d]) + i);
        for (int r = 0; r < kTileWY; r++) load_row(s.ring[r], plane, nbx, nby, wy0 + r, wx0, tid);
        cp_async_commit();

        float2 acc[kT];         // (even k, odd k) partial sums: two packed FFMA2 per float4 pair
#pragma unroll
        for (int j = 0; j < kT; j++) acc[j] = make_float2(0.f, 0.f);

        for (int by = 0; by < P.win_by; by++) {
            cp_async_wait_all();
            __syncthreads();                                   // rows by .. by+15 are resident; slot of row by-1 is free
            if (by + 1 < P.win_by) {                           // prefetch the row the next step adds
                load_row(s.ring[(by + kTileWY) % kRing], plane, nbx, nby, wy0 + by + kTileWY, wx0, tid);
                cp_async_commit();
            }
            float4 wv[WBX];
#pragma unroll
            for (int bx = 0; bx < WBX; bx++)
                wv[bx] = *reinterpret_cast<const float4*>(&s.coef[(by * WBX + bx) * kHist + 4 * q]);
            const float* rowp = &s.ring[(by + ly) % kRing][4 * ((wy0 + by + ly) & 7) + xg * kT * kHist + 4 * q];
#pragma unroll
            for (int X = 0; X < kT + WBX - 1; X++) {
                const float4 hv = *reinterpret_cast<const float4*>(rowp + X * kHist);
#pragma unroll
                for (int bx = 0; bx < WBX; bx++) {
                    const int j = X - bx;
                    if (j >= 0 && j < kT) {
                        acc[j] = fma2(make_float2(hv.x, hv.y), make_float2(wv[bx].x, wv[bx].y), acc[j]);
                        acc[j] = fma2(make_float2(hv.z, hv.w), make_float2(wv[bx].z, wv[bx].w), acc[j]);
                    }
                }
            }
        }
        // ---- reduce the 9 k-partials through shared memory (aliases the ring)
        __syncthreads();
        float* red = &s.ring[0][0];                            // [9][16][33]: odd row stride, the window rows of a store hit different banks
#pragma unroll
        for (int j = 0; j < kT; j++) red[(q * kTileWY + ly) * kRedStride + xg * kT + j] = acc[j].x + acc[j].y;
        __syncthreads();
        for (int w = tid; w < kTileWY * kTileWX; w += kSvmThreads) {      // whole warps: 512 = 288 + 7 * 32
            const int wy_l = w / kTileWX, wx_l = w % kTileWX;
            float sc = 0.f;
#pragma unroll
            for (int k = 0; k < kSvmWarps; k++) sc += red[(k * kTileWY + wy_l) * kRedStride + wx_l];
            sc += P.bias[d];
            const int vx = wx0 + wx_l, vy = wy0 + wy_l;
            bool valid = vx < vwx && vy < vwy && (vx % P.stride_bx == 0) && (vy % P.stride_by == 0);
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
                        Candidate c; c.level = it.level | (d << 16) | (P.tag + (it.frame << 20)); c.x = ox; c.y = oy; c.score = sc;
                        P.cand[pos] = c;
                    }
                }
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
                const SvmParams& p, cudaStream_t st)
{
    if (nitems <= 0 || p.ndet <= 0) return;
    if (p.win_bx == 7) {
        cudaFuncSetAttribute(svm_kernel<7>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SvmSmem));
        svm_kernel<7><<<nitems, kSvmThreads, sizeof(SvmSmem), st>>>(fs, d_levels, d_items, d_hist, p);
    } else if (p.win_bx == 5) {
        cudaFuncSetAttribute(svm_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SvmSmem));
        svm_kernel<5><<<nitems, kSvmThreads, sizeof(SvmSmem), st>>>(fs, d_levels, d_items, d_hist, p);
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
