// group.cu -- cv::groupRectangles(rects, groupThreshold, eps) on the device.
//
// Reference: modules/objdetect/src/cascadedetect.cpp:64-174 (grouping), core/operations.hpp:481-569 (cv::partition),
// objdetect.hpp:121-134 (SimilarRects).  The host version in hog_host.cu stays the default for the handful of hits
// a frame normally has; this one takes over when a frame yields tens of thousands of candidates (low hit threshold,
// dense pyramids), where the pairwise predicate dominates the frame time on the host.
//
// Same result, bit for bit, as the reference on the same input order:
//   * classes = connected components of SimilarRects; cv::partition numbers them by first appearance, i.e. by the
//     smallest member index.  Here: lock-free union-find whose hooking always puts the larger root under the
//     smaller one, so a root IS the smallest member; an exclusive scan over the root flags gives the class number.
//   * candidate pairs: rects sorted by x (CUB radix sort; with the CUB prefix scan below the two library primitives of this file -- a
//     sort and a scan are not the subject here -- both outside the three pipeline kernels);
//     a warp owns one rect and sweeps the sorted order to the right while x_b - x_a <= eps (w_a + h_a) / 2, which
//     bounds SimilarRects' own delta from above (the host's sweep, hog_host.cu group_rectangles).
//   * class sums in int32 atomics (order-free, exact), mean with float s = 1.f / n and round-half-even, the
//     "inside a stronger class" filter per class against all classes, ordered compaction by a second scan.
#include "hog_internal.h"

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <algorithm>

namespace hogk {

namespace {

struct RectI { int x, y, w, h; };

__global__ void iota_keys_kernel(const RectI* __restrict__ r, int n, int* __restrict__ keys, int* __restrict__ idx)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { keys[i] = r[i].x; idx[i] = i; }
}

__global__ void gather_kernel(const RectI* __restrict__ r, const int* __restrict__ idx, int n, RectI* __restrict__ sorted,
                              int* __restrict__ parent)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { sorted[i] = r[idx[i]]; parent[i] = i; }
}

// parent[] is read with relaxed GPU-scope loads (L2, where the atomics live): parents only ever decrease, so a value that
// is stale by the time it is used is harmless
__device__ __forceinline__ int uf_load(const int* p)
{
    int v;
    asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ int uf_find(int* parent, int a)
{
    // halve the path on the way up
    int p = uf_load(parent + a);
    while (p != a) {
        const int g = uf_load(parent + p);
        if (g != p) atomicMin(&parent[a], g);
        a = p; p = g;
    }
    return a;
}

// The sweep finds every similar pair, and most of them already sit in one set: the two parents are fetched together and
// compared before anything else (after path halving most nodes point at their root), and the two climbs to the roots
// advance in step, so that their loads overlap (the kernel is bound by these dependent L2 round trips: ncu, 55 long-
// scoreboard stall cycles per issued instruction before this arrangement).
__device__ __forceinline__ void uf_unite(int* parent, int a, int b)
{
    int pa = uf_load(parent + a), pb = uf_load(parent + b);
    if (pa == pb) return;
    for (;;) {
        for (;;) {
            const bool ra = pa == a, rb = pb == b;
            if (ra && rb) break;
            const int ga = ra ? pa : uf_load(parent + pa);
            const int gb = rb ? pb : uf_load(parent + pb);
            if (!ra) { if (ga != pa) atomicMin(&parent[a], ga); a = pa; pa = ga; }
            if (!rb) { if (gb != pb) atomicMin(&parent[b], gb); b = pb; pb = gb; }
        }
        if (a == b) return;
        if (a > b) { const int t = a; a = b; b = t; }
        const int old = atomicCAS(&parent[b], b, a);
        if (old == b) return;                               // b was still a root: now under the smaller root a
        pa = a; pb = old;                                   // a was a root a moment ago; b has moved under `old`
    }
}

// SimilarRects::operator() (objdetect.hpp:126-133), double arithmetic as in the reference
__device__ __forceinline__ bool similar(const RectI& r1, const RectI& r2, double eps)
{
    const double delta = eps * (min(r1.w, r2.w) + min(r1.h, r2.h)) * 0.5;
    return abs(r1.x - r2.x) <= delta && abs(r1.y - r2.y) <= delta &&
           abs(r1.x + r1.w - r2.x - r2.w) <= delta && abs(r1.y + r1.h - r2.y - r2.h) <= delta;
}

// One warp per rect (sorted position a): lanes test positions a+1+lane, a+33+lane, ... until all are out of reach.
// COLLECT: the partners SimilarRects accepts are first gathered (ballot-compacted into a per-warp list in shared memory) and
// then united with all lanes busy -- twice as fast on ONE dense list (261 558 rects of a pass-all frame: 18.4 -> 9.0 ms for
// the whole grouping), where a rect has hundreds of partners; on the chunk lists of the batch path (64 segments of ~1 600
// rects, ~8 partners per rect) uniting on the spot is 9 % faster (460 against 501 us per 105 000 rects), so each kernel
// takes its own form.  (A THREAD per rect with batched unites was measured too: 828 us.)
constexpr int kUniteWarps = 8, kHits = 224;
constexpr int kSegCap = 8192, kSegThreads = 1024, kTabMax = 4096;       // unite_seg_smem_kernel: rects per segment it takes (24 B of shared memory each), threads per CTA

template <bool SEG, bool COLLECT>
__device__ __forceinline__ void unite_sweep(const RectI* __restrict__ sorted, const int* __restrict__ sorted_seg,
                                            const int* __restrict__ idx, int n, double eps, int* __restrict__ parent, int* hits)
{
    const int lane = threadIdx.x & 31;
    const int a = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (a >= n) return;                                  // the whole warp
    const RectI ra = sorted[a];
    const int ia = idx[a], sa = SEG ? sorted_seg[a] : 0;
    const double reach = eps * (ra.w + ra.h) * 0.5;
    int nh = 0;
    for (int b = a + 1 + lane; ; b += 32) {
        bool in_reach = false, sim = false;
        int ib = 0;
        if (b < n) {
            const RectI rb = sorted[b];
            in_reach = (!SEG || sorted_seg[b] == sa) && rb.x - ra.x <= reach;
            if (in_reach && similar(ra, rb, eps)) { sim = true; ib = idx[b]; }
        }
        if (!COLLECT) {
            if (sim) uf_unite(parent, ia, ib);
            if (!__any_sync(0xffffffffu, in_reach)) break;         // sorted by (segment,) x: nothing further right can be in reach
            continue;
        }
        const unsigned m = __ballot_sync(0xffffffffu, sim);
        if (sim) hits[nh + __popc(m & ((1u << lane) - 1u))] = ib;
        nh += __popc(m);
        const bool more = __any_sync(0xffffffffu, in_reach);
        if (nh > kHits - 32 || !more) {
            __syncwarp();
            for (int k = lane; k < nh; k += 32) uf_unite(parent, ia, hits[k]);
            __syncwarp();
            nh = 0;
        }
        if (!more) break;
    }
}

__global__ void __launch_bounds__(kUniteWarps * 32) unite_kernel(const RectI* __restrict__ sorted, const int* __restrict__ idx, int n,
                                                                 double eps, int* __restrict__ parent)
{
    __shared__ int hits[kUniteWarps][kHits];
    unite_sweep<false, true>(sorted, nullptr, idx, n, eps, parent, hits[threadIdx.x >> 5]);
}

__global__ void root_flags_kernel(int* __restrict__ parent, int n, int* __restrict__ root_of, int* __restrict__ is_root)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        const int r = uf_find(parent, i);
        root_of[i] = r;
        is_root[i] = r == i;
    } else if (i == n) is_root[n] = 0;          // the scan runs over n + 1 entries: its last output is the total
}

// exclusive scan of in[0 .. n] (n + 1 entries, in[n] == 0) into out[0 .. n]: out[n] is the total.  (cub::DeviceScan, like the
// sort a library primitive beside the subject; the single-CTA scan it replaces took 270 us for 100 000 entries.)
inline size_t scan_bytes(int n)
{
    size_t b = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, b, (const int*)nullptr, (int*)nullptr, n + 1);
    return b;
}
inline void exclusive_scan(void* tmp, size_t tmp_bytes, const int* in, int* out, int n, cudaStream_t st)
{
    cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, in, out, n + 1, st);
}

// sums[cls] += (x, y, w, h, 1) for the rect of every lane: the lanes of a warp that hold the same class (neighbours in
// reference order mostly do -- with a few big classes, 100 000 rects otherwise queue up on a handful of addresses) add up
// among themselves first and their first lane issues the five atomics.  Integer sums: exact in any order.
__device__ __forceinline__ void class_add(int* __restrict__ sums, int cls, const RectI& r, bool live)
{
    const unsigned act = __ballot_sync(0xffffffffu, live);
    if (!live) return;
    const unsigned peers = __match_any_sync(act, cls);
    const int x = __reduce_add_sync(peers, r.x), y = __reduce_add_sync(peers, r.y), w = __reduce_add_sync(peers, r.w),
              h = __reduce_add_sync(peers, r.h);
    if ((threadIdx.x & 31) == __ffs(peers) - 1) {
        int* s = sums + 5 * cls;
        atomicAdd(s + 0, x); atomicAdd(s + 1, y); atomicAdd(s + 2, w); atomicAdd(s + 3, h);
        atomicAdd(s + 4, __popc(peers));
    }
}

// rrects[cls] += rect, rweights[cls]++ (cascadedetect.cpp:86-95); sums = 5 ints per class (x, y, w, h, count)
__global__ void accumulate_kernel(const RectI* __restrict__ r, const int* __restrict__ root_of, const int* __restrict__ class_of_root,
                                  int n, int* __restrict__ sums)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = i < n;
    class_add(sums, live ? class_of_root[root_of[i]] : 0, live ? r[i] : RectI{0, 0, 0, 0}, live);
}

// mean rect of a class: Rect(saturate_cast<int>(r.x * s), ...) with float s = 1.f / n (cascadedetect.cpp:97-106)
__device__ __forceinline__ RectI class_mean(const int* __restrict__ s)
{
    const float f = 1.f / (float)s[4];
    return RectI{__float2int_rn((float)s[0] * f), __float2int_rn((float)s[1] * f), __float2int_rn((float)s[2] * f),
                 __float2int_rn((float)s[3] * f)};
}

// keep class i unless it is too weak or sits inside a stronger class (cascadedetect.cpp:112-161)
__global__ void filter_kernel(const int* __restrict__ sums, int nclasses, int thr, double eps, int* __restrict__ keep,
                              RectI* __restrict__ mean)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == nclasses) keep[i] = 0;
    if (i >= nclasses) return;
    const int n1 = sums[5 * i + 4];
    const RectI r1 = class_mean(sums + 5 * i);
    mean[i] = r1;
    int k = n1 > thr;
    for (int j = 0; k && j < nclasses; j++) {
        const int n2 = sums[5 * j + 4];
        if (j == i || n2 <= thr) continue;
        const RectI r2 = class_mean(sums + 5 * j);
        const int dx = __double2int_rn(r2.w * eps), dy = __double2int_rn(r2.h * eps);
        if (r1.x >= r2.x - dx && r1.y >= r2.y - dy && r1.x + r1.w <= r2.x + r2.w + dx && r1.y + r1.h <= r2.y + r2.h + dy &&
            (n2 > max(3, n1) || n1 < 3))
            k = 0;
    }
    keep[i] = k;
}

__global__ void emit_kernel(const RectI* __restrict__ mean, const int* __restrict__ keep, const int* __restrict__ pos, int nclasses,
                            RectI* __restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nclasses && keep[i]) out[pos[i]] = mean[i];
}

inline size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

}  // namespace

size_t group_workspace_bytes(int n)
{
    size_t sort_bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (const int*)nullptr, (int*)nullptr, (const int*)nullptr, (int*)nullptr, n);
    sort_bytes = std::max(sort_bytes, scan_bytes(n));
    // rects, sorted rects, means, out: 4 x 16 n;  keys x2, idx x2, parent, root_of, flags, scan, keep, pos: 10 x 4 (n + 1);
    // sums 20 n; counters
    return 4 * align256((size_t)16 * n) + 10 * align256((size_t)4 * (n + 1)) + align256((size_t)20 * n) + align256(sort_bytes) + 512;
}

// rects (host, reference order) -> grouped rects (host).  ws: device workspace of group_workspace_bytes(n).
// Returns the number of output rects, or -1 on a CUDA error.
int group_rectangles_device(const int* h_rects, int n, int thr, double eps, int* h_out, void* ws, cudaStream_t st)
{
    if (n <= 0) return 0;
    unsigned char* p = (unsigned char*)ws;
    auto take = [&](size_t bytes) { void* r = p; p += align256(bytes); return r; };
    RectI* d_rects = (RectI*)take((size_t)16 * n); RectI* d_sorted = (RectI*)take((size_t)16 * n);
    RectI* d_mean = (RectI*)take((size_t)16 * n); RectI* d_out = (RectI*)take((size_t)16 * n);
    int* d_keys = (int*)take((size_t)4 * n); int* d_keys2 = (int*)take((size_t)4 * n);
    int* d_idx = (int*)take((size_t)4 * n); int* d_idx2 = (int*)take((size_t)4 * n);
    int* d_parent = (int*)take((size_t)4 * n); int* d_root = (int*)take((size_t)4 * n);
    int* d_flag = (int*)take((size_t)4 * (n + 1)); int* d_scan = (int*)take((size_t)4 * (n + 1));
    int* d_keep = (int*)take((size_t)4 * (n + 1)); int* d_pos = (int*)take((size_t)4 * (n + 1));
    int* d_sums = (int*)take((size_t)20 * n);
    size_t sort_bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (const int*)nullptr, (int*)nullptr, (const int*)nullptr, (int*)nullptr, n);
    sort_bytes = std::max(sort_bytes, scan_bytes(n));
    void* d_sort = take(sort_bytes);

    const int T = 256, B = (n + T - 1) / T, B1 = (n + T) / T;
    cudaMemcpyAsync(d_rects, h_rects, (size_t)16 * n, cudaMemcpyHostToDevice, st);
    iota_keys_kernel<<<B, T, 0, st>>>(d_rects, n, d_keys, d_idx);
    cub::DeviceRadixSort::SortPairs(d_sort, sort_bytes, d_keys, d_keys2, d_idx, d_idx2, n, 0, 32, st);
    gather_kernel<<<B, T, 0, st>>>(d_rects, d_idx2, n, d_sorted, d_parent);
    unite_kernel<<<(int)(((long long)n * 32 + T - 1) / T), T, 0, st>>>(d_sorted, d_idx2, n, eps, d_parent);
    root_flags_kernel<<<B1, T, 0, st>>>(d_parent, n, d_root, d_flag);
    exclusive_scan(d_sort, sort_bytes, d_flag, d_scan, n, st);
    int nclasses = 0;
    cudaMemcpyAsync(&nclasses, d_scan + n, sizeof(int), cudaMemcpyDeviceToHost, st);
    cudaMemsetAsync(d_sums, 0, (size_t)20 * n, st);
    accumulate_kernel<<<B, T, 0, st>>>(d_rects, d_root, d_scan, n, d_sums);
    if (cudaStreamSynchronize(st) != cudaSuccess) return -1;
    const int BC = (nclasses + T) / T;            // one thread more than classes: it writes the scan's sentinel
    filter_kernel<<<BC, T, 0, st>>>(d_sums, nclasses, thr, eps, d_keep, d_mean);
    exclusive_scan(d_sort, sort_bytes, d_keep, d_pos, nclasses, st);
    emit_kernel<<<BC, T, 0, st>>>(d_mean, d_keep, d_pos, nclasses, d_out);
    int nout = 0;
    cudaMemcpyAsync(&nout, d_pos + nclasses, sizeof(int), cudaMemcpyDeviceToHost, st);
    if (cudaStreamSynchronize(st) != cudaSuccess) return -1;
    if (nout > 0) cudaMemcpyAsync(h_out, d_out, (size_t)16 * nout, cudaMemcpyDeviceToHost, st);
    if (cudaStreamSynchronize(st) != cudaSuccess || cudaGetLastError() != cudaSuccess) return -1;
    return nout;
}

// =====================================================================================================================
// Candidates of a whole chunk of frames -> rects, on the device (HOG_Impl::detectMultiScale's tail, cudaobjdetect/src/hog.cpp:
// 471-487, for every (frame, detector) of the chunk at once): the compacted candidate list never travels to the host.
//   1. sort by (segment = frame * ndet + detector, level, y, x): the reference's order (levels outermost, windows row-major);
//   2. rect = (cvRound(x * stride * scale), cvRound(y * stride * scale), cvRound(win * scale)) in double, as hog.cpp:473-481;
//   3. groupRectangles per segment with the kernels above: the x-sort key carries the segment in its upper half and the
//      sweep stops at a segment boundary, so rects of different segments never unite; classes are numbered by a global scan
//      over the root flags, which -- segments being contiguous in the sorted order -- numbers every segment's classes by
//      first appearance, contiguously; the containment filter scans only the classes of its own segment.
// No intermediate host synchronisation: kernels over classes are sized by n and read the class count from device memory.
namespace {

struct Cand { int level, x, y; float score; };      // = hogk::Candidate

// key = segment | level | y | x, each field as wide as the plan needs (CandGeom bits_*): the radix sort makes one pass per
// 8 key bits, and 1080p / VGA keys are 26 / 25 bits wide instead of the 52 of a fixed layout
__global__ void cand_keys_kernel(const Cand* __restrict__ c, int n, int ndet, int bx, int by, int bl,
                                 unsigned long long* __restrict__ keys, int* __restrict__ idx)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Cand k = c[i];
    const unsigned seg = (unsigned)(k.level >> 20) * ndet + ((k.level >> 16) & 15);
    keys[i] = ((unsigned long long)seg << (bl + by + bx)) | ((unsigned long long)(k.level & 0xffff) << (by + bx)) |
              ((unsigned long long)k.y << bx) | (unsigned)k.x;
    idx[i] = i;
}

__global__ void cand_rects_kernel(const Cand* __restrict__ c, const unsigned long long* __restrict__ keys, const int* __restrict__ order,
                                  int n, const __grid_constant__ CandGeom G, RectI* __restrict__ rects, int* __restrict__ seg,
                                  unsigned long long* __restrict__ xkeys, int* __restrict__ idx, int* __restrict__ seg_count, int count_now)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Cand k = c[order[i]];
    const double scale = G.scale[k.level & 0xffff];
    RectI r;
    r.x = __double2int_rn((double)(k.x * G.stride_x) * scale); r.y = __double2int_rn((double)(k.y * G.stride_y) * scale);
    r.w = __double2int_rn(G.win_w * scale); r.h = __double2int_rn(G.win_h * scale);
    rects[i] = r;
    const int sg = (int)(keys[i] >> (G.bits_x + G.bits_y + G.bits_level));
    seg[i] = sg;
    xkeys[i] = ((unsigned long long)sg << G.bits_rx) | (unsigned)r.x;    // 0 <= r.x < 2^bits_rx: window origins are inside the frame
    idx[i] = i;
    if (count_now) atomicAdd(&seg_count[sg], 1);                        // no grouping: every candidate is an output rect
}

__global__ void gather_seg_kernel(const RectI* __restrict__ r, const int* __restrict__ seg, const int* __restrict__ idx, int n,
                                  RectI* __restrict__ sorted, int* __restrict__ sorted_seg, int* __restrict__ parent)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { sorted[i] = r[idx[i]]; sorted_seg[i] = seg[idx[i]]; parent[i] = i; }
}

// (segments of at most kSegCap rects have been grouped in shared memory by unite_seg_smem_kernel)
__global__ void __launch_bounds__(kUniteWarps * 32) unite_seg_kernel(const RectI* __restrict__ sorted, const int* __restrict__ sorted_seg,
                                                                     const int* __restrict__ idx, int n, double eps, int* __restrict__ parent,
                                                                     const int* __restrict__ s0, const int* __restrict__ s1, int done_below)
{
    const int a = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (a >= n) return;
    const int sg = sorted_seg[a];
    if (s1[sg] - s0[sg] <= done_below) return;
    unite_sweep<true, false>(sorted, sorted_seg, idx, n, eps, parent, nullptr);
}

// [s0[s], s1[s]) = the rects of segment s (the same range in reference order and in x order: both sorts are segment-major)
__global__ void seg_bounds_kernel(const int* __restrict__ seg, int n, int* __restrict__ s0, int* __restrict__ s1)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int s = seg[i];
    if (i == 0 || seg[i - 1] != s) s0[s] = i;
    if (i == n - 1 || seg[i + 1] != s) s1[s] = i + 1;
}

// union-find on a parent array in SHARED memory (indices local to a segment); same hooking rule as uf_unite
__device__ __forceinline__ int ufs_find(int* parent, int a)
{
    int p = ((volatile int*)parent)[a];
    while (p != a) {
        const int g = ((volatile int*)parent)[p];
        if (g != p) atomicMin(&parent[a], g);
        a = p; p = g;
    }
    return a;
}
__device__ __forceinline__ void ufs_unite(int* parent, int a, int b)
{
    if (((volatile int*)parent)[a] == ((volatile int*)parent)[b]) return;
    for (;;) {
        a = ufs_find(parent, a); b = ufs_find(parent, b);
        if (a == b) return;
        if (a > b) { const int t = a; a = b; b = t; }
        if (atomicCAS(&parent[b], b, a) == b) return;
    }
}

// One CTA per segment of at most kSegCap rects: the segment's rects and its union-find live in shared memory (a rect or a
// parent costs ~30 cycles instead of an L2 round trip -- the sweep is a chain of such dependent reads), warps take the
// segment's rects in turn and sweep as unite_sweep does; at the end every rect's ROOT goes to the global parent array.
// Larger segments are left to unite_seg_kernel.

__global__ void __launch_bounds__(kSegThreads) unite_seg_smem_kernel(const RectI* __restrict__ sorted, const int* __restrict__ idx,
                                                                     const int* __restrict__ s0, const int* __restrict__ s1,
                                                                     double eps, int mmax, int* __restrict__ parent)
{
    extern __shared__ __align__(16) unsigned char seg_smem[];
    const int r0 = s0[blockIdx.x], r1 = s1[blockIdx.x], ns = r1 - r0;
    if (ns <= 1 || ns > kSegCap) return;
    // the segment's rects (x order), their reference positions and the union-find parents, all local to the segment
    RectI* srect = reinterpret_cast<RectI*>(seg_smem);
    int* sidx = reinterpret_cast<int*>(seg_smem + (size_t)16 * kSegCap);
    int* lparent = sidx + kSegCap;
    int* dtab = lparent + kSegCap;
    for (int i = threadIdx.x; i < ns; i += kSegThreads) { srect[i] = sorted[r0 + i]; sidx[i] = idx[r0 + i] - r0; lparent[i] = i; }
    // SimilarRects compares integer differences with delta = eps * (min w + min h) * 0.5 in double: |d| <= delta is
    // |d| <= floor(delta), so the doubles are spent once per possible sum m (the same expression, the same rounding)
    // and the sweep is integer arithmetic -- the FP64 chain was what an iteration waited for once the loads were local
    for (int m = threadIdx.x; m <= mmax; m += kSegThreads) dtab[m] = __double2int_rd(eps * m * 0.5);
    __syncthreads();
    const int lane = threadIdx.x & 31;
    for (int a = threadIdx.x >> 5; a < ns; a += kSegThreads / 32) {
        const RectI ra = srect[a];
        const int ia = sidx[a];
        const int reach = dtab[ra.w + ra.h];
        for (int b = a + 1 + lane; ; b += 32) {
            bool in_reach = false;
            if (b < ns) {
                const RectI rb = srect[b];
                in_reach = rb.x - ra.x <= reach;
                if (in_reach) {
                    const int delta = dtab[min(ra.w, rb.w) + min(ra.h, rb.h)];
                    if (abs(ra.x - rb.x) <= delta && abs(ra.y - rb.y) <= delta && abs(ra.x + ra.w - rb.x - rb.w) <= delta &&
                        abs(ra.y + ra.h - rb.y - rb.h) <= delta)
                        ufs_unite(lparent, ia, sidx[b]);
                }
            }
            if (!__any_sync(0xffffffffu, in_reach)) break;
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < ns; i += kSegThreads) parent[r0 + i] = r0 + ufs_find(lparent, i);
}

// sums = 5 ints per class; class -> segment, and every segment's class range [seg_lo, seg_hi]
__global__ void accumulate_seg_kernel(const RectI* __restrict__ r, const int* __restrict__ seg, const int* __restrict__ root_of,
                                      const int* __restrict__ class_of_root, int n, int* __restrict__ sums, int* __restrict__ class_seg,
                                      int* __restrict__ seg_lo, int* __restrict__ seg_hi)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = i < n;
    const int cls = live ? class_of_root[root_of[i]] : 0;
    class_add(sums, cls, live ? r[i] : RectI{0, 0, 0, 0}, live);
    if (!live) return;
    if (root_of[i] == i) {
        class_seg[cls] = seg[i];
        atomicMin(&seg_lo[seg[i]], cls); atomicMax(&seg_hi[seg[i]], cls);
    }
}

__global__ void filter_seg_kernel(const int* __restrict__ sums, const int* __restrict__ nclasses_p, const int* __restrict__ class_seg,
                                  const int* __restrict__ seg_lo, const int* __restrict__ seg_hi, int thr, double eps,
                                  int* __restrict__ keep, RectI* __restrict__ mean, int* __restrict__ seg_count, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n) return;
    if (i >= *nclasses_p) { keep[i] = 0; return; }          // incl. the scan's sentinel keep[n] (nclasses <= n)
    const int n1 = sums[5 * i + 4];
    const RectI r1 = class_mean(sums + 5 * i);
    mean[i] = r1;
    const int sg = class_seg[i];
    int k = n1 > thr;
    for (int j = seg_lo[sg]; k && j <= seg_hi[sg]; j++) {
        const int n2 = sums[5 * j + 4];
        if (j == i || n2 <= thr) continue;
        const RectI r2 = class_mean(sums + 5 * j);
        const int dx = __double2int_rn(r2.w * eps), dy = __double2int_rn(r2.h * eps);
        if (r1.x >= r2.x - dx && r1.y >= r2.y - dy && r1.x + r1.w <= r2.x + r2.w + dx && r1.y + r1.h <= r2.y + r2.h + dy &&
            (n2 > max(3, n1) || n1 < 3))
            k = 0;
    }
    keep[i] = k;
    if (k) atomicAdd(&seg_count[sg], 1);
}

__global__ void emit_seg_kernel(const RectI* __restrict__ mean, const int* __restrict__ keep, const int* __restrict__ pos, int n,
                                RectI* __restrict__ out, int* __restrict__ total)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && keep[i]) out[pos[i]] = mean[i];
    if (i == 0) *total = pos[n];                 // number of output rects of the chunk
}

__global__ void fill_kernel(int* __restrict__ p, int n, int v)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

size_t sort64_bytes(int n)
{
    size_t b = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, b, (const unsigned long long*)nullptr, (unsigned long long*)nullptr, (const int*)nullptr,
                                    (int*)nullptr, n);
    return b;
}

}  // namespace

size_t cand_group_workspace_bytes(int n, int nseg)
{
    return 4 * align256((size_t)16 * n) + 2 * align256((size_t)8 * n) + 12 * align256((size_t)4 * (n + 1)) + align256((size_t)20 * n) +
           align256(std::max(sort64_bytes(n), scan_bytes(n))) + 6 * align256((size_t)4 * (nseg + 1)) + 1024;
}

// d_cand: n candidates of the chunk in device memory (any order).  On return *d_out_p points at the output rects in the
// workspace (segment-major, reference order inside a segment) and *d_hdr_p at [total, count of segment 0, 1, ...] (ints),
// both valid once `st` has drained.  Nothing is synchronised here.  Returns 0, or -1 on a CUDA error.
int cand_group_device(const void* d_cand, int n, int nseg, int ndet, const CandGeom& G, int thr, double eps, void* ws,
                      cudaStream_t st, const int** d_out_p, const int** d_hdr_p)
{
    unsigned char* p = (unsigned char*)ws;
    auto take = [&](size_t bytes) { void* r = p; p += align256(bytes); return r; };
    RectI* d_rects = (RectI*)take((size_t)16 * n); RectI* d_sorted = (RectI*)take((size_t)16 * n);
    RectI* d_mean = (RectI*)take((size_t)16 * n); RectI* d_out = (RectI*)take((size_t)16 * n);
    unsigned long long* d_k1 = (unsigned long long*)take((size_t)8 * n); unsigned long long* d_k2 = (unsigned long long*)take((size_t)8 * n);
    int* d_idx = (int*)take((size_t)4 * n); int* d_idx2 = (int*)take((size_t)4 * n);
    int* d_seg = (int*)take((size_t)4 * n); int* d_sseg = (int*)take((size_t)4 * n);
    int* d_parent = (int*)take((size_t)4 * n); int* d_root = (int*)take((size_t)4 * n);
    int* d_flag = (int*)take((size_t)4 * (n + 1)); int* d_scan = (int*)take((size_t)4 * (n + 1));
    int* d_keep = (int*)take((size_t)4 * (n + 1)); int* d_pos = (int*)take((size_t)4 * (n + 1));
    int* d_cseg = (int*)take((size_t)4 * n); int* d_order = (int*)take((size_t)4 * n);
    int* d_sums = (int*)take((size_t)20 * n);
    size_t sort_bytes = std::max(sort64_bytes(n), scan_bytes(n));
    void* d_sort = take(sort_bytes);
    int* d_hdr = (int*)take((size_t)4 * (nseg + 1));       // [total][count per segment]
    int* d_lo = (int*)take((size_t)4 * nseg); int* d_hi = (int*)take((size_t)4 * nseg);
    int* d_s0 = (int*)take((size_t)4 * nseg); int* d_s1 = (int*)take((size_t)4 * nseg);
    const Cand* c = (const Cand*)d_cand;
    const int T = 256, B = (n + T - 1) / T, B1 = (n + T) / T, BS = (nseg + T - 1) / T;
    int bseg = 1;
    while ((1 << bseg) < nseg) bseg++;
    const int bits1 = bseg + G.bits_level + G.bits_y + G.bits_x, bits2 = bseg + G.bits_rx;

    cudaMemsetAsync(d_hdr, 0, (size_t)4 * (nseg + 1), st);
    cand_keys_kernel<<<B, T, 0, st>>>(c, n, ndet, G.bits_x, G.bits_y, G.bits_level, d_k1, d_idx);
    cub::DeviceRadixSort::SortPairs(d_sort, sort_bytes, d_k1, d_k2, d_idx, d_order, n, 0, bits1, st);
    cand_rects_kernel<<<B, T, 0, st>>>(c, d_k2, d_order, n, G, d_rects, d_seg, d_k1, d_idx, d_hdr + 1, thr <= 0);
    *d_hdr_p = d_hdr;
    if (thr <= 0) {             // cv::groupRectangles returns the list untouched (cascadedetect.cpp:66-75)
        fill_kernel<<<1, 32, 0, st>>>(d_hdr, 1, n);
        *d_out_p = &d_rects[0].x;
        return cudaGetLastError() == cudaSuccess ? 0 : -1;
    }
    cub::DeviceRadixSort::SortPairs(d_sort, sort_bytes, d_k1, d_k2, d_idx, d_idx2, n, 0, bits2, st);
    gather_seg_kernel<<<B, T, 0, st>>>(d_rects, d_seg, d_idx2, n, d_sorted, d_sseg, d_parent);
    cudaMemsetAsync(d_s0, 0, (size_t)4 * nseg, st);
    cudaMemsetAsync(d_s1, 0, (size_t)4 * nseg, st);
    seg_bounds_kernel<<<B, T, 0, st>>>(d_seg, n, d_s0, d_s1);
    // the largest w + h a rect of this plan can have (cvRound(win * scale) of the last level): the size of the delta table
    double smax = 1.0;
    for (int l = 0; l < kMaxLevels; l++) smax = std::max(smax, G.scale[l]);
    const int mmax = (int)std::ceil(G.win_w * smax) + (int)std::ceil(G.win_h * smax) + 2;
    const bool in_smem = mmax < kTabMax;
    if (in_smem) {
        constexpr int smem = kSegCap * 24 + kTabMax * 4;
        static bool attr_set[64] = {};
        int dev = 0;
        cudaGetDevice(&dev);
        if (!attr_set[dev & 63])
            attr_set[dev & 63] = cudaFuncSetAttribute(unite_seg_smem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) == cudaSuccess;
        unite_seg_smem_kernel<<<nseg, kSegThreads, smem, st>>>(d_sorted, d_idx2, d_s0, d_s1, eps, mmax, d_parent);
    }
    unite_seg_kernel<<<(int)(((long long)n * 32 + T - 1) / T), T, 0, st>>>(d_sorted, d_sseg, d_idx2, n, eps, d_parent, d_s0, d_s1,
                                                                           in_smem ? kSegCap : 0);
    root_flags_kernel<<<B1, T, 0, st>>>(d_parent, n, d_root, d_flag);
    exclusive_scan(d_sort, sort_bytes, d_flag, d_scan, n, st);           // d_scan[n] = number of classes
    cudaMemsetAsync(d_sums, 0, (size_t)20 * n, st);
    fill_kernel<<<BS, T, 0, st>>>(d_lo, nseg, 0x7fffffff);
    fill_kernel<<<BS, T, 0, st>>>(d_hi, nseg, -1);
    accumulate_seg_kernel<<<B, T, 0, st>>>(d_rects, d_seg, d_root, d_scan, n, d_sums, d_cseg, d_lo, d_hi);
    filter_seg_kernel<<<B1, T, 0, st>>>(d_sums, d_scan + n, d_cseg, d_lo, d_hi, thr, eps, d_keep, d_mean, d_hdr + 1, n);
    exclusive_scan(d_sort, sort_bytes, d_keep, d_pos, n, st);
    emit_seg_kernel<<<B, T, 0, st>>>(d_mean, d_keep, d_pos, n, d_out, d_hdr);      // d_hdr[0] = number of output rects
    *d_out_p = &d_out[0].x;
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

}  // namespace hogk
