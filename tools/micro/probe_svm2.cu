// Variants of the SVM step (svm.cu) without its ring protocol, to see what keeps two warps per sub-partition from the FFMA2
// peak: MODE 0 = the kernel's step; 1 = coefficient rows loaded once per tile (no coefficient LDS in the steps); 2 = one
// histogram LDS per step instead of 22; 3 = both (FFMA2 only); 4 = the step with coefficient rotation (row t's registers
// serve window row B at step t + 1: 14 instead of 28 coefficient LDS per step).  WARPS 8 = two per sub-partition, 4 = one.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 tools/micro/probe_svm2.cu -o tools/micro/probe_svm2
#include <cstdio>
#include <cuda_runtime.h>

constexpr int kT = 16, kHist = 36, kRowFloats = 1376, kRows = 24, C = 7, kWinBy = 15;

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
// the same with the instruction order pinned (volatile): runs of FFMA2 that share the histogram operand stay together,
// so that the operand-reuse cache serves it
__device__ __forceinline__ float2 fma2v(float2 a, float2 b, float2 c)
{
    unsigned long long ra, rb, rc, rd;
    asm("mov.b64 %0, {%1,%2};" : "=l"(ra) : "f"(a.x), "f"(a.y));
    asm("mov.b64 %0, {%1,%2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
    asm("mov.b64 %0, {%1,%2};" : "=l"(rc) : "f"(c.x), "f"(c.y));
    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
    float2 d;
    asm("mov.b64 {%0,%1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
    return d;
}

template <int DEPTH>
__device__ __forceinline__ void stepv(float2 (&accA)[kT], float2 (&accB)[kT], float2 (&exA)[2], float2 (&exB)[2],
                                      const float* __restrict__ ca, const float* __restrict__ cb, const float* __restrict__ rowp, int q)
{
    float4 wa[C], wb[C];
#pragma unroll
    for (int bx = 0; bx < C; bx++) {
        wa[bx] = *reinterpret_cast<const float4*>(ca + bx * kHist + 4 * q);
        wb[bx] = *reinterpret_cast<const float4*>(cb + bx * kHist + 4 * q);
    }
    constexpr int NX = kT + C - 1;
    float4 hq[DEPTH];
#pragma unroll
    for (int d = 0; d < DEPTH; d++) hq[d] = *reinterpret_cast<const float4*>(rowp + d * kHist + 4 * q);
#pragma unroll
    for (int X = 0; X < NX; X++) {
        const float4 hv = hq[X % DEPTH];
        if (X + DEPTH < NX) hq[X % DEPTH] = *reinterpret_cast<const float4*>(rowp + (X + DEPTH) * kHist + 4 * q);
#pragma unroll
        for (int bx = 0; bx < C; bx++) {
            const int j = X - bx;
            if (j >= 0 && j < kT) {
                accA[j] = fma2v(make_float2(hv.x, hv.y), make_float2(wa[bx].x, wa[bx].y), accA[j]);
                accB[j] = fma2v(make_float2(hv.x, hv.y), make_float2(wb[bx].x, wb[bx].y), accB[j]);
            }
        }
#pragma unroll
        for (int bx = 0; bx < C; bx++) {
            const int j = X - bx;
            if (j >= 0 && j < kT) {
                accA[j] = fma2v(make_float2(hv.z, hv.w), make_float2(wa[bx].z, wa[bx].w), accA[j]);
                accB[j] = fma2v(make_float2(hv.z, hv.w), make_float2(wb[bx].z, wb[bx].w), accB[j]);
            }
        }
    }
    float4 wa9[C], wb9[C];
#pragma unroll
    for (int bx = 0; bx < C; bx++) {
        wa9[bx] = *reinterpret_cast<const float4*>(ca + bx * kHist + 32);
        wb9[bx] = *reinterpret_cast<const float4*>(cb + bx * kHist + 32);
    }
    const float* r8 = rowp + q * kHist + 32;
#pragma unroll
    for (int e = 0; e < 2; e++)
#pragma unroll
        for (int bx = 0; bx < C; bx++) {
            const float4 hv = *reinterpret_cast<const float4*>(r8 + (8 * e + bx) * kHist);
            exA[e] = fma2v(make_float2(hv.x, hv.y), make_float2(wa9[bx].x, wa9[bx].y), exA[e]);
            exB[e] = fma2v(make_float2(hv.x, hv.y), make_float2(wb9[bx].x, wb9[bx].y), exB[e]);
            exA[e] = fma2v(make_float2(hv.z, hv.w), make_float2(wa9[bx].z, wa9[bx].w), exA[e]);
            exB[e] = fma2v(make_float2(hv.z, hv.w), make_float2(wb9[bx].z, wb9[bx].w), exB[e]);
        }
}

template <bool LOADC, bool ONEH>
__device__ __forceinline__ void step(float2 (&accA)[kT], float2 (&accB)[kT], float2 (&exA)[2], float2 (&exB)[2],
                                     float4 (&wa)[C], float4 (&wb)[C], float4 (&wa9)[C], float4 (&wb9)[C],
                                     const float* __restrict__ ca, const float* __restrict__ cb, const float* __restrict__ rowp, int q)
{
    if (LOADC) {
#pragma unroll
        for (int bx = 0; bx < C; bx++) {
            wa[bx] = *reinterpret_cast<const float4*>(ca + bx * kHist + 4 * q);
            wb[bx] = *reinterpret_cast<const float4*>(cb + bx * kHist + 4 * q);
        }
    }
    float4 h1 = *reinterpret_cast<const float4*>(rowp + 4 * q);
#pragma unroll
    for (int X = 0; X < kT + C - 1; X++) {
        float4 hv;
        if (ONEH) { hv = h1; h1.x += 1.f; } else hv = *reinterpret_cast<const float4*>(rowp + X * kHist + 4 * q);
#pragma unroll
        for (int bx = 0; bx < C; bx++) {
            const int j = X - bx;
            if (j >= 0 && j < kT) {
                accA[j] = fma2(make_float2(hv.x, hv.y), make_float2(wa[bx].x, wa[bx].y), accA[j]);
                accB[j] = fma2(make_float2(hv.x, hv.y), make_float2(wb[bx].x, wb[bx].y), accB[j]);
            }
        }
#pragma unroll
        for (int bx = 0; bx < C; bx++) {
            const int j = X - bx;
            if (j >= 0 && j < kT) {
                accA[j] = fma2(make_float2(hv.z, hv.w), make_float2(wa[bx].z, wa[bx].w), accA[j]);
                accB[j] = fma2(make_float2(hv.z, hv.w), make_float2(wb[bx].z, wb[bx].w), accB[j]);
            }
        }
    }
    if (LOADC) {
#pragma unroll
        for (int bx = 0; bx < C; bx++) {
            wa9[bx] = *reinterpret_cast<const float4*>(ca + bx * kHist + 32);
            wb9[bx] = *reinterpret_cast<const float4*>(cb + bx * kHist + 32);
        }
    }
    const float* r8 = rowp + q * kHist + 32;
#pragma unroll
    for (int e = 0; e < 2; e++)
#pragma unroll
        for (int bx = 0; bx < C; bx++) {
            float4 hv;
            if (ONEH) { hv = h1; h1.y += 1.f; } else hv = *reinterpret_cast<const float4*>(r8 + (8 * e + bx) * kHist);
            exA[e] = fma2(make_float2(hv.x, hv.y), make_float2(wa9[bx].x, wa9[bx].y), exA[e]);
            exA[e] = fma2(make_float2(hv.z, hv.w), make_float2(wa9[bx].z, wa9[bx].w), exA[e]);
            exB[e] = fma2(make_float2(hv.x, hv.y), make_float2(wb9[bx].x, wb9[bx].y), exB[e]);
            exB[e] = fma2(make_float2(hv.z, hv.w), make_float2(wb9[bx].z, wb9[bx].w), exB[e]);
        }
}

// coefficient rotation: the step loads only row t (wa); wb is the previous step's wa.  Unrolled by two so that the
// two register sets swap roles without moves.
template <int MODE, int WARPS>
__global__ void __launch_bounds__(288, 1) probe(float* out, int tiles)
{
    extern __shared__ __align__(16) float sm[];
    float* ring = sm;
    float* coef = sm + kRows * kRowFloats;
    for (int i = threadIdx.x; i < kRows * kRowFloats + (kWinBy + 1) * C * kHist; i += 288) sm[i] = (i % 97) * 1e-3f;
    __syncthreads();
    if (threadIdx.x >= 32 * WARPS) return;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, q = lane & 7, pr = lane >> 3, cg = (w >> 2) & 1, band = w & 3;
    float total = 0.f;
    for (int tile = 0; tile < tiles; tile++) {
        float2 accA[kT], accB[kT], exA[2], exB[2];
        float4 wa[C], wb[C], wa9[C], wb9[C];
#pragma unroll
        for (int j = 0; j < kT; j++) accA[j] = accB[j] = make_float2(0.f, 0.f);
        exA[0] = exA[1] = exB[0] = exB[1] = make_float2(0.f, 0.f);
        if (MODE == 1 || MODE == 3) {
#pragma unroll
            for (int bx = 0; bx < C; bx++) {
                wa[bx] = *reinterpret_cast<const float4*>(coef + tile % 8 * C * kHist + bx * kHist + 4 * q);
                wb[bx] = *reinterpret_cast<const float4*>(coef + (tile % 8 + 1) * C * kHist + bx * kHist + 4 * q);
                wa9[bx] = *reinterpret_cast<const float4*>(coef + tile % 8 * C * kHist + bx * kHist + 32);
                wb9[bx] = *reinterpret_cast<const float4*>(coef + (tile % 8 + 1) * C * kHist + bx * kHist + 32);
            }
        }
        for (int t = 1; t < kWinBy; t++) {
            const float* rowp = ring + ((band * 2 + 2 * pr + t) % kRows) * kRowFloats + cg * kT * kHist;
            if constexpr (MODE >= 10) stepv<MODE - 10>(accA, accB, exA, exB, coef + t * C * kHist, coef + (t - 1) * C * kHist, rowp, q);
            else step<MODE == 0 || MODE == 2, MODE == 2 || MODE == 3>(accA, accB, exA, exB, wa, wb, wa9, wb9,
                coef + t * C * kHist, coef + (t - 1) * C * kHist, rowp, q);
            __syncwarp();
        }
#pragma unroll
        for (int j = 0; j < kT; j++) total += accA[j].x + accA[j].y + accB[j].x + accB[j].y;
        total += exA[0].x + exA[1].y + exB[0].x + exB[1].y;
    }
    out[blockIdx.x * 256 + threadIdx.x] = total;
}

template <int MODE, int WARPS>
void run(const char* name, float* d)
{
    const int smem = (kRows * kRowFloats + (kWinBy + 1) * C * kHist) * 4, tiles = 200;
    cudaFuncSetAttribute(probe<MODE, WARPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaFuncAttributes a; cudaFuncGetAttributes(&a, probe<MODE, WARPS>);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float ms = 0;
    for (int rep = 0; rep < 3; rep++) {
        cudaEventRecord(e0);
        probe<MODE, WARPS><<<148, 288, smem>>>(d, tiles);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
    }
    const double steps = tiles * (kWinBy - 1.0);
    const double cyc = ms * 1e-3 * 1.965e9 / steps;
    printf("%-52s %d warps %3d regs %8.3f ms  %6.0f cycles per step (FFMA2 pipe: %d)  %s\n", name, WARPS, a.numRegs, ms, cyc,
           1008 * WARPS / 4, cudaGetErrorString(cudaGetLastError()));
}

int main()
{
    float* d; cudaMalloc(&d, 148 * 256 * 4);
    run<0, 8>("the kernel's step", d);
    run<11, 8>("pinned order, 1 row ahead", d);
    run<12, 8>("pinned order, 2 rows ahead", d);
    run<13, 8>("pinned order, 3 rows ahead", d);
    run<0, 4>("the kernel's step", d);
    run<11, 4>("pinned order, 1 row ahead", d);
    run<12, 4>("pinned order, 2 rows ahead", d);
    run<13, 4>("pinned order, 3 rows ahead", d);
    return 0;
}
