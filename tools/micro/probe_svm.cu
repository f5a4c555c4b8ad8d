// What the SVM inner loop can reach WITHOUT its ring protocol: the step of svm.cu (22 histogram LDS.128 + 28 coefficient
// LDS.128 + 504 FFMA2 per lane and step, the kernel's own lane mapping and shared-memory strides) run by 8 warps per SM
// over fixed shared-memory rows, (a) under the register budget the kernel really gets (9 warps are granted 12 warps'
// registers: 168 per thread) and (b) with 255 registers (an 8-warp CTA).  Prints TFLOP/s against the FFMA2 peak.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 tools/micro/probe_svm.cu -o tools/micro/probe_svm
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

__device__ __forceinline__ void step(float2 (&accA)[kT], float2 (&accB)[kT], float2 (&exA)[2], float2 (&exB)[2],
                                     const float* __restrict__ ca, const float* __restrict__ cb, const float* __restrict__ rowp, int q)
{
    {
        float4 wa[C], wb[C];
#pragma unroll
        for (int bx = 0; bx < C; bx++) {
            wa[bx] = *reinterpret_cast<const float4*>(ca + bx * kHist + 4 * q);
            wb[bx] = *reinterpret_cast<const float4*>(cb + bx * kHist + 4 * q);
        }
#pragma unroll
        for (int X = 0; X < kT + C - 1; X++) {
            const float4 hv = *reinterpret_cast<const float4*>(rowp + X * kHist + 4 * q);
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
    }
    {
        float4 wa[C], wb[C];
#pragma unroll
        for (int bx = 0; bx < C; bx++) {
            wa[bx] = *reinterpret_cast<const float4*>(ca + bx * kHist + 32);
            wb[bx] = *reinterpret_cast<const float4*>(cb + bx * kHist + 32);
        }
        const float* r8 = rowp + q * kHist + 32;
#pragma unroll
        for (int e = 0; e < 2; e++)
#pragma unroll
            for (int bx = 0; bx < C; bx++) {
                const float4 hv = *reinterpret_cast<const float4*>(r8 + (8 * e + bx) * kHist);
                exA[e] = fma2(make_float2(hv.x, hv.y), make_float2(wa[bx].x, wa[bx].y), exA[e]);
                exA[e] = fma2(make_float2(hv.z, hv.w), make_float2(wa[bx].z, wa[bx].w), exA[e]);
                exB[e] = fma2(make_float2(hv.x, hv.y), make_float2(wb[bx].x, wb[bx].y), exB[e]);
                exB[e] = fma2(make_float2(hv.z, hv.w), make_float2(wb[bx].z, wb[bx].w), exB[e]);
            }
    }
}

template <int THREADS, bool SYNC>
__global__ void __launch_bounds__(THREADS, 1) probe(float* out, int tiles)
{
    extern __shared__ __align__(16) float sm[];
    float* ring = sm;                               // kRows rows of kRowFloats
    float* coef = sm + kRows * kRowFloats;          // (kWinBy + 1) rows of C * 36
    for (int i = threadIdx.x; i < kRows * kRowFloats + (kWinBy + 1) * C * kHist; i += THREADS) sm[i] = (i % 97) * 1e-3f;
    __syncthreads();
    if (threadIdx.x >= 256) return;                 // the ninth warp of the 288-thread variant idles
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, q = lane & 7, pr = lane >> 3, cg = w >> 2, band = w & 3;
    float total = 0.f;
    for (int tile = 0; tile < tiles; tile++) {
        float2 accA[kT], accB[kT], exA[2], exB[2];
#pragma unroll
        for (int j = 0; j < kT; j++) accA[j] = accB[j] = make_float2(0.f, 0.f);
        exA[0] = exA[1] = exB[0] = exB[1] = make_float2(0.f, 0.f);
        for (int t = 1; t < kWinBy; t++) {
            const float* rowp = ring + ((band * 2 + 2 * pr + t) % kRows) * kRowFloats + cg * kT * kHist;
            step(accA, accB, exA, exB, coef + t * C * kHist, coef + (t - 1) * C * kHist, rowp, q);
            if (SYNC) __syncwarp();
        }
#pragma unroll
        for (int j = 0; j < kT; j++) total += accA[j].x + accA[j].y + accB[j].x + accB[j].y;
        total += exA[0].x + exA[1].y + exB[0].x + exB[1].y;
    }
    out[blockIdx.x * 256 + threadIdx.x] = total;
}

template <int THREADS, bool SYNC>
void run(const char* name, float* d)
{
    const int smem = (kRows * kRowFloats + (kWinBy + 1) * C * kHist) * 4, tiles = 200;
    cudaFuncSetAttribute(probe<THREADS, SYNC>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaFuncAttributes a; cudaFuncGetAttributes(&a, probe<THREADS, SYNC>);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float ms = 0;
    for (int rep = 0; rep < 3; rep++) {
        cudaEventRecord(e0);
        probe<THREADS, SYNC><<<148, THREADS, smem>>>(d, tiles);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
    }
    const double ffma2 = 148.0 * 256 * tiles * (kWinBy - 1) * 504.0;     // per thread and step: 448 + 56
    const double tflops = ffma2 * 4 / (ms * 1e-3) / 1e12;
    printf("%-44s %3d regs  %8.3f ms  %6.2f TFLOP/s  (%.0f %% of the 74.5 TFLOP/s FP32 peak)  %s\n", name, a.numRegs, ms, tflops,
           100 * tflops / 74.5, cudaGetErrorString(cudaGetLastError()));
}

int main()
{
    float* d; cudaMalloc(&d, 148 * 256 * 4);
    run<288, true>("9 warps (168-register budget), syncwarp/step", d);
    run<256, true>("8 warps (255 registers), syncwarp/step", d);
    run<288, false>("9 warps (168-register budget), no sync", d);
    run<256, false>("8 warps (255 registers), no sync", d);
    return 0;
}
