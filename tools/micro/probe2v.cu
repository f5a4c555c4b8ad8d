// (probe2.cu with LANE-DEPENDENT coefficients: an FFMA2 then reads three vector operands, as in svm.cu, instead of two + a uniform one)
// FFMA2 / FFMA issue-rate probe at low occupancy: NACC independent accumulators per thread, the operand pattern of
// the SVM inner loop (acc[j] = fma(h, w[bx], acc[j]), h and w in registers).  Prints FMA-pipe efficiency for
// several CTA sizes (1 CTA per SM).
#include <cstdio>
#include <cuda_runtime.h>

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

template <int MODE, int NACC>
__global__ void probe(float* out, const float4* __restrict__ in, int iters)
{
    float2 acc[NACC];
#pragma unroll
    for (int i = 0; i < NACC; i++) acc[i] = make_float2(threadIdx.x + i, i);
    float4 w[7];
#pragma unroll
    for (int i = 0; i < 7; i++) w[i] = in[i + (threadIdx.x & 7) * 8];
    float4 h = in[8 + (threadIdx.x & 7)];
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int X = 0; X < NACC + 6; X++) {
            h.x += 1.0f;            // a new "histogram" value per X (one cheap op instead of a load)
#pragma unroll
            for (int bx = 0; bx < 7; bx++) {
                const int j = X - bx;
                if (j >= 0 && j < NACC) {
                    if (MODE == 1) {
                        acc[j] = fma2(make_float2(h.x, h.y), make_float2(w[bx].x, w[bx].y), acc[j]);
                        acc[j] = fma2(make_float2(h.z, h.w), make_float2(w[bx].z, w[bx].w), acc[j]);
                    } else {
                        acc[j].x = fmaf(h.x, w[bx].x, acc[j].x); acc[j].y = fmaf(h.y, w[bx].y, acc[j].y);
                        acc[j].x = fmaf(h.z, w[bx].z, acc[j].x); acc[j].y = fmaf(h.w, w[bx].w, acc[j].y);
                    }
                }
            }
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < NACC; i++) s += acc[i].x + acc[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE, int NACC>
void run(float* d, const float4* in, int threads, int ctas_per_sm)
{
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 2000;
    float ms = 0;
    for (int rep = 0; rep < 2; rep++) {
        cudaEventRecord(e0);
        probe<MODE, NACC><<<148 * ctas_per_sm, threads>>>(d, in, iters);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
    }
    const double fma_lane = 4.0 * NACC * 7 * iters * (double)threads * 148 * ctas_per_sm;   // scalar FMAs
    printf("%s NACC=%2d threads/SM=%4d: %.3f ms  %.1f TFLOP/s  (%.3f lane-FMA/clk/SM at 1.9 GHz, peak 128)\n",
           MODE ? "FFMA2" : "FFMA ", NACC, threads * ctas_per_sm, ms, 2 * fma_lane / ms * 1e-9, fma_lane / (ms * 1e-3) / 148 / 1.9e9);
}

int main()
{
    float* d; cudaMalloc(&d, 148 * 2048 * sizeof(float));
    float4* in; cudaMalloc(&in, 128 * sizeof(float4)); cudaMemset(in, 0, 128 * sizeof(float4));
    run<1, 32>(d, in, 128, 1); run<1, 32>(d, in, 256, 1); run<1, 32>(d, in, 384, 1); run<1, 32>(d, in, 512, 1);
    run<1, 16>(d, in, 128, 1); run<1, 16>(d, in, 256, 1); run<1, 16>(d, in, 512, 1); run<1, 16>(d, in, 1024, 1);
    return 0;
}
