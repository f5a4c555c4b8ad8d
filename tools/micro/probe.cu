// Micro-probes run once on the B200 box (results recorded in DESIGN.md section 6):
//  1. is sqrt.approx.ftz.f32(v) == sqrtf(v) for the 256 byte values? (decides LUT-free gamma correction)
//  2. FFMA vs FFMA2 issue throughput (independent chains, all warps resident)
#include <cstdio>
#include <cmath>
#include <cuda_runtime.h>

__global__ void sqrt_probe(float* out)
{
    float v = (float)threadIdx.x, r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v));
    out[threadIdx.x] = r;
}

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

template <int MODE>
__global__ void __launch_bounds__(256) fma_probe(float* out, int iters, float a, float b)
{
    float2 acc[8];
#pragma unroll
    for (int i = 0; i < 8; i++) acc[i] = make_float2(threadIdx.x + i, i);
    const float2 a2 = make_float2(a, a * 1.0001f), b2 = make_float2(b, b * 0.999f);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (MODE == 0) { acc[i].x = fmaf(acc[i].x, a2.x, b2.x); acc[i].y = fmaf(acc[i].y, a2.y, b2.y); }
            else acc[i] = fma2(acc[i], a2, b2);
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += acc[i].x + acc[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main()
{
    float* d; cudaMalloc(&d, 148 * 8 * 256 * sizeof(float));
    sqrt_probe<<<1, 256>>>(d);
    float h[256]; cudaMemcpy(h, d, sizeof h, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int i = 0; i < 256; i++) if (h[i] != sqrtf((float)i)) { bad++; if (bad < 5) printf("sqrt.approx(%d) = %.9g vs %.9g\n", i, h[i], sqrtf((float)i)); }
    printf("sqrt.approx.ftz mismatches over 0..255: %d\n", bad);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 20000;
    for (int mode = 0; mode < 2; mode++) {
        for (int rep = 0; rep < 2; rep++) {
            cudaEventRecord(e0);
            if (mode == 0) fma_probe<0><<<148 * 8, 256>>>(d, iters, 1.0001f, 0.5f);
            else fma_probe<1><<<148 * 8, 256>>>(d, iters, 1.0001f, 0.5f);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
        }
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        double flops = 2.0 * 16 * iters * 148.0 * 8 * 256;
        printf("%s: %.3f ms, %.1f TFLOP/s fp32\n", mode ? "FFMA2" : "FFMA ", ms, flops / ms * 1e-9);
    }
    return 0;
}
