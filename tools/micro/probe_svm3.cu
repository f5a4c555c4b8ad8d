// Probe of a different SVM scoring loop: lane = window (32 adjacent windows of one window row band), all of k in the lane,
// the coefficients warp-uniform (constant bank / uniform registers) so that an FFMA2 reads only two vector operands.
// A warp streams down block rows; 15 accumulators are live (one per window row that contains the block row), the row
// loop is unrolled by 15 so that slot <-> coefficient row is static.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 tools/micro/probe_svm3.cu -o tools/micro/probe_svm3
#include <cstdio>
#include <cuda_runtime.h>

constexpr int kHist = 36, kRowFloats = 1376, kRows = 16, kWinBy = 15, kWbx = 7, kF4 = kWbx * 9;   // 63 float4 per window row

__constant__ float4 c_coef[kWinBy * kF4];

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

template <int V, int UNROLL>
__device__ __forceinline__ void row(float2 (&acc)[kWinBy], const float* __restrict__ rowp)
{
#pragma unroll UNROLL
    for (int i = 0; i < kF4; i++) {
        const float4 hv = *reinterpret_cast<const float4*>(rowp + 4 * i);
#pragma unroll
        for (int s = 0; s < kWinBy; s++) {
            const int by = (V - s + kWinBy) % kWinBy;
            const float4 c = c_coef[by * kF4 + i];
            acc[s] = fma2(make_float2(hv.x, hv.y), make_float2(c.x, c.y), acc[s]);
        }
#pragma unroll
        for (int s = 0; s < kWinBy; s++) {
            const int by = (V - s + kWinBy) % kWinBy;
            const float4 c = c_coef[by * kF4 + i];
            acc[s] = fma2(make_float2(hv.z, hv.w), make_float2(c.z, c.w), acc[s]);
        }
    }
}

template <int UNROLL>
__global__ void __launch_bounds__(512, 1) probe(float* out, int rounds)
{
    extern __shared__ __align__(16) float sm[];
    for (int i = threadIdx.x; i < kRows * kRowFloats; i += blockDim.x) sm[i] = (i % 97) * 1e-3f;
    __syncthreads();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    float2 acc[kWinBy];
#pragma unroll
    for (int s = 0; s < kWinBy; s++) acc[s] = make_float2(0.f, 0.f);
    float total = 0.f;
    int r = w;
    for (int it = 0; it < rounds; it++) {
#define ROW(v) row<v, UNROLL>(acc, sm + ((r + v) % kRows) * kRowFloats + lane * kHist); total += acc[(v + 1) % kWinBy].x; acc[(v + 1) % kWinBy] = make_float2(0.f, 0.f);
        ROW(0) ROW(1) ROW(2) ROW(3) ROW(4) ROW(5) ROW(6) ROW(7) ROW(8) ROW(9) ROW(10) ROW(11) ROW(12) ROW(13) ROW(14)
#undef ROW
        r += 15;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = total;
}

template <int UNROLL>
void run(float* d, int warps)
{
    const int smem = kRows * kRowFloats * 4, rounds = 40;
    cudaFuncSetAttribute(probe<UNROLL>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaFuncAttributes a; cudaFuncGetAttributes(&a, probe<UNROLL>);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float ms = 0;
    for (int rep = 0; rep < 3; rep++) {
        cudaEventRecord(e0);
        probe<UNROLL><<<148, 32 * warps, smem>>>(d, rounds);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
    }
    const double ffma2 = 148.0 * 32 * warps * rounds * 15.0 * kF4 * 30;
    const double tflops = ffma2 * 4 / (ms * 1e-3) / 1e12;
    printf("unroll %d, %2d warps/SM: %3d regs  %8.3f ms  %6.2f TFLOP/s  (%.0f %% of the 74.5 TFLOP/s FP32 peak)  %s\n", UNROLL, warps,
           a.numRegs, ms, tflops, 100 * tflops / 74.5, cudaGetErrorString(cudaGetLastError()));
}

int main()
{
    float* d; cudaMalloc(&d, 148 * 512 * 4);
    float4* h = new float4[kWinBy * kF4];
    for (int i = 0; i < kWinBy * kF4; i++) h[i] = make_float4(1e-3f * i, 0.5f, -0.25f, 1e-4f * i);
    cudaMemcpyToSymbol(c_coef, h, sizeof(float4) * kWinBy * kF4);
    for (int warps : {4, 8, 12, 16}) { run<1>(d, warps); run<3>(d, warps); run<9>(d, warps); }
    return 0;
}
