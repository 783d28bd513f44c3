// Micro-benchmark: FP32 issue rates on sm_100a that decide the shape of the SIMT scan kernels.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_fma ubench_fma.cu && ./ubench_fma
#include <cstdio>
#include <cuda_runtime.h>

__global__ void k_ffma(float* out, float a, float b, int iters) {
    float acc[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = threadIdx.x * 0.001f + i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int i = 0; i < 16; ++i) acc[i] = fmaf(acc[i], a, b);
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_ffma2(float* out, float a, float b, int iters) {
    unsigned long long acc[8];
    unsigned long long av, bv;
    asm("mov.b64 %0, {%1, %1};" : "=l"(av) : "f"(a));
    asm("mov.b64 %0, {%1, %1};" : "=l"(bv) : "f"(b));
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        float x = threadIdx.x * 0.001f + i;
        asm("mov.b64 %0, {%1, %1};" : "=l"(acc[i]) : "f"(x));
    }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int i = 0; i < 8; ++i)
                asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(acc[i]) : "l"(av), "l"(bv));
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        float lo, hi;
        asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(acc[i]));
        s += lo + hi;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// FFMA fed by shared memory: per iteration NL LDS.128 (8 distinct float4 per warp) and NF FFMA
template <int NL, int NF>
__global__ void k_lds_ffma(float* out, int iters) {
    __shared__ float4 sm[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) sm[i] = make_float4(i, 1, 2, 3);
    __syncthreads();
    float acc[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = i;
    const int rg = threadIdx.x & 7;
    int base = rg;
    for (int it = 0; it < iters; ++it) {
        float4 v[NL];
#pragma unroll
        for (int l = 0; l < NL; ++l) v[l] = sm[(base + 8 * l) & 1023];
        base += 8 * NL;
#pragma unroll
        for (int f = 0; f < NF; ++f) {
            const float4 x = v[f % NL];
            const float m = (f & 3) == 0 ? x.x : (f & 3) == 1 ? x.y : (f & 3) == 2 ? x.z : x.w;
            acc[f & 15] = fmaf(acc[f & 15], m, x.x);
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <class F>
float time_ms(F f) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    f();
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    f();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    return ms;
}

int main() {
    float* out;
    cudaMalloc(&out, 148 * 8 * 1024 * sizeof(float));
    int clk;
    cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const double ghz = clk / 1e6;
    const int iters = 4000;
    for (int bs : {256, 512, 1024}) {
        const int grid = 148 * (1024 / bs) * 2;
        float ms = time_ms([&] { k_ffma<<<grid, bs>>>(out, 1.0001f, 0.5f, iters); });
        double fma = (double)grid * bs * iters * 128;
        printf("FFMA   block %4d: %.1f FMA/clk/SM (at %.2f GHz), %.1f TFLOP/s\n", bs,
               fma / (ms * 1e-3) / 148 / (ghz * 1e9), ghz, 2 * fma / (ms * 1e-3) / 1e12);
        ms = time_ms([&] { k_ffma2<<<grid, bs>>>(out, 1.0001f, 0.5f, iters); });
        printf("FFMA2  block %4d: %.1f FMA/clk/SM, %.1f TFLOP/s\n", bs,
               fma / (ms * 1e-3) / 148 / (ghz * 1e9), 2 * fma / (ms * 1e-3) / 1e12);
    }
    {
        const int bs = 512, grid = 148 * 4;
        float ms = time_ms([&] { k_lds_ffma<1, 16><<<grid, bs>>>(out, iters); });
        printf("LDS.128 x1 + FFMA x16: %.1f FMA/clk/SM, %.3f LDS.128/clk/SM\n",
               (double)grid * bs * iters * 16 / (ms * 1e-3) / 148 / (ghz * 1e9),
               (double)grid * bs / 32 * iters / (ms * 1e-3) / 148 / (ghz * 1e9));
        ms = time_ms([&] { k_lds_ffma<1, 32><<<grid, bs>>>(out, iters); });
        printf("LDS.128 x1 + FFMA x32: %.1f FMA/clk/SM, %.3f LDS.128/clk/SM\n",
               (double)grid * bs * iters * 32 / (ms * 1e-3) / 148 / (ghz * 1e9),
               (double)grid * bs / 32 * iters / (ms * 1e-3) / 148 / (ghz * 1e9));
        ms = time_ms([&] { k_lds_ffma<4, 4><<<grid, bs>>>(out, iters); });
        printf("LDS.128 x4 + FFMA x4 : %.1f FMA/clk/SM, %.3f LDS.128/clk/SM\n",
               (double)grid * bs * iters * 4 / (ms * 1e-3) / 148 / (ghz * 1e9),
               (double)grid * bs / 32 * iters * 4 / (ms * 1e-3) / 148 / (ghz * 1e9));
    }
    return 0;
}
