// Micro-benchmark: cost of LDS.128 on sm_100a by address pattern (per-SM cycles per warp instr).
#include <cstdio>
#include <cuda_runtime.h>

// mode 0: all lanes same address; 1: 8 distinct (lane&7); 2: 4 distinct (lane>>3);
// 3: 32 distinct consecutive; 4: 8 distinct by (lane>>2) [each group of 4 lanes same]
template <int W>
__global__ void k_lds(float* out, int iters, int mode) {
    __shared__ float4 sm[2048];
    for (int i = threadIdx.x; i < 2048; i += blockDim.x) sm[i] = make_float4(i, 1, 2, 3);
    __syncthreads();
    const int lane = threadIdx.x & 31;
    int idx = mode == 0 ? 0 : mode == 1 ? (lane & 7) : mode == 2 ? (lane >> 3) : mode == 3 ? lane : (lane >> 2);
    float4 s = make_float4(0, 0, 0, 0);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            float4 v;
            if (W == 16) v = sm[(idx + 32 * u) & 2047];
            else if (W == 8) { float2 t = reinterpret_cast<float2*>(sm)[(idx + 32 * u) & 4095]; v = make_float4(t.x, t.y, 0, 0); }
            else { float t = reinterpret_cast<float*>(sm)[(idx + 32 * u) & 8191]; v = make_float4(t, 0, 0, 0); }
            s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        }
        idx += 512;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s.x + s.y + s.z + s.w;
}

int main() {
    float* out;
    cudaMalloc(&out, 148 * 4 * 1024 * sizeof(float));
    int clk;
    cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const double ghz = clk / 1e6;
    const int iters = 2000, bs = 1024, grid = 148 * 2;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const char* names[5] = {"1 address", "8 distinct (lane&7)", "4 distinct (lane>>3)", "32 distinct", "8 distinct (lane>>2)"};
    for (int w : {16, 8, 4})
        for (int mode = 0; mode < 5; ++mode) {
            for (int rep = 0; rep < 2; ++rep) {
                cudaEventRecord(e0);
                if (w == 16) k_lds<16><<<grid, bs>>>(out, iters, mode);
                else if (w == 8) k_lds<8><<<grid, bs>>>(out, iters, mode);
                else k_lds<4><<<grid, bs>>>(out, iters, mode);
                cudaEventRecord(e1);
                cudaEventSynchronize(e1);
            }
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            double ninstr = (double)grid * bs / 32 * iters * 16 / 148; // warp instrs per SM
            printf("LDS.%-3d %-22s: %.2f cyc per warp-instr per SM\n", w * 8, names[mode],
                   ms * 1e-3 * ghz * 1e9 / ninstr);
        }
    return 0;
}
