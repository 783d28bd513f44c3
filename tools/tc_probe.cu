// tc_probe.cu -- hardware probe for the tcgen05 pieces of the tensor-core scan (run on a B200):
//   * kind::tf32 MMA with A in TMEM (tcgen05.st) and B in shared memory straight from the device
//     slab layout (K-major, no swizzle: SBO = 128 B, LBO = column stride of the TMA box)
//   * what the tensor core does with the low 13 mantissa bits of fp32 inputs (truncate / round)
//   * accuracy of the 3xTF32 split (hi*hi + lo*hi + hi*lo) against a double-precision product
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../hnsw_ivf_b200/csrc \
//        tc_probe.cu -o tc_probe   (see tools/Makefile)
#include "tc_ptx.cuh"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

using namespace b200ivf::ptx;

constexpr int M = 128, N = 128, K = 128, NF4 = K / 4;
constexpr int KC = 8;                      // float4 columns per TMA box
constexpr int BOX_BYTES = KC * 4 * 512;    // 16 KB
constexpr uint32_t LBO = 4 * 512, SBO = 128;

#define CK(x)                                                                          \
    do {                                                                               \
        cudaError_t e = (x);                                                           \
        if (e != cudaSuccess) {                                                        \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); \
            exit(1);                                                                   \
        }                                                                              \
    } while (0)

__device__ __forceinline__ float tf32_trunc(float x) {
    return __uint_as_float(__float_as_uint(x) & 0xffffe000u);
}
__device__ __forceinline__ float tf32_rn(float x) { // round to nearest (ties away), then truncate
    return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
}


// rows-major issue forms, BLK k-steps per elect block (bits 20..: see main): per 1 SS alone,
// 2 SS + TS, 3 SS(2N) + TS, 4 TS alone
template <int BLK>
__device__ __forceinline__ void issue_rows_major(int reps, int per, int nn, uint32_t tD, uint32_t tAlo,
                                                 uint32_t bhi, uint32_t blo, uint64_t* bars,
                                                 bool commits, bool lane0_loop) {
    const uint32_t idesc = make_idesc_tf32(M, nn);
    const uint32_t id2 = make_idesc_tf32(M, 2 * nn > 256 ? 256 : 2 * nn);
    const uint64_t d0 = make_smem_desc_kmajor_noswz(0, LBO, SBO);
    const bool me = lane0_loop ? (threadIdx.x & 31) == 0 : true;
    if (!me)
        return;
    for (int r = 0; r < reps; ++r)
        for (int g = 0; g < (K / 8) / BLK; ++g) {
            const uint32_t tDb = tD + ((r & 1) ? 128u : 0u);
            const uint64_t dhi = d0 + (uint64_t)((bhi + g * BLK * 2 * LBO) >> 4);
            const uint64_t dlo = d0 + (uint64_t)((blo + g * BLK * 2 * LBO) >> 4);
            if (lane0_loop || elect_one_sync()) {
#pragma unroll
                for (int ks = 0; ks < BLK; ++ks) {
                    const uint64_t kh = dhi + (uint64_t)(ks * ((2 * LBO) >> 4));
                    const uint64_t kl = dlo + (uint64_t)(ks * ((2 * LBO) >> 4));
                    if (per != 4)
                        mma_tf32_ss(tDb, kh, kl, per == 3 ? id2 : idesc, 1);
                    if (per != 1)
                        mma_tf32_ts(tDb, tAlo + (g * BLK + ks) * 8, kl, idesc, 1);
                }
                if (commits) {
                    mma_commit(&bars[2]);
                    mma_commit(&bars[3]);
                }
            }
            if (!lane0_loop)
                __syncwarp();
        }
}

// mode 1: D = A*B^T with raw fp32 bits (one MMA per k-step); mode 3: 3xTF32
__global__ void __launch_bounds__(256, 1)
        probe_kernel(const float* __restrict__ A, const __grid_constant__ CUtensorMap tmap,
                     float* __restrict__ D, int mode, int reps, int nmma, long long* cycles) {
    extern __shared__ __align__(1024) unsigned char smem[];
    float4* Bhi = reinterpret_cast<float4*>(smem);                   // [NF4][4 blocks][32 rows]
    float4* Blo = Bhi + NF4 * 128;
    uint64_t* bars = reinterpret_cast<uint64_t*>(Blo + NF4 * 128);   // [0] tma, [1] mma done
    uint32_t* tmem_base_s = reinterpret_cast<uint32_t*>(bars + 4);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        mbar_init(&bars[2], 0xfffff);
        mbar_init(&bars[3], 0xfffff);
        fence_barrier_init();
    }
    if (warp == 4) {
        tmem_alloc(tmem_base_s, 512);
        tmem_relinquish();
    }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem = *tmem_base_s;
    const uint32_t tD = tmem, tAhi = tmem + 256, tAlo = tmem + 384;

    if (warp == 4 && lane == 0) {
        mbar_expect_tx(&bars[0], (uint32_t)(NF4 * 128 * sizeof(float4)));
        for (int c = 0; c < NF4; c += KC)
            tma_load_3d(Bhi + c * 128, &tmap, 0, 0, c, &bars[0]);
    }
    if (warp < 4) {
        // A row `tid` -> TMEM lane tid, K consecutive columns
        const float* a = A + (size_t)tid * K;
        const uint32_t lane_addr = (uint32_t)(warp * 32) << 16;
        for (int c = 0; c < K; c += 8) {
            uint32_t hi[8], lo[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float x = a[c + i];
                if (mode == 1) {
                    hi[i] = __float_as_uint(x);
                    lo[i] = 0;
                } else {
                    const float h = tf32_trunc(x);
                    hi[i] = __float_as_uint(h);
                    lo[i] = __float_as_uint(tf32_rn(x - h));
                }
            }
            tmem_st_32x32b_x8(tAhi + lane_addr + c, hi);
            tmem_st_32x32b_x8(tAlo + lane_addr + c, lo);
        }
        tmem_st_wait();
    }
    // B lo = y - trunc(y), same layout
    mbar_wait(&bars[0], 0);
    for (int i = tid; i < NF4 * 128; i += 256) {
        const float4 y = Bhi[i];
        float4 l;
        l.x = tf32_rn(y.x - tf32_trunc(y.x));
        l.y = tf32_rn(y.y - tf32_trunc(y.y));
        l.z = tf32_rn(y.z - tf32_trunc(y.z));
        l.w = tf32_rn(y.w - tf32_trunc(y.w));
        Blo[i] = l;
    }
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();

    if (warp == 5 && reps > 0) {
        // the whole warp runs the loop, one elected lane issues (see scan_tcr.cu)
        // timing: `reps` passes over the K loop, nmma MMAs (N = nmma_n) per k-step, one commit
        // bit 8 of nmma: N = 256 (reads past Bhi into Blo); bits 16+: N given explicitly
        const int nn = ((nmma >> 16) & 0xff) ? ((nmma >> 16) & 0xff) : ((nmma & 0x100) ? 256 : 128);
        const int per = nmma & 0xff;
        const uint32_t idesc = make_idesc_tf32(M, nn);
        const uint32_t bhi = smem_u32(Bhi), blo = smem_u32(Blo);
        const long long t0 = clock64();
        if (nmma & 0x8000) {
            const int blk = (nmma >> 24) & 0xff ? (nmma >> 24) & 0xff : 4;
            const int n2 = (nmma >> 16) & 0xff;
            const bool l0 = (nmma & 0x1000) != 0, cm = (nmma & 0x200) != 0;
            if (blk == 2)
                issue_rows_major<2>(reps, per, n2, tD, tAlo, bhi, blo, bars, cm, l0);
            else if (blk == 4)
                issue_rows_major<4>(reps, per, n2, tD, tAlo, bhi, blo, bars, cm, l0);
            else if (blk == 8)
                issue_rows_major<8>(reps, per, n2, tD, tAlo, bhi, blo, bars, cm, l0);
            else
                issue_rows_major<16>(reps, per, n2, tD, tAlo, bhi, blo, bars, cm, l0);
        } else if (nmma & 0x4000) {
            // the kernels' issue shape: 12 MMAs of a stage (4 k-steps) straight-line under one elect
            const uint64_t d0 = make_smem_desc_kmajor_noswz(0, LBO, SBO);
            for (int r = 0; r < reps; ++r)
                for (int sg = 0; sg < K / 32; ++sg) {
                    const uint32_t tDb = tD + ((r & 1) ? 128u : 0u);
                    const uint64_t dhi = d0 + (uint64_t)((bhi + sg * 8 * LBO) >> 4);
                    const uint64_t dlo = d0 + (uint64_t)((blo + sg * 8 * LBO) >> 4);
                    if ((nmma & 0x8000) && elect_one_sync()) {
                        // rows-major forms, straight-line: per 1 SS alone, 2 SS + TS, 3 SS(2N) + TS,
                        // 4 TS alone
                        const uint32_t id2 = make_idesc_tf32(M, 2 * nn > 256 ? 256 : 2 * nn);
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) {
                            const uint64_t kh = dhi + (uint64_t)(ks * ((2 * LBO) >> 4));
                            const uint64_t kl = dlo + (uint64_t)(ks * ((2 * LBO) >> 4));
                            if (per != 4)
                                mma_tf32_ss(tDb, kh, kl, per == 3 ? id2 : idesc, 1);
                            if (per != 1)
                                mma_tf32_ts(tDb, tAlo + (sg * 4 + ks) * 8, kl, idesc, 1);
                        }
                        if (nmma & 0x200) {
                            mma_commit(&bars[2]);
                            mma_commit(&bars[3]);
                        }
                    } else if (!(nmma & 0x8000) && elect_one_sync()) {
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) {
                            const uint64_t kh = dhi + (uint64_t)(ks * ((2 * LBO) >> 4));
                            const uint64_t kl = dlo + (uint64_t)(ks * ((2 * LBO) >> 4));
                            mma_tf32_ts(tDb, tAlo + (sg * 4 + ks) * 8, kh, idesc, 1);
                            mma_tf32_ts(tDb, tAhi + (sg * 4 + ks) * 8, kl, idesc, 1);
                            mma_tf32_ts(tDb, tAhi + (sg * 4 + ks) * 8, kh, idesc, 1);
                        }
                        if (nmma & 0x200) {
                            mma_commit(&bars[2]);
                            mma_commit(&bars[3]);
                        }
                    }
                    __syncwarp();
                }
        } else
        for (int r = 0; r < reps; ++r)
            for (int ks = 0; ks < K / 8; ++ks) {
                const uint64_t dhi = make_smem_desc_kmajor_noswz(bhi + ks * 2 * LBO, LBO, SBO);
                const uint64_t dlo = make_smem_desc_kmajor_noswz(blo + ks * 2 * LBO, LBO, SBO);
                if (!elect_one_sync())
                    continue;
                if (nmma & 0x8000) {
                    // rows-major forms (scan_tcr.cu): A = the raw stage through a shared-memory
                    // descriptor (M = 128 rows), B = nn "queries" from the other region.
                    //   per 1: SS alone; per 2: SS (N = nn) + TS (A lo from tensor memory, N = nn);
                    //   per 3: SS with N = 2 nn (stacked) + TS N = nn
                    const uint32_t idn = make_idesc_tf32(M, nn);
                    const uint32_t id2 = make_idesc_tf32(M, 2 * nn > 256 ? 256 : 2 * nn);
                    const uint32_t tDb = tD + ((r & 1) ? 128u : 0u);
                    if (per == 4) { // TS alone
                        mma_tf32_ts(tDb, tAlo + ks * 8, dlo, idn, 1);
                    } else {
                        mma_tf32_ss(tDb, dhi, dlo, per == 3 ? id2 : idn, 1);
                        if (per > 1)
                            mma_tf32_ts(tDb, tAlo + ks * 8, dlo, idn, 1);
                    }
                } else if (nmma & 0x400) { // the real kernel's operand order, alternating D buffers
                    // 0x2000: alternate the accumulator every k-step (two dependency chains)
                    const uint32_t tDb =
                            tD + ((((nmma & 0x2000) ? ks : r) & 1) ? 128u : 0u);
                    mma_tf32_ts(tDb, tAlo + ks * 8, dhi, idesc, 1);
                    mma_tf32_ts(tDb, tAhi + ks * 8, dlo, idesc, 1);
                    mma_tf32_ts(tDb, tAhi + ks * 8, dhi, idesc, 1);
                } else {
                    mma_tf32_ts(tD, tAhi + ks * 8, dhi, idesc, 1);
                    if (per > 1)
                        mma_tf32_ts(tD, tAlo + ks * 8, dhi, idesc, 1);
                    if (per > 2)
                        mma_tf32_ts(tD, tAhi + ks * 8, dlo, idesc, 1);
                }
                if ((nmma & 0x1000) && (ks & 3) == 3)
                    tc_fence_after_sync();
                if ((nmma & 0x200) && (ks & 3) == 3) { // the real kernel's two commits per stage
                    mma_commit(&bars[2]);
                    mma_commit(&bars[3]);
                }
            }
        __syncwarp();
        const long long t1 = clock64();
        if (elect_one_sync())
            mma_commit(&bars[1]);
        __syncwarp();
        mbar_wait(&bars[1], 0);
        const long long t2 = clock64();
        if (lane == 0) {
            cycles[2 * blockIdx.x + 0] = t1 - t0;
            cycles[2 * blockIdx.x + 1] = t2 - t0;
        }
    } else if (warp == 5 && lane == 0) {
        const uint32_t idesc = make_idesc_tf32(M, N);
        const uint32_t bhi = smem_u32(Bhi), blo = smem_u32(Blo);
        for (int ks = 0; ks < K / 8; ++ks) {
            const uint64_t dhi = make_smem_desc_kmajor_noswz(bhi + ks * 2 * LBO, LBO, SBO);
            const uint64_t dlo = make_smem_desc_kmajor_noswz(blo + ks * 2 * LBO, LBO, SBO);
            mma_tf32_ts(tD, tAhi + ks * 8, dhi, idesc, ks > 0);
            if (mode == 3) {
                mma_tf32_ts(tD, tAlo + ks * 8, dhi, idesc, 1);
                mma_tf32_ts(tD, tAhi + ks * 8, dlo, idesc, 1);
            }
        }
        mma_commit(&bars[1]);
    }
    if (warp >= 6 && reps > 0 && (nmma & 0x800)) {
        // extra pollers, as the idle roles of the real kernel: spin until the MMAs are done
        while (!mbar_try_wait(&bars[1], 0)) {
        }
    }
    if (warp < 4) {
        mbar_wait(&bars[1], 0);
        tc_fence_after_sync();
        const uint32_t lane_addr = (uint32_t)(warp * 32) << 16;
        for (int c = 0; c < N; c += 32) {
            uint32_t r[32];
            tmem_ld_32x32b_x32(tD + lane_addr + c, r);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i)
                D[(size_t)tid * N + c + i] = __uint_as_float(r[i]);
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 4)
        tmem_dealloc(tmem, 512);
}

static float h_trunc(float x) {
    uint32_t u;
    memcpy(&u, &x, 4);
    u &= 0xffffe000u;
    memcpy(&x, &u, 4);
    return x;
}
static float h_rn(float x) {
    uint32_t u;
    memcpy(&u, &x, 4);
    u = (u + 0x1000u) & 0xffffe000u;
    memcpy(&x, &u, 4);
    return x;
}

int main() {
    std::vector<float> A((size_t)M * K), B((size_t)N * K), slab((size_t)N * NF4 * 4, 0.f);
    srand(7);
    for (auto& v : A)
        v = (float)rand() / RAND_MAX * 2.f - 1.f;
    for (auto& v : B)
        v = (float)rand() / RAND_MAX * 2.f - 1.f;
    for (int r = 0; r < N; ++r)
        for (int k = 0; k < K; ++k) {
            const size_t b = r / 32, ln = r % 32;
            slab[((b * NF4 + (k >> 2)) * 32 + ln) * 4 + (k & 3)] = B[(size_t)r * K + k];
        }
    float *dA, *dS, *dD;
    CK(cudaMalloc(&dA, A.size() * 4));
    CK(cudaMalloc(&dS, slab.size() * 4));
    CK(cudaMalloc(&dD, (size_t)M * N * 4));
    CK(cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dS, slab.data(), slab.size() * 4, cudaMemcpyHostToDevice));

    // tensor map: {128 floats, blocks, nf4} with strides {512 B per block... see below}
    using EncodeFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
    CUtensorMap tmap;
    const cuuint64_t nblocks = N / 32;
    const cuuint64_t gdim[3] = {128, nblocks, (cuuint64_t)NF4};
    const cuuint64_t gstride[2] = {(cuuint64_t)NF4 * 512, 512}; // bytes: dim1 = block, dim2 = column
    const cuuint32_t box[3] = {128, 4, KC};
    const cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = reinterpret_cast<EncodeFn>(fn)(
            &tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, dS, gdim, gstride, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
            CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        printf("cuTensorMapEncodeTiled failed: %d\n", (int)r);
        return 1;
    }
    const size_t smem_bytes = 2 * NF4 * 128 * sizeof(float4) + 64;
    static_assert(true, "");
    CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                            (int)smem_bytes));
    std::vector<float> D((size_t)M * N);
    for (int mode : {1, 3}) {
        CK(cudaMemset(dD, 0, D.size() * 4));
        probe_kernel<<<1, 256, smem_bytes>>>(dA, tmap, dD, mode, 0, 0, nullptr);
        CK(cudaGetLastError());
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
        double e_exact = 0, e_trunc = 0, e_rn = 0, e_f32 = 0, mag = 0;
        for (int i = 0; i < M; ++i)
            for (int j = 0; j < N; ++j) {
                double ex = 0, tr = 0, rn = 0;
                float f = 0.f;
                for (int k = 0; k < K; ++k) {
                    const float a = A[(size_t)i * K + k], b = B[(size_t)j * K + k];
                    ex += (double)a * b;
                    tr += (double)h_trunc(a) * h_trunc(b);
                    rn += (double)h_rn(a) * h_rn(b);
                    f = fmaf(a, b, f);
                }
                const double g = D[(size_t)i * N + j];
                e_exact = fmax(e_exact, fabs(g - ex));
                e_trunc = fmax(e_trunc, fabs(g - tr));
                e_rn = fmax(e_rn, fabs(g - rn));
                e_f32 = fmax(e_f32, fabs((double)f - ex));
                mag = fmax(mag, fabs(ex));
            }
        printf("mode %d: max|D-exact| %.3e  max|D-trunc| %.3e  max|D-rn| %.3e  "
               "(fp32 fma chain vs exact %.3e, max|exact| %.2f)\n",
               mode, e_exact, e_trunc, e_rn, e_f32, mag);
    }
    printf("D[0][0..3] = %g %g %g %g\n", D[0], D[1], D[2], D[3]);
    long long* dC;
    CK(cudaMalloc(&dC, 16 * 148));
    for (int grid : {1, 148})
        for (int nm : {(4 << 24) | (16 << 16) | 0x8201, (4 << 24) | (64 << 16) | 0x8201, (4 << 24) | (128 << 16) | 0x8201, (4 << 24) | (16 << 16) | 0x8204, (4 << 24) | (64 << 16) | 0x8204, (4 << 24) | (128 << 16) | 0x8204, (4 << 24) | (16 << 16) | 0x8202, (4 << 24) | (64 << 16) | 0x8202, (4 << 24) | (128 << 16) | 0x8202, (4 << 24) | (16 << 16) | 0x8203, (4 << 24) | (64 << 16) | 0x8203, (4 << 24) | (128 << 16) | 0x8203, (16 << 24) | (16 << 16) | 0x8201, (16 << 24) | (64 << 16) | 0x8201, (16 << 24) | (128 << 16) | 0x8201, (16 << 24) | (16 << 16) | 0x8204, (16 << 24) | (64 << 16) | 0x8204, (16 << 24) | (128 << 16) | 0x8204, (16 << 24) | (16 << 16) | 0x8202, (16 << 24) | (64 << 16) | 0x8202, (16 << 24) | (128 << 16) | 0x8202, (16 << 24) | (16 << 16) | 0x8203, (16 << 24) | (64 << 16) | 0x8203, (16 << 24) | (128 << 16) | 0x8203,
                       3, 0x403, 0x603, (128 << 16) | 0x403, (96 << 16) | 0x403, (48 << 16) | 0x403,
                       (16 << 16) | 0x403, (96 << 16) | 0x2403, (48 << 16) | 0x2403,
                       (16 << 16) | 0x2403, (128 << 16) | 0x4603, (96 << 16) | 0x4603, (64 << 16) | 0x4603, (48 << 16) | 0x4603,
                       (32 << 16) | 0x4603, (16 << 16) | 0x4603}) {
            const int reps = 64;
            probe_kernel<<<grid, 256, smem_bytes>>>(dA, tmap, dD, 3, reps, nm, dC);
            CK(cudaGetLastError());
            CK(cudaDeviceSynchronize());
            std::vector<long long> hc(2 * grid);
            CK(cudaMemcpy(hc.data(), dC, 16 * grid, cudaMemcpyDeviceToHost));
            const int pp = nm & 0xff;
            const int per_k = (nm & 0x8000) ? (pp == 1 || pp == 4 ? 1 : 2) : pp;
            if (grid == 148)
                continue;
            const int nmma = reps * (K / 8) * per_k;
            long long mn = hc[1], mx = hc[1];
            for (int b = 0; b < grid; ++b) {
                mn = std::min(mn, hc[2 * b + 1]);
                mx = std::max(mx, hc[2 * b + 1]);
            }
            printf("timing grid %d, flags 0x%x (N %d), %d MMA/k-step%s: %d MMAs, issue %lld cyc, complete min %lld max %lld "
                   "-> %.1f .. %.1f cyc/MMA\n",
                   grid, nm & 0xff00ffff, ((nm >> 16) & 0xff) ? ((nm >> 16) & 0xff) : 128, nm & 0xff, (nm & 0x200) ? " + 2 commits per 12" : "", nmma, hc[0], mn, mx,
                   (double)mn / nmma, (double)mx / nmma);
        }
    return 0;
}
