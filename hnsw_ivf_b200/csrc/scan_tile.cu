// scan_tile.cu -- K2 for lists that many queries of the batch probe: a CTA-tiled scan.
//
// Same contract as scan.cu (IVFFlatScanner::scan_codes faiss/IndexIVFFlat.cpp:155-179 applied to
// every (query, probe) pair, k best keys per (pair, row split) left in `part`), different shape:
// when a 10k-query batch probes 4096 lists, ~80 queries meet in every list, so the list rows are
// worth staging ONCE in shared memory and applying to a tile of TQ queries, instead of letting
// each warp re-stream them from L1/L2.
//
//  * Work item = (list, tile of <= TQ queries probing it, row split). One CTA of 8 warps.
//  * The query tile is transposed into shared memory once ([float4 column][query]).
//  * List rows arrive through a ring of NST stages filled by TMA (cp.async.bulk.tensor, completion
//    on an mbarrier). The device slab keeps each 32-row block dimension-interleaved (common.cuh);
//    seen as a 3-D tensor [block][float4 column][128 floats] one box of 16 blocks x KC columns is
//    one stage (16 KB) and ONE instruction by one thread; rows past the end of the slab and
//    columns past nf4 are zero-filled by the engine.
//  * Each thread owns 8 rows x 8 queries of fp32 accumulators; a warp covers 64 rows x 32
//    queries, the 8 warps of the CTA 512 rows. Lanes are arranged rows = lane >> 2, queries =
//    lane & 3, so that each quarter-warp of an LDS.128 reads at most 64 distinct bytes: measured
//    on B200 (tools/ubench_lds.cu) such a load costs 2.5 cycles of the SM's shared-memory pipe
//    against 4.0 when a quarter-warp reads 8 distinct float4. 16 LDS.128 feed 256 FMAs per
//    thread and column: ~60 % shared-memory pipe at full FMA rate.
//  * A stage is released by an atomic counter: the last of the 8 warps to finish it refills it.
//    The only CTA-wide barrier is one per 512-row chunk, in the candidate epilogue.
//  * Direct (x-y)^2 / x*y accumulation in fp32, dimension order 0..d-1 per accumulator.
//  * After the last column chunk of a row chunk, values that beat the query's current threshold
//    (strict compare, as the reference's C::cmp(simi[0], dis)) are appended to a small per-query
//    candidate buffer in shared memory; a full buffer is bitonic-sorted down to its k best, which
//    tightens the threshold. Nothing is ever dropped: a push that finds the buffer full is retried
//    after the compaction.
#include "common.cuh"
#include "kernels.h"

#include <cuda.h>

namespace b200ivf {

namespace {

constexpr int kTileThreads = 256;
constexpr int kTileKC = 2; // float4 columns per pipeline stage
constexpr int kTileRB = 16; // 32-row blocks per row chunk: two per warp

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile(
                "{\n\t"
                ".reg .pred p;\n\t"
                "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                "selp.u32 %0, 1, 0, p;\n\t"
                "}"
                : "=r"(done)
                : "r"(smem_u32(bar)), "r"(parity)
                : "memory");
    } while (!done);
}
// one TMA box {128 floats, KC columns, RB blocks} at (0, col, blk) -> shared, bytes counted on `bar`
__device__ __forceinline__ void tma_box_g2s(void* dst, const CUtensorMap* map, int col, int blk,
                                            uint64_t* bar) {
    asm volatile(
            "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
            "[%0], [%1, {%2, %3, %4}], [%5];" ::"r"(smem_u32(dst)),
            "l"(reinterpret_cast<uint64_t>(map)), "r"(0), "r"(col), "r"(blk), "r"(smem_u32(bar))
            : "memory");
}

template <int METRIC>
__device__ __forceinline__ float acc4(float acc, const float4& x, const float4& y) {
    if (METRIC == kMetricL2) {
        float t;
        t = x.x - y.x;
        acc = fmaf(t, t, acc);
        t = x.y - y.y;
        acc = fmaf(t, t, acc);
        t = x.z - y.z;
        acc = fmaf(t, t, acc);
        t = x.w - y.w;
        acc = fmaf(t, t, acc);
    } else {
        acc = fmaf(x.x, y.x, acc);
        acc = fmaf(x.y, y.y, acc);
        acc = fmaf(x.z, y.z, acc);
        acc = fmaf(x.w, y.w, acc);
    }
    return acc;
}

// One stage of the ring applied to the register tile. FULL: all 8 query slots are live and the
// stage holds KC columns -- no branch inside, so ptxas can software-pipeline the LDS.128s.
template <int METRIC, bool FULL>
__device__ __forceinline__ void tile_stage(float (&acc)[8][8], const float4* __restrict__ Yw,
                                           const float4* __restrict__ xp, int ncol, int jmax) {
#pragma unroll
    for (int c = 0; c < kTileKC; ++c) {
        if (FULL || c < ncol) {
            float4 y[8];
#pragma unroll
            for (int i = 0; i < 8; ++i)
                y[i] = Yw[((i >> 2) * kTileKC + c) * kBlockRows + 8 * (i & 3)];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (FULL || j < jmax) {
                    const float4 x = xp[c * kTileTQ + 4 * j];
#pragma unroll
                    for (int i = 0; i < 8; ++i)
                        acc[i][j] = acc4<METRIC>(acc[i][j], x, y[i]);
                }
            }
        }
    }
}

// Thread (rg = lane >> 2, qg = lane & 3) of warp w owns rows {rg + 8 i} (i < 8) of the 64 rows
// (two 32-row blocks) that warp w takes of every 512-row chunk, and queries {qg + 4 j} (j < 8).
template <int METRIC>
__global__ void __launch_bounds__(kTileThreads, 2)
        ivf_scan_tile_kernel(const TileScanArgs a, const __grid_constant__ CUtensorMap tmap) {
    constexpr int TQ = kTileTQ;
    constexpr int RB = kTileRB;
    constexpr int KC = kTileKC;
    constexpr int NW = kTileThreads / 32;
    constexpr int STAGE_F4 = RB * KC * kBlockRows;

    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;
    const int rg = lane >> 2, qg = lane & 3;

    const uint32_t item = blockIdx.x;
    if (item >= a.item_off[a.nlists])
        return;
    int lo = 0, hi = a.nlists;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (a.item_off[mid] <= item)
            lo = mid;
        else
            hi = mid;
    }
    const int l = lo;
    const int g0 = a.grp_start[l];
    const int cnt_l = a.grp_start[l + 1] - g0;
    const int nsub = (cnt_l + TQ - 1) / TQ;
    const uint32_t local = item - a.item_off[l];
    const int split = (int)(local / (uint32_t)nsub);
    const int sub = (int)(local % (uint32_t)nsub);
    const int q0 = g0 + sub * TQ;
    const int nqw = min(TQ, cnt_l - sub * TQ);

    const int len = a.list_len[l];
    const int nb = (len + kBlockRows - 1) / kBlockRows;
    const int b0 = (int)(((long long)nb * split) / a.nsplit);
    const int b1 = (int)(((long long)nb * (split + 1)) / a.nsplit);
    const int nf4 = a.nf4;
    const int nst = a.nstages;
    const int cap = a.cap;

    // shared-memory carve-up
    float4* Xs = reinterpret_cast<float4*>(smem_raw);                 // [nf4][TQ]
    float4* Ys = Xs + (size_t)nf4 * TQ;                               // [nst][RB][KC][32]
    uint64_t* bufs = reinterpret_cast<uint64_t*>(Ys + (size_t)nst * STAGE_F4); // [TQ][cap]
    uint64_t* full = bufs + (size_t)TQ * cap;                         // [4] mbarriers
    int* done_s = reinterpret_cast<int*>(full + 4);                   // [4] warps done per stage
    float* thr_s = reinterpret_cast<float*>(done_s + 4);              // [TQ]
    int* cnt_s = reinterpret_cast<int*>(thr_s + TQ);                  // [TQ]

    const int nchunks = (b1 - b0 + RB - 1) / RB;
    const int nkc = (nf4 + KC - 1) / KC;
    const int niter = nchunks * nkc;
    const uint32_t blk0 = a.list_blk0[l];

    // producer step (one thread): fill stage (it % nst) with column chunk (it % nkc) of row chunk
    // (it / nkc): one TMA box
    auto issue = [&](int it) {
        const int ci = it / nkc, ki = it - ci * nkc;
        const int st = it % nst;
        mbar_expect_tx(&full[st], (uint32_t)(STAGE_F4 * sizeof(float4)));
        tma_box_g2s(Ys + (size_t)st * STAGE_F4, &tmap, ki * KC, (int)blk0 + b0 + ci * RB,
                    &full[st]);
    };

    if (tid == 0) {
        for (int s = 0; s < 4; ++s) {
            mbar_init(&full[s], 1);
            done_s[s] = 0;
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int j = tid; j < TQ; j += kTileThreads) {
        thr_s[j] = j < nqw ? kNeutralScore : -__int_as_float(0x7f800000);
        cnt_s[j] = 0;
    }
    __syncthreads();
    if (tid == 0) {
        const int npre = min(nst, niter);
        for (int it = 0; it < npre; ++it)
            issue(it);
    }
    // query tile, transposed: Xs[c][j] = dims [4c, 4c+4) of the j-th query of the tile
    {
        const int d = a.d;
        const bool vec = (d & 3) == 0 && (reinterpret_cast<uintptr_t>(a.queries) & 15) == 0;
        for (int idx = tid; idx < TQ * nf4; idx += kTileThreads) {
            const int j = idx / nf4, c = idx - j * nf4;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (j < nqw) {
                const int q = a.pair_query ? a.pair_query[q0 + j] : (q0 + j);
                const float* src = a.queries + (size_t)q * d + 4 * c;
                if (vec) {
                    v = __ldg(reinterpret_cast<const float4*>(src));
                } else {
                    const int rem = d - 4 * c;
                    v.x = __ldg(src);
                    if (rem > 1)
                        v.y = __ldg(src + 1);
                    if (rem > 2)
                        v.z = __ldg(src + 2);
                    if (rem > 3)
                        v.w = __ldg(src + 3);
                }
            }
            Xs[(size_t)c * TQ + j] = v;
        }
    }
    __syncthreads();

    const int jmax = (nqw + 3) >> 2; // live query slots per thread (uniform over the CTA)
    const bool full_tile = jmax == 8;

    float acc[8][8];
    for (int it = 0; it < niter; ++it) {
        const int ci = it / nkc, ki = it - ci * nkc;
        const int c0 = ki * KC;
        const int ncol = min(KC, nf4 - c0);
        const int st = it % nst;
        const int bbw = b0 + ci * RB + 2 * warp; // first of this warp's two blocks
        if (ki == 0) {
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    acc[i][j] = 0.f;
        }
        mbar_wait(&full[st], (uint32_t)((it / nst) & 1));
        if (bbw < b1) { // warps whose rows lie past the end of the list skip the arithmetic
            const float4* Yw =
                    Ys + (size_t)st * STAGE_F4 + (size_t)(2 * warp) * KC * kBlockRows + rg;
            const float4* xp = Xs + (size_t)c0 * TQ + qg;
            if (full_tile && ncol == KC)
                tile_stage<METRIC, true>(acc, Yw, xp, ncol, jmax);
            else
                tile_stage<METRIC, false>(acc, Yw, xp, ncol, jmax);
        }
        // release the stage: the last of the 8 warps to finish refills it (no CTA-wide barrier)
        __syncwarp();
        if (lane == 0) {
            __threadfence_block();
            const int prev = atomicAdd(&done_s[st], 1);
            if (prev == NW - 1) {
                done_s[st] = 0;
                __threadfence_block();
                if (it + nst < niter)
                    issue(it + nst);
            }
        }

        if (ki == nkc - 1) {
            // ---- candidates of this row chunk ------------------------------------------------
            const uint32_t grow0 = (blk0 + (uint32_t)bbw) * kBlockRows + rg;
            unsigned long long pend = 0ull;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int qi = qg + 4 * j;
                const float t = thr_s[qi]; // -inf for padding queries
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int bb = bbw + (i >> 2);
                    const int row = bb * kBlockRows + rg + 8 * (i & 3);
                    const float s = (METRIC == kMetricL2 ? acc[i][j] : -acc[i][j]) + 0.f;
                    acc[i][j] = s;
                    if (bb < b1 && row < len && s < t) {
                        const int pos = atomicAdd(&cnt_s[qi], 1);
                        if (pos < cap)
                            bufs[(size_t)qi * cap + pos] = make_key(s, grow0 + 8 * i);
                        else
                            pend |= 1ull << (i * 8 + j);
                    }
                }
            }
            // a full buffer is compacted (threshold tightens), then the pushes that found it
            // full are retried; rare after the first chunk
            while (__syncthreads_or(pend != 0ull)) {
                for (int qi = warp; qi < nqw; qi += NW) {
                    if (cnt_s[qi] >= cap) {
                        KeyBufState s = warp_compact_keys(bufs + (size_t)qi * cap, cap, thr_s[qi],
                                                          a.k, cap, lane);
                        if (lane == 0) {
                            cnt_s[qi] = s.cnt;
                            thr_s[qi] = s.thr;
                        }
                    }
                }
                __syncthreads();
                if (pend) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int qi = qg + 4 * j;
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const unsigned long long bit = 1ull << (i * 8 + j);
                            if (pend & bit) {
                                bool keep = false;
                                if (acc[i][j] < thr_s[qi]) {
                                    const int pos = atomicAdd(&cnt_s[qi], 1);
                                    if (pos < cap)
                                        bufs[(size_t)qi * cap + pos] =
                                                make_key(acc[i][j], grow0 + 8 * i);
                                    else
                                        keep = true;
                                }
                                if (!keep)
                                    pend &= ~bit;
                            }
                        }
                    }
                }
            }
        }
    }

    // ---- the k best of every query of the tile -------------------------------------------------
    __syncthreads();
    for (int qi = warp; qi < nqw; qi += NW) {
        uint64_t* buf = bufs + (size_t)qi * cap;
        const int c = min(cnt_s[qi], cap);
        KeyBufState s = warp_compact_keys(buf, c, thr_s[qi], a.k, cap, lane);
        const int keep = min(s.cnt, a.k);
        uint64_t* out = a.part + ((size_t)(q0 + qi) * a.nsplit + split) * a.k;
        for (int i = lane; i < a.k; i += 32)
            out[i] = i < keep ? buf[i] : kSentinelKey;
    }
}

template <int METRIC>
cudaError_t launch_tile_one(uint32_t max_items, const TileScanArgs& a, cudaStream_t st) {
    const size_t smem = scan_tile_smem_bytes(a.nf4, a.cap, a.nstages);
    auto kern = ivf_scan_tile_kernel<METRIC>;
    cudaError_t e =
            cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess)
        return e;
    if (max_items == 0)
        return cudaSuccess;
    kern<<<max_items, kTileThreads, smem, st>>>(a, *reinterpret_cast<const CUtensorMap*>(a.tmap));
    count_launches(1);
    return cudaGetLastError();
}

} // namespace

// The slab as a rank-3 fp32 tensor {128 floats (one float4 column of a block), nf4, blocks} with a
// {128, KC, RB} box. libcuda is reached through the runtime (no link-time dependency on it).
cudaError_t make_slab_tensor_map(void* map_out, const void* slab, int nf4, uint64_t nblocks) {
    static_assert(sizeof(CUtensorMap) == kTensorMapBytes, "CUtensorMap size");
    using EncodeFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static EncodeFn encode = nullptr;
    if (!encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault,
                                                &qres);
        if (e != cudaSuccess)
            return e;
        if (qres != cudaDriverEntryPointSuccess || !fn)
            return cudaErrorNotSupported;
        encode = reinterpret_cast<EncodeFn>(fn);
    }
    const cuuint64_t gdim[3] = {128, (cuuint64_t)nf4, nblocks ? nblocks : 1};
    const cuuint64_t gstride[2] = {512, (cuuint64_t)nf4 * 512};
    const cuuint32_t box[3] = {128, (cuuint32_t)kTileKC, (cuuint32_t)kTileRB};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUresult r = encode(reinterpret_cast<CUtensorMap*>(map_out),
                              CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<void*>(slab), gdim,
                              gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                              CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? cudaSuccess : cudaErrorInvalidValue;
}

size_t scan_tile_smem_bytes(int nf4, int cap, int nstages) {
    const size_t tq = kTileTQ;
    return tq * nf4 * sizeof(float4) +
            (size_t)nstages * kTileRB * kTileKC * kBlockRows * sizeof(float4) +
            tq * cap * sizeof(uint64_t) + 4 * sizeof(uint64_t) + 4 * sizeof(int) + tq * 8;
}

cudaError_t launch_ivf_scan_tile(int metric, uint32_t max_items, const TileScanArgs& a,
                                 cudaStream_t st) {
    if (a.nstages < 2 || a.nstages > 4)
        return cudaErrorInvalidValue;
    if (metric == kMetricL2)
        return launch_tile_one<kMetricL2>(max_items, a, st);
    return launch_tile_one<kMetricInnerProduct>(max_items, a, st);
}

} // namespace b200ivf
