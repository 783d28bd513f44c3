// scan_tc.cu -- the tensor-core form of the scan: K1 (coarse quantizer) and K2 for lists that more
// than a handful of queries of the batch probe, for d <= 128.
//
// Same contract as scan.cu / scan_tile.cu (IVFFlatScanner::scan_codes,
// faiss/IndexIVFFlat.cpp:155-179, applied to every (query, probe) pair; for K1 knn_L2sqr /
// knn_inner_product over the centroid table, faiss/utils/distances.cpp:885-920,836-871): the best
// KP keys of every (pair, row split, epilogue group) are left in `part`. The scores this kernel
// produces are 3xTF32 products (x_hi*y_hi + x_lo*y_hi + x_hi*y_lo, fp32 accumulate): good to ~1e-6
// of |x||y|, which is enough to SELECT; the k-select (select.cu) recomputes the survivors'
// distances in plain fp32 in the reference's operation order before the final ordering, so the
// distances that leave the library never come from the tensor core.
//
// Work item = (list, tile of <= 128 queries probing it, row split); one CTA each.
//   D[128 queries x 128 rows] (TMEM, two buffers) += A[128 x 8] (TMEM) * B[128 rows x 8]^T (smem)
//  * A: the query tile, scaled by -1 (inner product) or -2 (L2), split into tf32 hi / lo parts and
//    written ONCE into tensor memory (tcgen05.st): 2 x d columns, lane = query.
//  * B hi: list rows straight from HBM by TMA. The slab's 32-row dimension-interleaved blocks
//    (common.cuh) are exactly the K-major no-swizzle UMMA operand layout (8 rows x 16 bytes = one
//    128-byte core matrix, SBO = 128 B); the TMA box {128 floats, 4 blocks, 8 columns} puts the
//    column stride at LBO = 2048 B. The tensor core ignores the low 13 mantissa bits of an fp32
//    operand (measured, tools/tc_probe.cu), so the raw fp32 rows ARE the hi part.
//  * B lo: y - trunc_tf32(y), rounded to tf32, computed by the split warps from the hi stage into a
//    twin stage (generic-proxy writes + fence.proxy.async).
//  * One thread issues the MMAs (3 per 8 dims); stages are released by tcgen05.commit.
//  * Epilogue: TWO groups of 4 warps (thread = query = TMEM lane); group g drains accumulator
//    buffer g, i.e. the tiles t = g (mod 2), so each group has two MMA tile times per tile.
//    Per 32 columns: tcgen05.ld, add the norms (L2: |y|^2 - 2xy + |x|^2), strict compare against
//    the thread's own threshold (the reference's C::cmp(simi[0], dis)) into a pass mask. Each
//    thread keeps its KP best (score, row) SORTED IN REGISTERS, so the threshold is always the
//    current KP-th best, as the reference's heap top is. Sparse chunks (steady state): passing
//    scores are staged in shared memory and every lane walks its OWN mask (ffs), so a warp
//    spends max-over-lanes(popc) insertions, not one per column some lane accepted. Dense chunks
//    (start of a list): the 32 scores are sorted by a register bitonic network and merged with
//    the KP best in two half-cleaner steps.
#include "common.cuh"
#include "kernels.h"
#include "tc_ptx.cuh"

#include <cstdio>
#include <cstdlib>

namespace b200ivf {

namespace {

using namespace ptx;

constexpr int kTcEpiWarps = 8;    // warps 0-3: epilogue group 0, warps 4-7: group 1
constexpr int kTcTmaWarp = 8;
constexpr int kTcMmaWarp = 9;
constexpr int kTcSplitWarp0 = 10; // warps 10-13
constexpr int kTcSplitThreads = 128;
constexpr int kTcThreads = (kTcSplitWarp0 + kTcSplitThreads / 32) * 32; // 448
constexpr int kTcKC = 8;          // float4 columns per stage (4 MMA k-steps)
constexpr int kTcTileBlocks = 4;  // 32-row blocks per MMA tile (N = 128)
constexpr int kTcTileRows = kTcTileBlocks * kBlockRows;
constexpr int kTcStageF4 = kTcKC * kTcTileBlocks * kBlockRows; // 1024 float4 = 16 KB
constexpr uint32_t kTcLBO = kTcTileBlocks * 512, kTcSBO = 128;
constexpr int kTcMaxStages = 12; // raw (hi) ring: deep, it hides the HBM latency
constexpr int kTcLoStages = 3;   // lo ring: shallow, it only decouples the split warps from the MMAs
constexpr uint32_t kTcColsD = 0, kTcColsAhi = 256, kTcColsAlo = 384;
constexpr int kTcDenseMin = 6;    // network path when some lane has >= this many passing columns

__device__ __forceinline__ float tf32_trunc(float x) {
    return __uint_as_float(__float_as_uint(x) & 0xffffe000u);
}
__device__ __forceinline__ float tf32_rn(float x) {
    return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
}
__device__ __forceinline__ float tf32_lo(float y) {
    return tf32_rn(y - tf32_trunc(y));
}

// ---- register sorting networks over (score, row) pairs ------------------------------------------
// compare-exchange: afterwards a <= b
__device__ __forceinline__ void ce(float& a, uint32_t& ia, float& b, uint32_t& ib) {
    const bool sw = b < a;
    const float lo = fminf(a, b), hi = fmaxf(a, b);
    const uint32_t il = sw ? ib : ia, ih = sw ? ia : ib;
    a = lo;
    b = hi;
    ia = il;
    ib = ih;
}
// keep the smaller of (a, b) in a
__device__ __forceinline__ void keep_min(float& a, uint32_t& ia, float b, uint32_t ib) {
    const bool sw = b < a;
    a = fminf(a, b);
    ia = sw ? ib : ia;
}
// bitonic sequence of N (power of two) -> ascending
template <int N>
__device__ __forceinline__ void bitonic_merge(float* s, uint32_t* id) {
#pragma unroll
    for (int stride = N >> 1; stride > 0; stride >>= 1) {
#pragma unroll
        for (int i = 0; i < N; ++i) {
            if ((i & stride) == 0)
                ce(s[i], id[i], s[i | stride], id[i | stride]);
        }
    }
}
// any sequence of N -> ascending
template <int N>
__device__ __forceinline__ void bitonic_sort(float* s, uint32_t* id) {
#pragma unroll
    for (int size = 2; size <= N; size <<= 1) {
#pragma unroll
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
#pragma unroll
            for (int i = 0; i < N; ++i) {
                const int j = i | stride;
                if ((i & stride) == 0) {
                    if ((i & size) == 0)
                        ce(s[i], id[i], s[j], id[j]);
                    else
                        ce(s[j], id[j], s[i], id[i]);
                }
            }
        }
    }
}

// Merge the 32 scores of a chunk (non-candidates already set to +inf) into the thread's sorted KP
// best. sc / id are clobbered.
template <int KP>
__device__ __forceinline__ void merge_chunk_network(float (&sc)[32], uint32_t (&id)[32],
                                                    float (&vs)[KP], uint32_t (&vr)[KP]) {
    bitonic_sort<16>(sc, id);
    bitonic_sort<16>(sc + 16, id + 16);
    if (KP == 16) {
        // the 16 smallest of the chunk as a bitonic sequence, then sorted
#pragma unroll
        for (int i = 0; i < 16; ++i)
            keep_min(sc[i], id[i], sc[31 - i], id[31 - i]);
        bitonic_merge<16>(sc, id);
        // the 16 smallest of (vs, chunk best)
#pragma unroll
        for (int i = 0; i < 16; ++i)
            keep_min(vs[i], vr[i], sc[15 - i], id[15 - i]);
        bitonic_merge<16>(vs, vr);
    } else {
        // [ascending 16, ascending 16] -> reverse the second half in place -> bitonic 32 -> sorted
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float ts = sc[16 + i];
            sc[16 + i] = sc[31 - i];
            sc[31 - i] = ts;
            const uint32_t ti = id[16 + i];
            id[16 + i] = id[31 - i];
            id[31 - i] = ti;
        }
        bitonic_merge<32>(sc, id);
#pragma unroll
        for (int i = 0; i < 32; ++i)
            keep_min(vs[i], vr[i], sc[31 - i], id[31 - i]);
        bitonic_merge<32>(vs, vr);
    }
}

// sorted insertion of one candidate (chain through one FMNMX per slot)
template <int KP>
__device__ __forceinline__ void insert_sorted(float sj, uint32_t rj, float (&vs)[KP],
                                              uint32_t (&vr)[KP]) {
#pragma unroll
    for (int u = 0; u < KP; ++u) {
        const bool lt = sj < vs[u];
        const float vmin = fminf(sj, vs[u]);
        const float vmax = fmaxf(sj, vs[u]);
        const uint32_t rmin = lt ? rj : vr[u];
        const uint32_t rmax = lt ? vr[u] : rj;
        vs[u] = vmin;
        vr[u] = rmin;
        sj = vmax;
        rj = rmax;
    }
}

struct TcSmem {
    float4* hi;        // [nst][kTcStageF4]
    float4* lo;        // [kTcLoStages][kTcStageF4]
    float* stg;        // [8 warps][32 columns][32 lanes] scores of the current chunk
    float* xnorm;      // [2][128] partial |x|^2 of the two halves of the query
    uint64_t* full;    // [kTcMaxStages] TMA landed
    uint64_t* empty;   // [kTcMaxStages] MMAs that read the raw stage are done
    uint64_t* lo_rdy;  // [4] lo stage written
    uint64_t* lo_empty;// [4] MMAs that read the lo stage are done
    uint64_t* d_full;  // [2] accumulator buffer complete
    uint64_t* d_empty; // [2] accumulator buffer drained by its epilogue group
    uint64_t* a_rdy;   // [1] query tile is in TMEM
    uint32_t* tmem;    // [1]
};

__device__ __forceinline__ TcSmem carve(unsigned char* raw, int nst) {
    TcSmem s;
    s.hi = reinterpret_cast<float4*>(raw);
    s.lo = s.hi + (size_t)nst * kTcStageF4;
    s.stg = reinterpret_cast<float*>(s.lo + (size_t)kTcLoStages * kTcStageF4);
    s.xnorm = s.stg + kTcEpiWarps * 32 * 32;
    s.full = reinterpret_cast<uint64_t*>(s.xnorm + 2 * kTcTQ);
    s.empty = s.full + kTcMaxStages;
    s.lo_rdy = s.empty + kTcMaxStages;
    s.lo_empty = s.lo_rdy + 4;
    s.d_full = s.lo_empty + 4;
    s.d_empty = s.d_full + 2;
    s.a_rdy = s.d_empty + 2;
    s.tmem = reinterpret_cast<uint32_t*>(s.a_rdy + 1);
    return s;
}

#define TC_TRACE(role, idx)                                              \
    do {                                                                 \
        if (tracing && (idx) < 64)                                       \
            a.trace[(role) * 64 + (idx)] = clock64();                    \
    } while (0)

template <int METRIC, int KP>
__global__ void __launch_bounds__(kTcThreads, 1)
        ivf_scan_tc_kernel(const TcScanArgs a, const __grid_constant__ CUtensorMap tmap) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;

    // ---- which item --------------------------------------------------------------------------
    const uint32_t item = blockIdx.x;
    if (item >= a.item_off[a.nlists])
        return;
    int lo_l = 0, hi_l = a.nlists;
    if (a.item_list) {
        lo_l = a.item_list[item]; // written by the plan: no dependent-load chain
    } else {
        while (hi_l - lo_l > 1) {
            const int mid = (lo_l + hi_l) >> 1;
            if (a.item_off[mid] <= item)
                lo_l = mid;
            else
                hi_l = mid;
        }
    }
    const int l = lo_l;
    const int g0 = a.grp_start[l];
    const int cnt_l = a.grp_start[l + 1] - g0;
    const int nsub = (cnt_l + kTcTQ - 1) / kTcTQ;
    const uint32_t local = item - a.item_off[l];
    const int split = (int)(local / (uint32_t)nsub);
    const int sub = (int)(local % (uint32_t)nsub);
    const int q0 = g0 + sub * kTcTQ;
    const int nqw = min(kTcTQ, cnt_l - sub * kTcTQ);

    const int len = a.list_len[l];
    const int nb = (len + kBlockRows - 1) / kBlockRows;
    const int b0 = (int)(((long long)nb * split) / a.nsplit);
    const int b1 = (int)(((long long)nb * (split + 1)) / a.nsplit);
    const uint32_t blk0 = a.list_blk0[l];
    const int nf4 = a.nf4;
    const int nst = a.nstages;
    const int ntiles = (b1 - b0 + kTcTileBlocks - 1) / kTcTileBlocks;
    const int nsg = (nf4 + kTcKC - 1) / kTcKC;  // stages per tile
    const int nks = (nf4 + 1) >> 1;             // MMA k-steps (8 dims each) per tile
    const int niter = ntiles * nsg;

    const TcSmem s = carve(smem_raw, nst);
    const bool tracing = a.trace != nullptr && item == a.trace_item;
    if (tid == 0) {
        TC_TRACE(0, 0);
        if (tracing) {
            a.trace[0 * 64 + 2] = ntiles;
            a.trace[0 * 64 + 3] = nqw;
        }
    }

    if (tid == 0) {
        for (int i = 0; i < kTcMaxStages; ++i) {
            mbar_init(&s.full[i], 1);
            mbar_init(&s.empty[i], 1);
        }
        for (int i = 0; i < 4; ++i) {
            mbar_init(&s.lo_rdy[i], kTcSplitThreads / 32);
            mbar_init(&s.lo_empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&s.d_full[i], 1);
            mbar_init(&s.d_empty[i], 4);
        }
        mbar_init(s.a_rdy, kTcEpiWarps);
        fence_barrier_init();
    }
    if (warp == kTcMmaWarp) {
        tmem_alloc(s.tmem, 512);
        tmem_relinquish();
    }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem = *s.tmem;
    if (tid == 0)
        TC_TRACE(0, 1);

    if (warp == kTcTmaWarp) {
        // ======================= TMA producer ===================================================
        if (lane == 0) {
            tma_prefetch_desc(&tmap);
            int st = 0;
            uint32_t ph = 0;
            for (int t = 0; t < ntiles; ++t) {
                const int blk = (int)blk0 + b0 + t * kTcTileBlocks;
                for (int sg = 0; sg < nsg; ++sg) {
                    if (t * nsg + sg >= nst)
                        mbar_wait(&s.empty[st], ph ^ 1u);
                    if ((a.dbg & 8) && t * nsg + sg >= nst) {
                        mbar_arrive(&s.full[st]); // experiment: recycle the stage without a load
                    } else {
                        mbar_expect_tx(&s.full[st], (uint32_t)(kTcStageF4 * sizeof(float4)));
                        tma_load_3d(s.hi + (size_t)st * kTcStageF4, &tmap, 0, blk, sg * kTcKC,
                                    &s.full[st]);
                    }
                    if (++st == nst) {
                        st = 0;
                        ph ^= 1u;
                    }
                }
                TC_TRACE(1, t);
            }
        }
    } else if (warp == kTcMmaWarp) {
        // ======================= MMA issuer =====================================================
        // The whole warp runs the loop (converged control flow, uniform registers); one elected
        // lane issues the tcgen05 instructions of a stage as one straight-line block.
        {
            const uint32_t idesc = make_idesc_tf32(kTcTQ, kTcTileRows);
            mbar_wait(s.a_rdy, 0);
            tc_fence_after_sync();
            if (lane == 0)
                TC_TRACE(0, 4);
            const uint64_t desc0 = make_smem_desc_kmajor_noswz(0, kTcLBO, kTcSBO);
            const uint32_t hi0 = smem_u32(s.hi), lo0 = smem_u32(s.lo);
            int st = 0, ls = 0;
            uint32_t lph = 0;
            constexpr uint64_t kstep = (2 * kTcLBO) >> 4; // 2 float4 columns per k-step
            for (int t = 0; t < ntiles; ++t) {
                const int b = t & 1;
                if (t >= 2) {
                    mbar_wait(&s.d_empty[b], (uint32_t)(((t >> 1) & 1) ^ 1));
                    tc_fence_after_sync();
                }
                const uint32_t tD = tmem + kTcColsD + (uint32_t)b * 128u;
                for (int sg = 0; sg < nsg; ++sg) {
                    // lo_rdy implies full: the split warps waited for the TMA before they wrote lo
                    mbar_wait(&s.lo_rdy[ls], lph);
                    tc_fence_after_sync();
                    // descriptors advance by 2 columns (2 * LBO bytes, in 16-byte units) per k-step
                    const uint64_t dhi =
                            desc0 + (uint64_t)((hi0 + (uint32_t)st * kTcStageF4 * 16u) >> 4);
                    const uint64_t dlo =
                            desc0 + (uint64_t)((lo0 + (uint32_t)ls * kTcStageF4 * 16u) >> 4);
                    const int ks0 = sg * (kTcKC / 2);
                    const int ksn = nks - ks0; // >= 1
                    const uint32_t ahi = tmem + kTcColsAhi + (uint32_t)ks0 * 8u;
                    const uint32_t alo = tmem + kTcColsAlo + (uint32_t)ks0 * 8u;
                    const uint32_t acc0 = sg != 0 ? 1u : 0u;
                    const bool last = sg == nsg - 1;
                    // straight line under the elect, every operand prepared before it: what the
                    // issuing lane executes between two tcgen05.mma paces this role
                    if (elect_one_sync()) {
                        if (ksn >= kTcKC / 2) {
#pragma unroll
                            for (int ks = 0; ks < kTcKC / 2; ++ks) {
                                // small terms first, the big product last
                                mma_tf32_ts(tD, alo + ks * 8, dhi + ks * kstep, idesc,
                                            ks ? 1u : acc0);
                                mma_tf32_ts(tD, ahi + ks * 8, dlo + ks * kstep, idesc, 1u);
                                mma_tf32_ts(tD, ahi + ks * 8, dhi + ks * kstep, idesc, 1u);
                            }
                        } else {
                            for (int ks = 0; ks < ksn; ++ks) {
                                mma_tf32_ts(tD, alo + ks * 8, dhi + ks * kstep, idesc,
                                            ks ? 1u : acc0);
                                mma_tf32_ts(tD, ahi + ks * 8, dlo + ks * kstep, idesc, 1u);
                                mma_tf32_ts(tD, ahi + ks * 8, dhi + ks * kstep, idesc, 1u);
                            }
                        }
                        mma_commit(&s.empty[st]);
                        mma_commit(&s.lo_empty[ls]);
                        if (last)
                            mma_commit(&s.d_full[b]);
                    }
                    __syncwarp();
                    if (++st == nst)
                        st = 0;
                    if (++ls == kTcLoStages) {
                        ls = 0;
                        lph ^= 1u;
                    }
                }
                if (lane == 0)
                    TC_TRACE(2, t);
            }
        }
    } else if (warp >= kTcSplitWarp0) {
        // ======================= split warps: lo = tf32(y - trunc(y)) ===========================
        const int t0 = tid - kTcSplitWarp0 * 32;
        int st = 0, ls = 0;
        uint32_t ph = 0, lph = 0;
        for (int it = 0; it < niter; ++it) {
            mbar_wait(&s.full[st], ph);
            if (it >= kTcLoStages)
                mbar_wait(&s.lo_empty[ls], lph ^ 1u);
            const float4* src = s.hi + (size_t)st * kTcStageF4;
            float4* dst = s.lo + (size_t)ls * kTcStageF4;
#pragma unroll
            for (int i = 0; i < kTcStageF4 / kTcSplitThreads && !(a.dbg & 4); i += 4) {
                float4 y[4];
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    y[u] = src[t0 + (i + u) * kTcSplitThreads];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    float4 r;
                    r.x = tf32_lo(y[u].x);
                    r.y = tf32_lo(y[u].y);
                    r.z = tf32_lo(y[u].z);
                    r.w = tf32_lo(y[u].w);
                    dst[t0 + (i + u) * kTcSplitThreads] = r;
                }
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0)
                mbar_arrive(&s.lo_rdy[ls]);
            if (tid == kTcSplitWarp0 * 32 && (it + 1) % nsg == 0)
                TC_TRACE(3, it / nsg);
            if (++st == nst) {
                st = 0;
                ph ^= 1u;
            }
            if (++ls == kTcLoStages) {
                ls = 0;
                lph ^= 1u;
            }
        }
    } else {
        // ======================= epilogue: 2 groups x 4 warps, thread = query ===================
        const int grp = warp >> 2;            // accumulator buffer / tile parity this group drains
        const int qi = tid & (kTcTQ - 1);     // query slot of the tile == TMEM lane
        const uint32_t lane_addr = (uint32_t)((warp & 3) * 32) << 16;
        const bool live = qi < nqw;
        float xn = 0.f;
        const int q = live ? (a.pair_query ? a.pair_query[q0 + qi] : (q0 + qi)) : -1;
        {
            // query -> TMEM (A operand): scaled, split into hi / lo, zero padded to 8 * nks dims;
            // group 0 writes the first half of the k-steps, group 1 the second
            const float* src = a.queries + (size_t)(q < 0 ? 0 : q) * a.d;
            const float scale = METRIC == kMetricL2 ? -2.f : -1.f;
            const int kh = (nks + 1) >> 1;
            const int c_lo = grp == 0 ? 0 : kh * 8, c_hi = grp == 0 ? kh * 8 : nks * 8;
            const bool vec = (a.d & 3) == 0 && (reinterpret_cast<uintptr_t>(a.queries) & 15) == 0;
            // all the loads of the thread's half first (<= 16 LDG.128 in flight), then the stores
            float4 xv[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const int c = c_lo + 4 * i;
                float4 u = make_float4(0.f, 0.f, 0.f, 0.f);
                if (live && c < c_hi) {
                    if (vec && c + 4 <= a.d) {
                        u = __ldg(reinterpret_cast<const float4*>(src + c));
                    } else {
                        u.x = c + 0 < a.d ? __ldg(src + c + 0) : 0.f;
                        u.y = c + 1 < a.d ? __ldg(src + c + 1) : 0.f;
                        u.z = c + 2 < a.d ? __ldg(src + c + 2) : 0.f;
                        u.w = c + 3 < a.d ? __ldg(src + c + 3) : 0.f;
                    }
                }
                xv[i] = u;
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int c = c_lo + 8 * i;
                if (c < c_hi) {
                    const float x[8] = {xv[2 * i].x,     xv[2 * i].y,     xv[2 * i].z,
                                        xv[2 * i].w,     xv[2 * i + 1].x, xv[2 * i + 1].y,
                                        xv[2 * i + 1].z, xv[2 * i + 1].w};
                    uint32_t hi[8], lo[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        xn = fmaf(x[j], x[j], xn);
                        const float xs = x[j] * scale;
                        const float h = tf32_trunc(xs);
                        hi[j] = __float_as_uint(h);
                        lo[j] = __float_as_uint(tf32_rn(xs - h));
                    }
                    tmem_st_32x32b_x8(tmem + kTcColsAhi + lane_addr + (uint32_t)c, hi);
                    tmem_st_32x32b_x8(tmem + kTcColsAlo + lane_addr + (uint32_t)c, lo);
                }
            }
            tmem_st_wait();
            tc_fence_before_sync();
            s.xnorm[grp * kTcTQ + qi] = xn;
            __syncwarp();
            if (lane == 0)
                mbar_arrive(s.a_rdy);
            named_bar_sync(1, kTcEpiWarps * 32);
            xn = s.xnorm[qi] + s.xnorm[kTcTQ + qi];
        }
        // the thread's KP best so far, ascending; empty slots hold FLT_MAX (heap_heapify's fill,
        // faiss/utils/Heap.h:333) and row ~0; dead lanes hold -inf so that nothing passes
        float vs[KP];
        uint32_t vr[KP];
#pragma unroll
        for (int j = 0; j < KP; ++j) {
            vs[j] = live ? kNeutralScore : -__int_as_float(0x7f800000);
            vr[j] = 0xffffffffu;
        }
        float* stg = s.stg + warp * 1024 + lane;
        // pruning bound of the query over ALL its work items (see TcScanArgs::qthr)
        volatile uint32_t* gq = (a.qthr != nullptr && live) ? a.qthr + q : nullptr;
        float thr_g = live ? kNeutralScore : -__int_as_float(0x7f800000);

        for (int t = grp; t < ntiles; t += 2) {
            if (gq)
                thr_g = fminf(thr_g, from_ordered_bits(*gq));
            const int tb = b0 + t * kTcTileBlocks; // first block of the tile, list-relative
            const int nvalid =
                    min(kTcTileRows, min((b1 - tb) * kBlockRows, len - tb * kBlockRows));
            const uint32_t grow0 = (blk0 + (uint32_t)tb) * kBlockRows;
            mbar_wait(&s.d_full[grp], (uint32_t)((t >> 1) & 1));
            tc_fence_after_sync();
            if ((tid & 127) == 0)
                TC_TRACE(4, t);
            const uint32_t tD = tmem + kTcColsD + (uint32_t)grp * 128u + lane_addr;
#pragma unroll 1
            for (int c0 = 0; c0 < kTcTileRows; c0 += 32) {
                uint32_t r[32];
                tmem_ld_32x32b_x32(tD + (uint32_t)c0, r);
                tmem_ld_wait();
                if (c0 == kTcTileRows - 32) {
                    // accumulator buffer drained: hand it back to the MMA thread
                    tc_fence_before_sync();
                    __syncwarp();
                    if (lane == 0)
                        mbar_arrive(&s.d_empty[grp]);
                }
                if (c0 >= nvalid || (a.dbg & 1))
                    continue;
                float sc[32];
                if (METRIC == kMetricL2) {
                    const float4* np =
                            reinterpret_cast<const float4*>(a.norms + grow0 + (uint32_t)c0);
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const float4 nv = __ldg(np + i);
                        sc[4 * i + 0] = (__uint_as_float(r[4 * i + 0]) + nv.x) + xn;
                        sc[4 * i + 1] = (__uint_as_float(r[4 * i + 1]) + nv.y) + xn;
                        sc[4 * i + 2] = (__uint_as_float(r[4 * i + 2]) + nv.z) + xn;
                        sc[4 * i + 3] = (__uint_as_float(r[4 * i + 3]) + nv.w) + xn;
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        sc[i] = __uint_as_float(r[i]) + 0.f;
                }
                float thr = fminf(vs[KP - 1], thr_g);
                unsigned mask = 0u;
#pragma unroll
                for (int i = 0; i < 32; ++i)
                    mask |= (sc[i] < thr) ? (1u << i) : 0u;
                const int nv_here = nvalid - c0; // columns of this chunk that are real rows
                if (nv_here < 32)
                    mask &= (1u << nv_here) - 1u;
                const int maxpop = __reduce_max_sync(0xffffffffu, __popc(mask));
                if (maxpop >= kTcDenseMin) {
                    // dense chunk: sort it in registers and merge
                    uint32_t id[32];
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        if (!((mask >> i) & 1u))
                            sc[i] = __int_as_float(0x7f800000);
                        id[i] = grow0 + (uint32_t)(c0 + i);
                    }
                    merge_chunk_network<KP>(sc, id, vs, vr);
                } else if (maxpop > 0) {
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        stg[i * 32] = sc[i];
                    // every lane walks its own mask; a lane only reads what it wrote itself
                    while (__any_sync(0xffffffffu, mask != 0u)) {
                        if (mask) {
                            const int j = __ffs(mask) - 1;
                            mask &= mask - 1u;
                            const float sj = stg[j * 32];
                            if (sj < thr) { // thr may have tightened since the mask was taken
                                insert_sorted<KP>(sj, grow0 + (uint32_t)(c0 + j), vs, vr);
                                thr = fminf(vs[KP - 1], thr_g);
                            }
                        }
                    }
                }
            }
            // publish this thread's KP-th best if it tightens the query's bound
            if (gq && vs[KP - 1] < thr_g) {
                thr_g = vs[KP - 1];
                atomicMin(a.qthr + q, ordered_bits(thr_g));
            }
            if ((tid & 127) == 0)
                TC_TRACE(6, t);
        }
        if ((tid & 127) == 0)
            TC_TRACE(5, grp);
        // ---- the KP best of this thread's query over this group's tiles ------------------------
        if (live) {
            uint64_t* out =
                    a.part + (((size_t)(q0 + qi) * a.nsplit + split) * 2 + grp) * KP;
#pragma unroll
            for (int j = 0; j < KP; ++j)
                out[j] = vr[j] == 0xffffffffu ? kSentinelKey : make_key(vs[j], vr[j]);
        }
    }

    tc_fence_before_sync();
    __syncthreads();
    if (warp == kTcMmaWarp)
        tmem_dealloc(tmem, 512);
    if (tid == 0)
        TC_TRACE(0, 5);
}

} // namespace

// The slab as a rank-3 fp32 tensor {128 floats (one float4 column of a block), blocks, nf4}: the
// column is the OUTER box dimension so that a box {128, 4 blocks, 8 columns} lands in shared memory
// as [column][128 rows][16 bytes] -- the K-major operand layout of the MMA.
cudaError_t make_slab_tensor_map_tc(void* map_out, const void* slab, int nf4, uint64_t nblocks) {
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
    const cuuint64_t gdim[3] = {128, nblocks ? nblocks : 1, (cuuint64_t)nf4};
    const cuuint64_t gstride[2] = {(cuuint64_t)nf4 * 512, 512};
    const cuuint32_t box[3] = {128, (cuuint32_t)kTcTileBlocks, (cuuint32_t)kTcKC};
    const cuuint32_t estr[3] = {1, 1, 1};
    // (tuning experiments: B200IVF_TMA_PROMO = 0 none, 1 64 B, 2 128 B, 3 256 B)
    static const int promo_i = getenv("B200IVF_TMA_PROMO") ? atoi(getenv("B200IVF_TMA_PROMO")) : 3;
    const CUtensorMapL2promotion promo = promo_i == 0 ? CU_TENSOR_MAP_L2_PROMOTION_NONE
            : promo_i == 1                            ? CU_TENSOR_MAP_L2_PROMOTION_L2_64B
            : promo_i == 2                            ? CU_TENSOR_MAP_L2_PROMOTION_L2_128B
                                                      : CU_TENSOR_MAP_L2_PROMOTION_L2_256B;
    const CUresult r = encode(reinterpret_cast<CUtensorMap*>(map_out),
                              CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<void*>(slab), gdim,
                              gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                              CU_TENSOR_MAP_SWIZZLE_NONE, promo,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? cudaSuccess : cudaErrorInvalidValue;
}

size_t scan_tc_smem_bytes(int nstages) {
    return (size_t)(nstages + kTcLoStages) * kTcStageF4 * sizeof(float4) +
            (size_t)kTcEpiWarps * 32 * 32 * sizeof(float) + 2 * kTcTQ * sizeof(float) +
            (kTcMaxStages * 2 + 4 * 2 + 2 * 2 + 1) * sizeof(uint64_t) + 16;
}

int scan_tc_kpart(int nf4, int k) {
    if (nf4 > 32 || nf4 < 1)
        return 0;
    if (k + 4 <= 16)
        return 16;
    if (k + 8 <= 32) // (k = 25..32 would leave the exact re-rank less than 8 spare keys: the
        return 32;   // rows-major kernel takes those with 64 kept)
    return 0;
}

template <int METRIC, int KP>
static cudaError_t launch_tc_one(uint32_t max_items, const TcScanArgs& a_in, cudaStream_t st) {
    TcScanArgs a = a_in;
    const int dbg = g_tc_dbg.load(std::memory_order_relaxed); // B200IVF_TC_DBG / "tc_dbg"
    a.dbg = a.nlists > 1 ? dbg : 0; // never for the coarse quantizer: the probes must stay real
    const size_t smem = scan_tc_smem_bytes(a.nstages);
    auto kern = ivf_scan_tc_kernel<METRIC, KP>;
    cudaError_t e =
            cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess)
        return e;
    if (max_items == 0)
        return cudaSuccess;
    static const char* tr = getenv("B200IVF_TC_TRACE");
    if (tr && a.nlists > 1) {
        static long long* buf = nullptr;
        if (!buf)
            cudaMalloc(&buf, 7 * 64 * sizeof(long long));
        cudaMemsetAsync(buf, 0, 7 * 64 * sizeof(long long), st);
        a.trace = buf;
        a.trace_item = (uint32_t)atoi(tr);
    }
    kern<<<max_items, kTcThreads, smem, st>>>(a, *reinterpret_cast<const CUtensorMap*>(a.tmap));
    count_launches(1);
    g_tc_launches.fetch_add(1, std::memory_order_relaxed);
    if (a.trace) {
        long long h[7 * 64];
        cudaStreamSynchronize(st);
        cudaMemcpy(h, a.trace, sizeof(h), cudaMemcpyDeviceToHost);
        const long long t0 = h[0];
        fprintf(stderr, "TCTRACE item %u ntiles %lld nq %lld: sync %lld a_rdy %lld end %lld\n",
                a.trace_item, h[2], h[3], h[1] - t0, h[4] - t0, h[5] - t0);
        const char* names[7] = {"", "tma", "mma", "split", "epi_start", "epi_final", "epi_end"};
        for (int r = 1; r < 7; ++r) {
            fprintf(stderr, "TCTRACE %-9s", names[r]);
            for (int i = 0; i < 24 && i < (r == 5 ? 2 : (int)h[2]); ++i)
                fprintf(stderr, " %lld", h[r * 64 + i] ? h[r * 64 + i] - t0 : -1);
            fprintf(stderr, "\n");
        }
    }
    return cudaGetLastError();
}

cudaError_t launch_ivf_scan_tc(int metric, uint32_t max_items, const TcScanArgs& a,
                               cudaStream_t st) {
    if (a.nstages < 2 || a.nstages > kTcMaxStages || a.nf4 > 32 ||
        scan_tc_smem_bytes(a.nstages) > 227 * 1024)
        return cudaErrorInvalidValue;
    if (a.kpart == 16)
        return metric == kMetricL2 ? launch_tc_one<kMetricL2, 16>(max_items, a, st)
                                   : launch_tc_one<kMetricInnerProduct, 16>(max_items, a, st);
    if (a.kpart == 32)
        return metric == kMetricL2 ? launch_tc_one<kMetricL2, 32>(max_items, a, st)
                                   : launch_tc_one<kMetricInnerProduct, 32>(max_items, a, st);
    return cudaErrorInvalidValue;
}

} // namespace b200ivf
