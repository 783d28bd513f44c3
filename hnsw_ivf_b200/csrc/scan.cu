// scan.cu -- K2, the inverted-list scan (and, pointed at the centroid table, the SIMT coarse
// quantizer). Replaces the reference's per-(query, probe) CPU loop
//   IVFFlatScanner::scan_codes          faiss/IndexIVFFlat.cpp:155-179
//   fvec_L2sqr / fvec_inner_product     faiss/utils/distances_simd.cpp:220-229,195-202
//   heap_replace_top on accept          faiss/utils/Heap.h:111-150
//
// Design (B200-first, not the reference's shape):
//  * Work is LIST-major: the probe table is inverted on device (plan.cu) so that all queries that
//    probe a list are applied to it together. One warp owns up to RQ of those queries and a row
//    range of the list; sibling warps of a CTA take the next sub-groups of the same rows, so a
//    list block is fetched from HBM once and re-served by L1/L2 to the other sub-groups.
//  * The slab is dimension-interleaved in 32-row blocks (common.cuh): lane = row, so every load is
//    one coalesced 512-byte LDG.128 per warp, distances need no cross-lane reduction at all, and
//    the query values are warp-uniform shared-memory broadcasts (1 wavefront per LDS.128).
//    Per warp step that is 4*RV + RQ L1/shared cycles against RV*RQ FMA issue cycles: with
//    RV=4, RQ=8 the FMA pipe, not the load path, is the on-chip limit; with few queries per list
//    the kernel degenerates into a pure HBM stream (RV*unroll 512-byte loads in flight per warp).
//  * Direct (x-y)^2 FMA chain for L2 (not the norm expansion), fp32 throughout.
//  * Accepts are rare after warm-up: a strict compare against a per-(warp, query) threshold
//    (the reference's C::cmp(simi[0], dis)), ballot, and a warp-collective append into a small
//    shared-memory buffer that is bitonic-sorted only when it overflows. Each (pair, split) leaves
//    its k best keys in `part` for K3.
#include "common.cuh"
#include "kernels.h"

namespace b200ivf {

namespace {

using PushState = KeyBufState;

// Warp-collective append of the lanes whose `pass` is set (m = their ballot, non-zero).
__device__ __noinline__ PushState push_candidates(uint64_t* buf, int cnt, float thr, int k, int cap,
                                                  bool pass, uint64_t key, unsigned m, int lane) {
    const int n = __popc(m);
    if (cnt + n > cap) {
        PushState s = warp_compact_keys(buf, cnt, thr, k, cap, lane);
        cnt = s.cnt;
        thr = s.thr;
    }
    if (pass)
        buf[cnt + __popc(m & ((1u << lane) - 1u))] = key;
    __syncwarp();
    return {cnt + n, thr};
}

template <int METRIC>
__device__ __forceinline__ float accumulate4(float acc, const float4& x, const float4& y) {
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

// FILT: the SearchParametersIVF form (faiss/IndexIVF.h:73-88) -- rows are kept only where the
// selector's row mask has their bit (IVFFlatScanner::scan_codes' use_sel arm,
// faiss/IndexIVFFlat.cpp:166) and, for max_codes, only the first pair_limit[slot] rows of the list
// count for that pair (scan_one_list's list_size_max, faiss/IndexIVF.cpp:545-548).
template <int METRIC, int RQ, int RV, bool FILT>
__global__ void __launch_bounds__(256) ivf_scan_kernel(const ScanArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int wpc = blockDim.x >> 5;
    const uint32_t item = blockIdx.x * (uint32_t)wpc + (uint32_t)warp;
    if (item >= a.item_off[a.nlists])
        return; // whole warp leaves together; the kernel has no block-wide barrier

    // which list: the last l with item_off[l] <= item
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
    const int nsub = (cnt_l + RQ - 1) / RQ;
    const uint32_t local = item - a.item_off[l];
    const int split = (int)(local / (uint32_t)nsub);
    const int sub = (int)(local % (uint32_t)nsub);
    const int q0 = g0 + sub * RQ;
    const int nqw = min(RQ, cnt_l - sub * RQ);

    const int len = a.list_len[l];
    const int nb = (len + kBlockRows - 1) / kBlockRows;
    const int b0 = (int)(((long long)nb * split) / a.nsplit);
    const int b1 = (int)(((long long)nb * (split + 1)) / a.nsplit);
    const int nf4 = a.nf4;

    const size_t wbytes = (size_t)RQ * nf4 * sizeof(float4) + (size_t)RQ * a.cap * sizeof(uint64_t);
    float4* qs = reinterpret_cast<float4*>(smem_raw + (size_t)warp * wbytes);
    uint64_t* bufs = reinterpret_cast<uint64_t*>(qs + (size_t)RQ * nf4);

    // stage this warp's queries (zero padded) in its private shared-memory strip
    {
        int myq = -1;
        if (lane < nqw)
            myq = a.pair_query ? a.pair_query[q0 + lane] : (q0 + lane);
        float* qsf = reinterpret_cast<float*>(qs);
        const int rowf = nf4 * 4;
#pragma unroll
        for (int j = 0; j < RQ; ++j) {
            const int q = __shfl_sync(0xffffffffu, myq, j);
            for (int dd = lane; dd < rowf; dd += 32) {
                float v = 0.f;
                if (q >= 0 && dd < a.d)
                    v = __ldg(a.queries + (size_t)q * a.d + dd);
                qsf[j * rowf + dd] = v;
            }
        }
        __syncwarp();
    }

    float thr[RQ];
    int cnt[RQ];
    int lim[RQ];
#pragma unroll
    for (int j = 0; j < RQ; ++j) {
        thr[j] = kNeutralScore;
        cnt[j] = 0;
        lim[j] = (FILT && a.pair_limit && j < nqw) ? a.pair_limit[q0 + j] : len;
    }

    const uint32_t blk0 = a.list_blk0[l];
    const float4* base = a.slab + (size_t)blk0 * nf4 * kBlockRows + lane;

    for (int b = b0; b < b1; b += RV) {
        float acc[RV][RQ];
        const float4* vp[RV];
#pragma unroll
        for (int v = 0; v < RV; ++v) {
            const int bb = min(b + v, b1 - 1);
            vp[v] = base + (size_t)bb * nf4 * kBlockRows;
#pragma unroll
            for (int j = 0; j < RQ; ++j)
                acc[v][j] = 0.f;
        }
        // loads in flight per lane = RV x unroll: 16 either way (a lone block per step, the
        // latency-bound small-batch shape, gets the deeper unroll)
#pragma unroll(RV == 1 ? 16 : 4)
        for (int c = 0; c < nf4; ++c) {
            float4 y[RV];
#pragma unroll
            for (int v = 0; v < RV; ++v)
                y[v] = __ldg(vp[v] + c * kBlockRows);
#pragma unroll
            for (int j = 0; j < RQ; ++j) {
                const float4 x = qs[j * nf4 + c];
#pragma unroll
                for (int v = 0; v < RV; ++v)
                    acc[v][j] = accumulate4<METRIC>(acc[v][j], x, y[v]);
            }
        }
#pragma unroll
        for (int v = 0; v < RV; ++v) {
            const int row = (b + v) * kBlockRows + lane;
            bool rowok = (b + v < b1) && (row < len);
            if (FILT && a.row_mask && b + v < b1)
                rowok = rowok && ((a.row_mask[blk0 + (uint32_t)(b + v)] >> lane) & 1u);
            const uint32_t grow = blk0 * kBlockRows + (uint32_t)row;
#pragma unroll
            for (int j = 0; j < RQ; ++j) {
                // smaller is better; "+ 0.f" folds -0 into +0 so equal scores have equal keys
                const float score = (METRIC == kMetricL2 ? acc[v][j] : -acc[v][j]) + 0.f;
                const bool pass =
                        rowok && (j < nqw) && (!FILT || row < lim[j]) && (score < thr[j]);
                const unsigned m = __ballot_sync(0xffffffffu, pass);
                if (m) {
                    PushState s = push_candidates(bufs + (size_t)j * a.cap, cnt[j], thr[j], a.k,
                                                  a.cap, pass, make_key(score, grow), m, lane);
                    cnt[j] = s.cnt;
                    thr[j] = s.thr;
                }
            }
        }
    }

    // leave the k best of every query in `part`
#pragma unroll
    for (int j = 0; j < RQ; ++j) {
        if (j < nqw) {
            uint64_t* buf = bufs + (size_t)j * a.cap;
            PushState s = warp_compact_keys(buf, cnt[j], thr[j], a.k, a.cap, lane);
            const int keep = min(s.cnt, a.k);
            uint64_t* out =
                    a.part + ((size_t)(q0 + j) * a.nsplit + split) * a.part_mul * a.k;
            for (int i = lane; i < a.k; i += 32)
                out[i] = i < keep ? buf[i] : kSentinelKey;
            if (a.part_mul == 2)
                for (int i = lane; i < a.k; i += 32)
                    out[a.k + i] = kSentinelKey;
        }
    }
}

template <int METRIC, int RQ, int RV, bool FILT = false>
cudaError_t launch_one(int wpc, uint32_t max_items, const ScanArgs& a, cudaStream_t st) {
    const size_t smem = scan_smem_bytes(RQ, wpc, a.nf4, a.cap);
    auto kern = ivf_scan_kernel<METRIC, RQ, RV, FILT>;
    if (smem > 48 * 1024) {
        cudaError_t e =
                cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess)
            return e;
    }
    const uint32_t grid = (max_items + wpc - 1) / wpc;
    if (grid == 0)
        return cudaSuccess;
    kern<<<grid, wpc * 32, smem, st>>>(a);
    count_launches(1);
    return cudaGetLastError();
}

} // namespace

size_t scan_smem_bytes(int rq, int warps_per_cta, int nf4, int cap) {
    return (size_t)warps_per_cta *
            ((size_t)rq * nf4 * sizeof(float4) + (size_t)rq * cap * sizeof(uint64_t));
}

cudaError_t launch_ivf_scan(int metric, int rq, int wpc, uint32_t max_items, const ScanArgs& a,
                            cudaStream_t st, int rv) {
    const bool filt = a.row_mask != nullptr || a.pair_limit != nullptr;
#define B200IVF_DISPATCH(M)                                    \
    if (filt) {                                                \
        switch (rq) {                                          \
            case 1:                                            \
                return launch_one<M, 1, 4, true>(wpc, max_items, a, st); \
            case 4:                                            \
                return launch_one<M, 4, 4, true>(wpc, max_items, a, st); \
            case 8:                                            \
                return launch_one<M, 8, 4, true>(wpc, max_items, a, st); \
            default:                                           \
                return cudaErrorInvalidValue;                  \
        }                                                      \
    }                                                          \
    switch (rq) {                                              \
        case 1:                                                \
            if (rv == 1)                                       \
                return launch_one<M, 1, 1>(wpc, max_items, a, st); \
            return launch_one<M, 1, 4>(wpc, max_items, a, st); \
        case 4:                                                \
            return launch_one<M, 4, 4>(wpc, max_items, a, st); \
        case 8:                                                \
            return launch_one<M, 8, 4>(wpc, max_items, a, st); \
        default:                                               \
            return cudaErrorInvalidValue;                      \
    }
    if (metric == kMetricL2) {
        B200IVF_DISPATCH(kMetricL2)
    } else {
        B200IVF_DISPATCH(kMetricInnerProduct)
    }
#undef B200IVF_DISPATCH
}

} // namespace b200ivf
