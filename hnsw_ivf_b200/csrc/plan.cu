// plan.cu -- inverts the probe table [n][nprobe] into per-list query groups so that the scan can be
// list-major. The reference walks probes per query (faiss/IndexIVF.cpp:612-622); the set of
// (query, list) pairs scanned is identical, only the traversal order changes (the result of a
// query does not depend on it: tests/test_lowlevel_ivf.cpp:398-549 pins exactly that property).
//
// Skip rules mirror scan_one_list (faiss/IndexIVF.cpp:512-529): key < 0 -> skipped, empty list ->
// skipped, key >= nlist -> error (counted in stats[3]; the host turns it into the reference's
// "Invalid key" failure). stats[1] counts lists visited (nlistv), stats[2] the distances computed
// (ndis, IndexIVF.cpp:624) -- also the algorithmic-bytes counter for the roofline.
#include "common.cuh"
#include "kernels.h"

#include <algorithm>

namespace b200ivf {

namespace {

__global__ void plan_count_kernel(const PlanArgs a) {
    const int64_t total = (int64_t)a.n * a.nprobe;
    unsigned long long ndis = 0, nvis = 0, bad = 0;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        int l = a.coarse_ids[i];
        if (l < 0)
            continue;
        if (l >= (a.nlist_total ? a.nlist_total : a.nlists)) {
            bad++;
            continue;
        }
        l -= a.list_lo;
        if (l < 0 || l >= a.nlists) // another shard's list
            continue;
        const int len = a.list_len[l];
        if (len == 0)
            continue;
        atomicAdd(&a.count[l], 1);
        ndis += (unsigned long long)(a.limit_qp ? a.limit_qp[i] : len);
        nvis++;
    }
    // warp-reduce the counters, one atomic per warp
    for (int o = 16; o > 0; o >>= 1) {
        ndis += __shfl_down_sync(0xffffffffu, ndis, o);
        nvis += __shfl_down_sync(0xffffffffu, nvis, o);
        bad += __shfl_down_sync(0xffffffffu, bad, o);
    }
    if ((threadIdx.x & 31) == 0 && a.stats) {
        if (nvis)
            atomicAdd(&a.stats[1], nvis);
        if (ndis)
            atomicAdd(&a.stats[2], ndis);
        if (bad)
            atomicAdd(&a.stats[3], bad);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0 && a.stats)
        atomicAdd(&a.stats[0], (unsigned long long)a.n);
}

// single-block exclusive scan over the lists: grp_start (pairs) and item_off (warp items).
// Every thread owns a contiguous run of lists: serial sums, one shuffle scan over the 1024 thread
// totals, then a second pass that writes the prefixes (two barriers in all; this kernel sits on
// the latency path of every small-batch search).
__device__ __forceinline__ void plan_list_items(const PlanArgs& a, int c, unsigned& it,
                                                unsigned& tl) {
    const bool tiled = a.tile_tq > 0 && c > a.tile_min;
    it = tiled ? 0u : (unsigned)((c + a.rq - 1) / a.rq) * (unsigned)a.nsplit;
    tl = tiled ? (unsigned)((c + a.tile_tq - 1) / a.tile_tq) * (unsigned)a.nsplit : 0u;
}

// The per-thread runs are read 16 lists at a time into registers (independent loads, all in flight
// together): the serial form paid one L2 round trip per list and took 35 us at nlist = 16384.
constexpr int kPlanChunk = 16;

__global__ void __launch_bounds__(1024) plan_scan_kernel(const PlanArgs a) {
    __shared__ int w_cnt[32];
    __shared__ unsigned w_itm[32], w_til[32];
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    const int per = (a.nlists + 1023) / 1024;
    const int l0 = min(t * per, a.nlists), l1 = min(l0 + per, a.nlists);
    int sc = 0;
    unsigned si = 0, sl = 0;
    for (int lb = l0; lb < l1; lb += kPlanChunk) {
        int c[kPlanChunk];
#pragma unroll
        for (int u = 0; u < kPlanChunk; ++u)
            c[u] = lb + u < l1 ? a.count[lb + u] : 0;
#pragma unroll
        for (int u = 0; u < kPlanChunk; ++u) {
            unsigned it, tl;
            plan_list_items(a, c[u], it, tl);
            sc += c[u];
            si += it;
            sl += tl;
        }
    }
    int ic = sc; // inclusive scans over the thread totals
    unsigned ii = si, il = sl;
    for (int o = 1; o < 32; o <<= 1) {
        const int vc = __shfl_up_sync(0xffffffffu, ic, o);
        const unsigned vi = __shfl_up_sync(0xffffffffu, ii, o);
        const unsigned vl = __shfl_up_sync(0xffffffffu, il, o);
        if (lane >= o) {
            ic += vc;
            ii += vi;
            il += vl;
        }
    }
    if (lane == 31) {
        w_cnt[w] = ic;
        w_itm[w] = ii;
        w_til[w] = il;
    }
    __syncthreads();
    if (w == 0) {
        const int tc = w_cnt[lane];
        const unsigned ti = w_itm[lane], tl = w_til[lane];
        int xc = tc;
        unsigned xi = ti, xl = tl;
        for (int o = 1; o < 32; o <<= 1) {
            const int vc = __shfl_up_sync(0xffffffffu, xc, o);
            const unsigned vi = __shfl_up_sync(0xffffffffu, xi, o);
            const unsigned vl = __shfl_up_sync(0xffffffffu, xl, o);
            if (lane >= o) {
                xc += vc;
                xi += vi;
                xl += vl;
            }
        }
        w_cnt[lane] = xc - tc; // exclusive prefix of the warp
        w_itm[lane] = xi - ti;
        w_til[lane] = xl - tl;
    }
    __syncthreads();
    int pc = w_cnt[w] + ic - sc;
    unsigned pi = w_itm[w] + ii - si, pl = w_til[w] + il - sl;
    for (int lb = l0; lb < l1; lb += kPlanChunk) {
        int c[kPlanChunk];
#pragma unroll
        for (int u = 0; u < kPlanChunk; ++u)
            c[u] = lb + u < l1 ? a.count[lb + u] : 0;
#pragma unroll
        for (int u = 0; u < kPlanChunk; ++u) {
            const int l = lb + u;
            if (l < l1) {
                unsigned it, tl;
                plan_list_items(a, c[u], it, tl);
                a.grp_start[l] = pc;
                a.item_off[l] = pi;
                if (a.item_off_tile)
                    a.item_off_tile[l] = pl;
                if (a.item_list_tile)
                    for (unsigned j = 0; j < tl; ++j)
                        a.item_list_tile[pl + j] = l;
                pc += c[u];
                pi += it;
                pl += tl;
            }
        }
    }
    if (t == 1023) { // the last thread's running prefix ends at the totals
        a.grp_start[a.nlists] = pc;
        a.item_off[a.nlists] = pi;
        if (a.item_off_tile)
            a.item_off_tile[a.nlists] = pl;
    }
}

__global__ void plan_scatter_kernel(const PlanArgs a) {
    const int64_t total = (int64_t)a.n * a.nprobe;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int l = a.coarse_ids[i] < 0 ? -1 : a.coarse_ids[i] - a.list_lo;
        int slot = -1;
        if (l >= 0 && l < a.nlists && a.list_len[l] > 0) {
            slot = a.grp_start[l] + atomicAdd(&a.fill[l], 1);
            a.pair_query[slot] = (int)(i / a.nprobe);
            if (a.limit_slot)
                a.limit_slot[slot] = a.limit_qp[i];
        }
        a.slot_of[i] = slot;
    }
}

// max_codes (faiss/IndexIVF.cpp:597-631, parallel_mode 0): a query walks its probes in order, a
// list counts for min(list_size, max_codes - nscan) rows (scan_one_list's list_size_max, :545-548)
// and the walk stops once nscan >= max_codes. One thread per query; probes after the stop are
// turned into -1 in the probe table (keys < 0 and empty lists count for 0 rows, :512-529).
__global__ void probe_limits_kernel(int32_t* coarse_ids, int n, int nprobe, int nlists,
                                    const int32_t* __restrict__ list_len, long long max_codes,
                                    int32_t* limit_qp, unsigned long long* query_ndis) {
    for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < n; q += gridDim.x * blockDim.x) {
        long long nscan = 0;
        bool stop = false;
        for (int p = 0; p < nprobe; ++p) {
            const int64_t i = (int64_t)q * nprobe + p;
            const int l = coarse_ids[i];
            int lim = 0;
            if (stop) {
                coarse_ids[i] = -1;
            } else if (l >= 0 && l < nlists) {
                const long long len = list_len[l];
                lim = (int)(len < max_codes - nscan ? len : max_codes - nscan);
                nscan += lim;
                stop = len > 0 && nscan >= max_codes;
            }
            if (limit_qp)
                limit_qp[i] = lim;
        }
        if (query_ndis)
            query_ndis[q] = (unsigned long long)nscan;
    }
}

// pair-major plan: see PlanArgs::vlen
__global__ void plan_pairs_kernel(const PlanArgs a) {
    const int total = a.n * a.nprobe;
    unsigned long long ndis = 0, nvis = 0, bad = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        int l = a.coarse_ids[i];
        int len = 0;
        uint32_t blk0 = 0;
        if (l >= (a.nlist_total ? a.nlist_total : a.nlists)) {
            bad++;
        } else if (l >= 0) {
            l -= a.list_lo;
            if (l >= 0 && l < a.nlists) { // (else: another shard's list)
                len = a.list_len[l];
                blk0 = a.list_blk0[l];
            }
        }
        a.vlen[i] = len;
        a.vblk0[i] = blk0;
        a.grp_start[i] = i;
        a.item_off[i] = (uint32_t)i * (uint32_t)a.nsplit;
        a.pair_query[i] = i / a.nprobe;
        a.slot_of[i] = len > 0 ? i : -1;
        if (len > 0) {
            ndis += (unsigned long long)len;
            nvis++;
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        ndis += __shfl_down_sync(0xffffffffu, ndis, o);
        nvis += __shfl_down_sync(0xffffffffu, nvis, o);
        bad += __shfl_down_sync(0xffffffffu, bad, o);
    }
    if ((threadIdx.x & 31) == 0 && a.stats) {
        if (nvis)
            atomicAdd(&a.stats[1], nvis);
        if (ndis)
            atomicAdd(&a.stats[2], ndis);
        if (bad)
            atomicAdd(&a.stats[3], bad);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        a.grp_start[total] = total;
        a.item_off[total] = (uint32_t)total * (uint32_t)a.nsplit;
        if (a.stats)
            atomicAdd(&a.stats[0], (unsigned long long)a.n);
    }
}

__global__ void plan_single_kernel(int32_t* grp_start, uint32_t* item_off, int n, uint32_t items) {
    grp_start[0] = 0;
    grp_start[1] = n;
    item_off[0] = 0;
    item_off[1] = items;
}

} // namespace

cudaError_t launch_plan_single(int32_t* grp_start, uint32_t* item_off, int n, uint32_t items,
                               cudaStream_t st) {
    plan_single_kernel<<<1, 1, 0, st>>>(grp_start, item_off, n, items);
    count_launches(1);
    return cudaGetLastError();
}

cudaError_t launch_probe_limits(int32_t* coarse_ids, int n, int nprobe, int nlists,
                                const int32_t* list_len, long long max_codes, int32_t* limit_qp,
                                unsigned long long* query_ndis, cudaStream_t st) {
    if (n == 0)
        return cudaSuccess;
    const int threads = 128;
    probe_limits_kernel<<<(n + threads - 1) / threads, threads, 0, st>>>(
            coarse_ids, n, nprobe, nlists, list_len, max_codes, limit_qp, query_ndis);
    count_launches(1);
    return cudaGetLastError();
}

cudaError_t launch_plan(const PlanArgs& a, cudaStream_t st) {
    cudaError_t e;
    if (a.vlen) { // pair-major: one launch
        const int total = a.n * a.nprobe;
        if (total == 0 || a.rq != 1 || a.tile_tq != 0 || a.limit_qp || !a.vblk0 || !a.list_blk0)
            return total == 0 ? cudaSuccess : cudaErrorInvalidValue;
        plan_pairs_kernel<<<(total + 255) / 256, 256, 0, st>>>(a);
        count_launches(1);
        return cudaGetLastError();
    }
    if ((e = cudaMemsetAsync(a.count, 0, sizeof(int32_t) * a.nlists, st)) != cudaSuccess)
        return e;
    if ((e = cudaMemsetAsync(a.fill, 0, sizeof(int32_t) * a.nlists, st)) != cudaSuccess)
        return e;
    const int64_t total = (int64_t)a.n * a.nprobe;
    if (total == 0)
        return cudaSuccess;
    const int threads = 256;
    const int grid = (int)std::min<int64_t>((total + threads - 1) / threads, 148 * 8);
    plan_count_kernel<<<grid, threads, 0, st>>>(a);
    count_launches(1);
    plan_scan_kernel<<<1, 1024, 0, st>>>(a);
    count_launches(1);
    plan_scatter_kernel<<<grid, threads, 0, st>>>(a);
    count_launches(1);
    return cudaGetLastError();
}

} // namespace b200ivf
