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
        const int l = a.coarse_ids[i];
        if (l < 0)
            continue;
        if (l >= a.nlists) {
            bad++;
            continue;
        }
        const int len = a.list_len[l];
        if (len == 0)
            continue;
        atomicAdd(&a.count[l], 1);
        ndis += (unsigned long long)len;
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

// single-block exclusive scan over the lists: grp_start (pairs) and item_off (warp items)
__global__ void plan_scan_kernel(const PlanArgs a) {
    __shared__ int s_cnt[1024];
    __shared__ unsigned s_itm[1024];
    __shared__ unsigned s_til[1024];
    __shared__ int carry_cnt;
    __shared__ unsigned carry_itm, carry_til;
    const int t = threadIdx.x;
    if (t == 0) {
        carry_cnt = 0;
        carry_itm = 0;
        carry_til = 0;
    }
    __syncthreads();
    for (int base = 0; base < a.nlists; base += 1024) {
        const int l = base + t;
        const int c = l < a.nlists ? a.count[l] : 0;
        const bool tiled = a.tile_tq > 0 && c > a.tile_min;
        const unsigned it = tiled ? 0u : (unsigned)((c + a.rq - 1) / a.rq) * (unsigned)a.nsplit;
        const unsigned tl =
                tiled ? (unsigned)((c + a.tile_tq - 1) / a.tile_tq) * (unsigned)a.nsplit : 0u;
        s_cnt[t] = c;
        s_itm[t] = it;
        s_til[t] = tl;
        __syncthreads();
        for (int o = 1; o < 1024; o <<= 1) { // Hillis-Steele inclusive scan
            int vc = 0;
            unsigned vi = 0, vt = 0;
            if (t >= o) {
                vc = s_cnt[t - o];
                vi = s_itm[t - o];
                vt = s_til[t - o];
            }
            __syncthreads();
            s_cnt[t] += vc;
            s_itm[t] += vi;
            s_til[t] += vt;
            __syncthreads();
        }
        if (l < a.nlists) {
            a.grp_start[l] = carry_cnt + s_cnt[t] - c;
            a.item_off[l] = carry_itm + s_itm[t] - it;
            if (a.item_off_tile)
                a.item_off_tile[l] = carry_til + s_til[t] - tl;
            if (a.item_list_tile) {
                const unsigned o = carry_til + s_til[t] - tl;
                for (unsigned j = 0; j < tl; ++j)
                    a.item_list_tile[o + j] = l;
            }
        }
        __syncthreads();
        if (t == 1023) {
            carry_cnt += s_cnt[t];
            carry_itm += s_itm[t];
            carry_til += s_til[t];
        }
        __syncthreads();
    }
    if (t == 0) {
        a.grp_start[a.nlists] = carry_cnt;
        a.item_off[a.nlists] = carry_itm;
        if (a.item_off_tile)
            a.item_off_tile[a.nlists] = carry_til;
    }
}

__global__ void plan_scatter_kernel(const PlanArgs a) {
    const int64_t total = (int64_t)a.n * a.nprobe;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int l = a.coarse_ids[i];
        int slot = -1;
        if (l >= 0 && l < a.nlists && a.list_len[l] > 0) {
            slot = a.grp_start[l] + atomicAdd(&a.fill[l], 1);
            a.pair_query[slot] = (int)(i / a.nprobe);
        }
        a.slot_of[i] = slot;
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

cudaError_t launch_plan(const PlanArgs& a, cudaStream_t st) {
    cudaError_t e;
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
