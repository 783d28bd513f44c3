// range.cu -- range search over the probed inverted lists: every stored vector whose distance to
// the query is inside the radius. Replaces, for the device path,
//   IndexIVF::range_search_preassigned      faiss/IndexIVF.cpp:1243-1399 (parallel_mode 0)
//   IVFFlatScanner::scan_codes_range        faiss/IndexIVFFlat.cpp:181-201
//   RangeSearchPartialResult::finalize      faiss/impl/AuxIndexStructures.cpp:134-190
// The reference appends results per query in probe order, then list order, into growable buffers
// and builds `lims` at the end. Here the result size is not known before the scan, so the scan
// runs twice with identical arithmetic: pass 1 counts per (query, probe) pair, an exclusive scan
// in (query, probe) order turns the counts into output offsets (= the reference's order), pass 2
// writes (distance, id) at those offsets. One warp per pair, lane = row of a 32-row block, so the
// position of a hit inside its pair is a ballot prefix: the output order is deterministic.
#include "common.cuh"
#include "kernels.h"

namespace b200ivf {

namespace {

template <int METRIC>
__device__ __forceinline__ float range_acc4(float acc, const float4& x, const float4& y) {
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

// FILL = false: pair_cnt[slot] = number of hits. FILL = true: hits written at
// lims[q] + pair_off[slot] + (position inside the pair).
template <int METRIC, bool FILL>
__global__ void __launch_bounds__(128) range_scan_kernel(const RangeArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int wpc = blockDim.x >> 5;
    const int64_t slot = (int64_t)blockIdx.x * wpc + warp;
    if (slot >= a.grp_start[a.nlists])
        return; // whole warp; no block-wide barrier below
    int lo = 0, hi = a.nlists; // the last list whose group starts at or before this slot
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (a.grp_start[mid] <= slot)
            lo = mid;
        else
            hi = mid;
    }
    const int l = lo;
    const int q = a.pair_query[slot];
    const int nf4 = a.nf4;
    float4* qs = reinterpret_cast<float4*>(smem_raw) + (size_t)warp * nf4;
    {
        float* qsf = reinterpret_cast<float*>(qs);
        for (int dd = lane; dd < nf4 * 4; dd += 32)
            qsf[dd] = dd < a.d ? __ldg(a.queries + (size_t)q * a.d + dd) : 0.f;
        __syncwarp();
    }
    const int len = a.list_len[l];
    const int nb = (len + kBlockRows - 1) / kBlockRows;
    const uint32_t blk0 = a.list_blk0[l];
    const float4* base = a.slab + (size_t)blk0 * nf4 * kBlockRows + lane;
    unsigned long long out0 = 0;
    if (FILL)
        out0 = a.lims[q] + a.pair_off[slot];
    int found = 0;
    for (int b = 0; b < nb; b += 2) {
        // two blocks per step: 2 x unroll 8 independent 512-byte loads in flight per warp
        const int bb1 = min(b + 1, nb - 1);
        const float4* p0 = base + (size_t)b * nf4 * kBlockRows;
        const float4* p1 = base + (size_t)bb1 * nf4 * kBlockRows;
        float acc0 = 0.f, acc1 = 0.f;
#pragma unroll 8
        for (int c = 0; c < nf4; ++c) {
            const float4 y0 = __ldg(p0 + c * kBlockRows);
            const float4 y1 = __ldg(p1 + c * kBlockRows);
            const float4 x = qs[c];
            acc0 = range_acc4<METRIC>(acc0, x, y0);
            acc1 = range_acc4<METRIC>(acc1, x, y1);
        }
#pragma unroll
        for (int v = 0; v < 2; ++v) {
            if (b + v >= nb)
                break;
            const float dis = v == 0 ? acc0 : acc1;
            const int row = (b + v) * kBlockRows + lane;
            bool in = row < len;
            if (a.row_mask)
                in = in && ((a.row_mask[blk0 + (uint32_t)(b + v)] >> lane) & 1u);
            // C::cmp(radius, dis): L2 keeps dis < radius, inner product keeps dis > radius
            in = in && (METRIC == kMetricL2 ? (a.radius > dis) : (a.radius < dis));
            const unsigned m = __ballot_sync(0xffffffffu, in);
            if (FILL && in) {
                const unsigned long long o = out0 + found + __popc(m & ((1u << lane) - 1u));
                a.out_dis[o] = dis;
                const size_t grow = (size_t)blk0 * kBlockRows + row;
                a.out_ids[o] = a.ids ? a.ids[grow] : (int64_t)grow;
            }
            found += __popc(m);
        }
    }
    if (!FILL && lane == 0)
        a.pair_cnt[slot] = found;
}

// One warp per query: offsets of its pairs inside the query's run (probe order), and the run length.
__global__ void range_offsets_kernel(const int32_t* __restrict__ slot_of,
                                     const int32_t* __restrict__ pair_cnt, int n, int nprobe,
                                     uint32_t* __restrict__ pair_off,
                                     unsigned long long* __restrict__ qcnt) {
    const int lane = threadIdx.x & 31;
    const int q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (q >= n)
        return;
    uint32_t carry = 0;
    for (int p0 = 0; p0 < nprobe; p0 += 32) {
        const int p = p0 + lane;
        int slot = -1;
        if (p < nprobe)
            slot = slot_of[(int64_t)q * nprobe + p];
        const uint32_t c = slot >= 0 ? (uint32_t)pair_cnt[slot] : 0u;
        uint32_t inc = c;
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o)
                inc += v;
        }
        if (slot >= 0)
            pair_off[slot] = carry + inc - c;
        carry += __shfl_sync(0xffffffffu, inc, 31);
    }
    if (lane == 0)
        qcnt[q] = carry;
}

// lims[0..n] = exclusive prefix of qcnt over the queries; single block of 1024 threads
__global__ void __launch_bounds__(1024) range_lims_kernel(const unsigned long long* __restrict__ qcnt,
                                                          int n, unsigned long long* __restrict__ lims) {
    __shared__ unsigned long long tot[1024];
    const int t = threadIdx.x;
    const int per = (n + 1023) / 1024;
    const int i0 = min(t * per, n), i1 = min(i0 + per, n);
    unsigned long long s = 0;
    for (int i = i0; i < i1; ++i)
        s += qcnt[i];
    tot[t] = s;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) { // inclusive Hillis-Steele scan of the thread totals
        const unsigned long long v = t >= o ? tot[t - o] : 0ull;
        __syncthreads();
        tot[t] += v;
        __syncthreads();
    }
    unsigned long long run = tot[t] - s;
    for (int i = i0; i < i1; ++i) {
        lims[i] = run;
        run += qcnt[i];
    }
    if (t == 1023)
        lims[n] = tot[1023];
}

// labels of a store_pairs search: slab row -> lo_build(list_no, offset) = list_no << 32 | offset
// (faiss/invlists/DirectMap.h:22-24, faiss/IndexIVFFlat.cpp:172)
__global__ void rows_to_pairs_kernel(int64_t* __restrict__ labels, int64_t n,
                                     const uint32_t* __restrict__ list_blk0, int nlists) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t row = labels[i];
        if (row < 0)
            continue;
        const uint32_t blk = (uint32_t)(row / kBlockRows);
        int lo = 0, hi = nlists; // the last list that starts at or before the block: empty lists
        while (hi - lo > 1) {    // share their start with the next non-empty one
            const int mid = (lo + hi) >> 1;
            if (list_blk0[mid] <= blk)
                lo = mid;
            else
                hi = mid;
        }
        const int64_t off = row - (int64_t)list_blk0[lo] * kBlockRows;
        labels[i] = ((int64_t)lo << 32) | off;
    }
}

} // namespace

cudaError_t launch_range_scan(int metric, bool fill, int64_t max_pairs, const RangeArgs& a,
                              cudaStream_t st) {
    if (max_pairs == 0)
        return cudaSuccess;
    const int wpc = 4;
    const size_t smem = (size_t)wpc * a.nf4 * sizeof(float4);
    if (smem > 48 * 1024)
        return cudaErrorInvalidValue; // d <= 3072
    const unsigned grid = (unsigned)((max_pairs + wpc - 1) / wpc);
    if (metric == kMetricL2) {
        if (fill)
            range_scan_kernel<kMetricL2, true><<<grid, wpc * 32, smem, st>>>(a);
        else
            range_scan_kernel<kMetricL2, false><<<grid, wpc * 32, smem, st>>>(a);
    } else {
        if (fill)
            range_scan_kernel<kMetricInnerProduct, true><<<grid, wpc * 32, smem, st>>>(a);
        else
            range_scan_kernel<kMetricInnerProduct, false><<<grid, wpc * 32, smem, st>>>(a);
    }
    count_launches(1);
    return cudaGetLastError();
}

cudaError_t launch_range_offsets(const int32_t* slot_of, const int32_t* pair_cnt, int n, int nprobe,
                                 uint32_t* pair_off, unsigned long long* qcnt,
                                 unsigned long long* lims, cudaStream_t st) {
    if (n == 0)
        return cudaSuccess;
    const int threads = 256;
    range_offsets_kernel<<<(int)(((int64_t)n * 32 + threads - 1) / threads), threads, 0, st>>>(
            slot_of, pair_cnt, n, nprobe, pair_off, qcnt);
    range_lims_kernel<<<1, 1024, 0, st>>>(qcnt, n, lims);
    count_launches(2);
    return cudaGetLastError();
}

cudaError_t launch_rows_to_pairs(int64_t* labels, int64_t n, const uint32_t* list_blk0, int nlists,
                                 cudaStream_t st) {
    if (n == 0)
        return cudaSuccess;
    const int threads = 256;
    const int grid = (int)((n + threads - 1) / threads < 148 * 8 ? (n + threads - 1) / threads
                                                                  : 148 * 8);
    rows_to_pairs_kernel<<<grid, threads, 0, st>>>(labels, n, list_blk0, nlists);
    count_launches(1);
    return cudaGetLastError();
}

} // namespace b200ivf
