// select.cu -- K3: fused per-query k-select over the partial candidate lists left by the scan, and
// the device-side merge of per-shard result tables.
//
// Replaces, for one query: the CPU result heap of search_preassigned (heap_heapify :607,
// heap_replace_top via scan_codes, heap_reorder :625 of faiss/IndexIVF.cpp) and, for the coarse
// quantizer, HeapBlockResultHandler::add_results/end_multiple (faiss/impl/ResultHandler.h:263-286).
// Output contract kept from heap_reorder (faiss/utils/Heap.h:420-450):
//   * best first; exact-distance ties ordered by cmp2: L2 -> ascending id, IP -> descending id
//     (the CMin heap pops (min dis, min id) first and writes it to the tail);
//   * fewer than k candidates: labels -1, distances FLT_MAX (L2) / -FLT_MAX (IP).
#include "common.cuh"
#include "kernels.h"

#include <algorithm>

namespace b200ivf {

namespace {


// E(q) of SelectArgs::err_coef, rounded up
__device__ __forceinline__ float rerank_err(const SelectArgs& a, float xn2) {
    const float ym2 = a.ymax2 ? *a.ymax2 : 0.f;
    const float e = a.metric == kMetricL2 ? a.err_coef * (xn2 + ym2)
                                          : a.err_coef * sqrtf(xn2 * ym2);
    return e * 1.000001f;
}
// the checked re-rank's verdict (SelectArgs::err_coef): true = the exact top k is proven
__device__ __forceinline__ bool rerank_proven(const SelectArgs& a, int q, float xn2, float bound,
                                              uint64_t kth_key) {
    if (kth_key == kSentinelKey)
        bound = fminf(bound, kNeutralScore); // fewer than k results: nothing may remain anywhere
    float b = bound;
    if (a.qthr)
        b = fminf(b, from_ordered_bits(a.qthr[q]));
    if (b >= kNeutralScore)
        return true; // nothing was left out
    if (kth_key == kSentinelKey)
        return false;
    // (rounded against the claim: the subtraction may only make the test harder to pass)
    return __fsub_rd(b, rerank_err(a, xn2)) > key_score(kth_key);
}
__device__ __forceinline__ void flag_query(const SelectArgs& a, int q) {
    if (a.flag_list && a.flag_count)
        a.flag_list[atomicAdd(a.flag_count, 1)] = q;
}

// One CTA per query. Shared memory: keys[M] then dis[k] then ids[k].
// Fast path (prefilter != 0): every partial list is sorted, so the kk-th smallest of the list HEADS
// is an upper bound T of the kk-th smallest key overall; only the keys <= T (a few dozen) are
// gathered and sorted. Falls back to the chunked full sort if more than kSelCap keys qualify.
constexpr int kSelCap = 512;
__global__ void select_kernel(const SelectArgs a, int M, int prefilter) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* keys = reinterpret_cast<uint64_t*>(smem_raw);
    int64_t* sid = reinterpret_cast<int64_t*>(keys + M);
    float* sdis = reinterpret_cast<float*>(sid + a.k);

    const int q = blockIdx.x;
    const int t = threadIdx.x;
    const int per_seg = a.nsplit * a.kpart; // keys per pair slot
    const int64_t total = (int64_t)a.nseg * per_seg;

    // keys kept between rounds: k, plus a margin when the scores are approximate (tensor-core scan)
    const int kk = a.slab ? a.k + a.rerank_extra : a.k;
    __shared__ int s_cnt;
    bool fast_done = false;
    uint64_t cut_key = kSentinelKey; // checked re-rank: every key left out so far was >= this
    if (prefilter) {
        // Pivot rule (the one select_warp_kernel uses): every partial list is sorted, so at least
        // r * m keys are <= the r-th smallest of the lists' m-th entries. With m = ceil(kk / L)
        // and r = ceil(kk / m) that is >= kk keys, whatever kk is against the number of lists
        // (kk > L happens for k = 100 at nprobe 32/64 with one row split: the heads alone bound
        // only L keys). No usable pivot (r > L after m was cut to kpart, or fewer than r lists
        // reach m entries): every real key passes and the chunked sort below takes over if they
        // do not fit.
        const int L = a.nseg * a.nsplit;
        const int Lp = pow2_ceil(L < 32 ? 32 : L);
        int m = (kk + L - 1) / L;
        if (m > a.kpart)
            m = a.kpart;
        const int r = (kk + m - 1) / m;
        for (int i = t; i < Lp; i += blockDim.x) {
            uint64_t key = kSentinelKey;
            if (i < L) {
                const int seg = i / a.nsplit;
                const int slot = a.slot_of ? a.slot_of[(int64_t)q * a.nseg + seg] : q;
                if (slot >= 0)
                    key = a.part[(int64_t)slot * per_seg + (int64_t)(i % a.nsplit) * a.kpart +
                                 (m - 1)];
            }
            keys[i] = key;
        }
        if (t == 0)
            s_cnt = 0;
        __syncthreads();
        block_sort_keys(keys, Lp);
        uint64_t T = r <= L ? keys[r - 1] : kSentinelKey;
        if (T == kSentinelKey)
            T = kSentinelKey - 1; // no bound: all real keys
        __syncthreads();
        for (int64_t c = t; c < total; c += blockDim.x) {
            const int seg = (int)(c / per_seg);
            const int slot = a.slot_of ? a.slot_of[(int64_t)q * a.nseg + seg] : q;
            if (slot >= 0) {
                const uint64_t key = a.part[(int64_t)slot * per_seg + (c % per_seg)];
                if (key <= T) {
                    const int pos = atomicAdd(&s_cnt, 1);
                    if (pos < kSelCap)
                        keys[pos] = key;
                }
            }
        }
        __syncthreads();
        const int cnt = s_cnt;
        // (cnt < kk under a real bound cannot happen; if it ever did, the full sort is the answer)
        if (cnt <= kSelCap && (cnt >= kk || T == kSentinelKey - 1)) {
            int P = pow2_ceil(cnt > kk ? cnt : kk + 1);
            if (P < 32)
                P = 32;
            for (int i = cnt + t; i < P; i += blockDim.x)
                keys[i] = kSentinelKey;
            __syncthreads();
            block_sort_keys(keys, P);
            fast_done = true;
            if (T != kSentinelKey - 1)
                cut_key = T; // keys > T were not gathered
        }
        __syncthreads();
    }
    int kept = 0; // keys[0..kept) hold the running best, sorted
    int64_t done = fast_done ? total : 0;
    while (done < total) {
        const int room = M - kept; // M >= k + 32, so there is always room
        const int64_t left = total - done;
        const int take = left < (int64_t)room ? (int)left : room;
        for (int i = t; i < room; i += blockDim.x) {
            uint64_t key = kSentinelKey;
            if (i < take) {
                const int64_t c = done + i;
                const int seg = (int)(c / per_seg);
                const int within = (int)(c % per_seg);
                const int slot = a.slot_of ? a.slot_of[(int64_t)q * a.nseg + seg] : q;
                if (slot >= 0)
                    key = a.part[(int64_t)slot * per_seg + within];
            }
            keys[kept + i] = key;
        }
        __syncthreads();
        block_sort_keys(keys, M);
        if (keys[kk] < cut_key)
            cut_key = keys[kk]; // what this round cuts off is >= its first cut key
        __syncthreads();        // (slot kk is overwritten by the next round's load)
        kept = kk;
        done += take;
    }

    // checked re-rank: the smallest approximate score among the keys that will NOT be re-scored
    float nb_bound = kNeutralScore;
    float xn2 = 0.f;
    int nre = kk; // keys to re-score (see select_warp_kernel)
    if (a.slab && a.err_coef > 0.f) {
        // (every path above leaves keys[0..] sorted with at least kk + 1 valid slots: the next key,
        // a sentinel if there is none; keys cut by the gather bound or by a chunked round were
        // all >= a key that is still in the array or >= the bound T)
        const float* xq = a.queries + (size_t)q * a.d;
        for (int j = 0; j < a.d; ++j) { // (every thread: the same value without a reduction)
            const float xv = __ldg(xq + j);
            xn2 = fmaf(xv, xv, xn2);
        }
        xn2 *= 1.0000005f;
        const uint64_t ak = keys[a.k - 1];
        if (ak != kSentinelKey) {
            const float lim = key_score(ak) + 2.f * rerank_err(a, xn2);
            int lo = a.k, hi = kk; // first index in [k, kk] whose key is outside the window
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (keys[mid] != kSentinelKey && key_score(keys[mid]) <= lim)
                    lo = mid + 1;
                else
                    hi = mid;
            }
            nre = lo;
        }
        const uint64_t nk = keys[nre] < cut_key ? keys[nre] : cut_key;
        if (nk != kSentinelKey)
            nb_bound = key_score(nk);
        __syncthreads();
        for (int i = nre + t; i < kk; i += blockDim.x)
            keys[i] = kSentinelKey;
        __syncthreads();
    }
    if (a.slab) {
        // exact re-rank of the survivors: plain fp32, dimension order 0..d-1 in one accumulator,
        // the arithmetic of scan.cu / scan_tile.cu (fvec_L2sqr / fvec_inner_product,
        // faiss/utils/distances_simd.cpp:220-229,195-202, up to summation order). The rows are
        // fetched by the whole block, 32 float4 columns at a time, so that the (sequential) sums
        // of the kk candidate threads never wait on global memory.
        const bool l2m = a.metric == kMetricL2;
        float4* ybuf = reinterpret_cast<float4*>(sdis + a.k + 4);  // [kk][32]
        ybuf = reinterpret_cast<float4*>((reinterpret_cast<uintptr_t>(ybuf) + 15) & ~(uintptr_t)15);
        float4* xbuf = ybuf + (size_t)kk * 32;                      // [32]
        const float* xp = a.queries + (size_t)q * a.d;
        float acc = 0.f;
        uint64_t mykey = kSentinelKey;
        if (t < kk)
            mykey = keys[t];
        for (int c0 = 0; c0 < a.nf4; c0 += 32) {
            const int nc = a.nf4 - c0 < 32 ? a.nf4 - c0 : 32;
            for (int idx = t; idx < kk * nc; idx += blockDim.x) {
                const int cand = idx / nc, c = idx - cand * nc;
                const uint64_t key = keys[cand];
                float4 y = make_float4(0.f, 0.f, 0.f, 0.f);
                if (key != kSentinelKey) {
                    const uint32_t row = key_row(key);
                    y = __ldg(a.slab + ((size_t)(row / kBlockRows) * a.nf4 + c0 + c) * kBlockRows +
                              (row % kBlockRows));
                }
                ybuf[cand * 32 + c] = y;
            }
            for (int c = t; c < nc; c += blockDim.x) {
                const int dd = 4 * (c0 + c);
                float4 x;
                x.x = __ldg(xp + dd);
                x.y = dd + 1 < a.d ? __ldg(xp + dd + 1) : 0.f;
                x.z = dd + 2 < a.d ? __ldg(xp + dd + 2) : 0.f;
                x.w = dd + 3 < a.d ? __ldg(xp + dd + 3) : 0.f;
                xbuf[c] = x;
            }
            __syncthreads();
            if (t < kk && mykey != kSentinelKey) {
                for (int c = 0; c < nc; ++c) {
                    const float4 y = ybuf[t * 32 + c];
                    const float4 x = xbuf[c];
                    if (l2m) {
                        float u;
                        u = x.x - y.x;
                        acc = fmaf(u, u, acc);
                        u = x.y - y.y;
                        acc = fmaf(u, u, acc);
                        u = x.z - y.z;
                        acc = fmaf(u, u, acc);
                        u = x.w - y.w;
                        acc = fmaf(u, u, acc);
                    } else {
                        acc = fmaf(x.x, y.x, acc);
                        acc = fmaf(x.y, y.y, acc);
                        acc = fmaf(x.z, y.z, acc);
                        acc = fmaf(x.w, y.w, acc);
                    }
                }
            }
            __syncthreads();
        }
        if (t < kk && mykey != kSentinelKey)
            keys[t] = make_key((l2m ? acc : -acc) + 0.f, key_row(mykey));
        const int P = pow2_ceil(kk);
        for (int i = kk + t; i < P; i += blockDim.x)
            keys[i] = kSentinelKey;
        __syncthreads();
        block_sort_keys(keys, P);
        if (a.err_coef > 0.f && t == 0) {
            if (!rerank_proven(a, q, xn2, nb_bound, keys[a.k - 1]))
                flag_query(a, q);
        }
    }

    // decode the k best, map rows to user ids
    const bool l2 = a.metric == kMetricL2;
    for (int i = t; i < a.k; i += blockDim.x) {
        const uint64_t key = keys[i];
        if (key == kSentinelKey) {
            sdis[i] = l2 ? kNeutralScore : -kNeutralScore;
            sid[i] = -1;
        } else {
            const float s = key_score(key);
            sdis[i] = l2 ? s : -s;
            const uint32_t row = key_row(key);
            sid[i] = a.ids ? a.ids[row] : (int64_t)row;
        }
    }
    __syncthreads();
    // exact-distance ties: order by id the way cmp2 does (ascending for L2, descending for IP)
    for (int i = t; i < a.k; i += blockDim.x) {
        const float di = sdis[i];
        const int64_t idi = sid[i];
        int pos = i;
        if (idi >= 0) {
            int lo = i, hi = i;
            while (lo > 0 && sdis[lo - 1] == di && sid[lo - 1] >= 0)
                --lo;
            while (hi + 1 < a.k && sdis[hi + 1] == di && sid[hi + 1] >= 0)
                ++hi;
            if (hi > lo) {
                int before = 0;
                for (int j = lo; j <= hi; ++j) {
                    const int64_t idj = sid[j];
                    const bool first = l2 ? (idj < idi) : (idj > idi);
                    // equal ids (duplicates): keep the original relative order
                    before += (first || (idj == idi && j < i)) ? 1 : 0;
                }
                pos = lo + before;
            }
        }
        const int64_t o = (int64_t)q * a.k + pos;
        a.out_dis[o] = di;
        if (a.out_ids64)
            a.out_ids64[o] = idi;
        if (a.out_ids32)
            a.out_ids32[o] = (int32_t)idi;
    }
}

// ---- warp-per-query k-select (k <= 64, <= 256 partial lists) --------------------------------------
// Same result as select_kernel, without a single block-wide barrier: the head-bound prefilter, the
// stream compaction (ballot / popc), the sorts (bitonic, in the warp's own shared-memory strip) and
// the exact re-rank (lane = candidate) are all warp-synchronous. 8 queries per CTA.
constexpr int kSelWarpCap = 256;
constexpr int kSelWarpK = 64;
constexpr int kSelWarpBytes = kSelWarpCap * 8 + kSelWarpK * 12;

template <bool L2M>
__device__ __forceinline__ float rerank_one(const SelectArgs& a, int q, uint32_t row) {
    const float4* yp = a.slab + (size_t)(row / kBlockRows) * a.nf4 * kBlockRows + (row % kBlockRows);
    const float* xp = a.queries + (size_t)q * a.d;
    const bool vec = (a.d & 3) == 0 && (reinterpret_cast<uintptr_t>(a.queries) & 15) == 0;
    float acc = 0.f;
    for (int c0 = 0; c0 < a.nf4; c0 += 8) {
        // the 8 row loads (one 16-byte gather each, L2 latency) go out together; the query values
        // are the same address in every lane (an L1 broadcast) and are fetched as they are used
        float4 y[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int c = c0 + u;
            y[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (c < a.nf4)
                y[u] = __ldg(yp + (size_t)c * kBlockRows);
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int c = c0 + u;
            if (c < a.nf4) {
                float4 x;
                if (vec) {
                    x = __ldg(reinterpret_cast<const float4*>(xp + 4 * c));
                } else {
                    const int dd = 4 * c;
                    x.x = __ldg(xp + dd);
                    x.y = dd + 1 < a.d ? __ldg(xp + dd + 1) : 0.f;
                    x.z = dd + 2 < a.d ? __ldg(xp + dd + 2) : 0.f;
                    x.w = dd + 3 < a.d ? __ldg(xp + dd + 3) : 0.f;
                }
                if (L2M) {
                    float t;
                    t = x.x - y[u].x;
                    acc = fmaf(t, t, acc);
                    t = x.y - y[u].y;
                    acc = fmaf(t, t, acc);
                    t = x.z - y[u].z;
                    acc = fmaf(t, t, acc);
                    t = x.w - y[u].w;
                    acc = fmaf(t, t, acc);
                } else {
                    acc = fmaf(x.x, y[u].x, acc);
                    acc = fmaf(x.y, y[u].y, acc);
                    acc = fmaf(x.z, y[u].z, acc);
                    acc = fmaf(x.w, y[u].w, acc);
                }
            }
        }
    }
    return (L2M ? acc : -acc) + 0.f;
}

__global__ void __launch_bounds__(256, 4) select_warp_kernel(const SelectArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31;
    const int w = threadIdx.x >> 5;
    const int q = blockIdx.x * (blockDim.x >> 5) + w;
    if (q >= a.n)
        return;
    uint64_t* keys = reinterpret_cast<uint64_t*>(smem_raw + (size_t)w * kSelWarpBytes);
    int64_t* sid = reinterpret_cast<int64_t*>(keys + kSelWarpCap);
    float* sdis = reinterpret_cast<float*>(sid + kSelWarpK);

    const int per_seg = a.nsplit * a.kpart;
    const int total = a.nseg * per_seg;
    const int kk = a.slab ? a.k + a.rerank_extra : a.k;
    const int L = a.nseg * a.nsplit;
    const unsigned lt_mask = (1u << lane) - 1u;

    // 1. a bound on the kk-th smallest key from one pivot per (sorted) partial list: at least
    // r * m keys are <= the r-th smallest of the lists' m-th entries. m = 1 (many lists): the
    // kk-th smallest head; few long lists (the coarse quantizer: 8 lists of 32 for kk = 40):
    // m = ceil(kk / L), which cuts the 256 keys to about kk before anything is sorted.
    uint64_t T = kSentinelKey - 1;
    if (L >= 2) {
        int m = (kk + L - 1) / L;
        if (m > a.kpart)
            m = a.kpart;
        const int r = (kk + m - 1) / m;
        const int Lp = pow2_ceil(L < 32 ? 32 : L);
        for (int i = lane; i < Lp; i += 32) {
            uint64_t key = kSentinelKey;
            if (i < L) {
                const int seg = i / a.nsplit;
                const int slot = a.slot_of ? a.slot_of[(int64_t)q * a.nseg + seg] : q;
                if (slot >= 0)
                    key = a.part[(int64_t)slot * per_seg + (int64_t)(i % a.nsplit) * a.kpart +
                                 (m - 1)];
            }
            keys[i] = key;
        }
        __syncwarp();
        warp_sort_keys(keys, Lp, lane);
        if (r <= L)
            T = keys[r - 1];
        if (T == kSentinelKey)
            T = kSentinelKey - 1;
        __syncwarp();
    }

    // 2. stream the partial lists, keep the keys <= T; keys[0..kept) sorted, then cnt unsorted
    int kept = 0, cnt = 0;
    // (segment, offset) of this lane's next key, advanced by 32 per load: no division per key
    int seg = lane / per_seg, off = lane - seg * per_seg;
    for (int c0 = 0; c0 < total; c0 += 256) {
        uint64_t kv[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            kv[u] = kSentinelKey;
            if (c0 + u * 32 + lane < total) {
                const int slot = a.slot_of ? a.slot_of[(int64_t)q * a.nseg + seg] : q;
                if (slot >= 0)
                    kv[u] = a.part[(int64_t)slot * per_seg + off];
            }
            off += 32;
            while (off >= per_seg) {
                off -= per_seg;
                ++seg;
            }
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            bool pass = kv[u] <= T;
            unsigned m = __ballot_sync(0xffffffffu, pass);
            if (m == 0u)
                continue;
            if (kept + cnt + __popc(m) > kSelWarpCap) {
                for (int i = kept + cnt + lane; i < kSelWarpCap; i += 32)
                    keys[i] = kSentinelKey;
                __syncwarp();
                warp_sort_keys(keys, kSelWarpCap, lane);
                kept = kept + cnt < kk ? kept + cnt : kk;
                cnt = 0;
                if (kept == kk && keys[kk - 1] < T)
                    T = keys[kk - 1];
                __syncwarp();
                pass = kv[u] <= T;
                m = __ballot_sync(0xffffffffu, pass);
            }
            if (pass)
                keys[kept + cnt + __popc(m & lt_mask)] = kv[u];
            cnt += __popc(m);
            __syncwarp();
        }
    }
    // 3. sort what was kept
    {
        const int tot = kept + cnt;
        int P = pow2_ceil(tot > kk ? tot : kk + 1); // (slot kk is read by the checked re-rank)
        if (P < 32)
            P = 32;
        for (int i = tot + lane; i < P; i += 32)
            keys[i] = kSentinelKey;
        __syncwarp();
        warp_sort_keys(keys, P, lane);
    }
    // 4. exact re-rank of the kk survivors (lane = candidate), see select_kernel
    float nb_bound = kNeutralScore; // checked re-rank: best approximate score left out
    float xn2 = 0.f;
    int nre = kk; // keys to re-score
    if (a.slab && a.err_coef > 0.f) {
        const float* xp = a.queries + (size_t)q * a.d;
        for (int j = lane; j < a.d; j += 32) {
            const float xv = __ldg(xp + j);
            xn2 = fmaf(xv, xv, xn2);
        }
        for (int o = 16; o > 0; o >>= 1)
            xn2 += __shfl_xor_sync(0xffffffffu, xn2, o);
        xn2 *= 1.0000005f; // (summation order differs from the scan's: err on the safe side)
        // Only the keys inside the error window of the k-th approximate score A can reach the
        // exact top k: k rows score <= A, so the exact k-th score is <= A + E, and a row with an
        // approximate score > A + 2 E scores > A + E exactly. Usually that is k keys and one or two
        // more; the cut is then proven by construction.
        const uint64_t ak = keys[a.k - 1];
        if (ak != kSentinelKey) {
            const float lim = key_score(ak) + 2.f * rerank_err(a, xn2);
            int c = 0;
            for (int i = lane; i < kk; i += 32)
                c += (keys[i] != kSentinelKey && key_score(keys[i]) <= lim) ? 1 : 0;
            for (int o = 16; o > 0; o >>= 1)
                c += __shfl_xor_sync(0xffffffffu, c, o);
            nre = c < a.k ? a.k : c;
        }
        const uint64_t nk = keys[nre]; // (P >= kk + 1 slots are valid: sentinel if none)
        if (nk != kSentinelKey)
            nb_bound = key_score(nk);
        if (T != kSentinelKey - 1)
            nb_bound = fminf(nb_bound, key_score(T)); // keys > T were never gathered
        __syncwarp();
        for (int i = nre + lane; i < kk; i += 32)
            keys[i] = kSentinelKey; // approximate scores must not meet exact ones
        __syncwarp();
    }
    if (a.slab) {
        for (int i = lane; i < nre; i += 32) {
            const uint64_t key = keys[i];
            if (key != kSentinelKey) {
                const uint32_t row = key_row(key);
                const float sc = a.metric == kMetricL2 ? rerank_one<true>(a, q, row)
                                                       : rerank_one<false>(a, q, row);
                keys[i] = make_key(sc, row);
            }
        }
        int P = pow2_ceil(kk);
        if (P < 32)
            P = 32;
        for (int i = kk + lane; i < P; i += 32)
            keys[i] = kSentinelKey;
        __syncwarp();
        warp_sort_keys(keys, P, lane);
        if (a.err_coef > 0.f) {
            if (lane == 0 && !rerank_proven(a, q, xn2, nb_bound, keys[a.k - 1]))
                flag_query(a, q);
        }
    }
    // 5. decode, order exact ties by id as cmp2 does, write
    const bool l2 = a.metric == kMetricL2;
    for (int i = lane; i < a.k; i += 32) {
        const uint64_t key = keys[i];
        if (key == kSentinelKey) {
            sdis[i] = l2 ? kNeutralScore : -kNeutralScore;
            sid[i] = -1;
        } else {
            const float sc = key_score(key);
            sdis[i] = l2 ? sc : -sc;
            const uint32_t row = key_row(key);
            sid[i] = a.ids ? a.ids[row] : (int64_t)row;
        }
    }
    __syncwarp();
    for (int i = lane; i < a.k; i += 32) {
        const float di = sdis[i];
        const int64_t idi = sid[i];
        int pos = i;
        if (idi >= 0) {
            int lo = i, hi = i;
            while (lo > 0 && sdis[lo - 1] == di && sid[lo - 1] >= 0)
                --lo;
            while (hi + 1 < a.k && sdis[hi + 1] == di && sid[hi + 1] >= 0)
                ++hi;
            if (hi > lo) {
                int before = 0;
                for (int j = lo; j <= hi; ++j) {
                    const int64_t idj = sid[j];
                    const bool first = l2 ? (idj < idi) : (idj > idi);
                    before += (first || (idj == idi && j < i)) ? 1 : 0;
                }
                pos = lo + before;
            }
        }
        const int64_t o = (int64_t)q * a.k + pos;
        a.out_dis[o] = di;
        if (a.out_ids64)
            a.out_ids64[o] = idi;
        if (a.out_ids32)
            a.out_ids32[o] = (int32_t)idi;
    }
}

// One warp per query: merge nshard sorted (n,k) tables by walking the shard heads. Same output as
// merge_knn_results (faiss/utils/Heap.cpp:158-232): equal distances come out in shard order (the
// reference's heap over the shard heads compares (distance, shard), :186-225: lower shard first
// for L2, higher first for inner product), exactly as merge_shards_rank_kernel below orders them.
__global__ void merge_shards_kernel(int metric, int nshard, int n, int k, const float* all_dis,
                                    const int64_t* all_ids, float* out_dis, int64_t* out_ids) {
    const int q = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (q >= n)
        return;
    const bool l2 = metric == kMetricL2;
    const size_t stride = (size_t)n * k;
    // lane s (and s+32, ...) owns the head pointer of shard s
    const int per = (nshard + 31) / 32;
    int head[8]; // up to 256 shards
    for (int i = 0; i < per && i < 8; ++i)
        head[i] = 0;
    for (int j = 0; j < k; ++j) {
        float best = 0.f;
        int64_t best_id = -1;
        int best_s = -1;
        for (int i = 0; i < per && i < 8; ++i) {
            const int s = lane + 32 * i;
            if (s < nshard && head[i] < k) {
                const size_t o = stride * s + (size_t)q * k + head[i];
                const int64_t id = all_ids[o];
                if (id >= 0) {
                    const float dv = all_dis[o];
                    bool better;
                    if (best_s < 0)
                        better = true;
                    else // equal distances: L2 keeps the lower shard (seen first), inner
                         // product takes the higher one (CMax::cmp2)
                        better = l2 ? (dv < best) : (dv >= best);
                    if (better) {
                        best = dv;
                        best_id = id;
                        best_s = s;
                    }
                }
            }
        }
        // warp arg-best
        for (int o = 16; o > 0; o >>= 1) {
            const float od = __shfl_xor_sync(0xffffffffu, best, o);
            const int64_t oid = __shfl_xor_sync(0xffffffffu, best_id, o);
            const int os = __shfl_xor_sync(0xffffffffu, best_s, o);
            bool take;
            if (os < 0)
                take = false;
            else if (best_s < 0)
                take = true;
            else if (od != best)
                take = l2 ? (od < best) : (od > best);
            else
                take = l2 ? os < best_s : os > best_s;
            if (take) {
                best = od;
                best_id = oid;
                best_s = os;
            }
        }
        if (best_s >= 0) {
            if ((best_s & 31) == lane)
                head[best_s >> 5]++;
        }
        if (lane == 0) {
            const size_t o = (size_t)q * k + j;
            if (best_s >= 0) {
                out_dis[o] = best;
                out_ids[o] = best_id;
            } else {
                out_dis[o] = l2 ? kNeutralScore : -kNeutralScore;
                out_ids[o] = -1;
            }
        }
    }
}

} // namespace

cudaError_t launch_select(const SelectArgs& a, cudaStream_t st) {
    if (a.n == 0)
        return cudaSuccess;
    const int64_t total = (int64_t)a.nseg * a.nsplit * a.kpart;
    const int kk = a.slab ? a.k + a.rerank_extra : a.k;
    // few queries with many keys each (small batches split over many warps): a warp per query
    // would walk its keys serially and leave the GPU empty; a wide CTA per query instead
    const bool few_wide = a.n <= 2 * 148 && total >= 1024;
    if (!few_wide && a.k <= kSelWarpK && kk <= kSelWarpK && (int64_t)a.nseg * a.nsplit <= 256 &&
        total < (1 << 30)) {
        const int wpb = 8;
        select_warp_kernel<<<(a.n + wpb - 1) / wpb, wpb * 32, wpb * kSelWarpBytes, st>>>(a);
        count_launches(1);
        return cudaGetLastError();
    }
    int M = pow2_ceil(kk + (int)std::min<int64_t>(std::max<int64_t>(total, 32), 2048));
    if (M < 64)
        M = 64;
    const int64_t L = (int64_t)a.nseg * a.nsplit;
    const int prefilter = (total > 256 && L >= 4 && L <= 2048) ? 1 : 0;
    if (prefilter)
        M = std::max(M, std::max(kSelCap, pow2_ceil((int)L)));
    size_t smem = (size_t)M * 8 + (size_t)a.k * 12 + 16;
    if (a.slab)
        smem += 32 + ((size_t)kk * 32 + 32) * sizeof(float4);
    int threads = std::min(512, std::max(64, M / 2));
    if (prefilter)
        threads = few_wide ? (int)std::min<int64_t>(1024, pow2_ceil((int)std::min<int64_t>(total / 4, 1024)))
                           : 128;
    if (a.slab) // the exact re-rank gives one thread to each of the kk survivors
        threads = std::max(threads, (kk + 31) / 32 * 32);
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(select_kernel,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess)
            return e;
    }
    select_kernel<<<a.n, threads, smem, st>>>(a, M, prefilter);
    count_launches(1);
    return cudaGetLastError();
}

// The same merge without its chain of dependent loads (the walk above pays one L2 round trip per
// output slot): one warp per query loads all nshard * k entries at once into shared memory, every
// entry finds its rank in the merged order as its own position plus, for every other shard, the
// number of that shard's entries that come before it (a binary search: shard tables are sorted),
// and the entries of rank < k go straight to their slots. nshard * k <= kMergeRankMax.
constexpr int kMergeRankMax = 512;
constexpr int kMergeRankWarps = 4;

// equal distances: ordered by shard number as the reference's heap over the shard heads orders
// them (cmp2 on (distance, shard), faiss/utils/Heap.cpp:186-225): lower shard first for L2, higher
// shard first for inner product; ids take no part, so the order inside a shard's own run of equal
// distances is whatever the shard delivered
__device__ __forceinline__ bool merge_before(bool l2, float da, int sa, float db, int sb) {
    if (da != db)
        return l2 ? (da < db) : (da > db);
    // CMin::cmp2 (L2) pops the LOWER shard first, CMax::cmp2 (inner product) the HIGHER one
    // (faiss/utils/ordered_key_value.h:51-53,74-76 applied to (distance, shard))
    return l2 ? sa < sb : sa > sb;
}

__global__ void __launch_bounds__(kMergeRankWarps * 32)
        merge_shards_rank_kernel(int metric, int nshard, int n, int k, const float* all_dis,
                                 const int64_t* all_ids, float* out_dis, int64_t* out_ids) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = blockIdx.x * kMergeRankWarps + w;
    if (q >= n)
        return;
    const bool l2 = metric == kMetricL2;
    const int E = nshard * k;
    int64_t* sid = reinterpret_cast<int64_t*>(smem_raw) + (size_t)w * E;
    float* sd = reinterpret_cast<float*>(reinterpret_cast<int64_t*>(smem_raw) +
                                         (size_t)kMergeRankWarps * E) + (size_t)w * E;
    const size_t stride = (size_t)n * k;
    for (int e = lane; e < E; e += 32) {
        const int s = e / k, pos = e - s * k;
        const size_t o = stride * s + (size_t)q * k + pos;
        sid[e] = all_ids[o];
        sd[e] = all_dis[o];
    }
    __syncwarp();
    int nvalid = 0;
    for (int e = lane; e < E; e += 32) {
        const int64_t id = sid[e];
        if (id < 0)
            continue;
        nvalid++;
        const int s = e / k, pos = e - s * k;
        const float dv = sd[e];
        int rank = pos;
        for (int s2 = 0; s2 < nshard; ++s2) {
            if (s2 == s)
                continue;
            // entries of shard s2 that come before this one: first index that does not
            int lo = 0, hi = k;
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                const int64_t im = sid[s2 * k + mid];
                if (im >= 0 && merge_before(l2, sd[s2 * k + mid], s2, dv, s))
                    lo = mid + 1;
                else
                    hi = mid;
            }
            rank += lo;
        }
        if (rank < k) {
            const size_t o = (size_t)q * k + rank;
            out_dis[o] = dv;
            out_ids[o] = id;
        }
    }
    for (int o = 16; o > 0; o >>= 1)
        nvalid += __shfl_xor_sync(0xffffffffu, nvalid, o);
    for (int j = nvalid + lane; j < k; j += 32) { // not enough results: the reference's padding
        out_dis[(size_t)q * k + j] = l2 ? kNeutralScore : -kNeutralScore;
        out_ids[(size_t)q * k + j] = -1;
    }
}

cudaError_t launch_merge_shards(int metric, int nshard, int n, int k, const float* all_dis,
                                const int64_t* all_ids, float* out_dis, int64_t* out_ids,
                                cudaStream_t st) {
    if (n == 0)
        return cudaSuccess;
    if (nshard > 256)
        return cudaErrorInvalidValue;
    if ((int64_t)nshard * k <= kMergeRankMax) {
        const size_t smem = (size_t)kMergeRankWarps * nshard * k * (sizeof(int64_t) + sizeof(float));
        merge_shards_rank_kernel<<<(n + kMergeRankWarps - 1) / kMergeRankWarps,
                                   kMergeRankWarps * 32, smem, st>>>(metric, nshard, n, k, all_dis,
                                                                     all_ids, out_dis, out_ids);
        count_launches(1);
        return cudaGetLastError();
    }
    const int wpb = 4;
    merge_shards_kernel<<<(n + wpb - 1) / wpb, wpb * 32, 0, st>>>(metric, nshard, n, k, all_dis,
                                                                 all_ids, out_dis, out_ids);
    count_launches(1);
    return cudaGetLastError();
}

} // namespace b200ivf
