// common.cuh -- shared device/host helpers for the B200 IVF-Flat search path.
//
// Selection everywhere works on one 64-bit key per candidate:
//     key = (ordered_bits(score) << 32) | row
// where score is "smaller is better" (L2: the squared distance; inner product: the negated
// product), ordered_bits() is the usual monotone float->uint map and row is the candidate's row in
// the device slab (or the centroid number for the coarse quantizer). One unsigned 64-bit compare is
// then the lexicographic (score, row) order, the same shape as the reference's cmp2
// (faiss/utils/ordered_key_value.h:51-53,74-76) with the slab row standing in for the id until the
// final pass (select.cu) re-orders exact ties by the user id.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace b200ivf {

constexpr int kMetricInnerProduct = 0; // faiss/MetricType.h:23
constexpr int kMetricL2 = 1;           // faiss/MetricType.h:24

constexpr int kBlockRows = 32;              // rows per interleaved block (one lane per row)
constexpr uint64_t kSentinelKey = ~0ull;    // sorts after every real candidate
constexpr float kNeutralScore = 3.402823466e+38f; // FLT_MAX: heap_heapify fill (Heap.h:333)

__host__ __device__ __forceinline__ uint32_t ordered_bits(float f) {
#ifdef __CUDA_ARCH__
    uint32_t u = __float_as_uint(f);
#else
    union {
        float f;
        uint32_t u;
    } c;
    c.f = f;
    uint32_t u = c.u;
#endif
    return u ^ ((uint32_t)((int32_t)u >> 31) | 0x80000000u);
}

__host__ __device__ __forceinline__ float from_ordered_bits(uint32_t o) {
    uint32_t u = (o & 0x80000000u) ? (o ^ 0x80000000u) : ~o;
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    union {
        float f;
        uint32_t u;
    } c;
    c.u = u;
    return c.f;
#endif
}

__host__ __device__ __forceinline__ uint64_t make_key(float score, uint32_t row) {
    return ((uint64_t)ordered_bits(score) << 32) | (uint64_t)row;
}
__host__ __device__ __forceinline__ float key_score(uint64_t key) {
    return from_ordered_bits((uint32_t)(key >> 32));
}
__host__ __device__ __forceinline__ uint32_t key_row(uint64_t key) {
    return (uint32_t)(key & 0xffffffffu);
}

__host__ __device__ __forceinline__ int pow2_ceil(int x) {
    int p = 1;
    while (p < x)
        p <<= 1;
    return p;
}

#ifdef __CUDACC__
// Ascending bitonic sort of n (power of two) keys in shared memory by ONE warp.
static __device__ __noinline__ void warp_sort_keys(uint64_t* a, int n, int lane) {
    for (int size = 2; size <= n; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int t = lane; t < (n >> 1); t += 32) {
                const int i = 2 * t - (t & (stride - 1));
                const int j = i + stride;
                const bool up = (i & size) == 0;
                const uint64_t x = a[i], y = a[j];
                if ((x > y) == up) {
                    a[i] = y;
                    a[j] = x;
                }
            }
            __syncwarp();
        }
    }
}

// Ascending bitonic sort of exactly 64 keys in shared memory by one warp, done in registers:
// lane holds elements `lane` and `lane + 32`; strides < 32 are shuffles, stride 32 is in-thread.
static __device__ __forceinline__ void warp_sort64_keys(uint64_t* a, int lane) {
    uint64_t v0 = a[lane], v1 = a[lane + 32];
#pragma unroll
    for (int size = 2; size <= 64; size <<= 1) {
#pragma unroll
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            if (stride == 32) {
                // elements lane and lane+32, ascending (size == 64)
                const uint64_t lo = v0 < v1 ? v0 : v1, hi = v0 < v1 ? v1 : v0;
                v0 = lo;
                v1 = hi;
            } else {
                const uint64_t p0 = __shfl_xor_sync(0xffffffffu, v0, stride);
                const uint64_t p1 = __shfl_xor_sync(0xffffffffu, v1, stride);
                const bool lower = (lane & stride) == 0; // this lane holds the lower index
                // element indices: e0 = lane, e1 = lane + 32
                const bool up0 = (lane & size) == 0;
                const bool up1 = ((lane + 32) & size) == 0;
                const bool take_min0 = lower == up0, take_min1 = lower == up1;
                v0 = take_min0 ? (v0 < p0 ? v0 : p0) : (v0 < p0 ? p0 : v0);
                v1 = take_min1 ? (v1 < p1 ? v1 : p1) : (v1 < p1 ? p1 : v1);
            }
        }
    }
    a[lane] = v0;
    a[lane + 32] = v1;
    __syncwarp();
}

// A per-query candidate buffer of `cap` keys (power of two) holding `cnt` entries: sort it and keep
// the k best; returns the new fill and the new acceptance threshold (the k-th best score once k
// candidates exist -- the reference's heap top, faiss/utils/Heap.h:111-150).
struct KeyBufState {
    int cnt;
    float thr;
};
static __device__ __noinline__ KeyBufState warp_compact_keys(uint64_t* buf, int cnt, float thr,
                                                             int k, int cap, int lane) {
    for (int i = cnt + lane; i < cap; i += 32)
        buf[i] = kSentinelKey;
    __syncwarp();
    if (cap == 64)
        warp_sort64_keys(buf, lane);
    else
        warp_sort_keys(buf, cap, lane);
    if (cnt >= k) {
        cnt = k;
        thr = key_score(buf[k - 1]);
    }
    return {cnt, thr};
}

// Same network run by a whole thread block.
__device__ __forceinline__ void block_sort_keys(uint64_t* a, int n) {
    for (int size = 2; size <= n; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int t = threadIdx.x; t < (n >> 1); t += blockDim.x) {
                const int i = 2 * t - (t & (stride - 1));
                const int j = i + stride;
                const bool up = (i & size) == 0;
                const uint64_t x = a[i], y = a[j];
                if ((x > y) == up) {
                    a[i] = y;
                    a[j] = x;
                }
            }
            __syncthreads();
        }
    }
}
#endif

// ---- device slab layout ------------------------------------------------------------------------
// A slab is a sequence of 32-row blocks. Block b holds rows 32b..32b+31 dimension-interleaved in
// float4 columns: float4 index ((b * nf4 + c) * 32 + lane) is dims [4c,4c+4) of row 32b+lane, so a
// warp that gives one lane to each row reads column c of a block as one 512-byte coalesced load.
// nf4 = ceil(d/4); dims d..4*nf4-1 and rows past the end of a list are zero.
__host__ __device__ __forceinline__ size_t slab_float_index(size_t row, int dim, int nf4) {
    const size_t b = row / kBlockRows;
    const int lane = (int)(row % kBlockRows);
    return ((b * nf4 + (dim >> 2)) * kBlockRows + lane) * 4 + (dim & 3);
}

} // namespace b200ivf
