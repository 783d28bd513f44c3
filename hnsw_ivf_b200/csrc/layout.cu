// layout.cu -- build-path kernels: put row-major fp32 vectors (the reference's code bytes,
// ArrayInvertedLists::codes, faiss/invlists/InvertedLists.h:243-245) into the device slab's
// interleaved 32-row blocks (common.cuh) and back.
#include "common.cuh"
#include "kernels.h"

#include <algorithm>

namespace b200ivf {

namespace {

// one warp per vector: lane strides over dims
__global__ void scatter_rows_kernel(const float* __restrict__ x, const int64_t* __restrict__ xids,
                                    const int64_t* __restrict__ dst_row, int64_t n, int d, int nf4,
                                    float* __restrict__ slab, int64_t* __restrict__ ids,
                                    int64_t id_base) {
    const int lane = threadIdx.x & 31;
    const int64_t w0 = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t i = w0; i < n; i += nw) {
        const int64_t row = dst_row[i];
        if (row < 0)
            continue;
        const float* xi = x + i * d;
        for (int dd = lane; dd < d; dd += 32)
            slab[slab_float_index((size_t)row, dd, nf4)] = xi[dd];
        if (lane == 0)
            ids[row] = xids ? xids[i] : id_base + i;
    }
}

// one CTA per list: whole blocks are position independent, so they move as plain float4 copies
__global__ void move_blocks_kernel(const float4* __restrict__ old_slab,
                                   const int64_t* __restrict__ old_ids,
                                   const uint32_t* __restrict__ old_blk0,
                                   const int32_t* __restrict__ old_len,
                                   const uint32_t* __restrict__ new_blk0, int nlists, int nf4,
                                   float4* __restrict__ new_slab, int64_t* __restrict__ new_ids) {
    for (int l = blockIdx.x; l < nlists; l += gridDim.x) {
        const int len = old_len[l];
        if (len == 0)
            continue;
        const int nb = (len + kBlockRows - 1) / kBlockRows;
        const size_t nvec = (size_t)nb * nf4 * kBlockRows;
        const float4* src = old_slab + (size_t)old_blk0[l] * nf4 * kBlockRows;
        float4* dst = new_slab + (size_t)new_blk0[l] * nf4 * kBlockRows;
        for (size_t i = threadIdx.x; i < nvec; i += blockDim.x)
            dst[i] = src[i];
        const int64_t* sid = old_ids + (size_t)old_blk0[l] * kBlockRows;
        int64_t* did = new_ids + (size_t)new_blk0[l] * kBlockRows;
        for (int i = threadIdx.x; i < len; i += blockDim.x)
            did[i] = sid[i];
    }
}

__global__ void gather_rows_kernel(const float* __restrict__ slab, int64_t row0, int64_t nrows,
                                   int d, int nf4, float* __restrict__ out) {
    const int64_t total = nrows * d;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / d;
        const int dd = (int)(i % d);
        out[i] = slab[slab_float_index((size_t)(row0 + r), dd, nf4)];
    }
}

__global__ void i64_to_i32_kernel(const int64_t* __restrict__ in, int32_t* __restrict__ out,
                                  int64_t n, int32_t lo, int32_t hi, unsigned long long* bad) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t v = in[i];
        int32_t o;
        if (v < 0) {
            o = -1;
        } else if (v >= hi || v < lo) {
            o = -1;
            if (bad)
                atomicAdd(bad, 1ull);
        } else {
            o = (int32_t)v;
        }
        out[i] = o;
    }
}

inline int grid_for(int64_t work, int threads) {
    return (int)std::max<int64_t>(1, std::min<int64_t>((work + threads - 1) / threads, 148 * 16));
}

} // namespace

cudaError_t launch_scatter_rows(const float* x, const int64_t* xids, const int64_t* dst_row,
                                int64_t n, int d, int nf4, float* slab, int64_t* ids,
                                int64_t id_base, cudaStream_t st) {
    if (n == 0)
        return cudaSuccess;
    scatter_rows_kernel<<<grid_for(n * 32, 256), 256, 0, st>>>(x, xids, dst_row, n, d, nf4, slab,
                                                               ids, id_base);
    count_launches(1);
    return cudaGetLastError();
}

cudaError_t launch_move_blocks(const float4* old_slab, const int64_t* old_ids,
                               const uint32_t* old_blk0, const int32_t* old_len,
                               const uint32_t* new_blk0, int nlists, int nf4, float4* new_slab,
                               int64_t* new_ids, cudaStream_t st) {
    if (nlists == 0)
        return cudaSuccess;
    move_blocks_kernel<<<std::min(nlists, 148 * 8), 256, 0, st>>>(
            old_slab, old_ids, old_blk0, old_len, new_blk0, nlists, nf4, new_slab, new_ids);
    count_launches(1);
    return cudaGetLastError();
}

// one thread per row: lane = row within its 32-row block, so each column read is coalesced
__global__ void row_norms_kernel(const float4* __restrict__ slab, int64_t nrows, int nf4,
                                 float* __restrict__ norms, float* __restrict__ max_out) {
    float mx = 0.f;
    for (int64_t row = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; row < nrows;
         row += (int64_t)gridDim.x * blockDim.x) {
        const float4* p =
                slab + (row / kBlockRows) * (int64_t)nf4 * kBlockRows + (row % kBlockRows);
        float acc = 0.f;
        for (int c = 0; c < nf4; ++c) {
            const float4 v = p[(int64_t)c * kBlockRows];
            acc = fmaf(v.x, v.x, acc);
            acc = fmaf(v.y, v.y, acc);
            acc = fmaf(v.z, v.z, acc);
            acc = fmaf(v.w, v.w, acc);
        }
        norms[row] = acc;
        mx = fmaxf(mx, acc);
    }
    if (max_out) { // (non-negative floats order like their bit patterns)
        for (int o = 16; o > 0; o >>= 1)
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        if ((threadIdx.x & 31) == 0 && mx > 0.f)
            atomicMax(reinterpret_cast<unsigned*>(max_out), __float_as_uint(mx));
    }
}

cudaError_t launch_row_norms(const float4* slab, int64_t nrows, int nf4, float* norms,
                             cudaStream_t st, float* max_out) {
    if (nrows == 0)
        return cudaSuccess;
    row_norms_kernel<<<grid_for(nrows, 256), 256, 0, st>>>(slab, nrows, nf4, norms, max_out);
    count_launches(1);
    return cudaGetLastError();
}

cudaError_t launch_gather_rows(const float* slab, int64_t row0, int64_t nrows, int d, int nf4,
                               float* out, cudaStream_t st) {
    if (nrows == 0)
        return cudaSuccess;
    gather_rows_kernel<<<grid_for(nrows * d, 256), 256, 0, st>>>(slab, row0, nrows, d, nf4, out);
    count_launches(1);
    return cudaGetLastError();
}

__global__ void fill_u32_kernel(uint32_t* p, int64_t n, uint32_t v) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x)
        p[i] = v;
}

cudaError_t launch_fill_u32(uint32_t* p, int64_t n, uint32_t v, cudaStream_t st) {
    if (n == 0)
        return cudaSuccess;
    fill_u32_kernel<<<grid_for(n, 256), 256, 0, st>>>(p, n, v);
    count_launches(1);
    return cudaGetLastError();
}

cudaError_t launch_i64_to_i32(const int64_t* in, int32_t* out, int64_t n, int32_t lo, int32_t hi,
                              unsigned long long* bad, cudaStream_t st) {
    if (n == 0)
        return cudaSuccess;
    i64_to_i32_kernel<<<grid_for(n, 256), 256, 0, st>>>(in, out, n, lo, hi, bad);
    count_launches(1);
    return cudaGetLastError();
}

namespace {
// one thread per (list, dimension): a strictly sequential sum over the list's rows, so that the
// result is the reference's (Clustering.cpp:183-204: c[j] += xi[j] in input order, then *= 1/n)
__global__ void list_means_kernel(const float* __restrict__ slab,
                                  const uint32_t* __restrict__ list_blk0,
                                  const int32_t* __restrict__ list_len, int d, int nf4,
                                  float* __restrict__ centroids) {
    const int l = blockIdx.x;
    const int len = list_len[l];
    const size_t row0 = (size_t)list_blk0[l] * kBlockRows;
    for (int dd = threadIdx.x; dd < d; dd += blockDim.x) {
        float c = 0.f;
        for (int r = 0; r < len; ++r)
            c += slab[slab_float_index(row0 + r, dd, nf4)];
        if (len > 0)
            c *= 1.f / (float)len;
        centroids[(size_t)l * d + dd] = c;
    }
}

__global__ void gather_by_index_kernel(const float* __restrict__ x, const int64_t* __restrict__ idx,
                                       int64_t n, int d, float* __restrict__ out) {
    const int64_t total = n * d;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / d;
        out[i] = x[idx[r] * d + (i - r * d)];
    }
}
} // namespace

cudaError_t launch_list_means(const float* slab, const uint32_t* list_blk0, const int32_t* list_len,
                              int nlists, int d, int nf4, float* centroids, cudaStream_t st) {
    if (nlists == 0)
        return cudaSuccess;
    const int threads = d >= 256 ? 256 : (d >= 128 ? 128 : 64);
    list_means_kernel<<<nlists, threads, 0, st>>>(slab, list_blk0, list_len, d, nf4, centroids);
    count_launches(1);
    return cudaGetLastError();
}

__global__ void gather_i64_kernel(const int64_t* __restrict__ in, const int64_t* __restrict__ idx,
                                  int64_t n, int64_t base, int64_t* __restrict__ out) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x)
        out[i] = in ? in[idx[i]] : base + idx[i];
}

cudaError_t launch_gather_i64(const int64_t* in, const int64_t* idx, int64_t n, int64_t base,
                              int64_t* out, cudaStream_t st) {
    if (n == 0)
        return cudaSuccess;
    const int threads = 256;
    const int grid = (int)std::min<int64_t>((n + threads - 1) / threads, 148 * 16);
    gather_i64_kernel<<<grid, threads, 0, st>>>(in, idx, n, base, out);
    count_launches(1);
    return cudaGetLastError();
}

cudaError_t launch_gather_by_index(const float* x, const int64_t* idx, int64_t n, int d, float* out,
                                   cudaStream_t st) {
    if (n == 0)
        return cudaSuccess;
    const int threads = 256;
    const int64_t total = n * d;
    const int grid = (int)std::min<int64_t>((total + threads - 1) / threads, 148 * 16);
    gather_by_index_kernel<<<grid, threads, 0, st>>>(x, idx, n, d, out);
    count_launches(1);
    return cudaGetLastError();
}

// ---- the flagged queries of a checked re-rank (SelectArgs::flag_list) ---------------------------
// rows idx[0..cnt) of a row-major table to a dense one (cnt is a host value: the caller has read
// the flag counter); W 4-byte words per row
__global__ void gather_rows_w_kernel(const uint32_t* __restrict__ in, const int32_t* __restrict__ idx,
                                     int64_t cnt, int W, uint32_t* __restrict__ out) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < cnt * W;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / W;
        out[i] = in[(int64_t)idx[r] * W + (i - r * W)];
    }
}
cudaError_t launch_gather_rows_w(const void* in, const int32_t* idx, int64_t cnt, int words,
                                 void* out, cudaStream_t st) {
    if (cnt == 0)
        return cudaSuccess;
    gather_rows_w_kernel<<<grid_for(cnt * words, 256), 256, 0, st>>>(
            static_cast<const uint32_t*>(in), idx, cnt, words, static_cast<uint32_t*>(out));
    count_launches(1);
    return cudaGetLastError();
}
// ... and the result rows of the redone queries back to their places
__global__ void scatter_results_kernel(const float* __restrict__ Df, const int64_t* __restrict__ If,
                                       const int32_t* __restrict__ idx, int64_t cnt, int k,
                                       float* __restrict__ D, int64_t* __restrict__ I) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < cnt * k;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / k;
        const int64_t o = (int64_t)idx[r] * k + (i - r * k);
        D[o] = Df[i];
        I[o] = If[i];
    }
}
cudaError_t launch_scatter_results(const float* Df, const int64_t* If, const int32_t* idx,
                                   int64_t cnt, int k, float* D, int64_t* I, cudaStream_t st) {
    if (cnt == 0)
        return cudaSuccess;
    scatter_results_kernel<<<grid_for(cnt * k, 256), 256, 0, st>>>(Df, If, idx, cnt, k, D, I);
    count_launches(1);
    return cudaGetLastError();
}

// rows of a dense table back to rows idx[0..cnt) of a larger one (`words` 4-byte words per row)
__global__ void scatter_rows_w_kernel(const uint32_t* __restrict__ in, const int32_t* __restrict__ idx,
                                      int64_t cnt, int W, uint32_t* __restrict__ out) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < cnt * W;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / W;
        out[(int64_t)idx[r] * W + (i - r * W)] = in[i];
    }
}
cudaError_t launch_scatter_rows_w(const void* in, const int32_t* idx, int64_t cnt, int words,
                                  void* out, cudaStream_t st) {
    if (cnt == 0)
        return cudaSuccess;
    scatter_rows_w_kernel<<<grid_for(cnt * words, 256), 256, 0, st>>>(
            static_cast<const uint32_t*>(in), idx, cnt, words, static_cast<uint32_t*>(out));
    count_launches(1);
    return cudaGetLastError();
}

} // namespace b200ivf
