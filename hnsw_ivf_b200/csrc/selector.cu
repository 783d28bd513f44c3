// selector.cu -- device evaluation of the IDSelector family into slab row masks (selector.h).
// One warp per 32-row block: lane r tests id[32*b + r], a ballot is the mask word.
#include "selector.h"

#include "common.cuh"
#include "index.h"
#include "kernels.h"

#include <algorithm>

namespace b200ivf {

namespace {

struct RangeTest {
    int64_t imin, imax;
    __device__ bool operator()(int64_t id) const {
        return id >= imin && id < imax;
    }
};
struct SortedSetTest {
    const int64_t* s;
    int64_t n;
    __device__ bool operator()(int64_t id) const {
        int64_t lo = 0, hi = n;
        while (lo < hi) {
            const int64_t mid = (lo + hi) >> 1;
            if (s[mid] < id)
                lo = mid + 1;
            else
                hi = mid;
        }
        return lo < n && s[lo] == id;
    }
};
struct BitmapTest {
    const uint8_t* bits;
    uint64_t nbytes;
    __device__ bool operator()(int64_t id) const {
        const uint64_t i = (uint64_t)id; // negative ids land beyond any bitmap (IDSelector.cpp:118)
        if ((i >> 3) >= nbytes)
            return false;
        return (bits[i >> 3] >> (i & 7)) & 1;
    }
};

template <class T>
__global__ void mask_eval_kernel(const int64_t* __restrict__ ids, int64_t nblocks,
                                 uint32_t* __restrict__ mask, T test) {
    const int lane = threadIdx.x & 31;
    const int64_t w0 = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t b = w0; b < nblocks; b += nw) {
        const bool in = test(ids[b * kBlockRows + lane]);
        const unsigned m = __ballot_sync(0xffffffffu, in);
        if (lane == 0)
            mask[b] = m;
    }
}

// op: 0 and, 1 or, 2 xor, 3 not (rhs unused), 4 fill ones
__global__ void mask_combine_kernel(uint32_t* __restrict__ a, const uint32_t* __restrict__ b,
                                    int64_t n, int op) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t x = a[i];
        uint32_t r;
        switch (op) {
            case 0:
                r = x & b[i];
                break;
            case 1:
                r = x | b[i];
                break;
            case 2:
                r = x ^ b[i];
                break;
            case 3:
                r = ~x;
                break;
            default:
                r = 0xffffffffu;
        }
        a[i] = r;
    }
}

template <class T>
void run_eval(const int64_t* ids, int64_t nblocks, uint32_t* mask, T test, cudaStream_t st) {
    if (nblocks == 0)
        return;
    const int threads = 256;
    const int grid = (int)std::min<int64_t>((nblocks * 32 + threads - 1) / threads, 148 * 16);
    mask_eval_kernel<T><<<grid, threads, 0, st>>>(ids, nblocks, mask, test);
    count_launches(1);
}

void run_combine(uint32_t* a, const uint32_t* b, int64_t n, int op, cudaStream_t st) {
    if (n == 0)
        return;
    const int threads = 256;
    const int grid = (int)std::min<int64_t>((n + threads - 1) / threads, 148 * 16);
    mask_combine_kernel<<<grid, threads, 0, st>>>(a, b, n, op);
    count_launches(1);
}

void check(cudaError_t e, const char* what) {
    if (e != cudaSuccess)
        throw B200Exception(std::string("CUDA error ") + std::to_string((int)e) + " (" +
                            cudaGetErrorString(e) + ") in " + what);
}

struct SetDevice {
    int prev = 0;
    explicit SetDevice(int dev) {
        cudaGetDevice(&prev);
        if (prev != dev)
            cudaSetDevice(dev);
    }
    ~SetDevice() {
        cudaSetDevice(prev);
    }
};

} // namespace

void IDSelectorRange::eval(const int64_t* ids, int64_t nblocks, uint32_t* mask,
                           cudaStream_t st) const {
    run_eval(ids, nblocks, mask, RangeTest{imin, imax}, st);
    check(cudaGetLastError(), "IDSelectorRange");
}

IDSelectorBatch::IDSelectorBatch(size_t n, const int64_t* indices, int dev) : device(dev) {
    if (n > 0 && !indices)
        throw B200Exception("IDSelectorBatch: null indices");
    sorted.assign(indices, indices + n);
    std::sort(sorted.begin(), sorted.end());
    sorted.erase(std::unique(sorted.begin(), sorted.end()), sorted.end());
    SetDevice g(device);
    check(cudaMalloc(&d_sorted, std::max<size_t>(sorted.size(), 1) * sizeof(int64_t)),
          "IDSelectorBatch");
    check(cudaMemcpy(d_sorted, sorted.data(), sorted.size() * sizeof(int64_t),
                     cudaMemcpyHostToDevice),
          "IDSelectorBatch");
}

IDSelectorBatch::~IDSelectorBatch() {
    if (d_sorted) {
        SetDevice g(device);
        cudaFree(d_sorted);
    }
}

bool IDSelectorBatch::is_member(int64_t id) const {
    return std::binary_search(sorted.begin(), sorted.end(), id);
}

void IDSelectorBatch::eval(const int64_t* ids, int64_t nblocks, uint32_t* mask,
                           cudaStream_t st) const {
    run_eval(ids, nblocks, mask, SortedSetTest{d_sorted, (int64_t)sorted.size()}, st);
    check(cudaGetLastError(), "IDSelectorBatch");
}

IDSelectorBitmap::IDSelectorBitmap(size_t n, const uint8_t* bitmap, int dev) : device(dev) {
    if (n > 0 && !bitmap)
        throw B200Exception("IDSelectorBitmap: null bitmap");
    bits.assign(bitmap, bitmap + n);
    SetDevice g(device);
    check(cudaMalloc(&d_bits, std::max<size_t>(n, 1)), "IDSelectorBitmap");
    check(cudaMemcpy(d_bits, bits.data(), n, cudaMemcpyHostToDevice), "IDSelectorBitmap");
}

IDSelectorBitmap::~IDSelectorBitmap() {
    if (d_bits) {
        SetDevice g(device);
        cudaFree(d_bits);
    }
}

bool IDSelectorBitmap::is_member(int64_t id) const {
    const uint64_t i = (uint64_t)id;
    if ((i >> 3) >= bits.size())
        return false;
    return (bits[i >> 3] >> (i & 7)) & 1;
}

void IDSelectorBitmap::eval(const int64_t* ids, int64_t nblocks, uint32_t* mask,
                            cudaStream_t st) const {
    run_eval(ids, nblocks, mask, BitmapTest{d_bits, (uint64_t)bits.size()}, st);
    check(cudaGetLastError(), "IDSelectorBitmap");
}

void IDSelectorNot::eval(const int64_t* ids, int64_t nblocks, uint32_t* mask,
                         cudaStream_t st) const {
    sel->eval(ids, nblocks, mask, st);
    run_combine(mask, nullptr, nblocks, 3, st);
    check(cudaGetLastError(), "IDSelectorNot");
}

void IDSelectorBinary::eval(const int64_t* ids, int64_t nblocks, uint32_t* mask,
                            cudaStream_t st) const {
    lhs->eval(ids, nblocks, mask, st);
    uint32_t* tmp = nullptr;
    check(cudaMallocAsync(&tmp, std::max<int64_t>(nblocks, 1) * sizeof(uint32_t), st),
          "IDSelector combine");
    try {
        rhs->eval(ids, nblocks, tmp, st);
    } catch (...) {
        cudaFreeAsync(tmp, st);
        throw;
    }
    run_combine(mask, tmp, nblocks, (int)op, st);
    check(cudaGetLastError(), "IDSelector combine");
    check(cudaFreeAsync(tmp, st), "IDSelector combine");
}

void IDSelectorAll::eval(const int64_t*, int64_t nblocks, uint32_t* mask, cudaStream_t st) const {
    run_combine(mask, nullptr, nblocks, 4, st);
    check(cudaGetLastError(), "IDSelectorAll");
}

} // namespace b200ivf
