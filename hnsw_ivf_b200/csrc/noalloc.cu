// noalloc.cu -- quantize() for callers that cannot allocate device memory (code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc). No counterpart in the
// reference.
//
// Same kernels as the small-batch coarse quantizer (the exact fp32 warp scan of scan.cu over the
// centroid table, then the k-select of select.cu), but their scratch -- the 4-word launch plan and
// the per-(query, row split) partial candidate lists -- lives in ONE file-scope __device__ array of
// kNoAllocScratchBytes (8 MiB) instead of the handle's grow-only cudaMalloc'd buffers. Every GPU
// that loads the library holds that array until the process exits.
//
// The array is shared by every call on a device, whatever the handle, and keeps the last call's
// contents. Nothing in it has to start at zero: the plan words and every partial-list entry the
// k-select reads are written by the call itself before they are read. Two calls that overlap on
// the GPU (different streams, or different handles on their own streams) RACE on it: callers
// serialise them.
#include "index.h"

#include "common.cuh"
#include "kernels.h"

#include <algorithm>
#include <mutex>
#include <string>

namespace b200ivf {

namespace {

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_scratch_mu;
void* g_scratch_ptr[kMaxDevices] = {};

// Device address of the array on the CURRENT device. In host code `g_noalloc_scratch` is only a
// stand-in for the symbol: the address a kernel may use comes from cudaGetSymbolAddress, once per
// device, with that device current.
void* scratch_on_current_device() {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices)
        throw B200Exception("quantize_noalloc: cannot identify the current CUDA device");
    std::lock_guard<std::mutex> lock(g_scratch_mu);
    if (!g_scratch_ptr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p)
            throw B200Exception(std::string("quantize_noalloc: cudaGetSymbolAddress failed: ") +
                                cudaGetErrorString(e));
        g_scratch_ptr[dev] = p;
    }
    return g_scratch_ptr[dev];
}

bool dev_ptr(const void* p) {
    cudaPointerAttributes at;
    if (!p || cudaPointerGetAttributes(&at, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

struct DevScope {
    int prev = 0;
    explicit DevScope(int dev) {
        cudaGetDevice(&prev);
        if (prev != dev && cudaSetDevice(dev) != cudaSuccess)
            throw B200Exception("quantize_noalloc: cudaSetDevice failed");
    }
    ~DevScope() {
        cudaSetDevice(prev);
    }
};

} // namespace

void IndexIVFFlatB200::quantize_noalloc(int64_t n, const float* x, int64_t kc, float* Dq,
                                        int64_t* Iq) const {
    std::lock_guard<std::mutex> lock(mu_);
    DevScope g(device_);
    if (!is_trained)
        throw B200Exception("index not trained");
    if (n == 0)
        return;
    if (n < 0 || n > kNoAllocMaxQueries)
        throw B200Exception("quantize_noalloc: n = " + std::to_string(n) + " exceeds the maximum of " +
                            std::to_string(kNoAllocMaxQueries) + " queries per call");
    if (kc <= 0 || kc > kNoAllocMaxProbe || kc > (int64_t)nlist)
        throw B200Exception("quantize_noalloc: nprobe = " + std::to_string(kc) +
                            " must be in [1, min(nlist, " + std::to_string(kNoAllocMaxProbe) + ")]");
    if (!dev_ptr(x) || !dev_ptr(Dq) || !dev_ptr(Iq))
        throw B200Exception("quantize_noalloc takes device buffers only (staging would allocate)");
    const CoarseView cv = coarse_view_();
    const int nf4 = cv.nf4;
    const int k = (int)kc;
    const int cap = pow2_ceil(k + 32);
    // queries per warp and warps per CTA that fit the shared-memory budget (index.cu's rule)
    const int max_rq = n <= 1 ? 1 : (n <= 4 ? 4 : 8);
    const int pref[9][2] = {{8, 4}, {4, 4}, {8, 2}, {4, 2}, {8, 1}, {4, 1}, {1, 4}, {1, 2}, {1, 1}};
    int rq = 0, wpc = 0;
    for (const auto& p : pref)
        if (p[0] <= max_rq && scan_smem_bytes(p[0], p[1], nf4, cap) <= (size_t)200 * 1024) {
            rq = p[0];
            wpc = p[1];
            break;
        }
    if (!rq)
        throw B200Exception("quantize_noalloc: dimension too large for the scan's shared memory");
    // scratch layout: [0, 256) the plan words, then part[n][S][k] keys
    const size_t part_off = 256;
    const size_t per_split = (size_t)n * k * sizeof(uint64_t);
    const int64_t ngroups = (n + rq - 1) / rq;
    const int64_t nblk = ((int64_t)nlist + kBlockRows - 1) / kBlockRows;
    int64_t S = (148 * 12 * 2 + ngroups - 1) / ngroups;
    S = std::min<int64_t>(S, std::max<int64_t>(1, nblk / 4));
    S = std::min<int64_t>(S, 64);
    S = std::min<int64_t>(S, (int64_t)((kNoAllocScratchBytes - part_off) / per_split));
    if (S < 1)
        throw B200Exception("quantize_noalloc: n * nprobe exceeds the fixed scratch array");
    unsigned char* scratch = static_cast<unsigned char*>(scratch_on_current_device());
    int32_t* grp_start = reinterpret_cast<int32_t*>(scratch);
    uint32_t* item_off = reinterpret_cast<uint32_t*>(scratch) + 2;
    uint64_t* part = reinterpret_cast<uint64_t*>(scratch + part_off);
    cudaStream_t st = stream_();
    auto check = [](cudaError_t e, const char* what) {
        if (e != cudaSuccess)
            throw B200Exception(std::string("quantize_noalloc: ") + what + ": " +
                                cudaGetErrorString(e));
    };
    check(launch_plan_single(grp_start, item_off, (int)n, (uint32_t)(ngroups * S), st), "plan");
    ScanArgs a;
    a.slab = static_cast<const float4*>(cv.slab);
    a.list_blk0 = cv.blk0;
    a.list_len = cv.len;
    a.grp_start = grp_start;
    a.pair_query = nullptr;
    a.item_off = item_off;
    a.nlists = 1;
    a.nsplit = (int)S;
    a.k = k;
    a.cap = cap;
    a.d = d;
    a.nf4 = nf4;
    a.queries = x;
    a.part = part;
    check(launch_ivf_scan(metric_type, rq, wpc, (uint32_t)(ngroups * S), a, st, 4), "scan");
    SelectArgs s;
    s.part = part;
    s.slot_of = nullptr;
    s.n = (int)n;
    s.nseg = 1;
    s.nsplit = (int)S;
    s.kpart = k;
    s.k = k;
    s.metric = metric_type;
    s.ids = nullptr;
    s.out_dis = Dq;
    s.out_ids64 = Iq;
    s.out_ids32 = nullptr;
    check(launch_select(s, st), "select");
}

} // namespace b200ivf
