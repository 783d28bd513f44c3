// index.cu -- host driver of the device-resident IVF-Flat index: owns the slab, plans launches,
// stages host buffers, keeps statistics. Control flow restated from the reference's
// IndexIVF::search (faiss/IndexIVF.cpp:303-397) / search_preassigned (:399-723) /
// IndexIVFFlat::add_core (faiss/IndexIVFFlat.cpp:46-91), with the OpenMP query slicing replaced by
// device-wide batches and the per-query heap by the scan/select kernels.
#include "index.h"

#include "common.cuh"
#include "kernels.h"
#include "selector.h"

#include <algorithm>
#include <climits>
#include <random>
#include <array>
#include <cstdlib>
#include <cstring>
#include <map>

namespace b200ivf {

std::atomic<unsigned long long> g_kernel_launches{0};
std::atomic<unsigned long long> g_tc_launches{0};
static std::atomic<int> g_force_nsplit{0}; // set_option("scan_nsplit"): row splits per list, 0 = auto
std::atomic<int> g_tc_dbg{getenv("B200IVF_TC_DBG") ? atoi(getenv("B200IVF_TC_DBG")) : 0};

#define CUDA_CHECK(expr)                                                                     \
    do {                                                                                     \
        cudaError_t e__ = (expr);                                                            \
        if (e__ != cudaSuccess) {                                                            \
            throw B200Exception(std::string("CUDA error ") + std::to_string((int)e__) + " (" + \
                                cudaGetErrorString(e__) + ") at " + __FILE__ + ":" +         \
                                std::to_string(__LINE__));                                   \
        }                                                                                    \
    } while (0)

#define B200_THROW_IF_NOT_MSG(cond, msg)                                                   \
    do {                                                                                   \
        if (!(cond))                                                                       \
            throw B200Exception(std::string("Error: '") + #cond + "' failed: " + (msg));   \
    } while (0)

void Buffer::reserve(size_t n) {
    if (n <= bytes)
        return;
    release();
    size_t want = std::max(n, bytes + bytes / 2);
    if (pinned_host)
        CUDA_CHECK(cudaMallocHost(&p, want));
    else
        CUDA_CHECK(cudaMalloc(&p, want));
    bytes = want;
}

void Buffer::release() {
    if (p) {
        if (pinned_host)
            cudaFreeHost(p);
        else
            cudaFree(p);
    }
    p = nullptr;
    bytes = 0;
}

namespace {

bool is_device_pointer(const void* p) {
    if (!p)
        return false;
    cudaPointerAttributes at;
    cudaError_t e = cudaPointerGetAttributes(&at, p);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

struct DeviceGuard {
    int prev = 0;
    explicit DeviceGuard(int dev) {
        CUDA_CHECK(cudaGetDevice(&prev));
        if (prev != dev)
            CUDA_CHECK(cudaSetDevice(dev));
    }
    ~DeviceGuard() {
        cudaSetDevice(prev);
    }
};

// shared-memory budget per CTA for the scan kernel (227 KB is the sm_100a limit)
constexpr size_t kScanSmemBudget = 200 * 1024;
// warps the scan should have available: 148 SMs x ~12 resident warps x 2 waves
constexpr int64_t kTargetScanItems = 148 * 12 * 2;
constexpr int64_t kTargetScanItemsSimt = 148 * 32 * 4;
constexpr int kScanRowsPerStep = 4; // RV of the scan kernel

struct ScanCfg {
    int rq, wpc, cap;
    int rv = 4; // 32-row blocks per warp step
};

// max_rq: queries per warp the caller can use (a warp computes all RQ accumulators whether or
// not the list has that many queries, so RQ follows how many queries meet in a list)
ScanCfg choose_scan_cfg(int nf4, int k, int max_rq = 8) {
    const int cap = pow2_ceil(k + 32);
    // preference order: many queries per warp first, then many warps per CTA
    const int pref[9][2] = {{8, 4}, {4, 4}, {8, 2}, {4, 2}, {8, 1}, {4, 1}, {1, 4}, {1, 2}, {1, 1}};
    for (const auto& p : pref)
        if (p[0] <= max_rq && scan_smem_bytes(p[0], p[1], nf4, cap) <= kScanSmemBudget)
            return {p[0], p[1], cap};
    if (scan_smem_bytes(1, 1, nf4, cap) <= 227 * 1024 - 1024)
        return {1, 1, cap};
    throw B200Exception("k too large for the scan kernel's shared-memory candidate buffer");
}

// the CTA-tiled scan (scan_tile.cu) takes the lists probed by more than kTileMinQueries queries
constexpr int kTileMinQueries = 8;
// the rows-major tensor-core kernel takes the whole scan when fewer queries than this meet in a list
constexpr double kTcrMaxAvgQueries = 48.0;
constexpr size_t kSmemPerSM = 233472, kSmemPerCtaMax = 232448, kSmemCtaReserve = 1024;

struct TileCfg {
    bool ok = false; // false: tiled kernel not usable for this (d, k)
    int nst = 0, cap = 0, tq = kTileTQ;
};

TileCfg choose_tile_cfg(int nf4, int k) {
    TileCfg c;
    c.cap = pow2_ceil(k + 32);
    // preference: two CTAs per SM with the deepest ring that fits, then one CTA per SM
    const int pref[6][2] = {{4, 2}, {3, 2}, {2, 2}, {4, 1}, {3, 1}, {2, 1}};
    for (const auto& p : pref) {
        const size_t sm = scan_tile_smem_bytes(nf4, c.cap, p[0]);
        if (sm <= kSmemPerCtaMax && (sm + kSmemCtaReserve) * p[1] <= kSmemPerSM) {
            c.ok = true;
            c.nst = p[0];
            return c;
        }
    }
    return c;
}

} // namespace

struct IndexIVFFlatB200::Impl {
    int nf4 = 0;
    size_t list_lo = 0, list_hi = 0;
    cudaStream_t own_stream = nullptr, stream = nullptr;

    // coarse quantizer: centroid table as a one-list slab
    Buffer cent_slab;
    Buffer cent_meta; // uint32 blk0[1] | int32 len[1]
    std::vector<float> h_centroids;

    Buffer cent_norms; // |c|^2 per centroid row (tensor-core scan, L2)
    // inverted lists
    Buffer slab, ids, list_blk0, list_len, norms;
    std::vector<uint32_t> h_blk0;
    std::vector<uint32_t> h_cap; // 32-row blocks reserved for each list (>= the blocks in use)
    std::vector<int32_t> h_len;
    size_t slab_blocks = 0;
    bool norms_dirty = false;    // rows were added since `norms` was computed
    int64_t nstored = 0;

    // scratch (grow only)
    Buffer item_list; // list number of every tensor-core work item (written by the plan)
    Buffer qthr; // per-query pruning bound shared by the tensor-core work items of one search
    // ONE-MMA filter mode of the rows-major scan (TcScanArgs::one_mma, SelectArgs::err_coef):
    // max |y|^2 over the slab, the queries whose result the checked re-rank could not prove, and
    // the scratch of their second, exact pass
    Buffer ymax2, flag_list, flag_count, host_flag, fb_x, fb_D, fb_I, fb_assign, fb_p32;
    Buffer cent_ymax2;                  // max |c|^2 over the centroid table (one-MMA coarse scan)
    Buffer pair_len, pair_blk0;         // pair-major plan (small batches): the "list" tables
    int coarse_level = 0;               // as one_level, for the coarse quantizer
    int64_t coarse_one_min_nlist = 16384; // "coarse_one_min_nlist"
    bool coarse_exact = false;          // the second, exact pass of a checked coarse scan is running
    // "tc_split" / B200IVF_TC_SPLIT: 1 = one MMA per k-step + checked re-rank (default),
    // 3 = the 3xTF32 split with a fixed re-rank margin (A/B runs)
    int tc_split = 1;
    // "one_margin": keys kept beyond k in that mode, at least (the error window of the k-th score
    // holds one or two rows beyond k on real data; what does not fit is redone exactly)
    int one_margin = 6;
    // queries per probed list below which the one-MMA rows-major kernel is preferred to the
    // queries-major 3xTF32 one (an item takes 64 queries: up to about one item per list)
    double one_max_avg_q = 72.0;
    double one_min_items = 2368.0; // "one_min_items": 8 waves of 2 x 148 CTAs
    // How well the error window fits the data is a property of the data (tight clusters far from
    // the origin crowd it): a handle whose second pass took more than 1/64 of a batch moves up a
    // level and stays there -- 0: one_margin, 1: a margin of 16 at least, 2: the 3xTF32 split
    int one_level = 0;
    unsigned long long fallback_queries = 0; // queries redone on the exact path so far
    Buffer coarse_ids, coarse_dis, part, count, fill, grp_start, item_off, item_off_tile,
            pair_query, slot_of,
            cplan, stage_x, stage_D, stage_I, stage_assign, dev_x, dev_D, dev_I, dev_assign64,
            dev_stats, host_stats;
    // SearchParametersIVF scratch: selector row mask, max_codes limits per (query, probe) and per
    // pair slot, per-query nscan (device + pinned host copy)
    Buffer row_mask, limit_qp, limit_slot, query_ndis, host_query_ndis;
    // range search scratch
    Buffer pair_cnt, pair_off, qcnt, lims, host_lims, range_dis, range_ids;
    // TMA views of the two slabs for the tiled scan (re-encoded whenever a slab moves)
    alignas(64) unsigned char cent_map[kTensorMapBytes] = {};
    alignas(64) unsigned char lists_map[kTensorMapBytes] = {};
    // ... and for the tensor-core scan (column-outer box)
    alignas(64) unsigned char cent_map_tc[kTensorMapBytes] = {};
    alignas(64) unsigned char lists_map_tc[kTensorMapBytes] = {};
    bool use_tc = true; // B200IVF_NO_TC=1 turns the tensor-core scan off (A/B runs)
    // which tensor-core kernel: 0 = by the number of queries that meet in a list (rows-major
    // scan_tcr.cu below kTcrMaxAvgQueries, queries-major scan_tc.cu above and for the coarse
    // quantizer), 1 / 2 = always queries-major / rows-major (B200IVF_TC_MODE, A/B runs)
    int tc_mode = 0;
    // Stage events of the last kEvSets search calls. A call whose buffers are all device pointers
    // returns as soon as its work is enqueued (include/b200ivf_c.h, set_stream): its event set stays
    // `pending` and is read out -- into last_ms, the sums and the stats -- when somebody asks
    // (last_timings, get_stats, ...) or when the ring comes round to it, by which time it is long
    // complete. Calls that stage host buffers synchronize anyway and read their set at once.
    static constexpr int kEvSets = 16;
    struct EvSet {
        cudaEvent_t ev[6] = {};
        bool pending = false;
        bool stages = false; // the per-stage events were recorded (not inside a graph replay)
    };
    EvSet evs[kEvSets];
    int ev_cur = 0;
    cudaEvent_t* ev = evs[0].ev; // the set of the call in progress / the last call
    bool async_outstanding = false;
    double last_ms[5] = {0, 0, 0, 0, 0};
    double sum_ms[5] = {0, 0, 0, 0, 0}; // since reset_stats
    unsigned long long sum_calls = 0;
    IndexIVFStatsB200 stats;

    // stage times of event set i -> last_ms (when it is the current call's), sums, stats
    void collect_set(int i, bool to_last) {
        EvSet& e = evs[i];
        float ms = 0.f;
        double v[5] = {0, 0, 0, 0, 0};
        if (e.stages)
            for (int s = 0; s < 4; ++s)
                if (cudaEventElapsedTime(&ms, e.ev[s], e.ev[s + 1]) == cudaSuccess)
                    v[s] = ms;
        if (cudaEventElapsedTime(&ms, e.ev[0], e.ev[5]) == cudaSuccess)
            v[4] = ms;
        cudaGetLastError();
        for (int s = 0; s < 5; ++s) {
            if (to_last)
                last_ms[s] += v[s];
            sum_ms[s] += v[s];
        }
        stats.quantization_time += v[0];
        stats.search_time += v[4];
        e.pending = false;
    }
    void begin_call() {
        ev_cur = (ev_cur + 1) % kEvSets;
        if (evs[ev_cur].pending) { // 16 calls old: complete long ago
            cudaEventSynchronize(evs[ev_cur].ev[5]);
            collect_set(ev_cur, false);
        }
        ev = evs[ev_cur].ev;
        for (double& v : last_ms)
            v = 0;
        sum_calls++;
    }

    // small-batch searches are launch-bound (a dozen short kernels): the kernel pipeline of one
    // (batch size, nprobe, k, preassigned?) shape is captured into a CUDA graph the second time the
    // shape is seen and replayed with one launch afterwards. A graph holds raw pointers and launch
    // shapes, so it is replayed only while `search_sig()` -- every buffer the pipeline touches plus
    // the sizes the launch plan depends on -- is what it was at capture. B200IVF_NO_GRAPH=1: off.
    struct GraphEntry {
        bool seen = false;
        uint64_t sig = 0;
        unsigned long long launches = 0; // kernels inside the graph (for b200_kernel_launches)
        cudaGraphExec_t exec = nullptr;
    };
    std::map<std::array<int64_t, 4>, GraphEntry> graphs;
    bool use_graph = true;
    bool use_pair_plan = true; // "use_pair_plan": pair-major plan for small batches (A/B runs)
    // average queries per probed list at or below which the SIMT warp kernel takes the whole scan
    // (B200IVF_SIMT_MAXQ, A/B runs)
    double simt_max_avg_q = 2.0;
    // ... and below which the rows-major tensor-core kernel is preferred to the queries-major one
    double tcr_max_avg_q = kTcrMaxAvgQueries;
    // ... and below which that kernel runs in its two-CTAs-per-SM shape
    double tcr_small_factor = 1.0;
    double tcr_small_max_avg_q = 1e9; // (cap for A/B runs; the rule is 3/4 of the small item size)
    unsigned long long graph_replays = 0;
    int coarse_s_override = 0;  // B200IVF_COARSE_S (read once, at construction) / "coarse_s"
    int coarse_tq = 0;          // "coarse_tq": queries per item of the rows-major coarse scan
    bool trace_sync = false;    // B200IVF_TC_TRACE set: no graphs, no asynchronous returns
    void drop_graphs() {
        for (auto& kv : graphs)
            if (kv.second.exec)
                cudaGraphExecDestroy(kv.second.exec);
        graphs.clear();
    }
    uint64_t search_sig() const {
        const Buffer* bufs[] = {&cent_slab, &cent_meta, &cent_norms, &slab, &ids, &list_blk0,
                                &list_len, &norms, &item_list, &qthr, &coarse_ids, &coarse_dis,
                                &part, &count, &fill, &grp_start, &item_off, &item_off_tile,
                                &pair_query, &slot_of, &cplan, &dev_x, &dev_D, &dev_I,
                                &dev_assign64, &dev_stats};
        uint64_t h = 1469598103934665603ull; // FNV-1a over the pointers and sizes
        auto mix = [&h](uint64_t v) {
            for (int i = 0; i < 8; ++i) {
                h ^= (v >> (8 * i)) & 0xff;
                h *= 1099511628211ull;
            }
        };
        for (const Buffer* b : bufs)
            mix((uint64_t)(uintptr_t)b->p);
        mix((uint64_t)nstored);
        mix((uint64_t)slab_blocks);
        mix((uint64_t)list_lo);
        mix((uint64_t)list_hi);
        mix((uint64_t)(uintptr_t)stream);
        return h;
    }

    Impl() {
        const char* e = getenv("B200IVF_NO_TC");
        use_tc = !(e && e[0] == '1');
        if (const char* sp = getenv("B200IVF_TC_SPLIT"))
            tc_split = atoi(sp) == 3 ? 3 : 1;
        const char* e2 = getenv("B200IVF_TC_MODE");
        tc_mode = e2 ? atoi(e2) : 0;
        const char* e3 = getenv("B200IVF_NO_GRAPH");
        use_graph = !(e3 && e3[0] == '1');
        if (const char* e4 = getenv("B200IVF_SIMT_MAXQ"))
            simt_max_avg_q = atof(e4);
        if (const char* e5 = getenv("B200IVF_COARSE_S")) // tuning experiments
            coarse_s_override = atoi(e5);
        trace_sync = getenv("B200IVF_TC_TRACE") != nullptr; // the trace read-out synchronizes
        stage_x.pinned_host = stage_D.pinned_host = stage_I.pinned_host = true;
        stage_assign.pinned_host = host_stats.pinned_host = host_flag.pinned_host = true;
        host_query_ndis.pinned_host = host_lims.pinned_host = true;
    }
};

IndexIVFFlatB200::IndexIVFFlatB200(size_t d_, size_t nlist_, int metric, int device)
        : d((int)d_), nlist(nlist_), metric_type(metric), impl_(nullptr), device_(device) {
    B200_THROW_IF_NOT_MSG(d_ > 0 && d_ < (1u << 20), "invalid dimension");
    B200_THROW_IF_NOT_MSG(nlist_ > 0 && nlist_ < (1u << 30), "invalid nlist");
    B200_THROW_IF_NOT_MSG(metric == kMetricL2 || metric == kMetricInnerProduct,
                          "metric type not supported"); // faiss/IndexIVFFlat.cpp:219
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        throw B200Exception(
                "no CUDA device: the B200 IVF-Flat path has no CPU fallback");
    }
    B200_THROW_IF_NOT_MSG(device >= 0 && device < ndev, "invalid CUDA device ordinal");
    DeviceGuard g(device_);
    impl_ = new Impl;
    impl_->nf4 = (d + 3) / 4;
    impl_->list_hi = nlist;
    CUDA_CHECK(cudaStreamCreateWithFlags(&impl_->own_stream, cudaStreamNonBlocking));
    impl_->stream = impl_->own_stream;
    for (auto& set : impl_->evs)
        for (auto& ev : set.ev)
            CUDA_CHECK(cudaEventCreate(&ev));
    impl_->h_blk0.assign(nlist, 0);
    impl_->h_cap.assign(nlist, 0);
    impl_->h_len.assign(nlist, 0);
    impl_->list_blk0.reserve(sizeof(uint32_t) * nlist);
    impl_->list_len.reserve(sizeof(int32_t) * nlist);
    CUDA_CHECK(cudaMemset(impl_->list_blk0.p, 0, sizeof(uint32_t) * nlist));
    CUDA_CHECK(cudaMemset(impl_->list_len.p, 0, sizeof(int32_t) * nlist));
    impl_->dev_stats.reserve(sizeof(unsigned long long) * 4);
    impl_->host_stats.reserve(sizeof(unsigned long long) * 4);
    CUDA_CHECK(cudaMemset(impl_->dev_stats.p, 0, sizeof(unsigned long long) * 4));
}

IndexIVFFlatB200::~IndexIVFFlatB200() {
    if (!impl_)
        return;
    int prev = 0;
    cudaGetDevice(&prev);
    cudaSetDevice(device_);
    cudaStreamSynchronize(impl_->stream);
    Buffer* bufs[] = {&impl_->item_list, &impl_->qthr, &impl_->cent_norms, &impl_->norms, &impl_->cent_slab, &impl_->cent_meta, &impl_->slab, &impl_->ids,
                      &impl_->list_blk0, &impl_->list_len, &impl_->coarse_ids, &impl_->coarse_dis,
                      &impl_->part, &impl_->count, &impl_->fill, &impl_->grp_start,
                      &impl_->item_off, &impl_->item_off_tile, &impl_->pair_query, &impl_->slot_of, &impl_->cplan,
                      &impl_->stage_x, &impl_->stage_D, &impl_->stage_I, &impl_->stage_assign,
                      &impl_->dev_x, &impl_->dev_D, &impl_->dev_I, &impl_->dev_assign64,
                      &impl_->dev_stats, &impl_->host_stats, &impl_->row_mask, &impl_->limit_qp,
                      &impl_->limit_slot, &impl_->query_ndis, &impl_->host_query_ndis,
                      &impl_->pair_cnt, &impl_->pair_off, &impl_->qcnt, &impl_->lims,
                      &impl_->host_lims, &impl_->range_dis, &impl_->range_ids};
    for (Buffer* b : bufs)
        b->release();
    impl_->drop_graphs();
    for (auto& set : impl_->evs)
        for (auto& ev : set.ev)
            if (ev)
                cudaEventDestroy(ev);
    if (impl_->own_stream)
        cudaStreamDestroy(impl_->own_stream);
    delete impl_;
    cudaSetDevice(prev);
}

cudaStream_t IndexIVFFlatB200::stream_() const {
    return impl_->stream;
}

IndexIVFFlatB200::CoarseView IndexIVFFlatB200::coarse_view_() const {
    const Impl& m = *impl_;
    return {m.cent_slab.p, m.cent_meta.as<uint32_t>(), m.cent_meta.as<int32_t>() + 1, m.nf4};
}

void IndexIVFFlatB200::set_stream(cudaStream_t s) {
    std::lock_guard<std::mutex> lock(mu_);
    DeviceGuard g(device_);
    finish_async_();
    CUDA_CHECK(cudaStreamSynchronize(impl_->stream));
    impl_->stream = s ? s : impl_->own_stream;
}

// ------------------------------------------------------------------------------------------------
// coarse quantizer table
// ------------------------------------------------------------------------------------------------
void IndexIVFFlatB200::set_centroids(const float* c) {
    std::lock_guard<std::mutex> lock(mu_);
    set_centroids_(c);
}

void IndexIVFFlatB200::set_centroids_(const float* c) {
    DeviceGuard g(device_);
    Impl& m = *impl_;
    finish_async_();
    m.drop_graphs();
    B200_THROW_IF_NOT_MSG(c != nullptr, "null centroids");
    const size_t nfl = nlist * (size_t)d;
    m.h_centroids.resize(nfl);
    const float* cd = c;
    if (is_device_pointer(c)) {
        // (on the handle's stream: device inputs are ordered with it, not with the legacy stream)
        CUDA_CHECK(cudaMemcpyAsync(m.h_centroids.data(), c, nfl * sizeof(float),
                                   cudaMemcpyDeviceToHost, m.stream));
        CUDA_CHECK(cudaStreamSynchronize(m.stream));
    } else {
        std::memcpy(m.h_centroids.data(), c, nfl * sizeof(float));
        m.dev_x.reserve(nfl * sizeof(float));
        CUDA_CHECK(cudaMemcpyAsync(m.dev_x.p, c, nfl * sizeof(float), cudaMemcpyHostToDevice,
                                   m.stream));
        cd = m.dev_x.as<float>();
    }
    const size_t nblk = (nlist + kBlockRows - 1) / kBlockRows;
    const size_t slab_bytes = nblk * m.nf4 * kBlockRows * sizeof(float4);
    m.cent_slab.reserve(slab_bytes);
    CUDA_CHECK(make_slab_tensor_map(m.cent_map, m.cent_slab.p, m.nf4, nblk));
    CUDA_CHECK(cudaMemsetAsync(m.cent_slab.p, 0, slab_bytes, m.stream));
    // identity placement: centroid j -> row j
    std::vector<int64_t> rows(nlist);
    for (size_t j = 0; j < nlist; ++j)
        rows[j] = (int64_t)j;
    m.dev_assign64.reserve(nlist * sizeof(int64_t));
    CUDA_CHECK(cudaMemcpyAsync(m.dev_assign64.p, rows.data(), nlist * sizeof(int64_t),
                               cudaMemcpyHostToDevice, m.stream));
    m.dev_I.reserve(nblk * kBlockRows * sizeof(int64_t)); // throw-away id column
    CUDA_CHECK(launch_scatter_rows(cd, nullptr, m.dev_assign64.as<int64_t>(), (int64_t)nlist, d,
                                   m.nf4, m.cent_slab.as<float>(), m.dev_I.as<int64_t>(), 0,
                                   m.stream));
    CUDA_CHECK(make_slab_tensor_map_tc(m.cent_map_tc, m.cent_slab.p, m.nf4, nblk));
    m.cent_ymax2.reserve(sizeof(float));
    CUDA_CHECK(cudaMemsetAsync(m.cent_ymax2.p, 0, sizeof(float), m.stream));
    m.cent_norms.reserve(nblk * kBlockRows * sizeof(float));
    CUDA_CHECK(launch_row_norms(m.cent_slab.as<float4>(), (int64_t)(nblk * kBlockRows), m.nf4,
                                m.cent_norms.as<float>(), m.stream, m.cent_ymax2.as<float>()));
    m.cent_meta.reserve(16);
    const uint32_t blk0 = 0;
    const int32_t len = (int32_t)nlist;
    CUDA_CHECK(cudaMemcpyAsync(m.cent_meta.p, &blk0, 4, cudaMemcpyHostToDevice, m.stream));
    CUDA_CHECK(cudaMemcpyAsync((char*)m.cent_meta.p + 4, &len, 4, cudaMemcpyHostToDevice,
                               m.stream));
    CUDA_CHECK(cudaStreamSynchronize(m.stream));
    is_trained = true;
}

void IndexIVFFlatB200::get_centroids(float* out) const {
    std::lock_guard<std::mutex> lock(mu_);
    B200_THROW_IF_NOT_MSG(is_trained, "index not trained");
    std::memcpy(out, impl_->h_centroids.data(), impl_->h_centroids.size() * sizeof(float));
}

void IndexIVFFlatB200::set_list_range(size_t lo, size_t hi) {
    std::lock_guard<std::mutex> lock(mu_);
    B200_THROW_IF_NOT_MSG(lo <= hi && hi <= nlist, "invalid list range");
    B200_THROW_IF_NOT_MSG(impl_->nstored == 0, "set_list_range must precede add");
    impl_->list_lo = lo;
    impl_->list_hi = hi;
}

void IndexIVFFlatB200::reset() {
    std::lock_guard<std::mutex> lock(mu_);
    reset_();
}

void IndexIVFFlatB200::reset_() {
    DeviceGuard g(device_);
    Impl& m = *impl_;
    finish_async_();
    CUDA_CHECK(cudaStreamSynchronize(m.stream));
    m.slab.release();
    m.ids.release();
    m.norms.release();
    m.slab_blocks = 0;
    m.nstored = 0;
    std::fill(m.h_blk0.begin(), m.h_blk0.end(), 0u);
    std::fill(m.h_cap.begin(), m.h_cap.end(), 0u);
    std::fill(m.h_len.begin(), m.h_len.end(), 0);
    m.norms_dirty = false;
    m.drop_graphs();
    CUDA_CHECK(cudaMemset(m.list_blk0.p, 0, sizeof(uint32_t) * nlist));
    CUDA_CHECK(cudaMemset(m.list_len.p, 0, sizeof(int32_t) * nlist));
    ntotal = 0;
}

size_t IndexIVFFlatB200::get_list_size(size_t l) const {
    std::lock_guard<std::mutex> lock(mu_);
    B200_THROW_IF_NOT_MSG(l < nlist, "invalid list number");
    return (size_t)impl_->h_len[l];
}

void IndexIVFFlatB200::get_list_ids(size_t l, int64_t* out) const {
    std::lock_guard<std::mutex> lock(mu_);
    DeviceGuard g(device_);
    B200_THROW_IF_NOT_MSG(l < nlist, "invalid list number");
    const Impl& m = *impl_;
    const size_t len = (size_t)m.h_len[l];
    if (len == 0)
        return;
    CUDA_CHECK(cudaStreamSynchronize(m.stream));
    CUDA_CHECK(cudaMemcpy(out, m.ids.as<int64_t>() + (size_t)m.h_blk0[l] * kBlockRows,
                          len * sizeof(int64_t), cudaMemcpyDeviceToHost));
}

void IndexIVFFlatB200::get_list_codes(size_t l, float* out) const {
    std::lock_guard<std::mutex> lock(mu_);
    DeviceGuard g(device_);
    B200_THROW_IF_NOT_MSG(l < nlist, "invalid list number");
    Impl& m = *impl_;
    const size_t len = (size_t)m.h_len[l];
    if (len == 0)
        return;
    m.dev_D.reserve(len * d * sizeof(float));
    CUDA_CHECK(launch_gather_rows(m.slab.as<float>(), (int64_t)m.h_blk0[l] * kBlockRows,
                                  (int64_t)len, d, m.nf4, m.dev_D.as<float>(), m.stream));
    CUDA_CHECK(cudaMemcpyAsync(out, m.dev_D.p, len * d * sizeof(float), cudaMemcpyDeviceToHost,
                               m.stream));
    CUDA_CHECK(cudaStreamSynchronize(m.stream));
}

// ------------------------------------------------------------------------------------------------
// search internals (all pointers are device pointers here)
// ------------------------------------------------------------------------------------------------
namespace {

inline int64_t ceil_div(int64_t a, int64_t b) {
    return (a + b - 1) / b;
}

} // namespace

// partial lists keep a few more than k entries (scan_tc_kpart), and the k-select re-scores
// k + kRerankExtra survivors exactly, whenever the tensor-core scan (approximate scores) takes part
constexpr int kRerankExtra = 8;
// ONE-MMA filter: |tf32(x) . trunc_tf32(y) - x . y| <= (2^-11 + 2^-10 + 2^-21) |x||y| from the two
// operand roundings (round-to-nearest for the query, the tensor core's truncation for the rows),
// plus 2^-13 |x||y| for everything else (fp32 accumulation inside the tensor core and in the exact
// chain, the norms of the L2 epilogue): E = kOneMmaErrCoef * |x||y| for the inner product and
// kOneMmaErrCoef * (|x|^2 + |y|^2) >= 2 kOneMmaErrCoef |x||y| for L2 (score = |y|^2 - 2 x.y + |x|^2)
constexpr float kOneMmaErrCoef = 0.00048828125f + 0.0009765625f + 0.0001220703125f + 4.8e-7f;
// searches of at most this many queries go through a captured CUDA graph
constexpr int64_t kGraphMaxBatch = 256;

// K1 as a scan of the centroid table: the kc nearest centroids of each query, best first, ties by
// centroid number as cmp2 orders them (faiss/IndexFlat.cpp:38-69 -> knn_L2sqr/knn_inner_product).
// With many queries the table is scanned on the tensor cores (scan_tc.cu, 3xTF32) and the
// survivors re-scored exactly in the k-select; with few, by the SIMT scan kernels.
void IndexIVFFlatB200::run_coarse_(const float* xd, int64_t n, int kc, int32_t* out_ids32,
                                   int64_t* out_ids64, float* out_dis, cudaStream_t st) const {
    Impl& m = *impl_;
    const int nf4 = m.nf4;
    const int kp_q = scan_tc_kpart(nf4, kc);   // queries-major kernel: d <= 128
    const int kp_r = scan_tcr_kpart(nf4, kc);  // rows-major kernel: d <= 1024
    const bool tc = m.use_tc && n > 32 && (kp_q > 0 || kp_r > 0) && !m.coarse_exact;
    // full 128-query tiles: the queries-major kernel where it exists
    const bool tcr = tc && (kp_q == 0 || (m.tc_mode == 2 && kp_r > 0));
    // Where the rows-major kernel takes the table (nprobe > 24, d > 128): its one-MMA filter form
    // with the checked re-rank (see search_impl_); queries it cannot prove are quantized again on
    // the exact SIMT kernels. Not while a CUDA graph is being captured (the check reads a counter
    // back).
    const int c_margin = m.coarse_level == 0 ? m.one_margin : std::max(m.one_margin, 16);
    const int kp_one = scan_tcr_kpart_one(nf4, kc, c_margin);
    // (long tables only: over the 32768 centroids of a C4 shard it takes the nprobe 64 scan from 4.3
    // to 3.2 ms; over C2's 4096 an item is a handful of tiles and the smaller query tiles of the
    // two-CTA shape cost more than the MMAs they save, 0.9 against 0.72 ms at nprobe 32). Above
    // d = 128 always: there the query operand is what limits the queries per item, and x_hi alone
    // is half of it (C3, d = 960, nprobe 16: 0.87 -> 0.54 ms)
    bool one = tcr && m.tc_split == 1 && m.coarse_level < 2 && kp_one > 0 && n >= 256 &&
            ((int64_t)nlist >= m.coarse_one_min_nlist || nf4 > 32);
    if (one) {
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(st, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone)
            one = false;
        cudaGetLastError();
    }
    const int kp = tc ? (tcr ? (one ? kp_one : kp_r) : kp_q) : kc;
    int tcr_tq = tcr ? scan_tcr_tq(nf4, kp, one) : 0;
    if (one) { // two CTAs per SM where the shape exists
        const int small = scan_tcr_tq_small(nf4, kp, true);
        if (small > 0)
            tcr_tq = small;
        if (m.coarse_tq > 0) // tuning experiments
            tcr_tq = std::max(16, std::min(m.coarse_tq, scan_tcr_tq(nf4, kp, true)) & ~15);
    }
    if (tcr && !one && m.coarse_tq > 0) { // tuning experiments
        const int small = scan_tcr_tq_small(nf4, kp);
        tcr_tq = m.coarse_tq <= small ? std::min(m.coarse_tq, small) : std::min(m.coarse_tq, tcr_tq);
        tcr_tq = std::max(16, tcr_tq & ~15);
    }
    const TileCfg tcfg = choose_tile_cfg(nf4, kp);
    // few queries: a warp carries no more query accumulators than there are queries
    const ScanCfg cfg = choose_scan_cfg(nf4, kp, n <= 1 ? 1 : (n <= 4 ? 4 : 8));
    const bool tiled = !tc && tcfg.ok && n > kTileMinQueries;
    const int64_t tq = tcr ? tcr_tq : tc ? kTcTQ : (tiled ? tcfg.tq : cfg.rq);
    const int64_t ngroups = ceil_div(n, tq);
    const int64_t nblk = ceil_div((int64_t)nlist, kBlockRows);
    // a lone query (or a handful) scanning the whole table is latency-bound: single-block row
    // splits with the deeper unroll (rv = 1) put four times as many warps on the table
    const bool fine = !tc && !tiled && cfg.rq == 1;
    const int64_t rows_step = tc ? 4 : (tiled ? 16 : (fine ? 1 : kScanRowsPerStep)); // blocks per step
    const int64_t target = tc ? 148 * 2 : (tiled ? 148 * 2 * 2 : kTargetScanItems);
    int64_t S = ceil_div(target, ngroups);
    S = std::max<int64_t>(1, std::min<int64_t>(S, std::max<int64_t>(1, nblk / rows_step)));
    S = std::min<int64_t>(S, fine ? 256 : ((!tc && !tiled) ? 128 : 64));
    if (tc) {
        // one CTA per SM: pick the split count whose (waves x tiles per item + per-item prologue,
        // worth ~3 tiles) is smallest, instead of a fixed item target (C2: 79 query tiles x 4
        // splits = 316 items = 2.1 waves of 148 ran as 3 waves; x 3 splits = 2 waves, 10 % faster)
        const int64_t T = ceil_div(nblk, rows_step);
        const int64_t smax = std::max<int64_t>(1, std::min<int64_t>(64, T));
        int64_t best = 1, best_cost = INT64_MAX;
        for (int64_t c = 1; c <= smax; ++c) {
            const int64_t cost = ceil_div(ngroups * c, 148) * (ceil_div(T, c) + 3);
            if (cost < best_cost) {
                best_cost = cost;
                best = c;
            }
        }
        S = best;
    }
    if (m.coarse_s_override > 0) // tuning experiments
        S = std::max<int64_t>(1, std::min<int64_t>(m.coarse_s_override,
                                                   std::max<int64_t>(1, nblk / rows_step)));
    const int pmul = (tc && !tcr) ? 2 : 1;
    m.part.reserve((size_t)n * S * pmul * kp * sizeof(uint64_t));
    m.cplan.reserve(16);
    int32_t* grp_start = m.cplan.as<int32_t>();
    uint32_t* item_off = m.cplan.as<uint32_t>() + 2;
    CUDA_CHECK(launch_plan_single(grp_start, item_off, (int)n, (uint32_t)(ngroups * S), st));
    if (tc) {
        TcScanArgs a;
        a.tmap = m.cent_map_tc;
        a.list_blk0 = m.cent_meta.as<uint32_t>();
        a.list_len = m.cent_meta.as<int32_t>() + 1;
        a.grp_start = grp_start;
        a.pair_query = nullptr;
        a.item_off = item_off;
        a.norms = m.cent_norms.as<float>();
        a.nlists = 1;
        a.nsplit = (int)S;
        a.kpart = kp;
        a.d = d;
        a.nf4 = nf4;
        a.nstages = kTcStages;
        a.queries = xd;
        a.part = m.part.as<uint64_t>();
        m.qthr.reserve((size_t)n * sizeof(uint32_t));
        CUDA_CHECK(launch_fill_u32(m.qthr.as<uint32_t>(), n, ordered_bits(kNeutralScore), st));
        a.qthr = m.qthr.as<uint32_t>();
        a.tq = tcr_tq;
        a.one_mma = one ? 1 : 0;
        if (tcr)
            CUDA_CHECK(launch_ivf_scan_tcr(metric_type, (uint32_t)(ngroups * S), a, st));
        else
            CUDA_CHECK(launch_ivf_scan_tc(metric_type, (uint32_t)(ngroups * S), a, st));
    } else if (tiled) {
        TileScanArgs a;
        a.tmap = m.cent_map;
        a.list_blk0 = m.cent_meta.as<uint32_t>();
        a.list_len = m.cent_meta.as<int32_t>() + 1;
        a.grp_start = grp_start;
        a.pair_query = nullptr;
        a.item_off = item_off;
        a.nlists = 1;
        a.nsplit = (int)S;
        a.k = kp;
        a.cap = tcfg.cap;
        a.d = d;
        a.nf4 = nf4;
        a.nstages = tcfg.nst;
        a.queries = xd;
        a.part = m.part.as<uint64_t>();
        CUDA_CHECK(launch_ivf_scan_tile(metric_type, (uint32_t)(ngroups * S), a, st));
    } else {
        ScanArgs a;
        a.slab = m.cent_slab.as<float4>();
        a.list_blk0 = m.cent_meta.as<uint32_t>();
        a.list_len = m.cent_meta.as<int32_t>() + 1;
        a.grp_start = grp_start;
        a.pair_query = nullptr;
        a.item_off = item_off;
        a.nlists = 1;
        a.nsplit = (int)S;
        a.k = kp;
        a.cap = cfg.cap;
        a.d = d;
        a.nf4 = nf4;
        a.queries = xd;
        a.part = m.part.as<uint64_t>();
        CUDA_CHECK(launch_ivf_scan(metric_type, cfg.rq, cfg.wpc, (uint32_t)(ngroups * S), a, st,
                                   fine ? 1 : 4));
    }
    SelectArgs s;
    s.part = m.part.as<uint64_t>();
    s.slot_of = nullptr;
    s.n = (int)n;
    s.nseg = 1;
    s.nsplit = (int)S * pmul;
    s.kpart = kp;
    s.k = kc;
    s.metric = metric_type;
    s.ids = nullptr;
    s.out_dis = out_dis;
    s.out_ids64 = out_ids64;
    s.out_ids32 = out_ids32;
    if (tc) {
        s.slab = m.cent_slab.as<float4>();
        s.queries = xd;
        s.d = d;
        s.nf4 = nf4;
        s.rerank_extra = std::min(kRerankExtra, kp - kc);
        if (one) {
            s.rerank_extra = std::min(std::max(c_margin, kc / 4), kp - kc);
            s.err_coef = kOneMmaErrCoef;
            s.ymax2 = m.cent_ymax2.as<float>();
            s.qthr = m.qthr.as<uint32_t>();
            m.flag_list.reserve((size_t)n * sizeof(int32_t));
            m.flag_count.reserve(sizeof(int32_t));
            m.host_flag.reserve(sizeof(int32_t));
            CUDA_CHECK(cudaMemsetAsync(m.flag_count.p, 0, sizeof(int32_t), st));
            s.flag_list = m.flag_list.as<int32_t>();
            s.flag_count = m.flag_count.as<int32_t>();
        }
    }
    CUDA_CHECK(launch_select(s, st));
    if (one) {
        CUDA_CHECK(cudaMemcpyAsync(m.host_flag.p, m.flag_count.p, sizeof(int32_t),
                                   cudaMemcpyDeviceToHost, st));
        CUDA_CHECK(cudaStreamSynchronize(st));
        const int64_t nf = *m.host_flag.as<int32_t>();
        if (nf * 64 > n && m.coarse_level < 2)
            m.coarse_level++;
        if (nf > 0) {
            // the flagged queries again, exactly, and their rows back into the tables
            m.fallback_queries += (unsigned long long)nf;
            m.fb_x.reserve((size_t)nf * d * sizeof(float));
            m.fb_D.reserve((size_t)nf * kc * sizeof(float));
            m.fb_I.reserve((size_t)nf * kc * sizeof(int64_t));
            m.fb_p32.reserve((size_t)nf * kc * sizeof(int32_t));
            const int32_t* fl = m.flag_list.as<int32_t>();
            CUDA_CHECK(launch_gather_rows_w(xd, fl, nf, d, m.fb_x.p, st));
            m.coarse_exact = true;
            try {
                run_coarse_(m.fb_x.as<float>(), nf, kc, out_ids32 ? m.fb_p32.as<int32_t>() : nullptr,
                            out_ids64 ? m.fb_I.as<int64_t>() : nullptr, m.fb_D.as<float>(), st);
            } catch (...) {
                m.coarse_exact = false;
                throw;
            }
            m.coarse_exact = false;
            if (out_ids32)
                CUDA_CHECK(launch_scatter_rows_w(m.fb_p32.p, fl, nf, kc, out_ids32, st));
            if (out_ids64)
                CUDA_CHECK(launch_scatter_rows_w(m.fb_I.p, fl, nf, 2 * kc, out_ids64, st));
            if (out_dis)
                CUDA_CHECK(launch_scatter_rows_w(m.fb_D.p, fl, nf, kc, out_dis, st));
        }
    }
}

// device-side search of one batch: coarse ids given -> results written to device D/I
struct BatchPlan {
    ScanCfg cfg;
    int64_t S;
};

static BatchPlan plan_batch(int nf4, int64_t kpart, int64_t nb, int64_t nprobe,
                            int64_t nlists_owned, int64_t nstored, bool simt_only = false,
                            bool filtered = false, double own_frac = 1.0) {
    BatchPlan bp;
    // the (query, probe) pairs expected to fall into this handle's list range
    const int64_t pairs = std::max<int64_t>(1, (int64_t)((double)(nb * nprobe) * own_frac));
    // queries that meet in a probed list, on average (uniform-probe estimate)
    const double avg_q =
            (double)pairs / (double)std::max<int64_t>(1, std::min<int64_t>(nlists_owned, pairs));
    // when the warp kernel is the whole scan, its queries-per-warp follows the regime: a warp
    // pays for RQ accumulators per row whether the list has RQ queries or one
    const int max_rq = !simt_only ? 8 : (avg_q <= 1.5 ? 1 : (avg_q <= 4.0 ? 4 : 8));
    bp.cfg = choose_scan_cfg(nf4, (int)kpart, max_rq);
    const int64_t items_est =
            std::max<int64_t>(1, std::max(ceil_div(pairs, bp.cfg.rq), std::min(pairs, nlists_owned)));
    const int64_t avg_blocks =
            std::max<int64_t>(1, nstored / std::max<int64_t>(1, nlists_owned) / kBlockRows);
    // When the warp kernel is the whole scan (hardly any list shared between queries) it is a pure
    // stream and wants many more warps than lists: 4 waves of 32 warps per SM (C2, nprobe 1: 4096
    // probed lists as 4096 warps ran at 2.8 TB/s, split five ways at 4.4 TB/s). The mixed plans keep
    // the small target: their row splits multiply the tensor-core items too.
    static const int64_t simt_target = getenv("B200IVF_SCAN_TARGET") // tuning experiments
            ? std::max<int64_t>(1, atoll(getenv("B200IVF_SCAN_TARGET")))
            : kTargetScanItemsSimt;
    int64_t S = ceil_div(kTargetScanItems, items_est);
    if (simt_only) // ... as long as a warp keeps at least 12 blocks (384 rows) to stream: below
                   // that the per-warp costs (query staging, final sort) outweigh the extra warps
        S = std::max(S, std::min(ceil_div(simt_target, items_est), avg_blocks / 12));
    // few pairs (small batches): split down to single blocks so that the whole GPU streams
    if (simt_only && !filtered && bp.cfg.rq == 1 && S > avg_blocks / kScanRowsPerStep) {
        bp.cfg.rv = 1;
        S = std::max<int64_t>(1, std::min(S, avg_blocks));
    } else {
        S = std::max<int64_t>(1, std::min(S, std::max<int64_t>(1, avg_blocks / kScanRowsPerStep)));
    }
    bp.S = std::min<int64_t>(S, 32);
    if (const int f = g_force_nsplit.load(std::memory_order_relaxed)) // tuning experiments
        bp.S = std::max<int64_t>(1, std::min<int64_t>(f, std::max<int64_t>(1, avg_blocks / 4)));
    return bp;
}

void IndexIVFFlatB200::search(int64_t n, const float* x, int64_t k, float* D, int64_t* I,
                              const SearchParametersIVF* params,
                              QueryLatencyStats* per_query_stats) const {
    search_impl_(n, x, k, nullptr, D, I, false, params, per_query_stats, nullptr);
}

void IndexIVFFlatB200::search_preassigned(int64_t n, const float* x, int64_t k,
                                          const int64_t* assign, const float* /*centroid_dis*/,
                                          float* D, int64_t* I, bool store_pairs,
                                          const SearchParametersIVF* params,
                                          QueryLatencyStats* per_query_stats) const {
    B200_THROW_IF_NOT_MSG(assign != nullptr, "null assign table");
    search_impl_(n, x, k, assign, D, I, store_pairs, params, per_query_stats, nullptr);
}

// search_preassigned with the probe table as the device holds it: [n][nprobe] int32 on the device,
// -1 = no probe (the table the sharded index all-gathers). Everything must be device memory.
void IndexIVFFlatB200::search_probed_i32(int64_t n, const float* x, int64_t k,
                                         const int32_t* probes, float* D, int64_t* I) const {
    B200_THROW_IF_NOT_MSG(probes != nullptr, "null probe table");
    search_impl_(n, x, k, nullptr, D, I, false, nullptr, nullptr, probes);
}

void IndexIVFFlatB200::search_impl_(int64_t n, const float* x, int64_t k, const int64_t* assign,
                                    float* D, int64_t* I, bool store_pairs,
                                    const SearchParametersIVF* params,
                                    QueryLatencyStats* per_query, const int32_t* probes32) const {
    std::lock_guard<std::mutex> lock(mu_);
    DeviceGuard g(device_);
    Impl& m = *impl_;
    // argument checks of IndexIVF::search / search_preassigned (faiss/IndexIVF.cpp:310-318,410-450)
    B200_THROW_IF_NOT_MSG(k > 0, "k > 0");
    B200_THROW_IF_NOT_MSG(is_trained, "index not trained");
    const int64_t np = (int64_t)std::min(nlist, params ? params->nprobe : nprobe);
    B200_THROW_IF_NOT_MSG(np > 0, "nprobe > 0");
    B200_THROW_IF_NOT_MSG(k <= 2048 && np <= 2048, "k and nprobe are limited to 2048");
    const IDSelector* sel = params ? params->sel : nullptr;
    const size_t mc = params ? params->max_codes : max_codes;
    B200_THROW_IF_NOT_MSG(!(sel && store_pairs), "selector and store_pairs cannot be combined");
    // selector / max_codes: a row mask and per-pair row limits in the scan
    const bool filtered = sel != nullptr || mc > 0;
    // an invalid key met by an earlier asynchronous call surfaces here (its counter has landed in
    // the pinned stats copy by now, or will have by the next call)
    if (m.async_outstanding && m.host_stats.as<unsigned long long>()[3] != 0)
        finish_async_();
    m.begin_call();
    if (per_query) // IndexIVF::search_stats zeroes the table first (faiss/IndexIVF.cpp:737-739)
        std::fill(per_query, per_query + n, QueryLatencyStats());
    if (n == 0)
        return;
    B200_THROW_IF_NOT_MSG(x && D && I, "null buffer");
    ensure_norms_(m.stream);

    const bool x_dev = is_device_pointer(x), D_dev = is_device_pointer(D),
               I_dev = is_device_pointer(I);
    const bool a_dev = assign && is_device_pointer(assign);
    B200_THROW_IF_NOT_MSG(!probes32 || (x_dev && D_dev && I_dev),
                          "search_probed_i32 takes device buffers only");
    const int64_t nlists_owned = (int64_t)(m.list_hi - m.list_lo);
    // every buffer on the device: the call is ordered on the handle's stream and returns once its
    // work is enqueued (stage times, statistics and the invalid-key check are read out later)
    const bool async_ok = x_dev && D_dev && I_dev && (!assign || a_dev) && !per_query &&
            !m.trace_sync;

    // tensor-core scan for the lists probed by many queries (d <= 128, moderate k)
    const int kp_q = scan_tc_kpart(m.nf4, (int)std::min<int64_t>(k, 1 << 20));  // d <= 128
    const int kp_r = scan_tcr_kpart(m.nf4, (int)std::min<int64_t>(k, 1 << 20)); // d <= 1024
    // batch size: bound the partial-candidate scratch
    int64_t nb = std::min<int64_t>(n, 32768);
    // queries per probed list, on average: few -> rows-major kernel (its MMA time follows the
    // query count and it streams a list at any count), many -> queries-major (128-query tiles)
    // (a list-range shard sees only its share of the probes: pairs * owned / nlist)
    const double own_frac = (double)nlists_owned / (double)nlist;
    const double pairs_own = std::max(1.0, (double)std::min<int64_t>(n, 32768) * np * own_frac);
    const double avg_q = pairs_own / std::max(1.0, std::min((double)nlists_owned, pairs_own));
    // hardly any list shared between queries (small batches, huge nlist): nothing to multiply, the
    // scan is a pure stream -> the SIMT warp kernel alone (exact fp32, no re-rank needed)
    // A filtered search (selector / max_codes) keeps the list-major reuse: the rows-major
    // tensor-core kernel has the mask and row-limit arms (the queries-major one does not).
    // (not const: the exact second pass of the one-MMA mode switches them, see below)
    bool simt_only = (filtered && kp_r == 0) || !m.use_tc || avg_q <= m.simt_max_avg_q ||
            (kp_q == 0 && kp_r == 0);
    bool tc = !simt_only;
    const int one_margin = m.one_level == 0 ? m.one_margin : std::max(m.one_margin, 16);
    const int kp_one = scan_tcr_kpart_one(m.nf4, (int)std::min<int64_t>(k, 1 << 20), one_margin);
    // ... and it needs enough work items to keep two CTAs per SM busy for several waves: the first
    // wave runs with cold pruning bounds and with every prologue at once, the last one half empty
    // (C2 over 8 / 4 GPUs: 512 / 1024 lists per rank = 3.3 / 6.6 waves took 0.43 / 0.59 ms against
    // 0.33 / 0.57 ms for the queries-major kernel, and the counter read-back delayed the result
    // gather by 0.08 ms; over 2 GPUs the one-MMA kernel wins, 0.875 ms)
    // (a list probed by q queries makes ceil(q / 64) items: about q / 64 + 1/2 on average)
    const double est_items =
            std::min((double)nlists_owned, pairs_own) * std::max(1.0, avg_q / 64.0 + 0.5);
    const bool one_ok = m.tc_split == 1 && m.one_level < 2 && kp_one > 0 &&
            (est_items >= m.one_min_items || filtered || kp_q == 0);
    // Against the queries-major 3xTF32 kernel the one-MMA kernel wins by more the longer the
    // lists are (its per-item costs amortise over more tiles): C1, 977 rows per list, loses above
    // ~72 queries per list (0.46 against 0.35 ms at 80); C2, 2441 rows per list, wins at every
    // count measured (82 / 123 / 163 / 244 per list: 1.67 / 2.15 / 2.62 / 3.53 ms against 1.99 /
    // 2.36 / 2.96 / 3.99)
    const double rows_per_list = (double)m.nstored / (double)std::max<int64_t>(1, nlists_owned);
    const double rl = std::max(1.0, rows_per_list / 1024.0);
    const double one_cap = m.one_max_avg_q * rl * rl * rl;
    bool tcr = tc && (filtered || kp_q == 0 ||
                      ((kp_r > 0 || one_ok) &&
                       (m.tc_mode == 2 ||
                        (m.tc_mode == 0 &&
                         avg_q < (one_ok ? std::max(one_cap, m.tcr_max_avg_q)
                                         : m.tcr_max_avg_q)))));
    if (tcr && kp_r == 0 && !one_ok) // (filtered, k beyond what the rows-major kernel keeps)
        tcr = false, tc = false, simt_only = true;
    // ONE-MMA filter mode (rows-major kernel only): the tensor core computes tf32(x) . tf32(y)
    // alone -- a third of the MMAs, no operand split -- and the k-select proves, query by query,
    // that nothing outside the exactly re-scored set could have reached the top k
    // (SelectArgs::err_coef); the few queries it cannot prove are searched again on the exact
    // warp kernel before the call returns
    bool one = tc && tcr && one_ok;
    int64_t kp = tc ? (tcr ? (one ? kp_one : kp_r) : kp_q) : k; // entries per partial list
    // few queries per list: the two-CTAs-per-SM shape of the rows-major kernel (<= 32 queries per
    // item; a list probed by more is cut into several items)
    // (worth it while most lists fit one or two items: up to that shape's item size on average;
    // measured on C2: 30 queries per list 1.49 ms against 1.66, 41 per list 1.85 against 1.78)
    int tq_small = tcr ? scan_tcr_tq_small(m.nf4, (int)kp, one) : 0;
    if (avg_q > std::min(m.tcr_small_max_avg_q, m.tcr_small_factor * tq_small))
        tq_small = 0;
    const int tcr_tq = tcr ? (tq_small ? tq_small : scan_tcr_tq(m.nf4, (int)kp, one)) : 0;
    BatchPlan bp = plan_batch(m.nf4, kp, nb, np, nlists_owned, m.nstored, simt_only, filtered, own_frac);
    int pmul = (tc && !tcr) ? 2 : 1; // partial lists per (pair, split)
    const int tc_tq = tcr ? tcr_tq : kTcTQ;
    const int tc_min = tcr ? 0 : kTileMinQueries; // lists with more queries than this: tensor cores
    while (nb > 1 && (size_t)nb * np * bp.S * pmul * kp * sizeof(uint64_t) > ((size_t)2 << 30)) {
        nb = (nb + 1) / 2;
        bp = plan_batch(m.nf4, kp, nb, np, nlists_owned, m.nstored, simt_only, filtered, own_frac);
    }
    cudaStream_t st = m.stream;
    unsigned long long* dstats = m.dev_stats.as<unsigned long long>();
    const uint32_t* row_mask = nullptr;
    if (sel && m.slab_blocks > 0) { // one pass over the id column per search call
        m.row_mask.reserve(m.slab_blocks * sizeof(uint32_t));
        sel->eval(m.ids.as<int64_t>(), (int64_t)m.slab_blocks, m.row_mask.as<uint32_t>(), st);
        row_mask = m.row_mask.as<uint32_t>();
    }

    // the kernel pipeline of one batch; every pointer is a device pointer. `timed`: stage events
    // are recorded (not while the pipeline is being captured into a graph)
    bool count_stats = true; // (off for the second pass: its queries were counted by the first)
    bool want_pq = per_query != nullptr; // per-query nscan (search_stats): first pass only
    const int32_t* last_probe_tab = nullptr; // the probe table the last enqueue() searched
    auto enqueue = [&](int64_t ni, const float* xd, float* Dd, int64_t* Id, const int64_t* ad,
                       bool timed, const int32_t* p32 = nullptr) {
        const int64_t pairs = ni * np;
        bp = plan_batch(m.nf4, kp, ni, np, nlists_owned, m.nstored, simt_only, filtered, own_frac);
        // stage 1: probe table
        m.coarse_ids.reserve((size_t)pairs * sizeof(int32_t));
        m.coarse_dis.reserve((size_t)pairs * sizeof(float));
        const int32_t* probe_tab = m.coarse_ids.as<int32_t>();
        last_probe_tab = probe_tab;
        if (p32) {
            // a device table in the plan's own format: read in place unless the max_codes walk is
            // going to edit it (keys >= nlist are counted as invalid by the plan)
            if (mc > 0)
                CUDA_CHECK(cudaMemcpyAsync(m.coarse_ids.p, p32, (size_t)pairs * sizeof(int32_t),
                                           cudaMemcpyDeviceToDevice, st));
            else
                probe_tab = last_probe_tab = p32;
        } else if (ad) {
            CUDA_CHECK(launch_i64_to_i32(ad, m.coarse_ids.as<int32_t>(), pairs, 0, (int32_t)nlist,
                                         dstats + 3, st));
        } else {
            run_coarse_(xd, ni, (int)np, m.coarse_ids.as<int32_t>(), nullptr,
                        m.coarse_dis.as<float>(), st);
        }
        if (timed)
            CUDA_CHECK(cudaEventRecord(m.ev[1], st));
        if (mc > 0 || want_pq) {
            if (mc > 0) {
                m.limit_qp.reserve((size_t)pairs * sizeof(int32_t));
                m.limit_slot.reserve((size_t)pairs * sizeof(int32_t));
            }
            if (want_pq)
                m.query_ndis.reserve((size_t)ni * sizeof(unsigned long long));
            CUDA_CHECK(launch_probe_limits(
                    m.coarse_ids.as<int32_t>(), (int)ni, (int)np, (int)nlist,
                    m.list_len.as<int32_t>(),
                    mc > 0 ? (long long)std::min<size_t>(mc, (size_t)LLONG_MAX) : LLONG_MAX,
                    mc > 0 ? m.limit_qp.as<int32_t>() : nullptr,
                    want_pq ? m.query_ndis.as<unsigned long long>() : nullptr, st));
        }
        // stage 2: invert the probe table, scan the lists
        // small batches: pair-major plan (every pair its own work-item group: nothing to share
        // between a handful of queries, and no pass over the nlist per-list counters)
        const bool pair_major = simt_only && !filtered && bp.cfg.rq == 1 && pairs <= 4096 &&
                m.use_pair_plan;
        const size_t nl_plan = pair_major ? (size_t)pairs : nlist;
        m.count.reserve(nlist * sizeof(int32_t));
        m.fill.reserve(nlist * sizeof(int32_t));
        m.grp_start.reserve((std::max<size_t>(nlist, 4096) + 1) * sizeof(int32_t));
        m.item_off.reserve((std::max<size_t>(nlist, 4096) + 1) * sizeof(uint32_t));
        m.item_off_tile.reserve((nlist + 1) * sizeof(uint32_t));
        if (pair_major) {
            m.pair_len.reserve(4096 * sizeof(int32_t));
            m.pair_blk0.reserve(4096 * sizeof(uint32_t));
        }
        (void)nl_plan;
        TileCfg tcfg = choose_tile_cfg(m.nf4, (int)kp);
        if (tc || filtered || pair_major)
            tcfg.ok = false; // the tensor-core kernel takes the many-query lists instead (and a
                             // filtered search stays in the warp kernel, which knows the mask)
        m.pair_query.reserve((size_t)pairs * sizeof(int32_t));
        m.slot_of.reserve((size_t)pairs * sizeof(int32_t));
        m.part.reserve((size_t)pairs * bp.S * pmul * kp * sizeof(uint64_t));
        {
            // debugging aid: B200IVF_POISON=1 fills the partial-candidate scratch with keys whose
            // row is far outside the slab (any entry a scan kernel fails to write then faults in
            // the k-select), =2 with sentinels
            static const int poison = getenv("B200IVF_POISON") ? atoi(getenv("B200IVF_POISON")) : 0;
            if (poison)
                CUDA_CHECK(cudaMemsetAsync(m.part.p, poison == 1 ? 0x7f : 0xff,
                                           (size_t)pairs * bp.S * pmul * kp * sizeof(uint64_t), st));
        }
        PlanArgs pa;
        pa.coarse_ids = probe_tab;
        pa.n = (int)ni;
        pa.nprobe = (int)np;
        // the plan and the scans cover this handle's list range only (per-list arrays in local
        // numbering): a shard of a 262144-list index does not walk the other shards' lists
        pa.nlists = (int)nlists_owned;
        pa.list_lo = (int)m.list_lo;
        pa.nlist_total = (int)nlist;
        pa.list_len = m.list_len.as<int32_t>() + m.list_lo;
        pa.count = m.count.as<int32_t>();
        pa.fill = m.fill.as<int32_t>();
        pa.grp_start = m.grp_start.as<int32_t>();
        pa.item_off = m.item_off.as<uint32_t>();
        pa.pair_query = m.pair_query.as<int32_t>();
        pa.slot_of = m.slot_of.as<int32_t>();
        pa.rq = bp.cfg.rq;
        pa.nsplit = (int)bp.S;
        pa.tile_tq = tc ? tc_tq : (tcfg.ok ? tcfg.tq : 0);
        pa.tile_min = tc ? tc_min : kTileMinQueries;
        pa.item_off_tile = m.item_off_tile.as<uint32_t>();
        const int64_t max_tc_items =
                (pairs / tc_tq + std::min<int64_t>(pairs / (tc_min + 1), nlists_owned)) * bp.S;
        if (tc) {
            m.item_list.reserve((size_t)(max_tc_items + 1) * sizeof(int32_t));
            pa.item_list_tile = m.item_list.as<int32_t>();
        }
        pa.stats = count_stats ? dstats : nullptr;
        if (pair_major) {
            pa.vlen = m.pair_len.as<int32_t>();
            pa.vblk0 = m.pair_blk0.as<uint32_t>();
            pa.list_blk0 = m.list_blk0.as<uint32_t>() + m.list_lo;
        }
        if (mc > 0) {
            pa.limit_qp = m.limit_qp.as<int32_t>();
            pa.limit_slot = m.limit_slot.as<int32_t>();
        }
        CUDA_CHECK(launch_plan(pa, st));
        if (timed)
            CUDA_CHECK(cudaEventRecord(m.ev[2], st));
        ScanArgs sa;
        sa.slab = m.slab.as<float4>();
        sa.list_blk0 = m.list_blk0.as<uint32_t>() + m.list_lo;
        sa.list_len = m.list_len.as<int32_t>() + m.list_lo;
        sa.grp_start = pa.grp_start;
        sa.pair_query = pa.pair_query;
        sa.item_off = pa.item_off;
        sa.nlists = (int)nlists_owned;
        if (pair_major) { // the pairs are the lists
            sa.list_blk0 = m.pair_blk0.as<uint32_t>();
            sa.list_len = m.pair_len.as<int32_t>();
            sa.nlists = (int)pairs;
        }
        sa.nsplit = (int)bp.S;
        sa.k = (int)kp;
        sa.cap = bp.cfg.cap;
        sa.d = d;
        sa.nf4 = m.nf4;
        sa.queries = xd;
        sa.part = m.part.as<uint64_t>();
        sa.part_mul = pmul;
        sa.row_mask = row_mask;
        sa.pair_limit = mc > 0 ? m.limit_slot.as<int32_t>() : nullptr;
        const int64_t max_items = pair_major
                ? pairs * bp.S
                : (ceil_div(pairs, bp.cfg.rq) + std::min<int64_t>(pairs, nlists_owned)) * bp.S;
        if (!tcr) // (every probed list goes to the rows-major tensor-core kernel otherwise)
            CUDA_CHECK(launch_ivf_scan(metric_type, bp.cfg.rq, bp.cfg.wpc, (uint32_t)max_items, sa,
                                       st, bp.cfg.rv));
        if (tc) {
            // lists probed by more than kTileMinQueries queries of this batch: tensor cores
            TcScanArgs ta;
            ta.tmap = m.lists_map_tc;
            ta.list_blk0 = sa.list_blk0;
            ta.list_len = sa.list_len;
            ta.grp_start = pa.grp_start;
            ta.pair_query = pa.pair_query;
            ta.item_off = pa.item_off_tile;
            ta.norms = m.norms.as<float>();
            ta.nlists = (int)nlists_owned;
            ta.nsplit = (int)bp.S;
            ta.kpart = (int)kp;
            ta.d = d;
            ta.nf4 = m.nf4;
            ta.nstages = kTcStages;
            ta.queries = xd;
            ta.part = sa.part;
            m.qthr.reserve((size_t)ni * sizeof(uint32_t));
            CUDA_CHECK(launch_fill_u32(m.qthr.as<uint32_t>(), ni, ordered_bits(kNeutralScore), st));
            ta.qthr = m.qthr.as<uint32_t>();
            ta.item_list = m.item_list.as<int32_t>();
            ta.tq = tcr_tq;
            if (tcr) {
                ta.row_mask = row_mask;
                ta.pair_limit = mc > 0 ? m.limit_slot.as<int32_t>() : nullptr;
                ta.one_mma = one ? 1 : 0;
            }
            if (tcr)
                CUDA_CHECK(launch_ivf_scan_tcr(metric_type, (uint32_t)max_tc_items, ta, st));
            else
                CUDA_CHECK(launch_ivf_scan_tc(metric_type, (uint32_t)max_tc_items, ta, st));
        } else if (tcfg.ok && ni > kTileMinQueries) {
            // lists probed by more than kTileMinQueries queries of this batch
            TileScanArgs ta;
            ta.tmap = m.lists_map;
            ta.list_blk0 = sa.list_blk0;
            ta.list_len = sa.list_len;
            ta.grp_start = pa.grp_start;
            ta.pair_query = pa.pair_query;
            ta.item_off = pa.item_off_tile;
            ta.nlists = (int)nlists_owned;
            ta.nsplit = (int)bp.S;
            ta.k = (int)kp;
            ta.cap = tcfg.cap;
            ta.d = d;
            ta.nf4 = m.nf4;
            ta.nstages = tcfg.nst;
            ta.queries = xd;
            ta.part = sa.part;
            const int64_t max_tiles =
                    (pairs / tcfg.tq + std::min<int64_t>(pairs / (kTileMinQueries + 1), nlists_owned)) *
                    bp.S;
            CUDA_CHECK(launch_ivf_scan_tile(metric_type, (uint32_t)max_tiles, ta, st));
        }
        if (timed)
            CUDA_CHECK(cudaEventRecord(m.ev[3], st));
        // stage 3: per-query k-select, row -> id
        SelectArgs se;
        se.part = m.part.as<uint64_t>();
        se.slot_of = pa.slot_of;
        se.n = (int)ni;
        se.nseg = (int)np;
        se.nsplit = (int)bp.S * pmul;
        se.kpart = (int)kp;
        se.k = (int)k;
        se.metric = metric_type;
        se.ids = store_pairs ? nullptr : m.ids.as<int64_t>(); // rows first, pairs below
        if (tc) {
            se.slab = m.slab.as<float4>();
            se.queries = xd;
            se.d = d;
            se.nf4 = m.nf4;
            se.rerank_extra = (int)std::min<int64_t>(kRerankExtra, kp - k);
            if (one) {
                // checked re-rank: everything inside the error window of the k-th score is
                // re-scored (the window holds a handful of rows: a margin of max(16, k / 4))
                se.rerank_extra =
                        (int)std::min<int64_t>(std::max<int64_t>(one_margin, k / 4), kp - k);
                se.err_coef = kOneMmaErrCoef;
                se.ymax2 = m.ymax2.as<float>();
                se.qthr = m.qthr.as<uint32_t>();
                se.flag_list = m.flag_list.as<int32_t>();
                se.flag_count = m.flag_count.as<int32_t>();
            }
        }
        se.out_dis = Dd;
        se.out_ids64 = Id;
        se.out_ids32 = nullptr;
        CUDA_CHECK(launch_select(se, st));
        if (store_pairs) // label = lo_build(list_no, offset), faiss/IndexIVFFlat.cpp:172
            CUDA_CHECK(launch_rows_to_pairs(Id, ni * k, m.list_blk0.as<uint32_t>(), (int)nlist, st));
    };
    // per-query figures of IndexIVF::search_stats (faiss/IndexIVF.cpp:727-862): the coarse time is
    // amortised over the batch as the reference does (:767-771); a batched kernel has no per-query
    // scan time, so the batch's plan + scan + select time is shared out by each query's nscan
    auto fill_per_query = [&](int64_t i0, int64_t ni, const double* ms_before) {
        if (!per_query)
            return;
        const unsigned long long* qn = m.host_query_ndis.as<unsigned long long>();
        double tot = 0;
        for (int64_t i = 0; i < ni; ++i)
            tot += (double)qn[i];
        const double coarse_us = (m.last_ms[0] - ms_before[0]) * 1000.0 / (double)ni;
        const double scan_us = ((m.last_ms[1] - ms_before[1]) + (m.last_ms[2] - ms_before[2]) +
                                (m.last_ms[3] - ms_before[3])) * 1000.0;
        for (int64_t i = 0; i < ni; ++i) {
            QueryLatencyStats& o = per_query[i0 + i];
            o.quantization_us = coarse_us;
            o.list_scan_us = tot > 0 ? scan_us * (double)qn[i] / tot : scan_us / (double)ni;
            o.total_us = o.quantization_us + o.list_scan_us;
        }
    };

    // ---- small batches: one CUDA-graph launch per search (see Impl::GraphEntry) ----
    const bool legacy_stream = st == nullptr || st == cudaStreamLegacy || st == cudaStreamPerThread;
    if (m.use_graph && n <= kGraphMaxBatch && !legacy_stream && !filtered && !store_pairs &&
        !per_query && !m.trace_sync && !probes32 && !one) {
        const int64_t pairs = n * np;
        m.dev_x.reserve((size_t)n * d * sizeof(float));
        m.dev_D.reserve((size_t)n * k * sizeof(float));
        m.dev_I.reserve((size_t)n * k * sizeof(int64_t));
        if (assign)
            m.dev_assign64.reserve((size_t)pairs * sizeof(int64_t));
        // Host buffers go through the handle's pinned staging areas: a copy between pageable
        // memory and the device makes the driver stage it and (device -> host) block the calling
        // thread per copy; from pinned memory every copy is a plain enqueue and the call blocks
        // once, in the stream synchronize at the end. These calls are latency-bound: it matters.
        const size_t xbytes = (size_t)n * d * sizeof(float), abytes = (size_t)pairs * sizeof(int64_t);
        const size_t Dbytes = (size_t)n * k * sizeof(float), Ibytes = (size_t)n * k * sizeof(int64_t);
        const void* xsrc = x;
        if (!x_dev) {
            m.stage_x.reserve(xbytes);
            std::memcpy(m.stage_x.p, x, xbytes);
            xsrc = m.stage_x.p;
        }
        const void* asrc = assign;
        if (assign && !a_dev) {
            m.stage_assign.reserve(abytes);
            std::memcpy(m.stage_assign.p, assign, abytes);
            asrc = m.stage_assign.p;
        }
        if (!D_dev)
            m.stage_D.reserve(Dbytes);
        if (!I_dev)
            m.stage_I.reserve(Ibytes);
        CUDA_CHECK(cudaEventRecord(m.ev[0], st));
        CUDA_CHECK(cudaMemcpyAsync(m.dev_x.p, xsrc, xbytes, cudaMemcpyDefault, st));
        if (assign)
            CUDA_CHECK(cudaMemcpyAsync(m.dev_assign64.p, asrc, abytes, cudaMemcpyDefault, st));
        const int64_t* ad = assign ? m.dev_assign64.as<int64_t>() : nullptr;
        Impl::GraphEntry& ge = m.graphs[{n, np, k, assign ? 1 : 0}];
        uint64_t sig = m.search_sig();
        if (ge.exec && ge.sig != sig) {
            cudaGraphExecDestroy(ge.exec);
            ge.exec = nullptr;
        }
        bool ran = false;
        if (!ge.exec && ge.seen) {
            // second sighting of the shape: capture (all scratch was sized by the first, eager run)
            const unsigned long long l0 = g_kernel_launches.load();
            cudaGraph_t graph = nullptr;
            CUDA_CHECK(cudaStreamBeginCapture(st, cudaStreamCaptureModeRelaxed));
            try {
                enqueue(n, m.dev_x.as<float>(), m.dev_D.as<float>(), m.dev_I.as<int64_t>(), ad, false);
            } catch (...) {
                cudaStreamEndCapture(st, &graph);
                if (graph)
                    cudaGraphDestroy(graph);
                cudaGetLastError();
                throw;
            }
            CUDA_CHECK(cudaStreamEndCapture(st, &graph));
            ge.launches = g_kernel_launches.load() - l0;
            g_kernel_launches.fetch_sub(ge.launches); // nothing ran yet
            const uint64_t sig2 = m.search_sig();
            if (sig2 == sig) {
                cudaError_t e = cudaGraphInstantiate(&ge.exec, graph, 0);
                cudaGraphDestroy(graph);
                CUDA_CHECK(e);
                ge.sig = sig;
            } else { // a scratch buffer moved while capturing: the graph is stale already
                cudaGraphDestroy(graph);
                sig = sig2;
            }
        }
        if (ge.exec) {
            CUDA_CHECK(cudaGraphLaunch(ge.exec, st));
            g_kernel_launches.fetch_add(ge.launches);
            m.graph_replays++;
            ran = true;
        }
        if (!ran) {
            ge.seen = true;
            enqueue(n, m.dev_x.as<float>(), m.dev_D.as<float>(), m.dev_I.as<int64_t>(), ad, true);
        }
        CUDA_CHECK(cudaEventRecord(m.ev[4], st));
        CUDA_CHECK(cudaMemcpyAsync(D_dev ? (void*)D : m.stage_D.p, m.dev_D.p, Dbytes,
                                   cudaMemcpyDefault, st));
        CUDA_CHECK(cudaMemcpyAsync(I_dev ? (void*)I : m.stage_I.p, m.dev_I.p, Ibytes,
                                   cudaMemcpyDefault, st));
        CUDA_CHECK(cudaEventRecord(m.ev[5], st));
        CUDA_CHECK(cudaMemcpyAsync(m.host_stats.p, dstats, 4 * sizeof(unsigned long long),
                                   cudaMemcpyDeviceToHost, st));
        m.evs[m.ev_cur].stages = !ran;
        if (async_ok) {
            m.evs[m.ev_cur].pending = true;
            m.async_outstanding = true;
            return;
        }
        CUDA_CHECK(cudaStreamSynchronize(st));
        if (!D_dev)
            std::memcpy(D, m.stage_D.p, Dbytes);
        if (!I_dev)
            std::memcpy(I, m.stage_I.p, Ibytes);
        m.collect_set(m.ev_cur, true);
    } else
    for (int64_t i0 = 0; i0 < n; i0 += nb) {
        const int64_t ni = std::min(nb, n - i0);
        const int64_t pairs = ni * np;
        CUDA_CHECK(cudaEventRecord(m.ev[0], st));
        // queries
        const float* xd = x + i0 * d;
        if (!x_dev) {
            m.dev_x.reserve((size_t)ni * d * sizeof(float));
            CUDA_CHECK(cudaMemcpyAsync(m.dev_x.p, xd, (size_t)ni * d * sizeof(float),
                                       cudaMemcpyHostToDevice, st));
            xd = m.dev_x.as<float>();
        }
        float* Dd = D + i0 * k;
        int64_t* Id = I + i0 * k;
        if (!D_dev) {
            m.dev_D.reserve((size_t)ni * k * sizeof(float));
            Dd = m.dev_D.as<float>();
        }
        if (!I_dev) {
            m.dev_I.reserve((size_t)ni * k * sizeof(int64_t));
            Id = m.dev_I.as<int64_t>();
        }
        const int64_t* ad = assign ? assign + i0 * np : nullptr;
        if (assign && !a_dev) {
            m.dev_assign64.reserve((size_t)pairs * sizeof(int64_t));
            CUDA_CHECK(cudaMemcpyAsync(m.dev_assign64.p, ad, (size_t)pairs * sizeof(int64_t),
                                       cudaMemcpyHostToDevice, st));
            ad = m.dev_assign64.as<int64_t>();
        }
        double ms_before[5];
        std::copy(m.last_ms, m.last_ms + 5, ms_before);
        if (one) {
            m.flag_list.reserve((size_t)ni * sizeof(int32_t));
            m.flag_count.reserve(sizeof(int32_t));
            m.host_flag.reserve(sizeof(int32_t));
            CUDA_CHECK(cudaMemsetAsync(m.flag_count.p, 0, sizeof(int32_t), st));
        }
        enqueue(ni, xd, Dd, Id, ad, true, probes32 ? probes32 + i0 * np : nullptr);
        if (one) {
            // the queries the checked re-rank could not prove: search them again, exactly (same
            // probe table, the SIMT warp kernel, no statistics: they were counted already)
            CUDA_CHECK(cudaMemcpyAsync(m.host_flag.p, m.flag_count.p, sizeof(int32_t),
                                       cudaMemcpyDeviceToHost, st));
            CUDA_CHECK(cudaStreamSynchronize(st));
            const int64_t nf = *m.host_flag.as<int32_t>();
            if (nf * 64 > ni && ni >= 256 && m.one_level < 2)
                m.one_level++; // (takes effect with the next call)
            if (nf > 0) {
                m.fallback_queries += (unsigned long long)nf;
                m.fb_x.reserve((size_t)nf * d * sizeof(float));
                m.fb_D.reserve((size_t)nf * k * sizeof(float));
                m.fb_I.reserve((size_t)nf * k * sizeof(int64_t));
                m.fb_p32.reserve((size_t)nf * np * sizeof(int32_t));
                const int32_t* fl = m.flag_list.as<int32_t>();
                CUDA_CHECK(launch_gather_rows_w(xd, fl, nf, d, m.fb_x.p, st));
                CUDA_CHECK(launch_gather_rows_w(last_probe_tab, fl, nf, (int)np, m.fb_p32.p, st));
                const bool s_simt = simt_only, s_tc = tc, s_tcr = tcr;
                const int64_t s_kp = kp;
                const int s_pmul = pmul;
                simt_only = true;
                tc = tcr = one = false;
                kp = k;
                pmul = 1;
                count_stats = false;
                want_pq = false;
                try {
                    enqueue(nf, m.fb_x.as<float>(), m.fb_D.as<float>(), m.fb_I.as<int64_t>(),
                            nullptr, false, m.fb_p32.as<int32_t>());
                } catch (...) {
                    simt_only = s_simt, tc = s_tc, tcr = s_tcr, one = true, kp = s_kp, pmul = s_pmul;
                    count_stats = true;
                    want_pq = per_query != nullptr;
                    throw;
                }
                simt_only = s_simt, tc = s_tc, tcr = s_tcr, one = true, kp = s_kp, pmul = s_pmul;
                count_stats = true;
                want_pq = per_query != nullptr;
                CUDA_CHECK(launch_scatter_results(m.fb_D.as<float>(), m.fb_I.as<int64_t>(), fl, nf,
                                                  (int)k, Dd, Id, st));
            }
        }
        CUDA_CHECK(cudaEventRecord(m.ev[4], st));
        if (per_query) {
            m.host_query_ndis.reserve((size_t)ni * sizeof(unsigned long long));
            CUDA_CHECK(cudaMemcpyAsync(m.host_query_ndis.p, m.query_ndis.p,
                                       (size_t)ni * sizeof(unsigned long long),
                                       cudaMemcpyDeviceToHost, st));
        }
        if (!D_dev)
            CUDA_CHECK(cudaMemcpyAsync(D + i0 * k, Dd, (size_t)ni * k * sizeof(float),
                                       cudaMemcpyDeviceToHost, st));
        if (!I_dev)
            CUDA_CHECK(cudaMemcpyAsync(I + i0 * k, Id, (size_t)ni * k * sizeof(int64_t),
                                       cudaMemcpyDeviceToHost, st));
        CUDA_CHECK(cudaEventRecord(m.ev[5], st));
        CUDA_CHECK(cudaMemcpyAsync(m.host_stats.p, dstats, 4 * sizeof(unsigned long long),
                                   cudaMemcpyDeviceToHost, st));
        m.evs[m.ev_cur].stages = true;
        if (async_ok && i0 + ni >= n) { // (earlier batches of a long call are read as they go)
            m.evs[m.ev_cur].pending = true;
            m.async_outstanding = true;
            return;
        }
        CUDA_CHECK(cudaStreamSynchronize(st));
        m.collect_set(m.ev_cur, true);
        fill_per_query(i0, ni, ms_before);
    }
    read_counters_();
}

// counters of the searches completed so far (the stream must be idle): nq / nlist / ndis, and the
// reference's "Invalid key" failure (faiss/IndexIVF.cpp:520-524) if a probe table held one
void IndexIVFFlatB200::read_counters_() const {
    Impl& m = *impl_;
    unsigned long long* hs = m.host_stats.as<unsigned long long>();
    m.stats.nq = (size_t)hs[0];
    m.stats.nlist = (size_t)hs[1];
    m.stats.ndis = (size_t)hs[2];
    if (hs[3] != 0) {
        hs[3] = 0;
        CUDA_CHECK(cudaMemset(m.dev_stats.as<unsigned long long>() + 3, 0,
                              sizeof(unsigned long long)));
        throw B200Exception("Invalid key in search_preassigned (key >= nlist)");
    }
}

// Bring the handle up to date with the asynchronous calls still in flight: wait for the stream,
// read their stage events, counters and deferred errors.
void IndexIVFFlatB200::finish_async_() const {
    Impl& m = *impl_;
    if (!m.async_outstanding)
        return;
    CUDA_CHECK(cudaStreamSynchronize(m.stream));
    for (int j = 1; j <= Impl::kEvSets; ++j) { // oldest first, the current call's set last
        const int i = (m.ev_cur + j) % Impl::kEvSets;
        if (m.evs[i].pending)
            m.collect_set(i, i == m.ev_cur);
    }
    m.async_outstanding = false;
    read_counters_();
}

void IndexIVFFlatB200::quantize(int64_t n, const float* x, int64_t kc, float* Dq,
                                int64_t* Iq) const {
    std::lock_guard<std::mutex> lock(mu_);
    DeviceGuard g(device_);
    Impl& m = *impl_;
    B200_THROW_IF_NOT_MSG(is_trained, "index not trained");
    B200_THROW_IF_NOT_MSG(kc > 0 && kc <= 2048, "nprobe must be in [1, 2048]");
    if (n == 0)
        return;
    const bool x_dev = is_device_pointer(x), D_dev = is_device_pointer(Dq),
               I_dev = is_device_pointer(Iq);
    cudaStream_t st = m.stream;
    const int64_t nb = std::min<int64_t>(n, 32768);
    for (int64_t i0 = 0; i0 < n; i0 += nb) {
        const int64_t ni = std::min(nb, n - i0);
        const float* xd = x + i0 * d;
        if (!x_dev) {
            m.dev_x.reserve((size_t)ni * d * sizeof(float));
            CUDA_CHECK(cudaMemcpyAsync(m.dev_x.p, xd, (size_t)ni * d * sizeof(float),
                                       cudaMemcpyHostToDevice, st));
            xd = m.dev_x.as<float>();
        }
        float* Dd = Dq + i0 * kc;
        int64_t* Id = Iq + i0 * kc;
        if (!D_dev) {
            m.dev_D.reserve((size_t)ni * kc * sizeof(float));
            Dd = m.dev_D.as<float>();
        }
        if (!I_dev) {
            m.dev_I.reserve((size_t)ni * kc * sizeof(int64_t));
            Id = m.dev_I.as<int64_t>();
        }
        run_coarse_(xd, ni, (int)kc, nullptr, Id, Dd, st);
        if (!D_dev)
            CUDA_CHECK(cudaMemcpyAsync(Dq + i0 * kc, Dd, (size_t)ni * kc * sizeof(float),
                                       cudaMemcpyDeviceToHost, st));
        if (!I_dev)
            CUDA_CHECK(cudaMemcpyAsync(Iq + i0 * kc, Id, (size_t)ni * kc * sizeof(int64_t),
                                       cudaMemcpyDeviceToHost, st));
        // all-device calls stay asynchronous on the handle's stream (the next batch only reuses
        // scratch in stream order); anything staged through host buffers must complete here
        if (!(x_dev && D_dev && I_dev))
            CUDA_CHECK(cudaStreamSynchronize(st));
    }
}

void IndexIVFFlatB200::quantize_i32(int64_t n, const float* x, int64_t kc, int32_t* Iq,
                                    float* Dq) const {
    std::lock_guard<std::mutex> lock(mu_);
    DeviceGuard g(device_);
    Impl& m = *impl_;
    B200_THROW_IF_NOT_MSG(is_trained, "index not trained");
    B200_THROW_IF_NOT_MSG(kc > 0 && kc <= 2048, "nprobe must be in [1, 2048]");
    if (n == 0)
        return;
    B200_THROW_IF_NOT_MSG(is_device_pointer(x) && is_device_pointer(Iq) && is_device_pointer(Dq),
                          "quantize_i32 takes device buffers only");
    for (int64_t i0 = 0; i0 < n; i0 += 32768) {
        const int64_t ni = std::min<int64_t>(32768, n - i0);
        run_coarse_(x + i0 * d, ni, (int)kc, Iq + i0 * kc, nullptr, Dq + i0 * kc, m.stream);
    }
}

void IndexIVFFlatB200::set_option(const char* name, double value) {
    std::lock_guard<std::mutex> lock(mu_);
    DeviceGuard g(device_);
    Impl& m = *impl_;
    B200_THROW_IF_NOT_MSG(name != nullptr, "null option name");
    const std::string n(name);
    CUDA_CHECK(cudaStreamSynchronize(m.stream));
    m.drop_graphs(); // a captured pipeline embodies the old choice
    if (n == "simt_max_avg_q")
        m.simt_max_avg_q = value;
    else if (n == "tcr_max_avg_q")
        m.tcr_max_avg_q = value;
    else if (n == "coarse_tq")
        m.coarse_tq = (int)value;
    else if (n == "coarse_s")
        m.coarse_s_override = (int)value;
    else if (n == "tc_split") { // 1 = one MMA + checked re-rank (rows-major kernel), 3 = 3xTF32
        m.tc_split = value == 1 ? 1 : 3;
        m.one_level = 0;
    }
    else if (n == "coarse_one_min_nlist")
        m.coarse_one_min_nlist = (int64_t)value;
    else if (n == "one_min_items")
        m.one_min_items = value;
    else if (n == "one_max_avg_q")
        m.one_max_avg_q = value;
    else if (n == "one_margin") {
        m.one_margin = std::max(1, (int)value);
        m.one_level = 0;
    }
    else if (n == "tcr_small_factor")
        m.tcr_small_factor = value;
    else if (n == "tcr_small_max_avg_q")
        m.tcr_small_max_avg_q = value;
    else if (n == "tc_mode")
        m.tc_mode = (int)value;
    else if (n == "use_tc")
        m.use_tc = value != 0;
    else if (n == "use_pair_plan")
        m.use_pair_plan = value != 0;
    else if (n == "use_graph")
        m.use_graph = value != 0;
    else if (n == "scan_nsplit") // timing experiments; process-wide
        g_force_nsplit.store((int)value);
    else if (n == "tc_dbg") // timing experiments (kernels.h); process-wide
        g_tc_dbg.store((int)value);
    else if (n == "cp_niter") // ClusteringParameters of the quantizer (faiss/Clustering.h:23-64),
        cp_niter = std::max(0, (int)value); // index.cp.niter in the reference
    else if (n == "cp_max_points_per_centroid")
        cp_max_points_per_centroid = std::max(1, (int)value);
    else if (n == "cp_seed")
        cp_seed = (int)value;
    else
        throw B200Exception("unknown option '" + n + "'");
}

IndexIVFStatsB200 IndexIVFFlatB200::get_stats() const {
    std::lock_guard<std::mutex> lock(mu_);
    DeviceGuard g(device_);
    finish_async_();
    return impl_->stats;
}

void IndexIVFFlatB200::reset_stats() {
    std::lock_guard<std::mutex> lock(mu_);
    DeviceGuard g(device_);
    try {
        finish_async_();
    } catch (const B200Exception&) { // (a deferred invalid-key report: the counters go anyway)
    }
    CUDA_CHECK(cudaStreamSynchronize(impl_->stream));
    CUDA_CHECK(cudaMemset(impl_->dev_stats.p, 0, 4 * sizeof(unsigned long long)));
    std::memset(impl_->host_stats.p, 0, 4 * sizeof(unsigned long long));
    impl_->stats = IndexIVFStatsB200();
    for (double& v : impl_->sum_ms)
        v = 0;
    impl_->sum_calls = 0;
}

void IndexIVFFlatB200::last_timings(double* ms5) const {
    std::lock_guard<std::mutex> lock(mu_);
    DeviceGuard g(device_);
    finish_async_();
    for (int i = 0; i < 5; ++i)
        ms5[i] = impl_->last_ms[i];
}

unsigned long long IndexIVFFlatB200::exact_redo_count() const {
    std::lock_guard<std::mutex> lock(mu_);
    return impl_->fallback_queries;
}

unsigned long long IndexIVFFlatB200::stage_time_sums(double* ms5) const {
    std::lock_guard<std::mutex> lock(mu_);
    DeviceGuard g(device_);
    finish_async_();
    for (int i = 0; i < 5; ++i)
        ms5[i] = impl_->sum_ms[i];
    return impl_->sum_calls;
}

void IndexIVFFlatB200::synchronize() const {
    std::lock_guard<std::mutex> lock(mu_);
    DeviceGuard g(device_);
    if (impl_->async_outstanding)
        finish_async_();
    else
        CUDA_CHECK(cudaStreamSynchronize(impl_->stream));
}

void IndexIVFFlatB200::get_list_sizes(int64_t* out) const {
    std::lock_guard<std::mutex> lock(mu_);
    for (size_t l = 0; l < nlist; ++l)
        out[l] = (int64_t)impl_->h_len[l];
}

// ------------------------------------------------------------------------------------------------
// build path: IndexIVF::add_with_ids (faiss/IndexIVF.cpp:191-241) -> IndexIVFFlat::add_core
// (faiss/IndexIVFFlat.cpp:46-91). Assignment = top-1 of the coarse quantizer on device; placement
// keeps input order inside every list (each reference list is appended by exactly one thread).
// ------------------------------------------------------------------------------------------------
void IndexIVFFlatB200::add_with_ids(int64_t n, const float* x, const int64_t* xids) {
    add_core(n, x, xids, nullptr);
}

void IndexIVFFlatB200::add_core(int64_t n, const float* x, const int64_t* xids,
                                const int64_t* assign) {
    std::lock_guard<std::mutex> lock(mu_);
    add_core_(n, x, xids, assign);
}

void IndexIVFFlatB200::add_core_(int64_t n, const float* x, const int64_t* xids,
                                 const int64_t* assign) {
    DeviceGuard g(device_);
    Impl& m = *impl_;
    B200_THROW_IF_NOT_MSG(is_trained, "index not trained"); // IndexIVFFlat.cpp:52
    if (n == 0)
        return;
    finish_async_();
    B200_THROW_IF_NOT_MSG(x != nullptr, "null vectors");
    cudaStream_t st = m.stream;
    const bool x_dev = is_device_pointer(x);
    const bool ids_dev = xids && is_device_pointer(xids);
    const bool a_dev = assign && is_device_pointer(assign);
    const int64_t chunk = 1 << 20;

    // pass 1: list number of every vector, on the host
    std::vector<int64_t> h_assign((size_t)n);
    if (assign) {
        if (a_dev) {
            CUDA_CHECK(cudaMemcpyAsync(h_assign.data(), assign, (size_t)n * sizeof(int64_t),
                                       cudaMemcpyDeviceToHost, st));
            CUDA_CHECK(cudaStreamSynchronize(st));
        } else
            std::memcpy(h_assign.data(), assign, (size_t)n * sizeof(int64_t));
    } else {
        for (int64_t i0 = 0; i0 < n; i0 += chunk) {
            const int64_t ni = std::min(chunk, n - i0);
            const float* xd = x + i0 * d;
            if (!x_dev) {
                m.dev_x.reserve((size_t)ni * d * sizeof(float));
                CUDA_CHECK(cudaMemcpyAsync(m.dev_x.p, xd, (size_t)ni * d * sizeof(float),
                                           cudaMemcpyHostToDevice, st));
                xd = m.dev_x.as<float>();
            }
            m.dev_I.reserve((size_t)ni * sizeof(int64_t));
            m.dev_D.reserve((size_t)ni * sizeof(float));
            // quantizer->assign == search with k = 1 (faiss/Index.cpp:84-88)
            for (int64_t j0 = 0; j0 < ni; j0 += 32768) {
                const int64_t nj = std::min<int64_t>(32768, ni - j0);
                run_coarse_(xd + j0 * d, nj, 1, nullptr, m.dev_I.as<int64_t>() + j0,
                            m.dev_D.as<float>() + j0, st);
            }
            CUDA_CHECK(cudaMemcpyAsync(h_assign.data() + i0, m.dev_I.p,
                                       (size_t)ni * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
            CUDA_CHECK(cudaStreamSynchronize(st));
        }
    }

    // host: new list lengths and the destination row of every vector (stable within a list)
    std::vector<int32_t> new_len(m.h_len);
    std::vector<int64_t> dst((size_t)n);
    int64_t nadd = 0;
    for (int64_t i = 0; i < n; ++i) {
        const int64_t l = h_assign[(size_t)i];
        if (l < 0 || (size_t)l < m.list_lo || (size_t)l >= m.list_hi) {
            B200_THROW_IF_NOT_MSG(l < (int64_t)nlist, "list number out of range in add_core");
            dst[(size_t)i] = -1;
            continue;
        }
        B200_THROW_IF_NOT_MSG(new_len[(size_t)l] < INT32_MAX, "list too long");
        dst[(size_t)i] = new_len[(size_t)l]++; // offset in list for now
        nadd++;
    }
    // Lists that outgrow the blocks reserved for them (reserve_lists; none by default) force a new
    // slab: every list gets what it needs now, the others keep their reservation. When everything
    // fits, the rows are appended in place and nothing moves.
    bool fits = m.slab.p != nullptr || nadd == 0;
    for (size_t l = 0; l < nlist && fits; ++l)
        fits = ((size_t)new_len[l] + kBlockRows - 1) / kBlockRows <= (size_t)m.h_cap[l];
    if (!fits) {
        std::vector<uint32_t> new_cap(m.h_cap);
        for (size_t l = 0; l < nlist; ++l)
            new_cap[l] = std::max<uint32_t>(
                    new_cap[l], (uint32_t)(((size_t)new_len[l] + kBlockRows - 1) / kBlockRows));
        relayout_(new_cap);
    }
    for (int64_t i = 0; i < n; ++i) {
        const int64_t l = h_assign[(size_t)i];
        if (dst[(size_t)i] >= 0)
            dst[(size_t)i] += (int64_t)m.h_blk0[(size_t)l] * kBlockRows;
    }

    // device: new rows scattered into their lists
    for (int64_t i0 = 0; i0 < n && nadd > 0; i0 += chunk) {
        const int64_t ni = std::min(chunk, n - i0);
        const float* xd = x + i0 * d;
        if (!x_dev) {
            m.dev_x.reserve((size_t)ni * d * sizeof(float));
            CUDA_CHECK(cudaMemcpyAsync(m.dev_x.p, xd, (size_t)ni * d * sizeof(float),
                                       cudaMemcpyHostToDevice, st));
            xd = m.dev_x.as<float>();
        }
        const int64_t* idd = nullptr;
        if (xids) {
            idd = xids + i0;
            if (!ids_dev) {
                m.dev_I.reserve((size_t)ni * sizeof(int64_t));
                CUDA_CHECK(cudaMemcpyAsync(m.dev_I.p, idd, (size_t)ni * sizeof(int64_t),
                                           cudaMemcpyHostToDevice, st));
                idd = m.dev_I.as<int64_t>();
            }
        }
        m.dev_assign64.reserve((size_t)ni * sizeof(int64_t));
        CUDA_CHECK(cudaMemcpyAsync(m.dev_assign64.p, dst.data() + i0, (size_t)ni * sizeof(int64_t),
                                   cudaMemcpyHostToDevice, st));
        CUDA_CHECK(launch_scatter_rows(xd, idd, m.dev_assign64.as<int64_t>(), ni, d, m.nf4,
                                       m.slab.as<float>(), m.ids.as<int64_t>(), ntotal + i0, st));
        CUDA_CHECK(cudaStreamSynchronize(st)); // staging buffers are reused by the next chunk
    }
    CUDA_CHECK(cudaMemcpyAsync(m.list_len.p, new_len.data(), nlist * sizeof(int32_t),
                               cudaMemcpyHostToDevice, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    m.h_len.swap(new_len);
    m.nstored += nadd;
    m.norms_dirty = m.norms_dirty || nadd > 0; // |y|^2 of the new rows: before the next search
    m.drop_graphs();                           // captured launch shapes embody the old lists
    ntotal += n; // IndexIVFFlat.cpp:90: every input vector counts, skipped or not
}

// New slab with cap[l] 32-row blocks for list l (>= the blocks the list uses now); the rows already
// stored move block-wise to their new places.
void IndexIVFFlatB200::relayout_(const std::vector<uint32_t>& cap) {
    Impl& m = *impl_;
    cudaStream_t st = m.stream;
    std::vector<uint32_t> new_blk0(nlist);
    size_t blocks = 0;
    for (size_t l = 0; l < nlist; ++l) {
        new_blk0[l] = (uint32_t)blocks;
        blocks += cap[l];
        B200_THROW_IF_NOT_MSG(blocks * kBlockRows < ((size_t)1 << 32),
                              "more than 2^32 rows on one device");
    }
    Buffer nslab, nids, nblk0, nnorms;
    const size_t slab_bytes = blocks * m.nf4 * kBlockRows * sizeof(float4);
    nslab.reserve(std::max<size_t>(slab_bytes, 16));
    nids.reserve(std::max<size_t>(blocks * kBlockRows * sizeof(int64_t), 16));
    nblk0.reserve(nlist * sizeof(uint32_t));
    CUDA_CHECK(cudaMemsetAsync(nslab.p, 0, slab_bytes, st));
    CUDA_CHECK(cudaMemsetAsync(nids.p, 0xff, blocks * kBlockRows * sizeof(int64_t), st));
    CUDA_CHECK(cudaMemcpyAsync(nblk0.p, new_blk0.data(), nlist * sizeof(uint32_t),
                               cudaMemcpyHostToDevice, st));
    if (m.nstored > 0)
        CUDA_CHECK(launch_move_blocks(m.slab.as<float4>(), m.ids.as<int64_t>(),
                                      m.list_blk0.as<uint32_t>(), m.list_len.as<int32_t>(),
                                      nblk0.as<uint32_t>(), (int)nlist, m.nf4, nslab.as<float4>(),
                                      nids.as<int64_t>(), st));
    CUDA_CHECK(cudaMemcpyAsync(m.list_blk0.p, new_blk0.data(), nlist * sizeof(uint32_t),
                               cudaMemcpyHostToDevice, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    m.norms.release(); // (before the new one is allocated: they are slab-sized / d)
    nnorms.reserve(std::max<size_t>(blocks * kBlockRows * sizeof(float), 16));
    m.slab = std::move(nslab);
    m.ids = std::move(nids);
    m.norms = std::move(nnorms);
    CUDA_CHECK(make_slab_tensor_map(m.lists_map, m.slab.p, m.nf4, blocks));
    CUDA_CHECK(make_slab_tensor_map_tc(m.lists_map_tc, m.slab.p, m.nf4, blocks));
    m.h_blk0.swap(new_blk0);
    m.h_cap = cap;
    m.slab_blocks = blocks;
    m.norms_dirty = true;
    m.drop_graphs();
}

// Room for sizes[l] rows in every list of this handle's range before the vectors arrive (the
// reference's InvertedLists::resize / GpuIndexIVF::reserveMemory, faiss/gpu/GpuIndexIVF.h:83-86,
// in list-wise form): later adds that stay within it append in place -- no second slab, no moves.
void IndexIVFFlatB200::reserve_lists(const int64_t* sizes) {
    std::lock_guard<std::mutex> lock(mu_);
    DeviceGuard g(device_);
    Impl& m = *impl_;
    B200_THROW_IF_NOT_MSG(sizes != nullptr, "null sizes");
    std::vector<int64_t> hs(nlist);
    if (is_device_pointer(sizes))
    {
        CUDA_CHECK(cudaMemcpyAsync(hs.data(), sizes, nlist * sizeof(int64_t),
                                   cudaMemcpyDeviceToHost, m.stream));
        CUDA_CHECK(cudaStreamSynchronize(m.stream));
    }
    else
        std::memcpy(hs.data(), sizes, nlist * sizeof(int64_t));
    std::vector<uint32_t> cap(m.h_cap);
    for (size_t l = m.list_lo; l < m.list_hi; ++l) {
        B200_THROW_IF_NOT_MSG(hs[l] >= 0 && hs[l] < INT32_MAX, "invalid list size");
        cap[l] = std::max<uint32_t>(cap[l], (uint32_t)((hs[l] + kBlockRows - 1) / kBlockRows));
    }
    finish_async_();
    CUDA_CHECK(cudaStreamSynchronize(m.stream));
    relayout_(cap);
}

// |y|^2 of every slab row (the L2 epilogue of the tensor-core scans) after the lists changed
void IndexIVFFlatB200::ensure_norms_(cudaStream_t st) const {
    Impl& m = *impl_;
    if (!m.norms_dirty)
        return;
    m.ymax2.reserve(sizeof(float));
    CUDA_CHECK(cudaMemsetAsync(m.ymax2.p, 0, sizeof(float), st));
    if (m.slab_blocks > 0)
        CUDA_CHECK(launch_row_norms(m.slab.as<float4>(), (int64_t)(m.slab_blocks * kBlockRows),
                                    m.nf4, m.norms.as<float>(), st, m.ymax2.as<float>()));
    m.norms_dirty = false;
}

// ------------------------------------------------------------------------------------------------
// range search: IndexIVF::range_search (faiss/IndexIVF.cpp:1203-1241) -> range_search_preassigned
// (:1243-1399, parallel_mode 0: every probe of every query is scanned, max_codes is not consulted
// on that path) -> IVFFlatScanner::scan_codes_range (faiss/IndexIVFFlat.cpp:181-201). Results of a
// query come out in probe order, then list order, as RangeQueryResult::add appends them.
// ------------------------------------------------------------------------------------------------
void IndexIVFFlatB200::range_search(int64_t n, const float* x, float radius,
                                    RangeSearchResult* result,
                                    const SearchParametersIVF* params) const {
    range_impl_(n, x, radius, nullptr, result, false, params);
}

void IndexIVFFlatB200::range_search_preassigned(int64_t n, const float* x, float radius,
                                                const int64_t* assign, const float* /*cdis*/,
                                                RangeSearchResult* result, bool store_pairs,
                                                const SearchParametersIVF* params) const {
    B200_THROW_IF_NOT_MSG(assign != nullptr, "null assign table");
    range_impl_(n, x, radius, assign, result, store_pairs, params);
}

void IndexIVFFlatB200::range_impl_(int64_t n, const float* x, float radius, const int64_t* assign,
                                   RangeSearchResult* result, bool store_pairs,
                                   const SearchParametersIVF* params) const {
    std::lock_guard<std::mutex> lock(mu_);
    DeviceGuard g(device_);
    Impl& m = *impl_;
    B200_THROW_IF_NOT_MSG(is_trained, "index not trained");
    B200_THROW_IF_NOT_MSG(result != nullptr, "null result");
    B200_THROW_IF_NOT_MSG((int64_t)result->nq == n, "result->nq must equal the number of queries");
    const int64_t np = (int64_t)std::min(nlist, params ? params->nprobe : nprobe);
    B200_THROW_IF_NOT_MSG(np > 0, "nprobe > 0"); // faiss/IndexIVF.cpp:1255
    B200_THROW_IF_NOT_MSG(np <= 2048, "nprobe is limited to 2048");
    const IDSelector* sel = params ? params->sel : nullptr;
    B200_THROW_IF_NOT_MSG(!(sel && store_pairs), "selector and store_pairs cannot be combined");
    result->lims.assign((size_t)n + 1, 0);
    result->labels.clear();
    result->distances.clear();
    finish_async_();
    m.begin_call();
    if (n == 0)
        return;
    B200_THROW_IF_NOT_MSG(x != nullptr, "null buffer");
    cudaStream_t st = m.stream;
    unsigned long long* dstats = m.dev_stats.as<unsigned long long>();
    const bool x_dev = is_device_pointer(x);
    const bool a_dev = assign && is_device_pointer(assign);
    const uint32_t* row_mask = nullptr;
    if (sel && m.slab_blocks > 0) {
        m.row_mask.reserve(m.slab_blocks * sizeof(uint32_t));
        sel->eval(m.ids.as<int64_t>(), (int64_t)m.slab_blocks, m.row_mask.as<uint32_t>(), st);
        row_mask = m.row_mask.as<uint32_t>();
    }
    int64_t nb = std::min<int64_t>(n, 32768);
    while (nb > 1 && nb * np > ((int64_t)1 << 26))
        nb = (nb + 1) / 2;
    size_t base = 0;
    for (int64_t i0 = 0; i0 < n; i0 += nb) {
        const int64_t ni = std::min(nb, n - i0);
        const int64_t pairs = ni * np;
        CUDA_CHECK(cudaEventRecord(m.ev[0], st));
        const float* xd = x + i0 * d;
        if (!x_dev) {
            m.dev_x.reserve((size_t)ni * d * sizeof(float));
            CUDA_CHECK(cudaMemcpyAsync(m.dev_x.p, xd, (size_t)ni * d * sizeof(float),
                                       cudaMemcpyHostToDevice, st));
            xd = m.dev_x.as<float>();
        }
        m.coarse_ids.reserve((size_t)pairs * sizeof(int32_t));
        m.coarse_dis.reserve((size_t)pairs * sizeof(float));
        if (assign) {
            const int64_t* ad = assign + i0 * np;
            if (!a_dev) {
                m.dev_assign64.reserve((size_t)pairs * sizeof(int64_t));
                CUDA_CHECK(cudaMemcpyAsync(m.dev_assign64.p, ad, (size_t)pairs * sizeof(int64_t),
                                           cudaMemcpyHostToDevice, st));
                ad = m.dev_assign64.as<int64_t>();
            }
            CUDA_CHECK(launch_i64_to_i32(ad, m.coarse_ids.as<int32_t>(), pairs, 0, (int32_t)nlist,
                                         dstats + 3, st));
        } else {
            run_coarse_(xd, ni, (int)np, m.coarse_ids.as<int32_t>(), nullptr,
                        m.coarse_dis.as<float>(), st);
        }
        CUDA_CHECK(cudaEventRecord(m.ev[1], st));
        // plan with one query per item and no row split: item == pair slot
        m.count.reserve(nlist * sizeof(int32_t));
        m.fill.reserve(nlist * sizeof(int32_t));
        m.grp_start.reserve((nlist + 1) * sizeof(int32_t));
        m.item_off.reserve((nlist + 1) * sizeof(uint32_t));
        m.pair_query.reserve((size_t)pairs * sizeof(int32_t));
        m.slot_of.reserve((size_t)pairs * sizeof(int32_t));
        PlanArgs pa;
        pa.coarse_ids = m.coarse_ids.as<int32_t>();
        pa.n = (int)ni;
        pa.nprobe = (int)np;
        pa.nlists = (int)nlist;
        pa.list_len = m.list_len.as<int32_t>();
        pa.count = m.count.as<int32_t>();
        pa.fill = m.fill.as<int32_t>();
        pa.grp_start = m.grp_start.as<int32_t>();
        pa.item_off = m.item_off.as<uint32_t>();
        pa.pair_query = m.pair_query.as<int32_t>();
        pa.slot_of = m.slot_of.as<int32_t>();
        pa.rq = 1;
        pa.nsplit = 1;
        pa.tile_tq = 0;
        pa.tile_min = 0;
        pa.item_off_tile = nullptr;
        pa.stats = dstats;
        CUDA_CHECK(launch_plan(pa, st));
        m.pair_cnt.reserve((size_t)pairs * sizeof(int32_t));
        m.pair_off.reserve((size_t)pairs * sizeof(uint32_t));
        m.qcnt.reserve((size_t)ni * sizeof(unsigned long long));
        m.lims.reserve((size_t)(ni + 1) * sizeof(unsigned long long));
        m.host_lims.reserve((size_t)(ni + 1) * sizeof(unsigned long long));
        RangeArgs ra;
        ra.slab = m.slab.as<float4>();
        ra.list_blk0 = m.list_blk0.as<uint32_t>();
        ra.list_len = m.list_len.as<int32_t>();
        ra.grp_start = pa.grp_start;
        ra.pair_query = pa.pair_query;
        ra.nlists = (int)nlist;
        ra.d = d;
        ra.nf4 = m.nf4;
        ra.queries = xd;
        ra.radius = radius;
        ra.row_mask = row_mask;
        ra.ids = store_pairs ? nullptr : m.ids.as<int64_t>();
        ra.pair_cnt = m.pair_cnt.as<int32_t>();
        CUDA_CHECK(launch_range_scan(metric_type, false, pairs, ra, st));
        CUDA_CHECK(launch_range_offsets(pa.slot_of, m.pair_cnt.as<int32_t>(), (int)ni, (int)np,
                                        m.pair_off.as<uint32_t>(),
                                        m.qcnt.as<unsigned long long>(),
                                        m.lims.as<unsigned long long>(), st));
        CUDA_CHECK(cudaMemcpyAsync(m.host_lims.p, m.lims.p,
                                   (size_t)(ni + 1) * sizeof(unsigned long long),
                                   cudaMemcpyDeviceToHost, st));
        CUDA_CHECK(cudaMemcpyAsync(m.host_stats.p, dstats, 4 * sizeof(unsigned long long),
                                   cudaMemcpyDeviceToHost, st));
        CUDA_CHECK(cudaStreamSynchronize(st));
        const unsigned long long* hl = m.host_lims.as<unsigned long long>();
        const size_t total = (size_t)hl[ni];
        if (total > 0) {
            m.range_dis.reserve(total * sizeof(float));
            m.range_ids.reserve(total * sizeof(int64_t));
            ra.pair_off = m.pair_off.as<uint32_t>();
            ra.lims = m.lims.as<unsigned long long>();
            ra.out_dis = m.range_dis.as<float>();
            ra.out_ids = m.range_ids.as<int64_t>();
            CUDA_CHECK(launch_range_scan(metric_type, true, pairs, ra, st));
            if (store_pairs)
                CUDA_CHECK(launch_rows_to_pairs(m.range_ids.as<int64_t>(), (int64_t)total,
                                                m.list_blk0.as<uint32_t>(), (int)nlist, st));
            result->distances.resize(base + total);
            result->labels.resize(base + total);
            CUDA_CHECK(cudaMemcpyAsync(result->distances.data() + base, m.range_dis.p,
                                       total * sizeof(float), cudaMemcpyDeviceToHost, st));
            CUDA_CHECK(cudaMemcpyAsync(result->labels.data() + base, m.range_ids.p,
                                       total * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
        }
        CUDA_CHECK(cudaEventRecord(m.ev[5], st));
        CUDA_CHECK(cudaStreamSynchronize(st));
        for (int64_t i = 0; i < ni; ++i)
            result->lims[(size_t)(i0 + i)] = base + (size_t)hl[i];
        base += total;
        {
            float ms;
            double v[5] = {0, 0, 0, 0, 0};
            CUDA_CHECK(cudaEventElapsedTime(&ms, m.ev[0], m.ev[1]));
            v[0] = ms;
            CUDA_CHECK(cudaEventElapsedTime(&ms, m.ev[1], m.ev[5]));
            v[2] = ms;
            CUDA_CHECK(cudaEventElapsedTime(&ms, m.ev[0], m.ev[5]));
            v[4] = ms;
            for (int s = 0; s < 5; ++s) {
                m.last_ms[s] += v[s];
                m.sum_ms[s] += v[s];
            }
            m.stats.quantization_time += v[0];
            m.stats.search_time += v[4];
        }
        const unsigned long long* hs = m.host_stats.as<unsigned long long>();
        if (hs[3] != 0) {
            CUDA_CHECK(cudaMemset(dstats + 3, 0, sizeof(unsigned long long)));
            throw B200Exception("Invalid key in range_search_preassigned (key >= nlist)");
        }
    }
    result->lims[(size_t)n] = base;
    const unsigned long long* hs = m.host_stats.as<unsigned long long>();
    m.stats.nq = (size_t)hs[0];
    m.stats.nlist = (size_t)hs[1];
    m.stats.ndis = (size_t)hs[2];
}

// ------------------------------------------------------------------------------------------------
// training: IndexIVF::train (faiss/IndexIVF.cpp:1404-1423) -> Level1Quantizer::train_q1 (:66-127,
// quantizer_trains_alone == 0) -> Clustering::train_encoded (faiss/Clustering.cpp:296-600) with the
// flat quantizer of the index's own metric as the assignment index. The host keeps the reference's
// control flow and random choices (std::mt19937 subsample and initial picks, split_clusters); the
// device does the work: assignment = K1 with k = 1, update = the add path (rows grouped per list
// in input order) followed by list_means_kernel, which sums a list's rows in that order -- the
// order compute_centroids uses (:183-204) -- so equal assignments give bit-equal centroids.
// ------------------------------------------------------------------------------------------------
namespace {

// faiss::rand_perm (faiss/utils/random.cpp:184-194) over RandomGenerator == std::mt19937
std::vector<int64_t> ref_rand_perm(size_t n, int64_t seed) {
    std::vector<int64_t> perm(n);
    for (size_t i = 0; i < n; ++i)
        perm[i] = (int64_t)i;
    std::mt19937 mt((unsigned int)seed);
    for (size_t i = 0; i + 1 < n; ++i) {
        const size_t i2 = i + (size_t)(mt() % (unsigned long)(int)(n - i));
        std::swap(perm[i], perm[i2]);
    }
    return perm;
}

// Clustering::post_process_centroids with spherical set (faiss/Clustering.cpp:52-56,
// fvec_renorm_L2 faiss/utils/distances.cpp:95-105)
void renorm_rows(std::vector<float>& c, size_t k, size_t d) {
    for (size_t i = 0; i < k; ++i) {
        float* xi = c.data() + i * d;
        float nr = 0.f;
        for (size_t j = 0; j < d; ++j)
            nr += xi[j] * xi[j];
        if (nr > 0) {
            const float inv_nr = (float)(1.0 / sqrtf(nr));
            for (size_t j = 0; j < d; ++j)
                xi[j] *= inv_nr;
        }
    }
}

// split_clusters (faiss/Clustering.cpp:232-276): an empty cluster takes a copy of a big one, chosen
// with probability proportional to its size, both perturbed symmetrically by 1/1024
int split_clusters_host(size_t d, size_t k, size_t n, std::vector<float>& hassign,
                        std::vector<float>& centroids) {
    int nsplit = 0;
    if (n <= k)
        return 0; // (the reference's pick loop cannot terminate in that case)
    std::mt19937 rng(1234);
    const double eps = 1 / 1024.;
    for (size_t ci = 0; ci < k; ++ci) {
        if (hassign[ci] != 0)
            continue;
        size_t cj;
        for (cj = 0; true; cj = (cj + 1) % k) {
            const float p = (float)((hassign[cj] - 1.0) / (float)(n - k));
            const float r = rng() / float(rng.max());
            if (r < p)
                break;
        }
        std::memcpy(centroids.data() + ci * d, centroids.data() + cj * d, sizeof(float) * d);
        for (size_t j = 0; j < d; ++j) {
            float& a = centroids[ci * d + j];
            float& b = centroids[cj * d + j];
            if (j % 2 == 0) {
                a = (float)(a * (1 + eps));
                b = (float)(b * (1 - eps));
            } else {
                a = (float)(a * (1 - eps));
                b = (float)(b * (1 + eps));
            }
        }
        hassign[ci] = hassign[cj] / 2;
        hassign[cj] -= hassign[ci];
        nsplit++;
    }
    return nsplit;
}

} // namespace

void IndexIVFFlatB200::train(int64_t n, const float* x) {
    std::lock_guard<std::mutex> lock(mu_);
    DeviceGuard g(device_);
    Impl& m = *impl_;
    if (is_trained) // "IVF quantizer does not need training", faiss/IndexIVF.cpp:72-74
        return;
    B200_THROW_IF_NOT_MSG(x != nullptr, "null training vectors");
    B200_THROW_IF_NOT_MSG(n >= (int64_t)nlist, // faiss/Clustering.cpp:302-306
                          "Number of training points should be at least as large as number of clusters");
    B200_THROW_IF_NOT_MSG(m.nstored == 0 && ntotal == 0, "train needs an empty index");
    cudaStream_t st = m.stream;
    const size_t k = nlist, dd = (size_t)d;
    const bool x_dev = is_device_pointer(x);
    finish_async_();

    if ((size_t)n == k) {
        // "just copy training set to clusters" (faiss/Clustering.cpp:346-368): the training
        // vectors, in input order, ARE the centroids -- no permutation, no iterations, and no
        // spherical renormalisation either (the reference returns before post_process_centroids)
        set_centroids_(x);
        return;
    }

    // subsample_training_set (faiss/Clustering.cpp:86-133): the first k * max_points_per_centroid
    // entries of a seeded permutation
    int64_t nx = n;
    std::vector<int64_t> pick;
    if ((size_t)n > k * (size_t)cp_max_points_per_centroid) {
        pick = ref_rand_perm((size_t)n, cp_seed);
        nx = (int64_t)(k * (size_t)cp_max_points_per_centroid);
        pick.resize((size_t)nx);
    }
    Buffer xt, didx, dcent;
    const float* xt_dev = x;
    auto cleanup = [&] {
        xt.release();
        didx.release();
        dcent.release();
    };
    try {
        if (!pick.empty() || !x_dev) {
            xt.reserve((size_t)nx * dd * sizeof(float));
            if (x_dev) {
                didx.reserve((size_t)nx * sizeof(int64_t));
                CUDA_CHECK(cudaMemcpyAsync(didx.p, pick.data(), (size_t)nx * sizeof(int64_t),
                                           cudaMemcpyHostToDevice, st));
                CUDA_CHECK(launch_gather_by_index(x, didx.as<int64_t>(), nx, d, xt.as<float>(), st));
                CUDA_CHECK(cudaStreamSynchronize(st));
            } else if (pick.empty()) {
                CUDA_CHECK(cudaMemcpyAsync(xt.p, x, (size_t)nx * dd * sizeof(float),
                                           cudaMemcpyHostToDevice, st));
                CUDA_CHECK(cudaStreamSynchronize(st));
            } else {
                std::vector<float> hx((size_t)nx * dd);
                for (int64_t i = 0; i < nx; ++i)
                    std::memcpy(hx.data() + (size_t)i * dd, x + (size_t)pick[(size_t)i] * dd,
                                dd * sizeof(float));
                CUDA_CHECK(cudaMemcpyAsync(xt.p, hx.data(), hx.size() * sizeof(float),
                                           cudaMemcpyHostToDevice, st));
                CUDA_CHECK(cudaStreamSynchronize(st));
            }
            xt_dev = xt.as<float>();
        }
        // initial centroids: k random training points (faiss/Clustering.cpp:425-441)
        std::vector<int64_t> perm = ref_rand_perm((size_t)nx, (int64_t)cp_seed + 1);
        perm.resize(k);
        didx.reserve(k * sizeof(int64_t));
        dcent.reserve(k * dd * sizeof(float));
        CUDA_CHECK(cudaMemcpyAsync(didx.p, perm.data(), k * sizeof(int64_t), cudaMemcpyHostToDevice,
                                   st));
        CUDA_CHECK(launch_gather_by_index(xt_dev, didx.as<int64_t>(), (int64_t)k, d,
                                          dcent.as<float>(), st));
        std::vector<float> cent(k * dd);
        CUDA_CHECK(cudaMemcpyAsync(cent.data(), dcent.p, k * dd * sizeof(float),
                                   cudaMemcpyDeviceToHost, st));
        CUDA_CHECK(cudaStreamSynchronize(st));
        const bool spherical = metric_type == kMetricInnerProduct; // faiss/IndexIVF.cpp:179-182
        if (spherical)
            renorm_rows(cent, k, dd);

        // Lloyd iterations over the whole list range, whatever shard this handle serves
        const size_t keep_lo = m.list_lo, keep_hi = m.list_hi;
        m.list_lo = 0;
        m.list_hi = nlist;
        try {
            std::vector<float> hassign(k);
            for (int it = 0; it < cp_niter; ++it) {
                set_centroids_(cent.data());
                reset_();
                add_core_(nx, xt_dev, nullptr, nullptr); // assign (K1, k = 1) + group rows by list
                CUDA_CHECK(launch_list_means(m.slab.as<float>(), m.list_blk0.as<uint32_t>(),
                                             m.list_len.as<int32_t>(), (int)nlist, d, m.nf4,
                                             dcent.as<float>(), st));
                CUDA_CHECK(cudaMemcpyAsync(cent.data(), dcent.p, k * dd * sizeof(float),
                                           cudaMemcpyDeviceToHost, st));
                CUDA_CHECK(cudaStreamSynchronize(st));
                for (size_t l = 0; l < k; ++l)
                    hassign[l] = (float)m.h_len[l];
                split_clusters_host(dd, k, (size_t)nx, hassign, cent);
                if (spherical)
                    renorm_rows(cent, k, dd);
            }
        } catch (...) {
            m.list_lo = keep_lo;
            m.list_hi = keep_hi;
            is_trained = false;
            throw;
        }
        reset_();
        m.list_lo = keep_lo;
        m.list_hi = keep_hi;
        set_centroids_(cent.data()); // quantizer->add(nlist, centroids); is_trained = true
    } catch (...) {
        cleanup();
        throw;
    }
    cleanup();
}

void merge_knn_results(int metric, int nshard, int64_t n, int64_t k, const float* all_D,
                       const int64_t* all_I, float* D, int64_t* I, int device, cudaStream_t st) {
    B200_THROW_IF_NOT_MSG(nshard > 0 && nshard <= 256, "nshard must be in [1, 256]");
    B200_THROW_IF_NOT_MSG(k > 0, "k > 0");
    if (n == 0)
        return;
    DeviceGuard g(device);
    const size_t cnt = (size_t)nshard * n * k;
    const bool in_dev = is_device_pointer(all_D) && is_device_pointer(all_I);
    const bool out_dev = is_device_pointer(D) && is_device_pointer(I);
    Buffer bD, bI, oD, oI;
    const float* dD = all_D;
    const int64_t* dI = all_I;
    if (!in_dev) {
        bD.reserve(cnt * sizeof(float));
        bI.reserve(cnt * sizeof(int64_t));
        CUDA_CHECK(cudaMemcpyAsync(bD.p, all_D, cnt * sizeof(float), cudaMemcpyHostToDevice, st));
        CUDA_CHECK(cudaMemcpyAsync(bI.p, all_I, cnt * sizeof(int64_t), cudaMemcpyHostToDevice, st));
        dD = bD.as<float>();
        dI = bI.as<int64_t>();
    }
    float* rD = D;
    int64_t* rI = I;
    if (!out_dev) {
        oD.reserve((size_t)n * k * sizeof(float));
        oI.reserve((size_t)n * k * sizeof(int64_t));
        rD = oD.as<float>();
        rI = oI.as<int64_t>();
    }
    CUDA_CHECK(launch_merge_shards(metric, nshard, (int)n, (int)k, dD, dI, rD, rI, st));
    if (in_dev && out_dev)
        return; // asynchronous on `st`, nothing was staged
    if (!out_dev) {
        CUDA_CHECK(cudaMemcpyAsync(D, rD, (size_t)n * k * sizeof(float), cudaMemcpyDeviceToHost, st));
        CUDA_CHECK(cudaMemcpyAsync(I, rI, (size_t)n * k * sizeof(int64_t), cudaMemcpyDeviceToHost,
                                   st));
    }
    CUDA_CHECK(cudaStreamSynchronize(st));
    bD.release();
    bI.release();
    oD.release();
    oI.release();
}

} // namespace b200ivf
