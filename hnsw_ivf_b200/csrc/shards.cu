// shards.cu -- IndexShardsIVFFlatB200: one rank of the list-sharded index (shards.h).
//
// The two exchange steps of IndexShardsIVF::search (faiss/IndexShardsIVF.cpp:158-243) -- the common
// coarse quantization handed to every shard (:175-212) and the per-shard result tables collected
// for merge_knn_results (:224-242) -- are NCCL all-gathers on the shard's own stream:
//   1. the coarse quantizer is split by QUERY (every rank holds the whole centroid table): rank r
//      quantizes queries [r * n / R, (r + 1) * n / R) and the [n][nprobe] int32 probe table is
//      all-gathered in place, so every rank ends with the table one GPU would have computed;
//   2. every rank scans the (query, probe) pairs whose list it owns (each pair on exactly one
//      rank) and k-selects its (n, k) table straight into its slot of the gather buffers;
//   3. one grouped all-gather of the distance and label tables, then merge_shards_*_kernel.
// Nothing in between touches the host. NCCL is bound at run time (dlopen of libnccl.so.2: the copy
// the process already has -- torch's, say -- or the system's), so the library has no link-time
// dependency on it and single-GPU users never load it.
#include "shards.h"

#include "common.cuh"
#include "kernels.h"

#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstring>
#include <string>
#include <vector>

namespace b200ivf {

namespace {

#define SH_CUDA_CHECK(expr)                                                                   \
    do {                                                                                      \
        cudaError_t e__ = (expr);                                                             \
        if (e__ != cudaSuccess) {                                                             \
            throw B200Exception(std::string("CUDA error ") + std::to_string((int)e__) + " (" + \
                                cudaGetErrorString(e__) + ") at " + __FILE__ + ":" +          \
                                std::to_string(__LINE__));                                    \
        }                                                                                     \
    } while (0)

struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t,
                              cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t,
                              cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t,
                         cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;

    template <class F>
    void bind(F& fn, const char* name) {
        fn = reinterpret_cast<F>(dlsym(lib, name));
        if (!fn)
            throw B200Exception(std::string("NCCL symbol not found: ") + name);
    }
    NcclApi() {
        for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
            lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (lib)
                break;
        }
        if (!lib)
            throw B200Exception("libnccl.so.2 not found: the sharded index needs NCCL");
        bind(GetUniqueId, "ncclGetUniqueId");
        bind(CommInitRank, "ncclCommInitRank");
        bind(CommDestroy, "ncclCommDestroy");
        bind(AllGather, "ncclAllGather");
        bind(Broadcast, "ncclBroadcast");
        bind(Send, "ncclSend");
        bind(Recv, "ncclRecv");
        bind(GroupStart, "ncclGroupStart");
        bind(GroupEnd, "ncclGroupEnd");
        bind(GetErrorString, "ncclGetErrorString");
    }
    static NcclApi& get() {
        static NcclApi api;
        return api;
    }
};

#define SH_NCCL_CHECK(expr)                                                                  \
    do {                                                                                     \
        ncclResult_t r__ = (expr);                                                           \
        if (r__ != ncclSuccess) {                                                            \
            throw B200Exception(std::string("NCCL error ") + std::to_string((int)r__) + " (" + \
                                NcclApi::get().GetErrorString(r__) + ") at " + __FILE__ +    \
                                ":" + std::to_string(__LINE__));                             \
        }                                                                                    \
    } while (0)

bool on_device(const void* p) {
    if (!p)
        return false;
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

struct DevGuard {
    int prev = 0;
    explicit DevGuard(int dev) {
        SH_CUDA_CHECK(cudaGetDevice(&prev));
        if (prev != dev)
            SH_CUDA_CHECK(cudaSetDevice(dev));
    }
    ~DevGuard() {
        cudaSetDevice(prev);
    }
};

constexpr int kPhases = 7; // bcast, coarse, probe gather, local search, result gather, merge, total

} // namespace

static_assert(sizeof(ncclUniqueId) == kNcclUniqueIdBytes, "ncclUniqueId size");

void nccl_get_unique_id(void* out128) {
    ncclUniqueId id;
    SH_NCCL_CHECK(NcclApi::get().GetUniqueId(&id));
    std::memcpy(out128, &id, sizeof(id));
}

struct IndexShardsIVFFlatB200::Impl {
    ncclComm_t comm = nullptr;
    // grow-only device scratch
    Buffer xq, probes, cdis, all_D, all_I, out_D, out_I;
    Buffer send_x, send_ids, send_assign, recv_x, recv_ids, recv_assign, perm, a64, adis, cnt_dev;
    // phase events of the last kSets searches (read out lazily: see IndexIVFFlatB200::Impl::EvSet)
    static constexpr int kSets = 16;
    struct EvSet {
        cudaEvent_t ev[kPhases] = {};
        bool pending = false;
    };
    EvSet evs[kSets];
    int cur = 0;
    double sum_ms[kPhases] = {0, 0, 0, 0, 0, 0, 0};
    unsigned long long calls = 0;

    void collect(int i) {
        EvSet& e = evs[i];
        float ms = 0.f;
        for (int s = 0; s + 1 < kPhases; ++s)
            if (cudaEventElapsedTime(&ms, e.ev[s], e.ev[s + 1]) == cudaSuccess)
                sum_ms[s] += ms;
        if (cudaEventElapsedTime(&ms, e.ev[0], e.ev[kPhases - 1]) == cudaSuccess)
            sum_ms[kPhases - 1] += ms;
        cudaGetLastError();
        e.pending = false;
    }
    EvSet& next_set() {
        cur = (cur + 1) % kSets;
        if (evs[cur].pending) {
            cudaEventSynchronize(evs[cur].ev[kPhases - 1]);
            collect(cur);
        }
        calls++;
        return evs[cur];
    }
    void collect_all() {
        for (int j = 1; j <= kSets; ++j) {
            const int i = (cur + j) % kSets;
            if (evs[i].pending) {
                cudaEventSynchronize(evs[i].ev[kPhases - 1]);
                collect(i);
            }
        }
    }
};

IndexShardsIVFFlatB200::IndexShardsIVFFlatB200(size_t d, size_t nlist, int metric, int device,
                                               int rank, int world, const void* nccl_unique_id)
        : impl_(nullptr), rank_(rank), world_(world), device_(device) {
    if (world < 1 || rank < 0 || rank >= world)
        throw B200Exception("invalid rank / world size");
    if (world > 256)
        throw B200Exception("at most 256 shards");
    local_.reset(new IndexIVFFlatB200(d, nlist, metric, device));
    // list range of the rank: the arithmetic of faiss/gpu/GpuCloner.cpp:304-305
    list_lo_ = (size_t)rank * nlist / (size_t)world;
    list_hi_ = (size_t)(rank + 1) * nlist / (size_t)world;
    local_->set_list_range(list_lo_, list_hi_);
    DevGuard g(device_);
    impl_ = new Impl;
    try {
        for (auto& set : impl_->evs)
            for (auto& ev : set.ev)
                SH_CUDA_CHECK(cudaEventCreate(&ev));
        if (world > 1) {
            if (!nccl_unique_id)
                throw B200Exception("null NCCL unique id");
            ncclUniqueId id;
            std::memcpy(&id, nccl_unique_id, sizeof(id));
            SH_NCCL_CHECK(NcclApi::get().CommInitRank(&impl_->comm, world, id, rank));
        }
    } catch (...) {
        delete impl_;
        impl_ = nullptr;
        throw;
    }
}

IndexShardsIVFFlatB200::~IndexShardsIVFFlatB200() {
    if (!impl_)
        return;
    int prev = 0;
    cudaGetDevice(&prev);
    cudaSetDevice(device_);
    cudaStreamSynchronize(local_->stream_());
    if (impl_->comm)
        NcclApi::get().CommDestroy(impl_->comm);
    for (auto& set : impl_->evs)
        for (auto& ev : set.ev)
            if (ev)
                cudaEventDestroy(ev);
    delete impl_;
    local_.reset();
    cudaSetDevice(prev);
}

void IndexShardsIVFFlatB200::set_centroids(const float* c, int root) {
    std::lock_guard<std::mutex> lock(mu_);
    DevGuard g(device_);
    Impl& m = *impl_;
    if (root < 0 || world_ == 1) {
        local_->set_centroids(c);
        return;
    }
    if (root >= world_)
        throw B200Exception("invalid root rank");
    cudaStream_t st = local_->stream_();
    const size_t nfl = local_->nlist * (size_t)local_->d;
    m.send_x.reserve(nfl * sizeof(float));
    if (rank_ == root) {
        if (!c)
            throw B200Exception("null centroids on the root rank");
        SH_CUDA_CHECK(cudaMemcpyAsync(m.send_x.p, c, nfl * sizeof(float), cudaMemcpyDefault, st));
    }
    SH_NCCL_CHECK(NcclApi::get().Broadcast(m.send_x.p, m.send_x.p, nfl, ncclFloat, root, m.comm,
                                           st));
    SH_CUDA_CHECK(cudaStreamSynchronize(st));
    local_->set_centroids(m.send_x.as<float>());
}

void IndexShardsIVFFlatB200::train(int64_t n, const float* x, int root) {
    if (root < 0 || root >= world_)
        throw B200Exception("train needs a root rank");
    std::vector<float> cent;
    if (rank_ == root) {
        local_->train(n, x); // Lloyd over all the lists, whatever range the rank serves
        cent.resize(local_->nlist * (size_t)local_->d);
        local_->get_centroids(cent.data());
    }
    if (world_ > 1)
        set_centroids(rank_ == root ? cent.data() : nullptr, root);
}

// ------------------------------------------------------------------------------------------------
// distributed add
// ------------------------------------------------------------------------------------------------
void IndexShardsIVFFlatB200::add_core(int64_t n, const float* x, const int64_t* xids,
                                      const int64_t* assign) {
    std::lock_guard<std::mutex> lock(mu_);
    DevGuard g(device_);
    Impl& m = *impl_;
    if (n < 0)
        throw B200Exception("negative vector count");
    if (n > 0 && !x)
        throw B200Exception("null vectors");
    if (!local_->is_trained)
        throw B200Exception("index not trained");
    cudaStream_t st = local_->stream_();
    NcclApi& nc = NcclApi::get();
    const int d = local_->d;
    const int R = world_;

    // how many vectors every rank brings: ids of id-less vectors continue from ntotal in rank order
    std::vector<long long> n_of(R, 0);
    n_of[rank_] = n;
    if (R > 1) {
        m.cnt_dev.reserve((size_t)R * R * sizeof(long long));
        long long* cd = m.cnt_dev.as<long long>();
        SH_CUDA_CHECK(cudaMemcpyAsync(cd + rank_, &n_of[rank_], sizeof(long long),
                                      cudaMemcpyHostToDevice, st));
        SH_NCCL_CHECK(nc.AllGather(cd + rank_, cd, 1, ncclInt64, m.comm, st));
        SH_CUDA_CHECK(cudaMemcpyAsync(n_of.data(), cd, R * sizeof(long long),
                                      cudaMemcpyDeviceToHost, st));
        SH_CUDA_CHECK(cudaStreamSynchronize(st));
    }
    long long id_base = ntotal_, n_all = 0, n_max = 0;
    for (int r = 0; r < R; ++r) {
        if (r < rank_)
            id_base += n_of[r];
        n_all += n_of[r];
        n_max = std::max(n_max, n_of[r]);
    }
    if (R == 1) {
        local_->add_core(n, x, xids, assign);
        ntotal_ += n_all;
        return;
    }

    // rounds of at most kChunk vectors per rank; every rank runs the same number of rounds
    const int64_t kChunk = (int64_t)1 << 22;
    const bool x_dev = on_device(x), ids_dev = xids && on_device(xids),
               a_dev = assign && on_device(assign);
    std::vector<int64_t> h_assign, h_perm;
    std::vector<long long> sendcnt(R), mat((size_t)R * R);
    for (int64_t i0 = 0; i0 < n_max; i0 += kChunk) {
        const int64_t ni = std::max<int64_t>(0, std::min<int64_t>(kChunk, n - i0));
        // 1. this round's vectors on the device, their list numbers on device and host
        const float* xd = nullptr;
        if (ni > 0) {
            xd = x + i0 * d;
            if (!x_dev) {
                m.xq.reserve((size_t)ni * d * sizeof(float));
                SH_CUDA_CHECK(cudaMemcpyAsync(m.xq.p, xd, (size_t)ni * d * sizeof(float),
                                              cudaMemcpyHostToDevice, st));
                xd = m.xq.as<float>();
            }
            m.a64.reserve((size_t)ni * sizeof(int64_t));
            if (assign) {
                SH_CUDA_CHECK(cudaMemcpyAsync(m.a64.p, assign + i0, (size_t)ni * sizeof(int64_t),
                                              a_dev ? cudaMemcpyDeviceToDevice
                                                    : cudaMemcpyHostToDevice, st));
            } else { // quantizer->assign == search with k = 1 (faiss/Index.cpp:84-88)
                m.adis.reserve((size_t)ni * sizeof(float));
                for (int64_t j0 = 0; j0 < ni; j0 += 32768) {
                    const int64_t nj = std::min<int64_t>(32768, ni - j0);
                    local_->run_coarse_(xd + j0 * d, nj, 1, nullptr, m.a64.as<int64_t>() + j0,
                                        m.adis.as<float>() + j0, st);
                }
            }
            h_assign.resize((size_t)ni);
            SH_CUDA_CHECK(cudaMemcpyAsync(h_assign.data(), m.a64.p, (size_t)ni * sizeof(int64_t),
                                          cudaMemcpyDeviceToHost, st));
            SH_CUDA_CHECK(cudaStreamSynchronize(st));
        }
        // 2. stable bucketing by owner rank (list l belongs to the rank r with
        //    r*nlist/R <= l < (r+1)*nlist/R); vectors with list -1 are dropped as add_core drops them
        std::fill(sendcnt.begin(), sendcnt.end(), 0);
        const size_t nlist = local_->nlist;
        auto owner_of = [&](int64_t l) -> int {
            int r = (int)(((unsigned long long)l * (unsigned long long)R) / nlist);
            while (r + 1 < R && (size_t)(r + 1) * nlist / R <= (size_t)l)
                ++r;
            while (r > 0 && (size_t)r * nlist / R > (size_t)l)
                --r;
            return r;
        };
        for (int64_t i = 0; i < ni; ++i) {
            const int64_t l = h_assign[(size_t)i];
            if (l >= (int64_t)nlist)
                throw B200Exception("list number out of range in add_core");
            if (l >= 0)
                sendcnt[owner_of(l)]++;
        }
        std::vector<long long> soff(R + 1, 0);
        for (int r = 0; r < R; ++r)
            soff[r + 1] = soff[r] + sendcnt[r];
        const int64_t nsend = soff[R];
        h_perm.resize((size_t)std::max<int64_t>(nsend, 1));
        {
            std::vector<long long> fill(soff.begin(), soff.end() - 1);
            for (int64_t i = 0; i < ni; ++i) {
                const int64_t l = h_assign[(size_t)i];
                if (l >= 0)
                    h_perm[(size_t)fill[owner_of(l)]++] = i;
            }
        }
        // 3. everybody's send counts -> what this rank receives from whom
        long long* cd = m.cnt_dev.as<long long>();
        SH_CUDA_CHECK(cudaMemcpyAsync(cd + (size_t)rank_ * R, sendcnt.data(), R * sizeof(long long),
                                      cudaMemcpyHostToDevice, st));
        SH_NCCL_CHECK(nc.AllGather(cd + (size_t)rank_ * R, cd, (size_t)R, ncclInt64, m.comm, st));
        SH_CUDA_CHECK(cudaMemcpyAsync(mat.data(), cd, (size_t)R * R * sizeof(long long),
                                      cudaMemcpyDeviceToHost, st));
        // 4. pack rows / ids / list numbers in owner order
        if (nsend > 0) {
            m.perm.reserve((size_t)nsend * sizeof(int64_t));
            m.send_x.reserve((size_t)nsend * d * sizeof(float));
            m.send_ids.reserve((size_t)nsend * sizeof(int64_t));
            m.send_assign.reserve((size_t)nsend * sizeof(int64_t));
            SH_CUDA_CHECK(cudaMemcpyAsync(m.perm.p, h_perm.data(), (size_t)nsend * sizeof(int64_t),
                                          cudaMemcpyHostToDevice, st));
            SH_CUDA_CHECK(launch_gather_by_index(xd, m.perm.as<int64_t>(), nsend, d,
                                                 m.send_x.as<float>(), st));
            const int64_t* idsrc = nullptr;
            if (xids) {
                idsrc = xids + i0;
                if (!ids_dev) {
                    m.recv_ids.reserve((size_t)ni * sizeof(int64_t)); // (free until step 5)
                    SH_CUDA_CHECK(cudaMemcpyAsync(m.recv_ids.p, idsrc, (size_t)ni * sizeof(int64_t),
                                                  cudaMemcpyHostToDevice, st));
                    idsrc = m.recv_ids.as<int64_t>();
                }
            }
            SH_CUDA_CHECK(launch_gather_i64(idsrc, m.perm.as<int64_t>(), nsend, id_base + i0,
                                            m.send_ids.as<int64_t>(), st));
            SH_CUDA_CHECK(launch_gather_i64(m.a64.as<int64_t>(), m.perm.as<int64_t>(), nsend, 0,
                                            m.send_assign.as<int64_t>(), st));
        }
        SH_CUDA_CHECK(cudaStreamSynchronize(st)); // `mat` is on the host; recv_ids may be reused
        std::vector<long long> roff(R + 1, 0);
        for (int r = 0; r < R; ++r)
            roff[r + 1] = roff[r] + mat[(size_t)r * R + rank_];
        const int64_t nrecv = roff[R];
        m.recv_x.reserve((size_t)std::max<int64_t>(nrecv, 1) * d * sizeof(float));
        m.recv_ids.reserve((size_t)std::max<int64_t>(nrecv, 1) * sizeof(int64_t));
        m.recv_assign.reserve((size_t)std::max<int64_t>(nrecv, 1) * sizeof(int64_t));
        // 5. the exchange: one grouped all-to-all of the three columns
        SH_NCCL_CHECK(nc.GroupStart());
        for (int r = 0; r < R; ++r) {
            const size_t sc = (size_t)sendcnt[r], so = (size_t)soff[r];
            const size_t rc = (size_t)mat[(size_t)r * R + rank_], ro = (size_t)roff[r];
            if (sc > 0) {
                SH_NCCL_CHECK(nc.Send(m.send_x.as<float>() + so * d, sc * d, ncclFloat, r, m.comm,
                                      st));
                SH_NCCL_CHECK(nc.Send(m.send_ids.as<int64_t>() + so, sc, ncclInt64, r, m.comm, st));
                SH_NCCL_CHECK(nc.Send(m.send_assign.as<int64_t>() + so, sc, ncclInt64, r, m.comm,
                                      st));
            }
            if (rc > 0) {
                SH_NCCL_CHECK(nc.Recv(m.recv_x.as<float>() + ro * d, rc * d, ncclFloat, r, m.comm,
                                      st));
                SH_NCCL_CHECK(nc.Recv(m.recv_ids.as<int64_t>() + ro, rc, ncclInt64, r, m.comm, st));
                SH_NCCL_CHECK(nc.Recv(m.recv_assign.as<int64_t>() + ro, rc, ncclInt64, r, m.comm,
                                      st));
            }
        }
        SH_NCCL_CHECK(nc.GroupEnd());
        // 6. what arrived goes into this rank's lists (source-rank order, then input order)
        if (nrecv > 0)
            local_->add_core(nrecv, m.recv_x.as<float>(), m.recv_ids.as<int64_t>(),
                             m.recv_assign.as<int64_t>());
        else
            SH_CUDA_CHECK(cudaStreamSynchronize(st));
    }
    ntotal_ += n_all;
}

// ------------------------------------------------------------------------------------------------
// search
// ------------------------------------------------------------------------------------------------
void IndexShardsIVFFlatB200::search(int64_t n, const float* x, int64_t k, float* D, int64_t* I,
                                    int root) const {
    std::lock_guard<std::mutex> lock(mu_);
    DevGuard g(device_);
    Impl& m = *impl_;
    if (k <= 0)
        throw B200Exception("Error: 'k > 0' failed");
    if (!local_->is_trained)
        throw B200Exception("index not trained");
    if (root >= world_)
        throw B200Exception("invalid root rank");
    if (n == 0)
        return;
    if (!D || !I)
        throw B200Exception("null buffer");
    const bool x_dev = on_device(x), D_dev = on_device(D), I_dev = on_device(I);
    const bool reads_x = root < 0 || rank_ == root;
    if (reads_x && !x)
        throw B200Exception("null buffer");
    if (world_ == 1) {
        local_->search(n, x, k, D, I);
        return;
    }
    NcclApi& nc = NcclApi::get();
    cudaStream_t st = local_->stream_();
    const int d = local_->d;
    const int R = world_;
    const int64_t np = (int64_t)std::min(local_->nlist, local_->nprobe);
    if (np <= 0)
        throw B200Exception("Error: 'nprobe > 0' failed");

    const int64_t nbmax = std::min<int64_t>(n, 32768);
    for (int64_t i0 = 0; i0 < n; i0 += nbmax) {
        const int64_t ni = std::min(nbmax, n - i0);
        Impl::EvSet& es = m.next_set();
        SH_CUDA_CHECK(cudaEventRecord(es.ev[0], st));
        // ---- queries on the device of every rank -------------------------------------------------
        const float* xd = x ? x + i0 * d : nullptr;
        if (!(x && x_dev)) {
            m.xq.reserve((size_t)ni * d * sizeof(float));
            if (reads_x)
                SH_CUDA_CHECK(cudaMemcpyAsync(m.xq.p, xd, (size_t)ni * d * sizeof(float),
                                              cudaMemcpyHostToDevice, st));
            xd = m.xq.as<float>();
        }
        if (root >= 0) // in place: root sends its buffer, the others receive into theirs
            SH_NCCL_CHECK(nc.Broadcast(xd, const_cast<float*>(xd), (size_t)ni * d, ncclFloat, root,
                                       m.comm, st));
        SH_CUDA_CHECK(cudaEventRecord(es.ev[1], st));
        // ---- the common coarse quantization, split by query ---------------------------------------
        const int64_t chunk = (ni + R - 1) / R;
        const int64_t lo = std::min<int64_t>((int64_t)rank_ * chunk, ni);
        const int64_t hi = std::min<int64_t>(lo + chunk, ni);
        m.probes.reserve((size_t)R * chunk * np * sizeof(int32_t));
        m.cdis.reserve((size_t)chunk * np * sizeof(float));
        int32_t* mine = m.probes.as<int32_t>() + (size_t)rank_ * chunk * np;
        if (hi - lo < chunk) // rows past the end of the batch: no probes
            SH_CUDA_CHECK(launch_fill_u32(reinterpret_cast<uint32_t*>(mine + (hi - lo) * np),
                                          (chunk - (hi - lo)) * np, 0xffffffffu, st));
        if (hi > lo)
            local_->run_coarse_(xd + lo * d, hi - lo, (int)np, mine, nullptr, m.cdis.as<float>(),
                                st);
        SH_CUDA_CHECK(cudaEventRecord(es.ev[2], st));
        SH_NCCL_CHECK(nc.AllGather(mine, m.probes.p, (size_t)chunk * np, ncclInt32, m.comm, st));
        SH_CUDA_CHECK(cudaEventRecord(es.ev[3], st));
        // ---- this rank's lists: plan, scan, k-select into its slot of the gather buffers ----------
        m.all_D.reserve((size_t)R * ni * k * sizeof(float));
        m.all_I.reserve((size_t)R * ni * k * sizeof(int64_t));
        float* Dl = m.all_D.as<float>() + (size_t)rank_ * ni * k;
        int64_t* Il = m.all_I.as<int64_t>() + (size_t)rank_ * ni * k;
        local_->search_probed_i32(ni, xd, k, m.probes.as<int32_t>(), Dl, Il);
        SH_CUDA_CHECK(cudaEventRecord(es.ev[4], st));
        // ---- collect the per-shard tables, merge (faiss/IndexShardsIVF.cpp:224-242) ---------------
        SH_NCCL_CHECK(nc.GroupStart());
        SH_NCCL_CHECK(nc.AllGather(Dl, m.all_D.p, (size_t)ni * k, ncclFloat, m.comm, st));
        SH_NCCL_CHECK(nc.AllGather(Il, m.all_I.p, (size_t)ni * k, ncclInt64, m.comm, st));
        SH_NCCL_CHECK(nc.GroupEnd());
        SH_CUDA_CHECK(cudaEventRecord(es.ev[5], st));
        float* Do = D + i0 * k;
        int64_t* Io = I + i0 * k;
        if (!D_dev) {
            m.out_D.reserve((size_t)ni * k * sizeof(float));
            Do = m.out_D.as<float>();
        }
        if (!I_dev) {
            m.out_I.reserve((size_t)ni * k * sizeof(int64_t));
            Io = m.out_I.as<int64_t>();
        }
        SH_CUDA_CHECK(launch_merge_shards(local_->metric_type, R, (int)ni, (int)k,
                                          m.all_D.as<float>(), m.all_I.as<int64_t>(), Do, Io, st));
        SH_CUDA_CHECK(cudaEventRecord(es.ev[6], st));
        es.pending = true;
        if (!D_dev)
            SH_CUDA_CHECK(cudaMemcpyAsync(D + i0 * k, Do, (size_t)ni * k * sizeof(float),
                                          cudaMemcpyDeviceToHost, st));
        if (!I_dev)
            SH_CUDA_CHECK(cudaMemcpyAsync(I + i0 * k, Io, (size_t)ni * k * sizeof(int64_t),
                                          cudaMemcpyDeviceToHost, st));
        // results in host memory must have landed; a further batch reuses the scratch in stream
        // order but is kept simple. All-device single-batch calls return here, work enqueued.
        if (!D_dev || !I_dev || i0 + ni < n)
            SH_CUDA_CHECK(cudaStreamSynchronize(st));
    }
}

unsigned long long IndexShardsIVFFlatB200::phase_time_sums(double* ms7) const {
    std::lock_guard<std::mutex> lock(mu_);
    DevGuard g(device_);
    impl_->collect_all();
    for (int i = 0; i < kPhases; ++i)
        ms7[i] = impl_->sum_ms[i];
    return impl_->calls;
}

void IndexShardsIVFFlatB200::reset_phase_times() {
    std::lock_guard<std::mutex> lock(mu_);
    DevGuard g(device_);
    impl_->collect_all();
    for (double& v : impl_->sum_ms)
        v = 0;
    impl_->calls = 0;
}

void IndexShardsIVFFlatB200::synchronize() const {
    std::lock_guard<std::mutex> lock(mu_);
    local_->synchronize();
}

} // namespace b200ivf
