// shards.h -- one rank of a list-sharded IVF-Flat index: the multi-GPU form of the path, behind the
// same C-ABI. Behavioural model: IndexShardsIVF (faiss/IndexShardsIVF.h:21-44,
// faiss/IndexShardsIVF.cpp:158-243) -- a common coarse quantizer, one search_preassigned per shard,
// merge_knn_results -- with shards that hold list RANGES (shard_type 4 / SUBSET_TYPE_INVLIST:
// faiss/gpu/GpuCloner.cpp:303-314, faiss/invlists/InvertedLists.cpp:168-176).
//
// One process per GPU. The reference fans out with host threads and merges on the CPU; here every
// rank drives its own device and the two exchange steps of the path (probe table, per-shard result
// tables) are NCCL all-gathers over NVLink on the handle's stream, followed by the device merge --
// no host round trip inside a search.
#pragma once

#include "index.h"

#include <memory>

namespace b200ivf {

constexpr int kNcclUniqueIdBytes = 128; // sizeof(ncclUniqueId)

// ncclGetUniqueId: rank 0 calls it and hands the bytes to the other ranks by whatever channel the
// application has (MPI, a socket, torch.distributed's store, ...)
void nccl_get_unique_id(void* out128);

class IndexShardsIVFFlatB200 {
  public:
    // Collective over the `world` ranks that share `nccl_unique_id`. The rank's sub-index holds the
    // lists [rank * nlist / world, (rank + 1) * nlist / world) on `device`.
    IndexShardsIVFFlatB200(size_t d, size_t nlist, int metric, int device, int rank, int world,
                           const void* nccl_unique_id);
    ~IndexShardsIVFFlatB200();
    IndexShardsIVFFlatB200(const IndexShardsIVFFlatB200&) = delete;
    IndexShardsIVFFlatB200& operator=(const IndexShardsIVFFlatB200&) = delete;

    int rank() const {
        return rank_;
    }
    int world() const {
        return world_;
    }
    size_t list_lo() const {
        return list_lo_;
    }
    size_t list_hi() const {
        return list_hi_;
    }
    // the shard of this rank (IndexShards::at, faiss/IndexShards.h): nprobe, stats, options, lists
    IndexIVFFlatB200* local() const {
        return local_.get();
    }
    int64_t ntotal() const {
        return ntotal_;
    }

    // Every call below is COLLECTIVE: all ranks make it, in the same order.

    // root >= 0: rank `root` supplies the table, the others receive it (their `c` may be null);
    // root < 0: every rank passes the same table
    void set_centroids(const float* c, int root);
    // IndexShardsIVF::train (faiss/IndexShardsIVF.cpp:102-135): the common quantizer is trained
    // once -- on rank `root`, with its n vectors -- and the centroids are shared
    void train(int64_t n, const float* x, int root);
    // IndexShardsIVF::add_with_ids (faiss/IndexShardsIVF.cpp:137-154), distributed: each rank
    // brings ITS OWN n vectors (any n, 0 allowed); they are assigned by the common quantizer where
    // they are and shipped to the rank that owns their list. xids null: ids continue from ntotal,
    // rank 0's vectors first. `assign` (optional) skips the quantizer.
    void add_core(int64_t n, const float* x, const int64_t* xids, const int64_t* assign);

    // IndexShardsIVF::search. root < 0: every rank passes the same n queries; root >= 0: only
    // rank root's x is read and the queries are broadcast (the other ranks' x must be a device
    // buffer to receive them, or null). Every rank gets the merged (D, I). With device buffers
    // the call returns once its work is enqueued on the shard's stream.
    void search(int64_t n, const float* x, int64_t k, float* D, int64_t* I, int root) const;

    // device time of the phases of the search calls since reset_phase_times, summed:
    // {query broadcast, coarse quantizer (this rank's slice), probe-table all-gather,
    //  local plan + scan + k-select, result all-gather, merge, total}; returns the number of calls
    unsigned long long phase_time_sums(double* ms7) const;
    void reset_phase_times();
    void synchronize() const;

  private:
    struct Impl;
    std::unique_ptr<IndexIVFFlatB200> local_;
    Impl* impl_;
    int rank_, world_, device_;
    size_t list_lo_, list_hi_;
    int64_t ntotal_ = 0;
    mutable std::mutex mu_;
};

} // namespace b200ivf
