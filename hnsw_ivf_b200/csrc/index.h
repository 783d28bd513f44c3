// index.h -- host-side C++ object behind the C-ABI: a device-resident IndexIVFFlat with its flat
// coarse quantizer. Mirrors the reference's interface for this one path (same names, argument
// meaning and error behaviour): faiss/IndexIVFFlat.h:26-70, faiss/IndexIVF.h:44-96,303-337,
// faiss/Index.h:77-148.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

namespace b200ivf {

// the FaissException of this library (C-ABI code -2, c_api/error_c.h:26)
struct B200Exception : std::runtime_error {
    using std::runtime_error::runtime_error;
};

struct IndexIVFStatsB200 { // faiss/IndexIVF.h:567-580
    size_t nq = 0, nlist = 0, ndis = 0, nheap_updates = 0;
    double quantization_time = 0, search_time = 0;
};

// per-query figures of the fork's IndexIVF::search_stats (faiss/IndexIVF.h:28-32), microseconds
struct QueryLatencyStats {
    double total_us = 0.0;
    double quantization_us = 0.0;
    double list_scan_us = 0.0;
};

struct IDSelector;          // selector.h
struct SearchParametersIVF; // selector.h

// faiss::RangeSearchResult (faiss/impl/AuxIndexStructures.h:30-58): query i has results
// lims[i]..lims[i+1] in labels / distances; host memory owned by the object
struct RangeSearchResult {
    size_t nq = 0;
    std::vector<size_t> lims; // [nq + 1]
    std::vector<int64_t> labels;
    std::vector<float> distances;
    explicit RangeSearchResult(size_t nq_) : nq(nq_), lims(nq_ + 1, 0) {}
};

// grow-only device / pinned-host buffer; owns its allocation (move-only), so a throw between an
// allocation and its hand-over (an out-of-memory CUDA_CHECK in add_core_, say) frees it
struct Buffer {
    void* p = nullptr;
    size_t bytes = 0;
    bool pinned_host = false;
    Buffer() = default;
    ~Buffer() {
        release();
    }
    Buffer(const Buffer&) = delete;
    Buffer& operator=(const Buffer&) = delete;
    Buffer(Buffer&& o) noexcept : p(o.p), bytes(o.bytes), pinned_host(o.pinned_host) {
        o.p = nullptr;
        o.bytes = 0;
    }
    Buffer& operator=(Buffer&& o) noexcept {
        if (this != &o) {
            release();
            p = o.p;
            bytes = o.bytes;
            pinned_host = o.pinned_host;
            o.p = nullptr;
            o.bytes = 0;
        }
        return *this;
    }
    void reserve(size_t n);
    void release();
    template <class T>
    T* as() const {
        return reinterpret_cast<T*>(p);
    }
};

constexpr int64_t kNoAllocMaxQueries = 4096;
constexpr int64_t kNoAllocMaxProbe = 64;
constexpr size_t kNoAllocScratchBytes = (size_t)8 << 20; // per GPU that loads the library

class IndexIVFFlatB200 {
  public:
    IndexIVFFlatB200(size_t d, size_t nlist, int metric, int device);
    ~IndexIVFFlatB200();
    IndexIVFFlatB200(const IndexIVFFlatB200&) = delete;
    IndexIVFFlatB200& operator=(const IndexIVFFlatB200&) = delete;

    // ---- fields with the reference's names (faiss/Index.h:81-90, faiss/IndexIVF.h:44,95) ----
    int d;
    size_t nlist;
    int metric_type;
    size_t nprobe = 1;
    int64_t ntotal = 0;
    bool is_trained = false;
    size_t max_codes = 0; // faiss/IndexIVF.h:96 (0 = no limit)
    // Level1Quantizer::cp as IndexIVF sets it up (faiss/IndexIVF.cpp:55, :179-182; Clustering.h:23-64)
    int cp_niter = 10;
    int cp_seed = 1234;
    int cp_max_points_per_centroid = 256;

    void set_centroids(const float* c);
    void get_centroids(float* out) const;
    void set_list_range(size_t lo, size_t hi);
    void add_with_ids(int64_t n, const float* x, const int64_t* xids);
    void add_core(int64_t n, const float* x, const int64_t* xids, const int64_t* assign);
    void reset();
    // room for sizes[l] rows in list l ([nlist] entries, host or device; entries outside this
    // handle's list range are ignored): adds that stay within it append in place
    void reserve_lists(const int64_t* sizes);
    // sizes of all nlist lists at once (0 outside this handle's range), host array
    void get_list_sizes(int64_t* out) const;
    size_t get_list_size(size_t l) const;
    void get_list_ids(size_t l, int64_t* out) const;
    void get_list_codes(size_t l, float* out) const;

    // IndexIVF::train -> Level1Quantizer::train_q1 -> Clustering::train with the flat quantizer
    // as the assignment index (faiss/IndexIVF.cpp:66-127, faiss/Clustering.cpp:296-600); a no-op
    // when the quantizer already holds nlist centroids (:72-74). The index must be empty.
    void train(int64_t n, const float* x);

    // params (nullptr: the handle's nprobe / max_codes, no selector) and per_query_stats (nullptr:
    // not wanted; HOST array of n entries) are the reference's SearchParametersIVF
    // (faiss/IndexIVF.h:73-88) and the fork's QueryLatencyStats (search_stats, IndexIVF.cpp:727)
    void search(int64_t n, const float* x, int64_t k, float* D, int64_t* I,
                const SearchParametersIVF* params = nullptr,
                QueryLatencyStats* per_query_stats = nullptr) const;
    void search_preassigned(int64_t n, const float* x, int64_t k, const int64_t* assign,
                            const float* centroid_dis, float* D, int64_t* I, bool store_pairs,
                            const SearchParametersIVF* params = nullptr,
                            QueryLatencyStats* per_query_stats = nullptr) const;
    // IndexIVF::range_search / range_search_preassigned (faiss/IndexIVF.cpp:1203-1399)
    void range_search(int64_t n, const float* x, float radius, RangeSearchResult* result,
                      const SearchParametersIVF* params = nullptr) const;
    void range_search_preassigned(int64_t n, const float* x, float radius, const int64_t* assign,
                                  const float* centroid_dis, RangeSearchResult* result,
                                  bool store_pairs = false,
                                  const SearchParametersIVF* params = nullptr) const;
    void quantize(int64_t n, const float* x, int64_t nprobe, float* Dq, int64_t* Iq) const;
    // search_preassigned with a DEVICE probe table of int32 (-1 = no probe); device buffers only
    void search_probed_i32(int64_t n, const float* x, int64_t k, const int32_t* probes, float* D,
                           int64_t* I) const;
    // quantize() for callers that may not allocate (noalloc.cu): device buffers only, scratch in a
    // fixed file-scope __device__ array shared by every call; n <= kNoAllocMaxQueries,
    // nprobe <= kNoAllocMaxProbe. No cudaMalloc / cudaFree on this path. Asynchronous on the
    // handle's stream.
    void quantize_noalloc(int64_t n, const float* x, int64_t nprobe, float* Dq, int64_t* Iq) const;
    // quantizer->search into a device int32 table (device buffers only, asynchronous)
    void quantize_i32(int64_t n, const float* x, int64_t nprobe, int32_t* Iq, float* Dq) const;

    IndexIVFStatsB200 get_stats() const;
    void reset_stats();
    void last_timings(double* ms5) const;
    // the same five stage times summed over the search calls since reset_stats; returns how many
    unsigned long long stage_time_sums(double* ms5) const;
    // queries searched a second time on the exact kernel because the checked re-rank of the
    // one-MMA filter mode ("tc_split" = 1) could not prove their result, since construction
    unsigned long long exact_redo_count() const;
    // wait for the asynchronous calls in flight; a deferred failure (invalid key) is thrown here
    void synchronize() const;
    void set_stream(cudaStream_t s);
    // kernel-selection knobs (the ParameterSpace of this path): "simt_max_avg_q" / "tcr_max_avg_q"
    // (queries per probed list below which the exact warp kernel / the rows-major tensor-core
    // kernel takes the scan), "tc_mode" (0 auto, 1 queries-major, 2 rows-major), "use_tc",
    // "use_graph". Results do not depend on them.
    void set_option(const char* name, double value);

    int device() const {
        return device_;
    }

  private:
    friend class IndexShardsIVFFlatB200; // shards.cu: shares the handle's stream and scratch
    struct Impl;
    void set_centroids_(const float* c);
    void reset_();
    void add_core_(int64_t n, const float* x, const int64_t* xids, const int64_t* assign);
    void relayout_(const std::vector<uint32_t>& cap);
    void ensure_norms_(cudaStream_t st) const;
    void finish_async_() const;
    void read_counters_() const;
    cudaStream_t stream_() const; // the stream every call of this handle is ordered on
    // the centroid table as the scan kernels read it (one-list slab + its blk0 / len words)
    struct CoarseView {
        const void* slab;
        const uint32_t* blk0;
        const int32_t* len;
        int nf4;
    };
    CoarseView coarse_view_() const;
    void search_impl_(int64_t n, const float* x, int64_t k, const int64_t* assign, float* D,
                      int64_t* I, bool store_pairs, const SearchParametersIVF* params,
                      QueryLatencyStats* per_query, const int32_t* probes32) const;
    void range_impl_(int64_t n, const float* x, float radius, const int64_t* assign,
                     RangeSearchResult* result, bool store_pairs,
                     const SearchParametersIVF* params) const;
    // K1: the kc nearest centroids of n device-resident queries (ids as int32 and/or int64)
    void run_coarse_(const float* xd, int64_t n, int kc, int32_t* out_ids32, int64_t* out_ids64,
                     float* out_dis, cudaStream_t st) const;
    Impl* impl_;
    int device_;
    mutable std::mutex mu_;
};

// merge of per-shard tables (faiss/utils/Heap.cpp:158-232) on `device`
void merge_knn_results(int metric, int nshard, int64_t n, int64_t k, const float* all_D,
                       const int64_t* all_I, float* D, int64_t* I, int device, cudaStream_t st);

// The reference's index file format for this index type ("IwFl" + "IxF2"/"IxFI" + "ilar"):
// faiss/impl/index_read.cpp:782-787 / index_write.cpp:635-638 (index_io.cu). The caller owns the
// returned index.
IndexIVFFlatB200* read_index(const char* fname, int io_flags, int device);
void write_index(const IndexIVFFlatB200* idx, const char* fname);

} // namespace b200ivf
