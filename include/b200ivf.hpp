// b200ivf.hpp -- header-only C++ face of the C-ABI (b200ivf_c.h) with the reference's own names, so
// that code written against faiss::IndexIVFFlat for this path compiles against the device index by
// changing the type:
//   faiss::Index::search(n, x, k, distances, labels, params)      faiss/Index.h:142-148
//   fields d, ntotal, is_trained, metric_type                     faiss/Index.h:81-90
//   nprobe, nlist, max_codes                                      faiss/IndexIVF.h:44,95,96
//   train / add / add_with_ids / reset / range_search             faiss/Index.h:107-170
//   SearchParametersIVF, IDSelectorRange / Batch / Bitmap / Not   faiss/IndexIVF.h:73-88,
//                                                                 faiss/impl/IDSelector.h
//   RangeSearchResult {nq, lims, labels, distances}               faiss/impl/AuxIndexStructures.h:30-58
//   the fork's search_stats + QueryLatencyStats                   faiss/IndexIVF.h:28-32,322-330
// Failures throw b200::Exception carrying the C-ABI's last error (the reference throws
// FaissException from the same checks). Nothing here computes: every call is one C-ABI call.
#pragma once

#include "b200ivf_c.h"

#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace b200 {

using idx_t = int64_t; // faiss/MetricType.h:50

enum MetricType { METRIC_INNER_PRODUCT = 0, METRIC_L2 = 1 }; // faiss/MetricType.h:23-24

struct Exception : std::runtime_error {
    int code;
    Exception(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};

inline void check(int rc) {
    if (rc != 0)
        throw Exception(rc, b200_get_last_error());
}

// ---- selectors (operands of Not / And / Or / XOr are borrowed, as in the reference) -------------
struct IDSelector {
    B200IDSelector* h = nullptr;
    IDSelector() = default;
    IDSelector(const IDSelector&) = delete;
    IDSelector& operator=(const IDSelector&) = delete;
    virtual ~IDSelector() {
        b200_IDSelector_free(h);
    }
    bool is_member(idx_t id) const {
        return b200_IDSelector_is_member(h, id) != 0;
    }
};
struct IDSelectorRange : IDSelector {
    IDSelectorRange(idx_t imin, idx_t imax) {
        check(b200_IDSelectorRange_new(&h, imin, imax));
    }
};
struct IDSelectorBatch : IDSelector {
    IDSelectorBatch(size_t n, const idx_t* indices, int device = 0) {
        check(b200_IDSelectorBatch_new(&h, n, indices, device));
    }
};
struct IDSelectorBitmap : IDSelector {
    IDSelectorBitmap(size_t n, const uint8_t* bitmap, int device = 0) {
        check(b200_IDSelectorBitmap_new(&h, n, bitmap, device));
    }
};
struct IDSelectorNot : IDSelector {
    explicit IDSelectorNot(const IDSelector* sel) {
        check(b200_IDSelectorNot_new(&h, sel->h));
    }
};
struct IDSelectorAnd : IDSelector {
    IDSelectorAnd(const IDSelector* lhs, const IDSelector* rhs) {
        check(b200_IDSelectorAnd_new(&h, lhs->h, rhs->h));
    }
};
struct IDSelectorOr : IDSelector {
    IDSelectorOr(const IDSelector* lhs, const IDSelector* rhs) {
        check(b200_IDSelectorOr_new(&h, lhs->h, rhs->h));
    }
};
struct IDSelectorXOr : IDSelector {
    IDSelectorXOr(const IDSelector* lhs, const IDSelector* rhs) {
        check(b200_IDSelectorXOr_new(&h, lhs->h, rhs->h));
    }
};

struct SearchParametersIVF { // faiss/IndexIVF.h:73-88
    size_t nprobe = 1;
    size_t max_codes = 0;
    IDSelector* sel = nullptr;
};

struct RangeSearchResult { // faiss/impl/AuxIndexStructures.h:30-58 (arrays owned by the object)
    size_t nq;
    std::vector<size_t> lims;
    std::vector<idx_t> labels;
    std::vector<float> distances;
    explicit RangeSearchResult(size_t nq_) : nq(nq_), lims(nq_ + 1, 0) {}
};

using QueryLatencyStats = B200QueryLatencyStats; // {total_us, quantization_us, list_scan_us}

// ---- the index --------------------------------------------------------------------------------
struct IndexIVFFlat {
    int d;
    size_t nlist;
    MetricType metric_type;
    size_t nprobe = 1;    // faiss/IndexIVF.h:95
    size_t max_codes = 0; // faiss/IndexIVF.h:96
    B200IndexIVFFlat* h = nullptr;

    // faiss::IndexIVFFlat(quantizer, d, nlist, metric) with a flat quantizer of the same metric
    IndexIVFFlat(size_t d_, size_t nlist_, MetricType metric = METRIC_L2, int device = 0)
            : d((int)d_), nlist(nlist_), metric_type(metric) {
        check(b200_IndexIVFFlat_new_with_metric(&h, d_, nlist_, (B200MetricType)metric, device));
    }
    // faiss::read_index(fname) for an IndexIVFFlat file of the reference
    IndexIVFFlat(const char* fname, int device) : d(0), nlist(0), metric_type(METRIC_L2) {
        check(b200_read_index_fname(fname, 0, device, &h));
        d = b200_IndexIVFFlat_d(h);
        nlist = b200_IndexIVFFlat_nlist(h);
        metric_type = (MetricType)b200_IndexIVFFlat_metric_type(h);
        nprobe = b200_IndexIVFFlat_nprobe(h);
    }
    IndexIVFFlat(const IndexIVFFlat&) = delete;
    IndexIVFFlat& operator=(const IndexIVFFlat&) = delete;
    ~IndexIVFFlat() {
        b200_IndexIVFFlat_free(h);
    }

    idx_t ntotal() const {
        return b200_IndexIVFFlat_ntotal(h);
    }
    bool is_trained() const {
        return b200_IndexIVFFlat_is_trained(h) != 0;
    }

    void train(idx_t n, const float* x) {
        check(b200_IndexIVFFlat_train(h, n, x));
    }
    void set_centroids(const float* centroids) { // quantizer->add(nlist, centroids)
        check(b200_IndexIVFFlat_set_centroids(h, centroids));
    }
    void add(idx_t n, const float* x) {
        check(b200_IndexIVFFlat_add_with_ids(h, n, x, nullptr));
    }
    void add_with_ids(idx_t n, const float* x, const idx_t* xids) {
        check(b200_IndexIVFFlat_add_with_ids(h, n, x, xids));
    }
    void reset() {
        check(b200_IndexIVFFlat_reset(h));
    }

    void search(idx_t n, const float* x, idx_t k, float* distances, idx_t* labels,
                const SearchParametersIVF* params = nullptr) const {
        if (!params) {
            b200_IndexIVFFlat_set_nprobe(h, nprobe);
            b200_IndexIVFFlat_set_max_codes(h, max_codes);
            check(b200_IndexIVFFlat_search(h, n, x, k, distances, labels));
            return;
        }
        Params p(*params);
        check(b200_IndexIVFFlat_search_with_params(h, n, x, k, p.h, distances, labels));
    }
    void search_preassigned(idx_t n, const float* x, idx_t k, const idx_t* assign,
                            const float* centroid_dis, float* distances, idx_t* labels,
                            bool store_pairs, const SearchParametersIVF* params = nullptr) const {
        SearchParametersIVF own;
        own.nprobe = nprobe;
        own.max_codes = max_codes;
        Params p(params ? *params : own);
        check(b200_IndexIVFFlat_search_preassigned_with_params(
                h, n, x, k, assign, centroid_dis, distances, labels, store_pairs ? 1 : 0, p.h));
    }
    // the fork's IndexIVF::search_stats
    void search_stats(idx_t n, const float* x, idx_t k, float* distances, idx_t* labels,
                      const SearchParametersIVF* params = nullptr,
                      QueryLatencyStats* per_query_stats = nullptr) const {
        SearchParametersIVF own;
        own.nprobe = nprobe;
        own.max_codes = max_codes;
        Params p(params ? *params : own);
        check(b200_IndexIVFFlat_search_stats(h, n, x, k, distances, labels, p.h, per_query_stats));
    }
    void range_search(idx_t n, const float* x, float radius, RangeSearchResult* result,
                      const SearchParametersIVF* params = nullptr) const {
        SearchParametersIVF own;
        own.nprobe = nprobe;
        Params p(params ? *params : own);
        B200RangeSearchResult* r = nullptr;
        check(b200_RangeSearchResult_new(&r, n));
        const int rc = b200_IndexIVFFlat_range_search(h, n, x, radius, r, p.h);
        if (rc == 0) {
            size_t* lims;
            idx_t* lab;
            float* dis;
            b200_RangeSearchResult_lims(r, &lims);
            b200_RangeSearchResult_labels(r, &lab, &dis);
            result->nq = (size_t)n;
            result->lims.assign(lims, lims + n + 1);
            result->labels.assign(lab, lab + lims[n]);
            result->distances.assign(dis, dis + lims[n]);
        }
        b200_RangeSearchResult_free(r);
        check(rc);
    }

  private:
    struct Params { // RAII over the C handle of one call's parameters
        B200SearchParametersIVF* h = nullptr;
        explicit Params(const SearchParametersIVF& p) {
            check(b200_SearchParametersIVF_new_with(&h, p.sel ? p.sel->h : nullptr, p.nprobe,
                                                    p.max_codes));
        }
        ~Params() {
            b200_SearchParametersIVF_free(h);
        }
    };
};

// ---- one rank of the list-sharded index: faiss::IndexShardsIVF (faiss/IndexShardsIVF.h:21-44) with
// one process per GPU instead of one thread per shard. Every method is collective over the ranks
// that share `nccl_unique_id` (rank 0 makes it with nccl_unique_id() and hands the 128 bytes to the
// others by whatever channel the application has).
inline void nccl_unique_id(unsigned char (&out)[B200_NCCL_UNIQUE_ID_BYTES]) {
    check(b200_nccl_get_unique_id(out));
}

struct IndexShardsIVFFlat {
    B200IndexShardsIVFFlat* h = nullptr;
    size_t nprobe = 1;

    IndexShardsIVFFlat(size_t d, size_t nlist, MetricType metric, int device, int rank, int world,
                       const void* nccl_unique_id128) {
        check(b200_IndexShardsIVFFlat_new(&h, d, nlist, (B200MetricType)metric, device, rank, world,
                                          nccl_unique_id128));
    }
    IndexShardsIVFFlat(const IndexShardsIVFFlat&) = delete;
    IndexShardsIVFFlat& operator=(const IndexShardsIVFFlat&) = delete;
    ~IndexShardsIVFFlat() {
        b200_IndexShardsIVFFlat_free(h);
    }
    int rank() const {
        return b200_IndexShardsIVFFlat_rank(h);
    }
    int world() const {
        return b200_IndexShardsIVFFlat_world(h);
    }
    idx_t ntotal() const {
        return b200_IndexShardsIVFFlat_ntotal(h);
    }
    // IndexShards::at (faiss/IndexShards.h): this rank's shard, borrowed (options, stats, lists)
    B200IndexIVFFlat* local() const {
        return b200_IndexShardsIVFFlat_local(h);
    }
    void train(idx_t n, const float* x, int root = 0) {
        check(b200_IndexShardsIVFFlat_train(h, n, x, root));
    }
    void set_centroids(const float* centroids, int root = -1) {
        check(b200_IndexShardsIVFFlat_set_centroids(h, centroids, root));
    }
    // each rank brings ITS OWN vectors; they travel to the rank that owns their list
    void add_with_ids(idx_t n, const float* x, const idx_t* xids) {
        check(b200_IndexShardsIVFFlat_add_with_ids(h, n, x, xids));
    }
    // root < 0: every rank passes the same queries; root >= 0: rank root's are broadcast
    void search(idx_t n, const float* x, idx_t k, float* distances, idx_t* labels,
                int root = -1) const {
        b200_IndexIVFFlat_set_nprobe(b200_IndexShardsIVFFlat_local(h), nprobe);
        check(b200_IndexShardsIVFFlat_search(h, n, x, k, distances, labels, root));
    }
    void synchronize() const {
        check(b200_IndexShardsIVFFlat_synchronize(h));
    }
};

} // namespace b200
