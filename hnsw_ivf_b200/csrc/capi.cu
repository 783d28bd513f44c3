// capi.cu -- the extern "C" layer declared in include/b200ivf_c.h. Same shape as the reference's C
// API: reinterpret_cast of an opaque handle plus try/catch mapped to int codes
// (c_api/macros_impl.h:22-55), thread-local last error (c_api/error_impl.cpp:15-27).
#include "../../include/b200ivf_c.h"

#include "index.h"
#include "kernels.h"
#include "selector.h"
#include "shards.h"

#include <exception>
#include <string>

using b200ivf::IndexIVFFlatB200;

namespace {
thread_local std::string g_last_error;

template <class F>
int guarded(F&& f) {
    try {
        f();
        return 0;
    } catch (const b200ivf::B200Exception& e) {
        g_last_error = e.what();
        return -2; // FAISS_EXCEPT
    } catch (const std::exception& e) {
        g_last_error = e.what();
        return -4; // STD_EXCEPT
    } catch (...) {
        g_last_error = "Unknown error";
        return -1; // UNKNOWN_EXCEPT
    }
}

inline IndexIVFFlatB200* H(B200IndexIVFFlat* p) {
    return reinterpret_cast<IndexIVFFlatB200*>(p);
}
inline const IndexIVFFlatB200* H(const B200IndexIVFFlat* p) {
    return reinterpret_cast<const IndexIVFFlatB200*>(p);
}
inline b200ivf::IndexShardsIVFFlatB200* SH(B200IndexShardsIVFFlat* p) {
    return reinterpret_cast<b200ivf::IndexShardsIVFFlatB200*>(p);
}
inline const b200ivf::IndexShardsIVFFlatB200* SH(const B200IndexShardsIVFFlat* p) {
    return reinterpret_cast<const b200ivf::IndexShardsIVFFlatB200*>(p);
}
inline const b200ivf::IDSelector* S(const B200IDSelector* p) {
    return reinterpret_cast<const b200ivf::IDSelector*>(p);
}
inline const b200ivf::SearchParametersIVF* P(const B200SearchParametersIVF* p) {
    return reinterpret_cast<const b200ivf::SearchParametersIVF*>(p);
}
inline b200ivf::RangeSearchResult* R(B200RangeSearchResult* p) {
    return reinterpret_cast<b200ivf::RangeSearchResult*>(p);
}
template <class T>
void put_selector(B200IDSelector** p_sel, T* sel) {
    if (!p_sel) {
        delete sel;
        throw b200ivf::B200Exception("null output handle");
    }
    *p_sel = reinterpret_cast<B200IDSelector*>(static_cast<b200ivf::IDSelector*>(sel));
}
static_assert(sizeof(B200QueryLatencyStats) == sizeof(b200ivf::QueryLatencyStats),
              "QueryLatencyStats layout");
} // namespace

extern "C" {

const char* b200_get_last_error(void) {
    return g_last_error.c_str();
}

int b200_IndexIVFFlat_new_with_metric(B200IndexIVFFlat** p_index, size_t d, size_t nlist,
                                      B200MetricType metric, int device) {
    return guarded([&] {
        if (!p_index)
            throw b200ivf::B200Exception("null output handle");
        *p_index = reinterpret_cast<B200IndexIVFFlat*>(
                new IndexIVFFlatB200(d, nlist, (int)metric, device));
    });
}

void b200_IndexIVFFlat_free(B200IndexIVFFlat* index) {
    delete H(index);
}

size_t b200_IndexIVFFlat_nlist(const B200IndexIVFFlat* index) {
    return H(index)->nlist;
}
size_t b200_IndexIVFFlat_nprobe(const B200IndexIVFFlat* index) {
    return H(index)->nprobe;
}
void b200_IndexIVFFlat_set_nprobe(B200IndexIVFFlat* index, size_t nprobe) {
    H(index)->nprobe = nprobe;
}
int b200_IndexIVFFlat_d(const B200IndexIVFFlat* index) {
    return H(index)->d;
}
b200_idx_t b200_IndexIVFFlat_ntotal(const B200IndexIVFFlat* index) {
    return H(index)->ntotal;
}
int b200_IndexIVFFlat_is_trained(const B200IndexIVFFlat* index) {
    return H(index)->is_trained ? 1 : 0;
}
B200MetricType b200_IndexIVFFlat_metric_type(const B200IndexIVFFlat* index) {
    return (B200MetricType)H(index)->metric_type;
}

int b200_IndexIVFFlat_set_centroids(B200IndexIVFFlat* index, const float* centroids) {
    return guarded([&] { H(index)->set_centroids(centroids); });
}
int b200_IndexIVFFlat_get_centroids(const B200IndexIVFFlat* index, float* out) {
    return guarded([&] { H(index)->get_centroids(out); });
}
int b200_IndexIVFFlat_set_list_range(B200IndexIVFFlat* index, size_t lo, size_t hi) {
    return guarded([&] { H(index)->set_list_range(lo, hi); });
}
int b200_IndexIVFFlat_add_with_ids(B200IndexIVFFlat* index, b200_idx_t n, const float* x,
                                   const b200_idx_t* xids) {
    return guarded([&] { H(index)->add_with_ids(n, x, xids); });
}
int b200_IndexIVFFlat_add_core(B200IndexIVFFlat* index, b200_idx_t n, const float* x,
                               const b200_idx_t* xids, const int64_t* precomputed_idx) {
    return guarded([&] {
        if (!precomputed_idx) // FAISS_THROW_IF_NOT(coarse_idx), faiss/IndexIVFFlat.cpp:53
            throw b200ivf::B200Exception("Error: 'coarse_idx' failed");
        H(index)->add_core(n, x, xids, precomputed_idx);
    });
}
int b200_IndexIVFFlat_reset(B200IndexIVFFlat* index) {
    return guarded([&] { H(index)->reset(); });
}

int b200_IndexIVFFlat_reserve_lists(B200IndexIVFFlat* index, const int64_t* sizes) {
    return guarded([&] { H(index)->reserve_lists(sizes); });
}
int b200_IndexIVFFlat_get_list_sizes(const B200IndexIVFFlat* index, int64_t* sizes_out) {
    return guarded([&] {
        if (!sizes_out)
            throw b200ivf::B200Exception("null output buffer");
        H(index)->get_list_sizes(sizes_out);
    });
}
size_t b200_IndexIVFFlat_get_list_size(const B200IndexIVFFlat* index, size_t list_no) {
    size_t r = 0;
    guarded([&] { r = H(index)->get_list_size(list_no); });
    return r;
}
int b200_IndexIVFFlat_invlists_get_ids(const B200IndexIVFFlat* index, size_t list_no,
                                       b200_idx_t* invlist) {
    return guarded([&] { H(index)->get_list_ids(list_no, invlist); });
}
int b200_IndexIVFFlat_invlists_get_codes(const B200IndexIVFFlat* index, size_t list_no,
                                         float* codes) {
    return guarded([&] { H(index)->get_list_codes(list_no, codes); });
}

int b200_IndexIVFFlat_search(const B200IndexIVFFlat* index, b200_idx_t n, const float* x,
                             b200_idx_t k, float* distances, b200_idx_t* labels) {
    return guarded([&] { H(index)->search(n, x, k, distances, labels); });
}
int b200_IndexIVFFlat_search_preassigned(const B200IndexIVFFlat* index, b200_idx_t n,
                                         const float* x, b200_idx_t k, const b200_idx_t* assign,
                                         const float* centroid_dis, float* distances,
                                         b200_idx_t* labels, int store_pairs) {
    return guarded([&] {
        if (!assign)
            throw b200ivf::B200Exception("null assign table");
        H(index)->search_preassigned(n, x, k, assign, centroid_dis, distances, labels,
                                     store_pairs != 0);
    });
}
int b200_IndexIVFFlat_quantize(const B200IndexIVFFlat* index, b200_idx_t n, const float* x,
                               b200_idx_t nprobe, float* centroid_dis, b200_idx_t* assign) {
    return guarded([&] { H(index)->quantize(n, x, nprobe, centroid_dis, assign); });
}

int b200_IndexIVFFlat_quantize_noalloc(const B200IndexIVFFlat* index, b200_idx_t n, const float* x,
                                       b200_idx_t nprobe, float* centroid_dis, b200_idx_t* assign) {
    return guarded([&] { H(index)->quantize_noalloc(n, x, nprobe, centroid_dis, assign); });
}

void b200_IndexIVFFlat_get_stats(const B200IndexIVFFlat* index, B200IndexIVFStats* stats) {
    b200ivf::IndexIVFStatsB200 s;
    guarded([&] { s = H(index)->get_stats(); }); // (a deferred failure stays in last_error)
    stats->nq = s.nq;
    stats->nlist = s.nlist;
    stats->ndis = s.ndis;
    stats->nheap_updates = s.nheap_updates;
    stats->quantization_time = s.quantization_time;
    stats->search_time = s.search_time;
}
void b200_IndexIVFFlat_reset_stats(B200IndexIVFFlat* index) {
    guarded([&] { H(index)->reset_stats(); });
}
void b200_IndexIVFFlat_last_timings(const B200IndexIVFFlat* index, double* ms5) {
    for (int i = 0; i < 5; ++i)
        ms5[i] = 0;
    guarded([&] { H(index)->last_timings(ms5); });
}
unsigned long long b200_IndexIVFFlat_stage_time_sums(const B200IndexIVFFlat* index, double* ms5) {
    unsigned long long r = 0;
    for (int i = 0; i < 5; ++i)
        ms5[i] = 0;
    guarded([&] { r = H(index)->stage_time_sums(ms5); });
    return r;
}
unsigned long long b200_IndexIVFFlat_exact_redo_count(const B200IndexIVFFlat* index) {
    unsigned long long r = 0;
    guarded([&] { r = H(index)->exact_redo_count(); });
    return r;
}
int b200_IndexIVFFlat_synchronize(const B200IndexIVFFlat* index) {
    return guarded([&] { H(index)->synchronize(); });
}
int b200_IndexIVFFlat_set_stream(B200IndexIVFFlat* index, void* cuda_stream) {
    return guarded([&] { H(index)->set_stream((cudaStream_t)cuda_stream); });
}

int b200_IndexIVFFlat_set_option(B200IndexIVFFlat* index, const char* name, double value) {
    return guarded([&] { H(index)->set_option(name, value); });
}

unsigned long long b200_tensor_core_launches(void) {
    return b200ivf::g_tc_launches.load();
}

int b200_debug_trap_info(unsigned long long* out8) {
    const unsigned long long* h = b200ivf::tc_trap_info_host();
    for (int i = 0; i < 8; ++i)
        out8[i] = h ? h[i] : 0ull;
    return h && h[0] ? 1 : 0;
}

unsigned long long b200_kernel_launches(void) {
    return b200ivf::g_kernel_launches.load();
}

int b200_merge_knn_results(B200MetricType metric, int nshard, b200_idx_t n, b200_idx_t k,
                           const float* all_distances, const b200_idx_t* all_labels,
                           float* distances, b200_idx_t* labels, int device, void* cuda_stream) {
    return guarded([&] {
        b200ivf::merge_knn_results((int)metric, nshard, n, k, all_distances, all_labels, distances,
                                   labels, device, (cudaStream_t)cuda_stream);
    });
}

int b200_IndexIVFFlat_train(B200IndexIVFFlat* index, b200_idx_t n, const float* x) {
    return guarded([&] { H(index)->train(n, x); });
}

int b200_IDSelectorRange_new(B200IDSelector** p_sel, b200_idx_t imin, b200_idx_t imax) {
    return guarded([&] { put_selector(p_sel, new b200ivf::IDSelectorRange(imin, imax)); });
}
int b200_IDSelectorBatch_new(B200IDSelector** p_sel, size_t n, const b200_idx_t* indices,
                             int device) {
    return guarded([&] { put_selector(p_sel, new b200ivf::IDSelectorBatch(n, indices, device)); });
}
int b200_IDSelectorBitmap_new(B200IDSelector** p_sel, size_t n, const uint8_t* bitmap,
                              int device) {
    return guarded([&] { put_selector(p_sel, new b200ivf::IDSelectorBitmap(n, bitmap, device)); });
}
int b200_IDSelectorNot_new(B200IDSelector** p_sel, const B200IDSelector* sel) {
    return guarded([&] {
        if (!sel)
            throw b200ivf::B200Exception("null selector");
        put_selector(p_sel, new b200ivf::IDSelectorNot(S(sel)));
    });
}
static int binary_selector_new(B200IDSelector** p_sel, b200ivf::IDSelectorBinary::Op op,
                               const B200IDSelector* lhs, const B200IDSelector* rhs) {
    return guarded([&] {
        if (!lhs || !rhs)
            throw b200ivf::B200Exception("null selector");
        put_selector(p_sel, new b200ivf::IDSelectorBinary(op, S(lhs), S(rhs)));
    });
}
int b200_IDSelectorAnd_new(B200IDSelector** p_sel, const B200IDSelector* lhs,
                           const B200IDSelector* rhs) {
    return binary_selector_new(p_sel, b200ivf::IDSelectorBinary::AND, lhs, rhs);
}
int b200_IDSelectorOr_new(B200IDSelector** p_sel, const B200IDSelector* lhs,
                          const B200IDSelector* rhs) {
    return binary_selector_new(p_sel, b200ivf::IDSelectorBinary::OR, lhs, rhs);
}
int b200_IDSelectorXOr_new(B200IDSelector** p_sel, const B200IDSelector* lhs,
                           const B200IDSelector* rhs) {
    return binary_selector_new(p_sel, b200ivf::IDSelectorBinary::XOR, lhs, rhs);
}
int b200_IDSelector_is_member(const B200IDSelector* sel, b200_idx_t id) {
    return S(sel)->is_member(id) ? 1 : 0;
}
void b200_IDSelector_free(B200IDSelector* sel) {
    delete reinterpret_cast<b200ivf::IDSelector*>(sel);
}

int b200_SearchParametersIVF_new(B200SearchParametersIVF** p_sp) {
    return b200_SearchParametersIVF_new_with(p_sp, nullptr, 1, 0); // IndexIVF.h:74-76 defaults
}
int b200_SearchParametersIVF_new_with(B200SearchParametersIVF** p_sp, B200IDSelector* sel,
                                      size_t nprobe, size_t max_codes) {
    return guarded([&] {
        if (!p_sp)
            throw b200ivf::B200Exception("null output handle");
        auto* sp = new b200ivf::SearchParametersIVF;
        sp->sel = S(sel);
        sp->nprobe = nprobe;
        sp->max_codes = max_codes;
        *p_sp = reinterpret_cast<B200SearchParametersIVF*>(sp);
    });
}
void b200_SearchParametersIVF_free(B200SearchParametersIVF* sp) {
    delete reinterpret_cast<b200ivf::SearchParametersIVF*>(sp);
}
const B200IDSelector* b200_SearchParametersIVF_sel(const B200SearchParametersIVF* sp) {
    return reinterpret_cast<const B200IDSelector*>(P(sp)->sel);
}
size_t b200_SearchParametersIVF_nprobe(const B200SearchParametersIVF* sp) {
    return P(sp)->nprobe;
}
void b200_SearchParametersIVF_set_nprobe(B200SearchParametersIVF* sp, size_t nprobe) {
    reinterpret_cast<b200ivf::SearchParametersIVF*>(sp)->nprobe = nprobe;
}
size_t b200_SearchParametersIVF_max_codes(const B200SearchParametersIVF* sp) {
    return P(sp)->max_codes;
}
void b200_SearchParametersIVF_set_max_codes(B200SearchParametersIVF* sp, size_t max_codes) {
    reinterpret_cast<b200ivf::SearchParametersIVF*>(sp)->max_codes = max_codes;
}

size_t b200_IndexIVFFlat_max_codes(const B200IndexIVFFlat* index) {
    return H(index)->max_codes;
}
void b200_IndexIVFFlat_set_max_codes(B200IndexIVFFlat* index, size_t max_codes) {
    H(index)->max_codes = max_codes;
}

int b200_IndexIVFFlat_search_with_params(const B200IndexIVFFlat* index, b200_idx_t n,
                                         const float* x, b200_idx_t k,
                                         const B200SearchParametersIVF* params, float* distances,
                                         b200_idx_t* labels) {
    return guarded([&] { H(index)->search(n, x, k, distances, labels, P(params)); });
}
int b200_IndexIVFFlat_search_preassigned_with_params(
        const B200IndexIVFFlat* index, b200_idx_t n, const float* x, b200_idx_t k,
        const b200_idx_t* assign, const float* centroid_dis, float* distances, b200_idx_t* labels,
        int store_pairs, const B200SearchParametersIVF* params) {
    return guarded([&] {
        H(index)->search_preassigned(n, x, k, assign, centroid_dis, distances, labels,
                                     store_pairs != 0, P(params));
    });
}
int b200_IndexIVFFlat_search_stats(const B200IndexIVFFlat* index, b200_idx_t n, const float* x,
                                   b200_idx_t k, float* distances, b200_idx_t* labels,
                                   const B200SearchParametersIVF* params,
                                   B200QueryLatencyStats* per_query_stats) {
    return guarded([&] {
        H(index)->search(n, x, k, distances, labels, P(params),
                         reinterpret_cast<b200ivf::QueryLatencyStats*>(per_query_stats));
    });
}

int b200_RangeSearchResult_new(B200RangeSearchResult** p_rsr, b200_idx_t nq) {
    return guarded([&] {
        if (!p_rsr || nq < 0)
            throw b200ivf::B200Exception("invalid RangeSearchResult arguments");
        *p_rsr = reinterpret_cast<B200RangeSearchResult*>(
                new b200ivf::RangeSearchResult((size_t)nq));
    });
}
void b200_RangeSearchResult_free(B200RangeSearchResult* rsr) {
    delete R(rsr);
}
size_t b200_RangeSearchResult_nq(const B200RangeSearchResult* rsr) {
    return reinterpret_cast<const b200ivf::RangeSearchResult*>(rsr)->nq;
}
void b200_RangeSearchResult_lims(B200RangeSearchResult* rsr, size_t** lims) {
    *lims = R(rsr)->lims.data();
}
void b200_RangeSearchResult_labels(B200RangeSearchResult* rsr, b200_idx_t** labels,
                                   float** distances) {
    *labels = R(rsr)->labels.data();
    *distances = R(rsr)->distances.data();
}
int b200_IndexIVFFlat_range_search(const B200IndexIVFFlat* index, b200_idx_t n, const float* x,
                                   float radius, B200RangeSearchResult* result,
                                   const B200SearchParametersIVF* params) {
    return guarded([&] { H(index)->range_search(n, x, radius, R(result), P(params)); });
}
int b200_IndexIVFFlat_range_search_preassigned(const B200IndexIVFFlat* index, b200_idx_t n,
                                               const float* x, float radius,
                                               const b200_idx_t* assign, const float* centroid_dis,
                                               B200RangeSearchResult* result, int store_pairs,
                                               const B200SearchParametersIVF* params) {
    return guarded([&] {
        H(index)->range_search_preassigned(n, x, radius, assign, centroid_dis, R(result),
                                           store_pairs != 0, P(params));
    });
}

int b200_nccl_get_unique_id(void* out128) {
    return guarded([&] {
        if (!out128)
            throw b200ivf::B200Exception("null output buffer");
        b200ivf::nccl_get_unique_id(out128);
    });
}
int b200_IndexShardsIVFFlat_new(B200IndexShardsIVFFlat** p_index, size_t d, size_t nlist,
                                B200MetricType metric, int device, int rank, int world,
                                const void* nccl_unique_id128) {
    return guarded([&] {
        if (!p_index)
            throw b200ivf::B200Exception("null output handle");
        *p_index = reinterpret_cast<B200IndexShardsIVFFlat*>(new b200ivf::IndexShardsIVFFlatB200(
                d, nlist, (int)metric, device, rank, world, nccl_unique_id128));
    });
}
void b200_IndexShardsIVFFlat_free(B200IndexShardsIVFFlat* index) {
    delete SH(index);
}
B200IndexIVFFlat* b200_IndexShardsIVFFlat_local(B200IndexShardsIVFFlat* index) {
    return reinterpret_cast<B200IndexIVFFlat*>(SH(index)->local());
}
int b200_IndexShardsIVFFlat_rank(const B200IndexShardsIVFFlat* index) {
    return SH(index)->rank();
}
int b200_IndexShardsIVFFlat_world(const B200IndexShardsIVFFlat* index) {
    return SH(index)->world();
}
b200_idx_t b200_IndexShardsIVFFlat_ntotal(const B200IndexShardsIVFFlat* index) {
    return SH(index)->ntotal();
}
int b200_IndexShardsIVFFlat_set_centroids(B200IndexShardsIVFFlat* index, const float* centroids,
                                          int root) {
    return guarded([&] { SH(index)->set_centroids(centroids, root); });
}
int b200_IndexShardsIVFFlat_train(B200IndexShardsIVFFlat* index, b200_idx_t n, const float* x,
                                  int root) {
    return guarded([&] { SH(index)->train(n, x, root); });
}
int b200_IndexShardsIVFFlat_add_with_ids(B200IndexShardsIVFFlat* index, b200_idx_t n,
                                         const float* x, const b200_idx_t* xids) {
    return guarded([&] { SH(index)->add_core(n, x, xids, nullptr); });
}
int b200_IndexShardsIVFFlat_add_core(B200IndexShardsIVFFlat* index, b200_idx_t n, const float* x,
                                     const b200_idx_t* xids, const int64_t* precomputed_idx) {
    return guarded([&] {
        if (!precomputed_idx && n > 0)
            throw b200ivf::B200Exception("Error: 'coarse_idx' failed");
        SH(index)->add_core(n, x, xids, precomputed_idx);
    });
}
int b200_IndexShardsIVFFlat_search(const B200IndexShardsIVFFlat* index, b200_idx_t n,
                                   const float* x, b200_idx_t k, float* distances,
                                   b200_idx_t* labels, int root) {
    return guarded([&] { SH(index)->search(n, x, k, distances, labels, root); });
}
unsigned long long b200_IndexShardsIVFFlat_phase_time_sums(const B200IndexShardsIVFFlat* index,
                                                           double* ms7) {
    unsigned long long r = 0;
    for (int i = 0; i < 7; ++i)
        ms7[i] = 0;
    guarded([&] { r = SH(index)->phase_time_sums(ms7); });
    return r;
}
void b200_IndexShardsIVFFlat_reset_phase_times(B200IndexShardsIVFFlat* index) {
    guarded([&] { SH(index)->reset_phase_times(); });
}
int b200_IndexShardsIVFFlat_synchronize(const B200IndexShardsIVFFlat* index) {
    return guarded([&] { SH(index)->synchronize(); });
}

int b200_write_index_fname(const B200IndexIVFFlat* index, const char* fname) {
    return guarded([&] { b200ivf::write_index(H(index), fname); });
}

int b200_read_index_fname(const char* fname, int io_flags, int device, B200IndexIVFFlat** p_out) {
    return guarded([&] {
        if (!p_out)
            throw b200ivf::B200Exception("null output handle");
        *p_out = reinterpret_cast<B200IndexIVFFlat*>(b200ivf::read_index(fname, io_flags, device));
    });
}

} // extern "C"
