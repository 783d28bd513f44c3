// oracle/ref_shim.cpp -- TEST INFRASTRUCTURE, not product code.
//
// A thin C-ABI over the UNMODIFIED reference CPU library (faiss 1.10 fork, compiled from
// /root/reference by oracle/Makefile into oracle/_ref/libfaiss_ref.a). It exists so that the
// Python tests and bench.py's cpu_baseline / --impl reference legs can drive the reference's own
// IndexIVFFlat through ctypes. Nothing here re-implements reference logic: every function is a
// direct call into the reference's public C++ API:
//   faiss::IndexFlat            faiss/IndexFlat.h:23      (coarse quantizer)
//   faiss::IndexIVFFlat         faiss/IndexIVFFlat.h:26
//   IndexIVF::search            faiss/IndexIVF.cpp:303
//   IndexIVF::search_preassigned faiss/IndexIVF.cpp:399
//   merge_knn_results           faiss/utils/Heap.cpp:158
//   knn_L2sqr/knn_inner_product faiss/utils/distances.cpp:885,836
//   write_index / read_index    faiss/impl/index_write.cpp:391, faiss/impl/index_read.cpp:593
//   IDSelector*                 faiss/impl/IDSelector.h:19-170
//   IndexIVF::range_search*     faiss/IndexIVF.cpp:1203-1399
//   IndexIVF::search_stats      faiss/IndexIVF.cpp:727 (fork)
// Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may load this library.

#include <faiss/IndexFlat.h>
#include <faiss/IndexIVFFlat.h>
#include <faiss/impl/FaissAssert.h>
#include <faiss/impl/AuxIndexStructures.h>
#include <faiss/impl/FaissException.h>
#include <faiss/impl/IDSelector.h>
#include <faiss/index_io.h>
#include <faiss/invlists/InvertedLists.h>
#include <faiss/utils/Heap.h>
#include <faiss/utils/distances.h>
#include <faiss/utils/utils.h>

#include <omp.h>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

namespace {
thread_local std::string g_err;

struct RefIndex {
    std::unique_ptr<faiss::IndexFlat> quantizer;
    std::unique_ptr<faiss::IndexIVFFlat> index;
};

template <class F>
int guarded(F&& f) {
    try {
        f();
        return 0;
    } catch (const faiss::FaissException& e) {
        g_err = e.what();
        return -2;
    } catch (const std::exception& e) {
        g_err = e.what();
        return -4;
    } catch (...) {
        g_err = "unknown";
        return -1;
    }
}
} // namespace

extern "C" {

const char* ref_last_error() {
    return g_err.c_str();
}

void ref_set_omp_threads(int nt) {
    omp_set_num_threads(nt);
}

int ref_omp_max_threads() {
    return omp_get_max_threads();
}

double ref_getmillisecs() {
    return faiss::getmillisecs();
}

// metric: 0 = METRIC_INNER_PRODUCT, 1 = METRIC_L2 (faiss/MetricType.h:23-24)
void* ref_ivfflat_new(int64_t d, int64_t nlist, int metric) {
    auto* r = new RefIndex;
    auto mt = (faiss::MetricType)metric;
    r->quantizer.reset(new faiss::IndexFlat(d, mt));
    r->index.reset(new faiss::IndexIVFFlat(r->quantizer.get(), d, nlist, mt));
    r->index->own_fields = false;
    return r;
}

void ref_ivfflat_free(void* h) {
    delete (RefIndex*)h;
}

// install centroids without training: quantizer->add(nlist, c). IndexIVF::train then sees
// quantizer->is_trained && ntotal == nlist and skips k-means (faiss/IndexIVF.cpp:72-74).
int ref_ivfflat_set_centroids(void* h, const float* c) {
    auto* r = (RefIndex*)h;
    return guarded([&] {
        r->quantizer->reset();
        r->quantizer->add(r->index->nlist, c);
        r->index->is_trained = true;
    });
}

int ref_ivfflat_train(void* h, int64_t n, const float* x) {
    auto* r = (RefIndex*)h;
    return guarded([&] { r->index->train(n, x); });
}

int ref_ivfflat_get_centroids(void* h, float* out) {
    auto* r = (RefIndex*)h;
    return guarded([&] {
        std::memcpy(out, r->quantizer->get_xb(),
                    sizeof(float) * r->index->nlist * r->index->d);
    });
}

int ref_ivfflat_add_with_ids(void* h, int64_t n, const float* x, const int64_t* ids) {
    auto* r = (RefIndex*)h;
    return guarded([&] {
        if (ids)
            r->index->add_with_ids(n, x, ids);
        else
            r->index->add(n, x);
    });
}

int ref_ivfflat_add_core(void* h, int64_t n, const float* x, const int64_t* ids,
                         const int64_t* assign) {
    auto* r = (RefIndex*)h;
    return guarded([&] { r->index->add_core(n, x, ids, assign); });
}

// bulk install of one inverted list, bypassing assignment (InvertedLists::add_entries)
int ref_ivfflat_append_list(void* h, int64_t list_no, int64_t n, const int64_t* ids,
                            const float* codes) {
    auto* r = (RefIndex*)h;
    return guarded([&] {
        r->index->invlists->add_entries(list_no, n, ids, (const uint8_t*)codes);
        r->index->ntotal += n;
    });
}

int64_t ref_ivfflat_ntotal(void* h) {
    return ((RefIndex*)h)->index->ntotal;
}

int64_t ref_ivfflat_list_size(void* h, int64_t l) {
    return ((RefIndex*)h)->index->invlists->list_size(l);
}

int ref_ivfflat_get_list(void* h, int64_t l, float* codes, int64_t* ids) {
    auto* r = (RefIndex*)h;
    return guarded([&] {
        auto* il = r->index->invlists;
        size_t n = il->list_size(l);
        faiss::InvertedLists::ScopedCodes sc(il, l);
        faiss::InvertedLists::ScopedIds si(il, l);
        if (codes)
            std::memcpy(codes, sc.get(), n * il->code_size);
        if (ids)
            std::memcpy(ids, si.get(), n * sizeof(int64_t));
    });
}

int ref_ivfflat_quantize(void* h, int64_t n, const float* x, int64_t nprobe, float* Dq,
                         int64_t* Iq) {
    auto* r = (RefIndex*)h;
    return guarded([&] { r->quantizer->search(n, x, nprobe, Dq, Iq); });
}

int ref_ivfflat_search(void* h, int64_t n, const float* x, int64_t k, int64_t nprobe, float* D,
                       int64_t* I) {
    auto* r = (RefIndex*)h;
    return guarded([&] {
        r->index->nprobe = nprobe;
        r->index->search(n, x, k, D, I);
    });
}

int ref_ivfflat_search_preassigned(void* h, int64_t n, const float* x, int64_t k, int64_t nprobe,
                                   const int64_t* assign, const float* cdis, float* D,
                                   int64_t* I) {
    auto* r = (RefIndex*)h;
    return guarded([&] {
        r->index->nprobe = nprobe;
        r->index->search_preassigned(n, x, k, assign, cdis, D, I, false);
    });
}

// faiss::write_index(index, fname): the reference's own file of this index
int ref_ivfflat_write_index(void* h, const char* fname) {
    auto* r = (RefIndex*)h;
    return guarded([&] { faiss::write_index(r->index.get(), fname); });
}

// faiss::read_index(fname) -> handle (NULL on failure or if the file is not an IndexIVFFlat over
// an IndexFlat quantizer). out5 = {d, nlist, metric, nprobe, ntotal}.
void* ref_read_index(const char* fname, int64_t* out5) {
    RefIndex* r = nullptr;
    const int rc = guarded([&] {
        std::unique_ptr<faiss::Index> idx(faiss::read_index(fname));
        auto* ivf = dynamic_cast<faiss::IndexIVFFlat*>(idx.get());
        FAISS_THROW_IF_NOT_MSG(ivf, "not an IndexIVFFlat file");
        auto* q = dynamic_cast<faiss::IndexFlat*>(ivf->quantizer);
        FAISS_THROW_IF_NOT_MSG(q, "quantizer is not an IndexFlat");
        r = new RefIndex;
        ivf->own_fields = false; // the quantizer is owned by the RefIndex below
        r->quantizer.reset(q);
        idx.release();
        r->index.reset(ivf);
        out5[0] = ivf->d;
        out5[1] = (int64_t)ivf->nlist;
        out5[2] = (int64_t)ivf->metric_type;
        out5[3] = (int64_t)ivf->nprobe;
        out5[4] = ivf->ntotal;
    });
    return rc == 0 ? r : nullptr;
}

void ref_ivfflat_set_nprobe(void* h, int64_t nprobe) {
    ((RefIndex*)h)->index->nprobe = (size_t)nprobe;
}

void ref_ivfflat_set_parallel_mode(void* h, int pm) {
    ((RefIndex*)h)->index->parallel_mode = pm;
}

// layout: nq, nlist, ndis, nheap_updates (as double), quantization_time, search_time
void ref_stats_get(double* out6) {
    auto& s = faiss::indexIVF_stats;
    out6[0] = (double)s.nq;
    out6[1] = (double)s.nlist;
    out6[2] = (double)s.ndis;
    out6[3] = (double)s.nheap_updates;
    out6[4] = s.quantization_time;
    out6[5] = s.search_time;
}

void ref_stats_reset() {
    faiss::indexIVF_stats.reset();
}

int ref_knn(int metric, const float* x, const float* y, int64_t d, int64_t nx, int64_t ny,
            int64_t k, float* D, int64_t* I) {
    return guarded([&] {
        if (metric == 0)
            faiss::knn_inner_product(x, y, d, nx, ny, k, D, I);
        else
            faiss::knn_L2sqr(x, y, d, nx, ny, k, D, I);
    });
}

float ref_fvec_L2sqr(const float* x, const float* y, int64_t d) {
    return faiss::fvec_L2sqr(x, y, d);
}

float ref_fvec_inner_product(const float* x, const float* y, int64_t d) {
    return faiss::fvec_inner_product(x, y, d);
}

// IndexShardsIVF-style merge (faiss/IndexShardsIVF.cpp:224-242): L2 -> CMin<float,int>, IP -> CMax
int ref_merge_knn_results(int metric, int64_t n, int64_t k, int nshard, const float* all_D,
                          const int64_t* all_I, float* D, int64_t* I) {
    return guarded([&] {
        if (metric == 1)
            faiss::merge_knn_results<faiss::idx_t, faiss::CMin<float, int>>(
                    n, k, nshard, all_D, all_I, D, I);
        else
            faiss::merge_knn_results<faiss::idx_t, faiss::CMax<float, int>>(
                    n, k, nshard, all_D, all_I, D, I);
    });
}

// heap primitives exposed for known-answer checks of the restatement (faiss/utils/Heap.h)
void ref_heap_select(int metric, int64_t k, int64_t n, const float* vals, const int64_t* ids,
                     float* out_v, int64_t* out_i) {
    if (metric == 1) {
        using C = faiss::CMax<float, int64_t>;
        faiss::heap_heapify<C>(k, out_v, out_i);
        faiss::heap_addn<C>(k, out_v, out_i, vals, ids, n);
        faiss::heap_reorder<C>(k, out_v, out_i);
    } else {
        using C = faiss::CMin<float, int64_t>;
        faiss::heap_heapify<C>(k, out_v, out_i);
        faiss::heap_addn<C>(k, out_v, out_i, vals, ids, n);
        faiss::heap_reorder<C>(k, out_v, out_i);
    }
}

// ---- IDSelector family (faiss/impl/IDSelector.h); batch/bitmap inputs are copied so that the
// Python caller need not keep them alive ---------------------------------------------------------
struct RefSel {
    std::unique_ptr<faiss::IDSelector> sel;
    std::vector<uint8_t> bitmap;
};
void* ref_sel_range_new(int64_t imin, int64_t imax) {
    auto* r = new RefSel;
    r->sel.reset(new faiss::IDSelectorRange(imin, imax));
    return r;
}
void* ref_sel_batch_new(int64_t n, const int64_t* ids) {
    auto* r = new RefSel;
    r->sel.reset(new faiss::IDSelectorBatch(n, ids));
    return r;
}
void* ref_sel_bitmap_new(int64_t n, const uint8_t* bitmap) {
    auto* r = new RefSel;
    r->bitmap.assign(bitmap, bitmap + n);
    r->sel.reset(new faiss::IDSelectorBitmap(n, r->bitmap.data()));
    return r;
}
// kind: 3 not, 4 and, 5 or, 6 xor (operands borrowed)
void* ref_sel_combine_new(int kind, void* lhs, void* rhs) {
    auto* r = new RefSel;
    const faiss::IDSelector* a = ((RefSel*)lhs)->sel.get();
    const faiss::IDSelector* b = rhs ? ((RefSel*)rhs)->sel.get() : nullptr;
    if (kind == 3)
        r->sel.reset(new faiss::IDSelectorNot(a));
    else if (kind == 4)
        r->sel.reset(new faiss::IDSelectorAnd(a, b));
    else if (kind == 5)
        r->sel.reset(new faiss::IDSelectorOr(a, b));
    else
        r->sel.reset(new faiss::IDSelectorXOr(a, b));
    return r;
}
int ref_sel_is_member(void* s, int64_t id) {
    return ((RefSel*)s)->sel->is_member(id) ? 1 : 0;
}
void ref_sel_free(void* s) {
    delete (RefSel*)s;
}

// IndexIVF::search with SearchParametersIVF (faiss/IndexIVF.cpp:303-397)
int ref_ivfflat_search_params(void* h, int64_t n, const float* x, int64_t k, int64_t nprobe,
                              int64_t max_codes, void* sel, float* D, int64_t* I) {
    auto* r = (RefIndex*)h;
    return guarded([&] {
        faiss::SearchParametersIVF p;
        p.nprobe = nprobe;
        p.max_codes = max_codes;
        p.sel = sel ? ((RefSel*)sel)->sel.get() : nullptr;
        r->index->search(n, x, k, D, I, &p);
    });
}

// IndexIVF::search_preassigned with every trailing argument (faiss/IndexIVF.cpp:399-723)
int ref_ivfflat_search_preassigned_ex(void* h, int64_t n, const float* x, int64_t k,
                                      int64_t nprobe, const int64_t* assign, const float* cdis,
                                      float* D, int64_t* I, int store_pairs, int64_t max_codes,
                                      void* sel) {
    auto* r = (RefIndex*)h;
    return guarded([&] {
        faiss::SearchParametersIVF p;
        p.nprobe = nprobe;
        p.max_codes = max_codes;
        p.sel = sel ? ((RefSel*)sel)->sel.get() : nullptr;
        r->index->search_preassigned(n, x, k, assign, cdis, D, I, store_pairs != 0, &p);
    });
}

// the fork's IndexIVF::search_stats (faiss/IndexIVF.cpp:727-862); stats: n x {total_us,
// quantization_us, list_scan_us}
int ref_ivfflat_search_stats(void* h, int64_t n, const float* x, int64_t k, int64_t nprobe,
                             float* D, int64_t* I, double* stats3) {
    auto* r = (RefIndex*)h;
    static_assert(sizeof(faiss::QueryLatencyStats) == 3 * sizeof(double), "layout");
    return guarded([&] {
        r->index->nprobe = nprobe;
        r->index->search_stats(n, x, k, D, I, nullptr, (faiss::QueryLatencyStats*)stats3);
    });
}

// IndexIVF::range_search / range_search_preassigned (faiss/IndexIVF.cpp:1203-1399) into a
// faiss::RangeSearchResult the caller reads and frees through the three functions below
void* ref_ivfflat_range_search(void* h, int64_t n, const float* x, float radius, int64_t nprobe,
                               const int64_t* assign, int store_pairs, void* sel) {
    auto* r = (RefIndex*)h;
    faiss::RangeSearchResult* res = nullptr;
    const int rc = guarded([&] {
        std::unique_ptr<faiss::RangeSearchResult> rr(new faiss::RangeSearchResult(n));
        faiss::SearchParametersIVF p;
        p.nprobe = nprobe;
        p.sel = sel ? ((RefSel*)sel)->sel.get() : nullptr;
        if (assign) {
            std::vector<float> cd((size_t)n * nprobe, 0.f);
            r->index->range_search_preassigned(n, x, radius, assign, cd.data(), rr.get(),
                                               store_pairs != 0, &p);
        } else {
            r->index->range_search(n, x, radius, rr.get(), &p);
        }
        res = rr.release();
    });
    return rc == 0 ? res : nullptr;
}
const size_t* ref_range_lims(void* res) {
    return ((faiss::RangeSearchResult*)res)->lims;
}
const int64_t* ref_range_labels(void* res) {
    return ((faiss::RangeSearchResult*)res)->labels;
}
const float* ref_range_distances(void* res) {
    return ((faiss::RangeSearchResult*)res)->distances;
}
void ref_range_free(void* res) {
    delete (faiss::RangeSearchResult*)res;
}

// Level1Quantizer::cp knobs before train (faiss/Clustering.h:23-64)
void ref_ivfflat_set_cp(void* h, int niter, int seed, int max_points_per_centroid) {
    auto& cp = ((RefIndex*)h)->index->cp;
    cp.niter = niter;
    cp.seed = seed;
    cp.max_points_per_centroid = max_points_per_centroid;
}

} // extern "C"
