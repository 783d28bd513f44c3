/* oracle/ivfflat_oracle.c -- TEST INFRASTRUCTURE, not product code (see ivfflat_oracle.h).
 *
 * Plain-C restatement of the reference's CPU IndexIVFFlat::search path. Each function cites the
 * reference file:line it follows. Written from the behaviour of the reference, not copied: the
 * reference is templated C++ over comparator classes, this is C with a runtime metric flag.
 */
#include "ivfflat_oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <sys/time.h>

/* ---------------------------------------------------------------------------------------------
 * distance primitives: faiss/utils/distances_simd.cpp:194-230 (plain loops; the reference lets
 * the compiler reassociate them, so low-order bits are compiler-dependent there as well).
 * ------------------------------------------------------------------------------------------- */
float orc_fvec_inner_product(const float* x, const float* y, size_t d) {
    float res = 0.f;
    for (size_t i = 0; i < d; i++)
        res += x[i] * y[i];
    return res;
}

float orc_fvec_L2sqr(const float* x, const float* y, size_t d) {
    float res = 0.f;
    for (size_t i = 0; i < d; i++) {
        const float t = x[i] - y[i];
        res += t * t;
    }
    return res;
}

float orc_fvec_norm_L2sqr(const float* x, size_t d) {
    float res = 0.f;
    for (size_t i = 0; i < d; i++)
        res += x[i] * x[i];
    return res;
}

/* ---------------------------------------------------------------------------------------------
 * comparators: faiss/utils/ordered_key_value.h:41-84.
 * is_max=1 -> CMax (max-heap, used for L2 results); is_max=0 -> CMin (min-heap, IP results).
 * ------------------------------------------------------------------------------------------- */
static inline int c_cmp(int is_max, float a, float b) {
    return is_max ? (a > b) : (a < b);
}
static inline int c_cmp2(int is_max, float a1, float b1, int64_t a2, int64_t b2) {
    return is_max ? ((a1 > b1) || ((a1 == b1) && (a2 > b2)))
                  : ((a1 < b1) || ((a1 == b1) && (a2 < b2)));
}
static inline float c_neutral(int is_max) {
    return is_max ? FLT_MAX : -FLT_MAX;
}
static inline int metric_is_max(int metric) {
    return metric == ORC_METRIC_L2; /* faiss/IndexIVF.cpp:436-437 HeapForL2 = CMax */
}

/* ---------------------------------------------------------------------------------------------
 * binary heap on parallel arrays, worst element at slot 0: faiss/utils/Heap.h:43-150
 * ------------------------------------------------------------------------------------------- */
static void h_sift_from_top(int is_max, size_t k, float* v0, int64_t* i0, float val, int64_t id) {
    /* shared body of heap_pop (Heap.h:47-77) and heap_replace_top (Heap.h:111-150): walk down
     * from the root moving the cmp2-larger child up until (val,id) dominates both children. */
    float* v = v0 - 1; /* 1-based */
    int64_t* ix = i0 - 1;
    size_t i = 1;
    for (;;) {
        size_t c1 = i << 1, c2 = c1 + 1;
        if (c1 > k)
            break;
        size_t c = (c2 == k + 1 || c_cmp2(is_max, v[c1], v[c2], ix[c1], ix[c2])) ? c1 : c2;
        if (c_cmp2(is_max, val, v[c], id, ix[c]))
            break;
        v[i] = v[c];
        ix[i] = ix[c];
        i = c;
    }
    v[i] = val;
    ix[i] = id;
}

static void h_pop(int is_max, size_t k, float* v, int64_t* ix) {
    /* Heap.h:47-77: remove the root, re-insert the last element from the top */
    h_sift_from_top(is_max, k, v, ix, v[k - 1], ix[k - 1]);
}

static void h_push(int is_max, size_t k, float* v0, int64_t* i0, float val, int64_t id) {
    /* Heap.h:82-104: heap holds k-1 elements on entry, k on exit */
    float* v = v0 - 1;
    int64_t* ix = i0 - 1;
    size_t i = k;
    while (i > 1) {
        size_t f = i >> 1;
        if (!c_cmp2(is_max, val, v[f], id, ix[f]))
            break;
        v[i] = v[f];
        ix[i] = ix[f];
        i = f;
    }
    v[i] = val;
    ix[i] = id;
}

void orc_heap_heapify(int metric, size_t k, float* val, int64_t* ids) {
    /* Heap.h:316-338 with k0 = 0 */
    const int is_max = metric_is_max(metric);
    for (size_t i = 0; i < k; i++) {
        val[i] = c_neutral(is_max);
        ids[i] = -1;
    }
}

void orc_heap_replace_top(int metric, size_t k, float* val, int64_t* ids, float v, int64_t id) {
    h_sift_from_top(metric_is_max(metric), k, val, ids, v, id);
}

size_t orc_heap_addn(int metric, size_t k, float* val, int64_t* ids, const float* x,
                     const int64_t* xids, size_t n) {
    /* Heap.h:367-388: strict compare against the current worst */
    const int is_max = metric_is_max(metric);
    size_t nup = 0;
    for (size_t i = 0; i < n; i++) {
        if (c_cmp(is_max, val[0], x[i])) {
            h_sift_from_top(is_max, k, val, ids, x[i], xids ? xids[i] : (int64_t)i);
            nup++;
        }
    }
    return nup;
}

static size_t h_reorder(int is_max, size_t k, float* val, int64_t* ids) {
    /* Heap.h:420-450: pop worst-first into the tail, drop id == -1 slots, pad */
    size_t ii = 0;
    for (size_t i = 0; i < k; i++) {
        float v = val[0];
        int64_t id = ids[0];
        h_pop(is_max, k - i, val, ids);
        val[k - ii - 1] = v;
        ids[k - ii - 1] = id;
        if (id != -1)
            ii++;
    }
    size_t nel = ii;
    memmove(val, val + k - ii, ii * sizeof(*val));
    memmove(ids, ids + k - ii, ii * sizeof(*ids));
    for (; ii < k; ii++) {
        val[ii] = c_neutral(is_max);
        ids[ii] = -1;
    }
    return nel;
}

size_t orc_heap_reorder(int metric, size_t k, float* val, int64_t* ids) {
    return h_reorder(metric_is_max(metric), k, val, ids);
}

/* ---------------------------------------------------------------------------------------------
 * flat kNN (the coarse quantizer): faiss/utils/distances.cpp
 * ------------------------------------------------------------------------------------------- */
#define ORC_BLAS_THRESHOLD 20 /* distances.cpp:831 */
#define ORC_QUERY_BS 4096     /* distances.cpp:832 */
#define ORC_DB_BS 1024        /* distances.cpp:833 */

static void knn_seq(int metric, const float* x, const float* y, size_t d, size_t nx, size_t ny,
                    size_t k, float* vals, int64_t* ids) {
    /* exhaustive_{inner_product,L2sqr}_seq, distances.cpp:136-199, with the Heap handler
     * (ResultHandler.h:213-252). Top1 (k==1) and Reservoir (k>=100, ResultHandler.h:631-636)
     * select the same values and differ from the heap only in which of several equal values
     * survive. */
    const int is_max = metric_is_max(metric);
#pragma omp parallel for
    for (int64_t i = 0; i < (int64_t)nx; i++) {
        float* hv = vals + i * k;
        int64_t* hi = ids + i * k;
        orc_heap_heapify(metric, k, hv, hi);
        const float* xi = x + i * d;
        for (size_t j = 0; j < ny; j++) {
            float dis = metric == ORC_METRIC_L2 ? orc_fvec_L2sqr(xi, y + j * d, d)
                                                : orc_fvec_inner_product(xi, y + j * d, d);
            if (c_cmp(is_max, hv[0], dis))
                h_sift_from_top(is_max, k, hv, hi, dis, (int64_t)j);
        }
        h_reorder(is_max, k, hv, hi);
    }
}

static void knn_blas_form(int metric, const float* x, const float* y, size_t d, size_t nx,
                          size_t ny, size_t k, float* vals, int64_t* ids) {
    /* exhaustive_L2sqr_blas_default_impl distances.cpp:258-342 and
     * exhaustive_inner_product_blas :202-254. sgemm_ (third-party, un-vendored, unpinned BLAS) is
     * restated as a plain dot product: its accumulation order is unspecified, so coarse distance
     * BITS are "parity unpinned" (SURVEY.md 8c); the arithmetic FORM is what is kept:
     * dis = ||x||^2 + ||y||^2 - 2 x.y, negative -> 0 (:323-331). Blocking (4096 x 1024) does not
     * change any value, only the order candidates reach the heap (ascending j either way). */
    const int is_max = metric_is_max(metric);
    float* xn = NULL;
    float* yn = NULL;
    if (metric == ORC_METRIC_L2) {
        xn = (float*)malloc(sizeof(float) * nx);
        yn = (float*)malloc(sizeof(float) * ny);
        for (size_t i = 0; i < nx; i++)
            xn[i] = orc_fvec_norm_L2sqr(x + i * d, d);
        for (size_t j = 0; j < ny; j++)
            yn[j] = orc_fvec_norm_L2sqr(y + j * d, d);
    }
#pragma omp parallel for
    for (int64_t i = 0; i < (int64_t)nx; i++) {
        float* hv = vals + i * k;
        int64_t* hi = ids + i * k;
        orc_heap_heapify(metric, k, hv, hi);
        const float* xi = x + i * d;
        for (size_t j = 0; j < ny; j++) {
            float ip = orc_fvec_inner_product(xi, y + j * d, d);
            float dis = ip;
            if (metric == ORC_METRIC_L2) {
                dis = xn[i] + yn[j] - 2 * ip;
                if (dis < 0)
                    dis = 0;
            }
            if (c_cmp(is_max, hv[0], dis))
                h_sift_from_top(is_max, k, hv, hi, dis, (int64_t)j);
        }
        h_reorder(is_max, k, hv, hi);
    }
    free(xn);
    free(yn);
}

void orc_knn(int metric, const float* x, const float* y, size_t d, size_t nx, size_t ny, size_t k,
             float* vals, int64_t* ids) {
    /* Run_search_inner_product / Run_search_L2sqr, distances.cpp:790-823 */
    if (nx < ORC_BLAS_THRESHOLD)
        knn_seq(metric, x, y, d, nx, ny, k, vals, ids);
    else
        knn_blas_form(metric, x, y, d, nx, ny, k, vals, ids);
}

/* ---------------------------------------------------------------------------------------------
 * index object: ArrayInvertedLists (faiss/invlists/InvertedLists.h:243-283) + flat quantizer
 * ------------------------------------------------------------------------------------------- */
struct OrcIndex {
    size_t d, nlist;
    int metric;
    int64_t ntotal;
    float* centroids; /* [nlist][d], faiss/IndexFlat.h:69-72 */
    float** codes;    /* per list: [len][d] row-major fp32 == code bytes */
    int64_t** ids;
    size_t* len;
    size_t* cap;
};

OrcIndex* orc_index_new(size_t d, size_t nlist, int metric) {
    OrcIndex* ix = (OrcIndex*)calloc(1, sizeof(OrcIndex));
    ix->d = d;
    ix->nlist = nlist;
    ix->metric = metric;
    ix->centroids = (float*)calloc(nlist * d, sizeof(float));
    ix->codes = (float**)calloc(nlist, sizeof(float*));
    ix->ids = (int64_t**)calloc(nlist, sizeof(int64_t*));
    ix->len = (size_t*)calloc(nlist, sizeof(size_t));
    ix->cap = (size_t*)calloc(nlist, sizeof(size_t));
    return ix;
}

void orc_index_free(OrcIndex* ix) {
    if (!ix)
        return;
    for (size_t l = 0; l < ix->nlist; l++) {
        free(ix->codes[l]);
        free(ix->ids[l]);
    }
    free(ix->codes);
    free(ix->ids);
    free(ix->len);
    free(ix->cap);
    free(ix->centroids);
    free(ix);
}

void orc_index_set_centroids(OrcIndex* ix, const float* c) {
    memcpy(ix->centroids, c, sizeof(float) * ix->nlist * ix->d);
}

size_t orc_index_list_size(const OrcIndex* ix, size_t l) {
    return ix->len[l];
}
const float* orc_index_list_codes(const OrcIndex* ix, size_t l) {
    return ix->codes[l];
}
const int64_t* orc_index_list_ids(const OrcIndex* ix, size_t l) {
    return ix->ids[l];
}
int64_t orc_index_ntotal(const OrcIndex* ix) {
    return ix->ntotal;
}

void orc_index_add_core(OrcIndex* ix, int64_t n, const float* x, const int64_t* xids,
                        const int64_t* assign) {
    /* faiss/IndexIVFFlat.cpp:46-91: every list is owned by one thread, so entries of a list stay
     * in input order whatever the thread count; id = xids ? xids[i] : ntotal + i (:72);
     * list_no < 0 is skipped but still counted in ntotal (:90). */
    for (int64_t i = 0; i < n; i++) {
        int64_t l = assign[i];
        if (l < 0)
            continue;
        if (ix->len[l] == ix->cap[l]) {
            size_t nc = ix->cap[l] ? ix->cap[l] * 2 : 16;
            ix->codes[l] = (float*)realloc(ix->codes[l], nc * ix->d * sizeof(float));
            ix->ids[l] = (int64_t*)realloc(ix->ids[l], nc * sizeof(int64_t));
            ix->cap[l] = nc;
        }
        memcpy(ix->codes[l] + ix->len[l] * ix->d, x + i * ix->d, ix->d * sizeof(float));
        ix->ids[l][ix->len[l]] = xids ? xids[i] : ix->ntotal + i;
        ix->len[l]++;
    }
    ix->ntotal += n;
}

void orc_index_add_with_ids(OrcIndex* ix, int64_t n, const float* x, const int64_t* xids) {
    /* faiss/IndexIVF.cpp:191-241: coarse_idx = quantizer->assign(x) (k=1 search), processed in
     * blocks of 65536 (:205-221) - the block size only matters for seq-vs-BLAS when a block has
     * < 20 vectors. */
    const int64_t bs = 65536;
    int64_t* assign = (int64_t*)malloc(sizeof(int64_t) * (size_t)(n < bs ? n : bs));
    float* dis = (float*)malloc(sizeof(float) * (size_t)(n < bs ? n : bs));
    for (int64_t i0 = 0; i0 < n; i0 += bs) {
        int64_t ni = n - i0 < bs ? n - i0 : bs;
        orc_knn(ix->metric, x + i0 * ix->d, ix->centroids, ix->d, ni, ix->nlist, 1, dis, assign);
        orc_index_add_core(ix, ni, x + i0 * ix->d, xids ? xids + i0 : NULL, assign);
    }
    free(assign);
    free(dis);
}

void orc_index_quantize(const OrcIndex* ix, int64_t n, const float* x, size_t nprobe, float* Dq,
                        int64_t* Iq) {
    /* IndexFlat::search faiss/IndexFlat.cpp:38-69 */
    orc_knn(ix->metric, x, ix->centroids, ix->d, n, ix->nlist, nprobe, Dq, Iq);
}

static double now_ms(void) {
    struct timeval tv;
    gettimeofday(&tv, NULL);
    return tv.tv_sec * 1e3 + tv.tv_usec * 1e-3;
}

int orc_index_search_preassigned(const OrcIndex* ix, int64_t n, const float* x, int64_t k,
                                 size_t nprobe, const int64_t* keys, float* D, int64_t* I,
                                 OrcStats* stats) {
    /* faiss/IndexIVF.cpp:399-723, parallel_mode 0, store_pairs=false, no selector, max_codes=0.
     * Per query: heapify (:607), probes in the given order (:612-622), strict compare in
     * scan_codes (faiss/IndexIVFFlat.cpp:155-179), reorder (:625). */
    const int is_max = metric_is_max(ix->metric);
    const size_t d = ix->d;
    size_t nlistv = 0, ndis = 0, nheap = 0;
    int bad = 0;
    if (k <= 0 || nprobe == 0)
        return -2; /* FAISS_THROW_IF_NOT(k > 0), (nprobe > 0): IndexIVF.cpp:410,414 */
    if (nprobe > ix->nlist)
        nprobe = ix->nlist;
#pragma omp parallel for reduction(+ : nlistv, ndis, nheap) reduction(| : bad)
    for (int64_t i = 0; i < n; i++) {
        const float* xi = x + i * d;
        float* simi = D + i * k;
        int64_t* idxi = I + i * k;
        orc_heap_heapify(ix->metric, k, simi, idxi);
        size_t nscan = 0;
        for (size_t ik = 0; ik < nprobe; ik++) {
            int64_t key = keys[i * nprobe + ik];
            if (key < 0)
                continue; /* :516 */
            if ((size_t)key >= ix->nlist) {
                bad = 1; /* :520 throws */
                continue;
            }
            size_t ls = ix->len[key];
            if (ls == 0)
                continue; /* :527 */
            nlistv++;
            const float* yl = ix->codes[key];
            const int64_t* idl = ix->ids[key];
            for (size_t j = 0; j < ls; j++) {
                float dis = ix->metric == ORC_METRIC_INNER_PRODUCT
                        ? orc_fvec_inner_product(xi, yl + j * d, d)
                        : orc_fvec_L2sqr(xi, yl + j * d, d);
                if (c_cmp(is_max, simi[0], dis)) {
                    h_sift_from_top(is_max, k, simi, idxi, dis, idl[j]);
                    nheap++;
                }
            }
            nscan += ls;
        }
        ndis += nscan;
        h_reorder(is_max, k, simi, idxi);
    }
    if (stats) {
        stats->nq += n;
        stats->nlist += nlistv;
        stats->ndis += ndis;
        stats->nheap_updates += nheap;
    }
    return bad ? -2 : 0;
}

int orc_index_search(const OrcIndex* ix, int64_t n, const float* x, int64_t k, size_t nprobe,
                     int nt, float* D, int64_t* I, OrcStats* stats) {
    /* faiss/IndexIVF.cpp:303-397: nprobe clamp (:316), nt = min(threads, n) slices (:359), slice s
     * = [n*s/nt, n*(s+1)/nt) (:366-367); per slice quantize then search_preassigned (:321-356). */
    if (k <= 0)
        return -2;
    if (nprobe > ix->nlist)
        nprobe = ix->nlist;
    if (nprobe == 0)
        return -2;
    if (nt < 1)
        nt = 1;
    if (nt > n)
        nt = (int)n;
    int rc = 0;
    for (int s = 0; s < nt; s++) {
        int64_t i0 = n * s / nt, i1 = n * (s + 1) / nt;
        if (i1 <= i0)
            continue;
        int64_t ns = i1 - i0;
        int64_t* idx = (int64_t*)malloc(sizeof(int64_t) * ns * nprobe);
        float* cd = (float*)malloc(sizeof(float) * ns * nprobe);
        double t0 = now_ms();
        orc_index_quantize(ix, ns, x + i0 * ix->d, nprobe, cd, idx);
        double t1 = now_ms();
        int r = orc_index_search_preassigned(ix, ns, x + i0 * ix->d, k, nprobe, idx, D + i0 * k,
                                             I + i0 * k, stats);
        double t2 = now_ms();
        if (stats) {
            stats->quantization_time += t1 - t0;
            stats->search_time += t2 - t0;
        }
        if (r)
            rc = r;
        free(idx);
        free(cd);
    }
    return rc;
}

void orc_merge_knn_results(int metric, size_t n, size_t k, int nshard, const float* all_D,
                           const int64_t* all_I, float* D, int64_t* I) {
    /* faiss/utils/Heap.cpp:158-232 as called from faiss/IndexShardsIVF.cpp:224-242: a heap over
     * shard heads with the BEST value on top (L2 -> CMin<float,int>, IP -> CMax<float,int>), ties
     * between shards broken by shard number; inputs are per-shard sorted (n,k) tables, -1 ends a
     * shard's list; output padded with -1 / Crev::neutral(). */
    const int top_is_max = metric == ORC_METRIC_L2 ? 0 : 1;
    const size_t stride = n * k;
    if (k == 0)
        return;
    int* pointer = (int*)malloc(sizeof(int) * nshard);
    int64_t* shard_ids = (int64_t*)malloc(sizeof(int64_t) * nshard);
    float* heap_vals = (float*)malloc(sizeof(float) * nshard);
    for (size_t i = 0; i < n; i++) {
        const float* Din = all_D + i * k;
        const int64_t* Iin = all_I + i * k;
        size_t hs = 0;
        for (int s = 0; s < nshard; s++) {
            pointer[s] = 0;
            if (Iin[stride * s] >= 0)
                h_push(top_is_max, ++hs, heap_vals, shard_ids, Din[stride * s], s);
        }
        size_t j;
        for (j = 0; j < k && hs > 0; j++) {
            int s = (int)shard_ids[0];
            D[i * k + j] = heap_vals[0];
            I[i * k + j] = Iin[stride * s + pointer[s]];
            h_pop(top_is_max, hs--, heap_vals, shard_ids);
            pointer[s]++;
            if ((size_t)pointer[s] < k && Iin[stride * s + pointer[s]] >= 0)
                h_push(top_is_max, ++hs, heap_vals, shard_ids, Din[stride * s + pointer[s]], s);
        }
        for (; j < k; j++) {
            I[i * k + j] = -1;
            D[i * k + j] = c_neutral(!top_is_max);
        }
    }
    free(pointer);
    free(shard_ids);
    free(heap_vals);
}
