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
const float* orc_index_centroids(const OrcIndex* ix) {
    return ix->centroids;
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

/* ---------------------------------------------------------------------------------------------
 * IDSelector family: faiss/impl/IDSelector.h:19-170, faiss/impl/IDSelector.cpp:17-123
 * ------------------------------------------------------------------------------------------- */
struct OrcSel {
    int kind; /* ORC_SEL_* */
    int64_t imin, imax;    /* range */
    int64_t* set;          /* batch / array: sorted copy */
    size_t nset;
    uint8_t* bitmap;       /* bitmap copy, nbytes */
    size_t nbytes;
    const OrcSel *lhs, *rhs; /* not (lhs) / and / or / xor: borrowed */
};

static int cmp_i64(const void* a, const void* b) {
    int64_t x = *(const int64_t*)a, y = *(const int64_t*)b;
    return x < y ? -1 : (x > y ? 1 : 0);
}

OrcSel* orc_sel_range(int64_t imin, int64_t imax) {
    OrcSel* s = (OrcSel*)calloc(1, sizeof(OrcSel));
    s->kind = ORC_SEL_RANGE;
    s->imin = imin;
    s->imax = imax;
    return s;
}
OrcSel* orc_sel_batch(size_t n, const int64_t* ids) {
    OrcSel* s = (OrcSel*)calloc(1, sizeof(OrcSel));
    s->kind = ORC_SEL_BATCH;
    s->set = (int64_t*)malloc(sizeof(int64_t) * (n ? n : 1));
    memcpy(s->set, ids, sizeof(int64_t) * n);
    qsort(s->set, n, sizeof(int64_t), cmp_i64);
    s->nset = n;
    return s;
}
OrcSel* orc_sel_bitmap(size_t nbytes, const uint8_t* bitmap) {
    OrcSel* s = (OrcSel*)calloc(1, sizeof(OrcSel));
    s->kind = ORC_SEL_BITMAP;
    s->bitmap = (uint8_t*)malloc(nbytes ? nbytes : 1);
    memcpy(s->bitmap, bitmap, nbytes);
    s->nbytes = nbytes;
    return s;
}
OrcSel* orc_sel_combine(int kind, const OrcSel* lhs, const OrcSel* rhs) {
    OrcSel* s = (OrcSel*)calloc(1, sizeof(OrcSel));
    s->kind = kind;
    s->lhs = lhs;
    s->rhs = rhs;
    return s;
}
void orc_sel_free(OrcSel* s) {
    if (!s)
        return;
    free(s->set);
    free(s->bitmap);
    free(s);
}
int orc_sel_is_member(const OrcSel* s, int64_t id) {
    switch (s->kind) {
        case ORC_SEL_RANGE: /* IDSelector.cpp:20-22 */
            return id >= s->imin && id < s->imax;
        case ORC_SEL_BATCH: { /* IDSelector.cpp:102-108: set membership */
            size_t lo = 0, hi = s->nset;
            while (lo < hi) {
                size_t mid = (lo + hi) / 2;
                if (s->set[mid] < id)
                    lo = mid + 1;
                else
                    hi = mid;
            }
            return lo < s->nset && s->set[lo] == id;
        }
        case ORC_SEL_BITMAP: { /* IDSelector.cpp:117-123 */
            uint64_t i = (uint64_t)id;
            if ((i >> 3) >= s->nbytes)
                return 0;
            return (s->bitmap[i >> 3] >> (i & 7)) & 1;
        }
        case ORC_SEL_NOT: /* IDSelector.h:133-139 */
            return !orc_sel_is_member(s->lhs, id);
        case ORC_SEL_AND:
            return orc_sel_is_member(s->lhs, id) && orc_sel_is_member(s->rhs, id);
        case ORC_SEL_OR:
            return orc_sel_is_member(s->lhs, id) || orc_sel_is_member(s->rhs, id);
        case ORC_SEL_XOR:
            return orc_sel_is_member(s->lhs, id) ^ orc_sel_is_member(s->rhs, id);
    }
    return 0;
}

int orc_index_search_preassigned_ex(const OrcIndex* ix, int64_t n, const float* x, int64_t k,
                                    size_t nprobe, const int64_t* keys, float* D, int64_t* I,
                                    OrcStats* stats, size_t max_codes, const OrcSel* sel,
                                    int store_pairs) {
    /* faiss/IndexIVF.cpp:399-723, parallel_mode 0. Per query: heapify (:607), probes in the given
     * order (:612-622) each cut to max_codes - nscan rows (:545-548) until nscan >= max_codes
     * (:619-621), strict compare in scan_codes with the selector test first
     * (faiss/IndexIVFFlat.cpp:155-179), label = lo_build(list, j) when store_pairs (:172),
     * reorder (:625). */
    const int is_max = metric_is_max(ix->metric);
    const size_t d = ix->d;
    size_t nlistv = 0, ndis = 0, nheap = 0;
    int bad = 0;
    if (k <= 0 || nprobe == 0)
        return -2; /* FAISS_THROW_IF_NOT(k > 0), (nprobe > 0): IndexIVF.cpp:410,414 */
    if (sel && store_pairs)
        return -2; /* :428-430 */
    if (nprobe > ix->nlist)
        nprobe = ix->nlist;
    if (max_codes == 0)
        max_codes = (size_t)INT64_MAX; /* :453-455 */
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
            if (ls > max_codes - nscan)
                ls = max_codes - nscan; /* :546-548 */
            const float* yl = ix->codes[key];
            const int64_t* idl = ix->ids[key];
            for (size_t j = 0; j < ls; j++) {
                if (sel && !orc_sel_is_member(sel, idl[j]))
                    continue; /* IndexIVFFlat.cpp:166 */
                float dis = ix->metric == ORC_METRIC_INNER_PRODUCT
                        ? orc_fvec_inner_product(xi, yl + j * d, d)
                        : orc_fvec_L2sqr(xi, yl + j * d, d);
                if (c_cmp(is_max, simi[0], dis)) {
                    int64_t id = store_pairs ? (int64_t)(((uint64_t)key << 32) | (uint64_t)j)
                                             : idl[j];
                    h_sift_from_top(is_max, k, simi, idxi, dis, id);
                    nheap++;
                }
            }
            nscan += ls;
            if (nscan >= max_codes)
                break; /* :619-621 */
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

int orc_index_search_preassigned(const OrcIndex* ix, int64_t n, const float* x, int64_t k,
                                 size_t nprobe, const int64_t* keys, float* D, int64_t* I,
                                 OrcStats* stats) {
    return orc_index_search_preassigned_ex(ix, n, x, k, nprobe, keys, D, I, stats, 0, NULL, 0);
}

int orc_index_range_search_preassigned(const OrcIndex* ix, int64_t n, const float* x, float radius,
                                       size_t nprobe, const int64_t* keys, const OrcSel* sel,
                                       int store_pairs, size_t* lims, int64_t** labels_out,
                                       float** dis_out) {
    /* faiss/IndexIVF.cpp:1243-1399 (parallel_mode 0) + scan_codes_range
     * faiss/IndexIVFFlat.cpp:181-201: per query, probes in order, rows in list order, a row is
     * kept when C::cmp(radius, dis) (L2: radius > dis, IP: radius < dis). */
    const int is_max = metric_is_max(ix->metric);
    const size_t d = ix->d;
    if (nprobe == 0 || (sel && store_pairs))
        return -2;
    if (nprobe > ix->nlist)
        nprobe = ix->nlist;
    size_t cap = 1024, cnt = 0;
    int64_t* lab = (int64_t*)malloc(cap * sizeof(int64_t));
    float* dd = (float*)malloc(cap * sizeof(float));
    int bad = 0;
    for (int64_t i = 0; i < n; i++) {
        const float* xi = x + i * d;
        lims[i] = cnt;
        for (size_t ik = 0; ik < nprobe; ik++) {
            int64_t key = keys[i * nprobe + ik];
            if (key < 0)
                continue;
            if ((size_t)key >= ix->nlist) {
                bad = 1;
                continue;
            }
            const size_t ls = ix->len[key];
            const float* yl = ix->codes[key];
            const int64_t* idl = ix->ids[key];
            for (size_t j = 0; j < ls; j++) {
                if (sel && !orc_sel_is_member(sel, idl[j]))
                    continue;
                float dis = ix->metric == ORC_METRIC_INNER_PRODUCT
                        ? orc_fvec_inner_product(xi, yl + j * d, d)
                        : orc_fvec_L2sqr(xi, yl + j * d, d);
                if (c_cmp(is_max, radius, dis)) {
                    if (cnt == cap) {
                        cap *= 2;
                        lab = (int64_t*)realloc(lab, cap * sizeof(int64_t));
                        dd = (float*)realloc(dd, cap * sizeof(float));
                    }
                    lab[cnt] = store_pairs ? (int64_t)(((uint64_t)key << 32) | (uint64_t)j)
                                           : idl[j];
                    dd[cnt] = dis;
                    cnt++;
                }
            }
        }
    }
    lims[n] = cnt;
    *labels_out = lab;
    *dis_out = dd;
    return bad ? -2 : 0;
}

void orc_free(void* p) {
    free(p);
}

/* ---------------------------------------------------------------------------------------------
 * training: Clustering::train_encoded (faiss/Clustering.cpp:296-600) as IndexIVF drives it
 * (faiss/IndexIVF.cpp:66-127: the flat quantizer of the index's metric assigns; niter 10,
 * spherical for inner product :179-182). RandomGenerator is std::mt19937
 * (faiss/utils/random.cpp:37-49), restated here as MT19937 (Matsumoto & Nishimura 1998).
 * ------------------------------------------------------------------------------------------- */
typedef struct {
    uint32_t mt[624];
    int idx;
} Mt19937;

static void mt_seed(Mt19937* g, uint32_t seed) {
    g->mt[0] = seed;
    for (int i = 1; i < 624; i++)
        g->mt[i] = 1812433253u * (g->mt[i - 1] ^ (g->mt[i - 1] >> 30)) + (uint32_t)i;
    g->idx = 624;
}

static uint32_t mt_next(Mt19937* g) {
    if (g->idx >= 624) {
        for (int i = 0; i < 624; i++) {
            uint32_t y = (g->mt[i] & 0x80000000u) | (g->mt[(i + 1) % 624] & 0x7fffffffu);
            g->mt[i] = g->mt[(i + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
        }
        g->idx = 0;
    }
    uint32_t y = g->mt[g->idx++];
    y ^= y >> 11;
    y ^= (y << 7) & 0x9d2c5680u;
    y ^= (y << 15) & 0xefc60000u;
    y ^= y >> 18;
    return y;
}

/* faiss::rand_perm, faiss/utils/random.cpp:184-194 */
static void ref_rand_perm(int64_t* perm, size_t n, int64_t seed) {
    Mt19937 g;
    mt_seed(&g, (uint32_t)seed);
    for (size_t i = 0; i < n; i++)
        perm[i] = (int64_t)i;
    for (size_t i = 0; i + 1 < n; i++) {
        size_t i2 = i + (size_t)(mt_next(&g) % (uint32_t)(n - i));
        int64_t t = perm[i];
        perm[i] = perm[i2];
        perm[i2] = t;
    }
}

static void renorm_rows(float* c, size_t k, size_t d) { /* distances.cpp:95-105 */
    for (size_t i = 0; i < k; i++) {
        float* xi = c + i * d;
        float nr = orc_fvec_norm_L2sqr(xi, d);
        if (nr > 0) {
            const float inv_nr = (float)(1.0 / sqrtf(nr));
            for (size_t j = 0; j < d; j++)
                xi[j] *= inv_nr;
        }
    }
}

int orc_index_train(OrcIndex* ix, int64_t n, const float* x, int niter, int seed,
                    int max_points_per_centroid, double* objective_out) {
    const size_t k = ix->nlist, d = ix->d;
    if (n < (int64_t)k)
        return -2; /* Clustering.cpp:302-306 */
    int64_t nx = n;
    float* xt = NULL;
    const float* xs = x;
    if ((size_t)n == k) { /* "just copy training set to clusters", Clustering.cpp:346-368 */
        memcpy(ix->centroids, x, sizeof(float) * k * d);
        if (objective_out)
            *objective_out = 0;
        return 0;
    }
    if ((size_t)n > k * (size_t)max_points_per_centroid) { /* subsample_training_set :86-133 */
        int64_t* perm = (int64_t*)malloc(sizeof(int64_t) * (size_t)n);
        ref_rand_perm(perm, (size_t)n, seed);
        nx = (int64_t)(k * (size_t)max_points_per_centroid);
        xt = (float*)malloc(sizeof(float) * (size_t)nx * d);
        for (int64_t i = 0; i < nx; i++)
            memcpy(xt + i * d, x + perm[i] * d, sizeof(float) * d);
        free(perm);
        xs = xt;
    }
    float* cent = ix->centroids;
    { /* initial centroids: :425-441 */
        int64_t* perm = (int64_t*)malloc(sizeof(int64_t) * (size_t)nx);
        ref_rand_perm(perm, (size_t)nx, (int64_t)seed + 1);
        for (size_t i = 0; i < k; i++)
            memcpy(cent + i * d, xs + perm[i] * d, sizeof(float) * d);
        free(perm);
    }
    const int spherical = ix->metric == ORC_METRIC_INNER_PRODUCT;
    if (spherical)
        renorm_rows(cent, k, d);
    int64_t* assign = (int64_t*)malloc(sizeof(int64_t) * (size_t)nx);
    float* dis = (float*)malloc(sizeof(float) * (size_t)nx);
    float* hassign = (float*)malloc(sizeof(float) * k);
    double obj = 0;
    for (int it = 0; it < niter; it++) {
        const int64_t bs = 65536;
        for (int64_t i0 = 0; i0 < nx; i0 += bs) { /* index.search(nx, x, 1, dis, assign) :470-474 */
            int64_t ni = nx - i0 < bs ? nx - i0 : bs;
            orc_knn(ix->metric, xs + i0 * d, cent, d, ni, k, 1, dis + i0, assign + i0);
        }
        obj = 0;
        for (int64_t j = 0; j < nx; j++)
            obj += dis[j];
        /* compute_centroids :150-221: sums in input order, then * 1/count */
        memset(cent, 0, sizeof(float) * k * d);
        memset(hassign, 0, sizeof(float) * k);
        for (int64_t i = 0; i < nx; i++) {
            float* c = cent + assign[i] * d;
            const float* xi = xs + i * d;
            hassign[assign[i]] += 1.0f;
            for (size_t j = 0; j < d; j++)
                c[j] += xi[j];
        }
        for (size_t ci = 0; ci < k; ci++) {
            if (hassign[ci] == 0)
                continue;
            float norm = 1 / hassign[ci];
            for (size_t j = 0; j < d; j++)
                cent[ci * d + j] *= norm;
        }
        /* split_clusters :232-276 */
        if ((size_t)nx > k) {
            Mt19937 g;
            mt_seed(&g, 1234);
            for (size_t ci = 0; ci < k; ci++) {
                if (hassign[ci] != 0)
                    continue;
                size_t cj;
                for (cj = 0; 1; cj = (cj + 1) % k) {
                    float p = (float)((hassign[cj] - 1.0) / (float)(nx - (int64_t)k));
                    float r = mt_next(&g) / (float)4294967295u;
                    if (r < p)
                        break;
                }
                memcpy(cent + ci * d, cent + cj * d, sizeof(float) * d);
                for (size_t j = 0; j < d; j++) {
                    if (j % 2 == 0) {
                        cent[ci * d + j] = (float)(cent[ci * d + j] * (1 + 1 / 1024.));
                        cent[cj * d + j] = (float)(cent[cj * d + j] * (1 - 1 / 1024.));
                    } else {
                        cent[ci * d + j] = (float)(cent[ci * d + j] * (1 - 1 / 1024.));
                        cent[cj * d + j] = (float)(cent[cj * d + j] * (1 + 1 / 1024.));
                    }
                }
                hassign[ci] = hassign[cj] / 2;
                hassign[cj] -= hassign[ci];
            }
        }
        if (spherical)
            renorm_rows(cent, k, d);
    }
    if (objective_out)
        *objective_out = obj;
    free(assign);
    free(dis);
    free(hassign);
    free(xt);
    return 0;
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
