/* oracle/ivfflat_oracle.h -- TEST INFRASTRUCTURE, not product code.
 *
 * Plain-C restatement of the reference's CPU IndexIVFFlat search path (faiss 1.10 fork under
 * /root/reference). Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use
 * it, and only as the checker. Parity pinning: this restatement is checked against the compiled
 * reference itself (oracle/_ref/libref_ivfflat.so, see tests/test_oracle_vs_reference.py) and
 * against golden vectors generated from the compiled reference (tests/golden/, generator
 * tests/golden/make_golden.py). The reference holds no stored golden vectors of its own for this
 * path.
 */
#ifndef IVFFLAT_ORACLE_H
#define IVFFLAT_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_METRIC_INNER_PRODUCT 0 /* faiss/MetricType.h:23 */
#define ORC_METRIC_L2 1            /* faiss/MetricType.h:24 */

typedef struct OrcIndex OrcIndex;

typedef struct OrcStats { /* faiss/IndexIVF.h:567-580 */
    size_t nq, nlist, ndis, nheap_updates;
    double quantization_time, search_time;
} OrcStats;

/* faiss/utils/distances_simd.cpp:194-230 */
float orc_fvec_L2sqr(const float* x, const float* y, size_t d);
float orc_fvec_inner_product(const float* x, const float* y, size_t d);
float orc_fvec_norm_L2sqr(const float* x, size_t d);

/* faiss/utils/Heap.h:316-338, 111-150, 367-388, 420-450 (metric picks CMax(L2)/CMin(IP)) */
void orc_heap_heapify(int metric, size_t k, float* val, int64_t* ids);
void orc_heap_replace_top(int metric, size_t k, float* val, int64_t* ids, float v, int64_t id);
size_t orc_heap_addn(int metric, size_t k, float* val, int64_t* ids, const float* x,
                     const int64_t* xids, size_t n);
size_t orc_heap_reorder(int metric, size_t k, float* val, int64_t* ids);

/* faiss/utils/distances.cpp:836-920: knn over a flat table; nx < 20 takes the seq path
 * (:136-199), else the ||x||^2+||y||^2-2xy "BLAS" form (:258-342, :202-254). */
void orc_knn(int metric, const float* x, const float* y, size_t d, size_t nx, size_t ny, size_t k,
             float* vals, int64_t* ids);

OrcIndex* orc_index_new(size_t d, size_t nlist, int metric);
void orc_index_free(OrcIndex* ix);
void orc_index_set_centroids(OrcIndex* ix, const float* c);
size_t orc_index_list_size(const OrcIndex* ix, size_t l);
const float* orc_index_list_codes(const OrcIndex* ix, size_t l);
const int64_t* orc_index_list_ids(const OrcIndex* ix, size_t l);
int64_t orc_index_ntotal(const OrcIndex* ix);

/* faiss/IndexIVFFlat.cpp:46-91 (ids==NULL -> ntotal+i; assign<0 skipped) */
void orc_index_add_core(OrcIndex* ix, int64_t n, const float* x, const int64_t* xids,
                        const int64_t* assign);
/* faiss/IndexIVF.cpp:191-195,197-241: quantizer->assign then add_core */
void orc_index_add_with_ids(OrcIndex* ix, int64_t n, const float* x, const int64_t* xids);

/* quantizer->search: faiss/IndexFlat.cpp:38-69 */
void orc_index_quantize(const OrcIndex* ix, int64_t n, const float* x, size_t nprobe, float* Dq,
                        int64_t* Iq);
/* faiss/IndexIVF.cpp:399-723 (parallel_mode 0) + IVFFlatScanner::scan_codes
 * faiss/IndexIVFFlat.cpp:155-179. Returns 0, or -2 on an invalid key (>= nlist). */
int orc_index_search_preassigned(const OrcIndex* ix, int64_t n, const float* x, int64_t k,
                                 size_t nprobe, const int64_t* keys, float* D, int64_t* I,
                                 OrcStats* stats);
/* IDSelector family, faiss/impl/IDSelector.h:19-170 / IDSelector.cpp:17-123. batch/bitmap copy
 * their input; not/and/or/xor borrow their operands (rhs NULL for not). */
#define ORC_SEL_RANGE 0
#define ORC_SEL_BATCH 1
#define ORC_SEL_BITMAP 2
#define ORC_SEL_NOT 3
#define ORC_SEL_AND 4
#define ORC_SEL_OR 5
#define ORC_SEL_XOR 6
typedef struct OrcSel OrcSel;
OrcSel* orc_sel_range(int64_t imin, int64_t imax);
OrcSel* orc_sel_batch(size_t n, const int64_t* ids);
OrcSel* orc_sel_bitmap(size_t nbytes, const uint8_t* bitmap);
OrcSel* orc_sel_combine(int kind, const OrcSel* lhs, const OrcSel* rhs);
void orc_sel_free(OrcSel* s);
int orc_sel_is_member(const OrcSel* s, int64_t id);

/* search_preassigned with the SearchParametersIVF arms: max_codes (faiss/IndexIVF.cpp:545-548,
 * 619-621; 0 = no limit), selector (faiss/IndexIVFFlat.cpp:166), store_pairs (:172) */
int orc_index_search_preassigned_ex(const OrcIndex* ix, int64_t n, const float* x, int64_t k,
                                    size_t nprobe, const int64_t* keys, float* D, int64_t* I,
                                    OrcStats* stats, size_t max_codes, const OrcSel* sel,
                                    int store_pairs);
/* faiss/IndexIVF.cpp:1243-1399 + scan_codes_range faiss/IndexIVFFlat.cpp:181-201. lims: [n+1]
 * caller buffer; *labels_out / *dis_out are malloc'd (orc_free). */
int orc_index_range_search_preassigned(const OrcIndex* ix, int64_t n, const float* x, float radius,
                                       size_t nprobe, const int64_t* keys, const OrcSel* sel,
                                       int store_pairs, size_t* lims, int64_t** labels_out,
                                       float** dis_out);
void orc_free(void* p);
/* k-means of the coarse quantizer as IndexIVF::train runs it (faiss/IndexIVF.cpp:66-127 ->
 * faiss/Clustering.cpp:296-600); writes the centroids into the index. objective_out (may be
 * NULL): sum of the assignment distances of the last iteration (:486-489). */
int orc_index_train(OrcIndex* ix, int64_t n, const float* x, int niter, int seed,
                    int max_points_per_centroid, double* objective_out);
const float* orc_index_centroids(const OrcIndex* ix);

/* faiss/IndexIVF.cpp:303-397; nt = number of OpenMP slices to emulate (slice size decides the
 * seq-vs-BLAS coarse form, :358-368). */
int orc_index_search(const OrcIndex* ix, int64_t n, const float* x, int64_t k, size_t nprobe,
                     int nt, float* D, int64_t* I, OrcStats* stats);

/* faiss/utils/Heap.cpp:158-232 */
void orc_merge_knn_results(int metric, size_t n, size_t k, int nshard, const float* all_D,
                           const int64_t* all_I, float* D, int64_t* I);

#ifdef __cplusplus
}
#endif
#endif
