/* b200ivf_c.h -- C-ABI of the B200-native IVF-Flat search path.
 *
 * Drop-in boundary for ONE path of the reference (faiss 1.10 fork Quaternijkon/hnsw-ivf):
 * batched IndexIVFFlat::search(n, x, k, D, I) with a flat coarse quantizer, METRIC_L2 /
 * METRIC_INNER_PRODUCT, fp32. Conventions follow the reference's own C API:
 *   - opaque handles, created by *_new*, destroyed by *_free      (c_api/macros_impl.h:90-93)
 *   - every fallible call returns int: 0 OK, -1 unknown, -2 library exception, -4 std exception
 *                                                                  (c_api/error_c.h:19-30)
 *   - message of the last failure on this thread: b200_get_last_error()
 *                                                                  (c_api/error_impl.cpp:15-27)
 *   - idx_t = int64_t, metric enum values as faiss/MetricType.h:23-24
 * Plain pointers and sizes only. x / distances / labels may be HOST or DEVICE pointers (detected
 * with cudaPointerGetAttributes); host buffers are staged through pinned memory.
 * There is no CPU fallback: every entry point fails with -2 if no CUDA device is usable.
 */
#ifndef B200IVF_C_H
#define B200IVF_C_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef int64_t b200_idx_t; /* faiss idx_t, faiss/MetricType.h:50 */

typedef enum B200MetricType { /* == FaissMetricType, c_api/Index_c.h:28-41 */
    B200_METRIC_INNER_PRODUCT = 0,
    B200_METRIC_L2 = 1
} B200MetricType;

typedef struct B200IndexIVFFlat_H B200IndexIVFFlat;

/* layout identical to FaissIndexIVFStats, c_api/IndexIVF_c.h:162-169 */
typedef struct B200IndexIVFStats {
    size_t nq;
    size_t nlist;
    size_t ndis;
    size_t nheap_updates; /* always 0: the reference counts replacements of ONE heap fed in probe
                           * order (faiss/IndexIVFFlat.cpp:163,174-177); a list-major scan has
                           * no such order, the figure is not reproducible */
    double quantization_time; /* ms, device time of the coarse quantizer */
    double search_time;       /* ms, device time of the whole search */
} B200IndexIVFStats;

/* c_api/error_impl.cpp:15-27 faiss_get_last_error */
const char* b200_get_last_error(void);

/* faiss_IndexIVFFlat_new_with_metric, c_api/IndexIVFFlat_c.h:53-58. The flat quantizer
 * (IndexFlat, faiss/IndexFlat.h:23) is part of the object; `device` is the CUDA ordinal. */
int b200_IndexIVFFlat_new_with_metric(B200IndexIVFFlat** p_index, size_t d, size_t nlist,
                                      B200MetricType metric, int device);
/* faiss_IndexIVFFlat_free, c_api/IndexIVFFlat_c.h:27 */
void b200_IndexIVFFlat_free(B200IndexIVFFlat* index);

/* getters/setters, c_api/IndexIVFFlat_c.h:29-32 + c_api/Index_c.h:85-97 */
size_t b200_IndexIVFFlat_nlist(const B200IndexIVFFlat* index);
size_t b200_IndexIVFFlat_nprobe(const B200IndexIVFFlat* index);
void b200_IndexIVFFlat_set_nprobe(B200IndexIVFFlat* index, size_t nprobe);
int b200_IndexIVFFlat_d(const B200IndexIVFFlat* index);
b200_idx_t b200_IndexIVFFlat_ntotal(const B200IndexIVFFlat* index);
int b200_IndexIVFFlat_is_trained(const B200IndexIVFFlat* index);
B200MetricType b200_IndexIVFFlat_metric_type(const B200IndexIVFFlat* index);

/* Install trained centroids == quantizer->add(nlist, c) on an IndexFlat; afterwards
 * is_trained (faiss/IndexIVF.cpp:72-74,178). c: [nlist][d] fp32, host or device. */
int b200_IndexIVFFlat_set_centroids(B200IndexIVFFlat* index, const float* centroids);
int b200_IndexIVFFlat_get_centroids(const B200IndexIVFFlat* index, float* centroids_out);

/* Restrict this handle to the inverted lists [list_lo, list_hi): the list-range shard of
 * SUBSET_TYPE_INVLIST (faiss/invlists/InvertedLists.cpp:168-176, faiss/gpu/GpuCloner.cpp:303-314).
 * Centroids stay complete; vectors assigned elsewhere are dropped by add. Call before add. */
int b200_IndexIVFFlat_set_list_range(B200IndexIVFFlat* index, size_t list_lo, size_t list_hi);

/* faiss_Index_add_with_ids, c_api/Index_c.h:133-137 (xids NULL: ids ntotal..ntotal+n-1,
 * faiss/IndexIVFFlat.cpp:72) */
int b200_IndexIVFFlat_add_with_ids(B200IndexIVFFlat* index, b200_idx_t n, const float* x,
                                   const b200_idx_t* xids);
/* faiss_IndexIVFFlat_add_core, c_api/IndexIVFFlat_c.h:60-65 (precomputed_idx < 0 skips) */
int b200_IndexIVFFlat_add_core(B200IndexIVFFlat* index, b200_idx_t n, const float* x,
                               const b200_idx_t* xids, const int64_t* precomputed_idx);
/* faiss_Index_reset, c_api/Index_c.h:151 */
int b200_IndexIVFFlat_reset(B200IndexIVFFlat* index);

/* Room for sizes[l] rows in list l before the vectors arrive (sizes: [nlist] int64, host or
 * device; entries outside the handle's list range are ignored). Adds that stay within the
 * reservation append in place: no second slab, nothing moves. The list-wise form of
 * GpuIndexIVF::reserveMemory (faiss/gpu/GpuIndexIVF.h) / InvertedLists::resize
 * (faiss/invlists/InvertedLists.h:103). */
int b200_IndexIVFFlat_reserve_lists(B200IndexIVFFlat* index, const int64_t* sizes);
/* sizes of all nlist lists at once (InvertedLists::list_size, faiss/invlists/InvertedLists.h:56),
 * HOST array of nlist entries */
int b200_IndexIVFFlat_get_list_sizes(const B200IndexIVFFlat* index, int64_t* sizes_out);

/* faiss_IndexIVF_get_list_size, c_api/IndexIVF_c.h:123 */
size_t b200_IndexIVFFlat_get_list_size(const B200IndexIVFFlat* index, size_t list_no);
/* faiss_IndexIVF_invlists_get_ids, c_api/IndexIVF_c.h:148-151 (host buffer) */
int b200_IndexIVFFlat_invlists_get_ids(const B200IndexIVFFlat* index, size_t list_no,
                                       b200_idx_t* invlist);
/* InvertedLists::get_codes as fp32 rows (faiss/invlists/InvertedLists.h:62), host buffer */
int b200_IndexIVFFlat_invlists_get_codes(const B200IndexIVFFlat* index, size_t list_no,
                                         float* codes);

/* faiss_Index_search, c_api/Index_c.h:108-114 == IndexIVF::search faiss/IndexIVF.cpp:303 */
int b200_IndexIVFFlat_search(const B200IndexIVFFlat* index, b200_idx_t n, const float* x,
                             b200_idx_t k, float* distances, b200_idx_t* labels);
/* faiss_IndexIVF_search_preassigned, c_api/IndexIVF_c.h:112-121; store_pairs != 0 returns
 * list_no << 32 | offset as labels (faiss/IndexIVFFlat.cpp:172);
 * nprobe is the handle's nprobe; centroid_dis may be NULL (unused by IVFFlat,
 * faiss/IndexIVFFlat.cpp:143) */
int b200_IndexIVFFlat_search_preassigned(const B200IndexIVFFlat* index, b200_idx_t n,
                                         const float* x, b200_idx_t k, const b200_idx_t* assign,
                                         const float* centroid_dis, float* distances,
                                         b200_idx_t* labels, int store_pairs);
/* quantizer->search (faiss/IndexFlat.cpp:38-69): the nprobe nearest centroids of each query */
int b200_IndexIVFFlat_quantize(const B200IndexIVFFlat* index, b200_idx_t n, const float* x,
                               b200_idx_t nprobe, float* centroid_dis, b200_idx_t* assign);

/* b200_IndexIVFFlat_quantize for callers that cannot allocate device memory (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc). It has NO counterpart in the
 * reference. x, centroid_dis, assign must be
 * DEVICE pointers; n <= 4096 and nprobe <= min(nlist, 64), anything larger fails with -2 and a
 * message naming the limit. The call makes no cudaMalloc / cudaFree: its scratch (launch plan and
 * partial candidate lists) is a fixed 8 MiB file-scope __device__ array that every GPU which loads
 * the library holds until exit. ALL calls on a device share that array: calls that overlap on the
 * GPU race on it -- serialise them. Work is enqueued on the handle's stream (set_stream). */
int b200_IndexIVFFlat_quantize_noalloc(const B200IndexIVFFlat* index, b200_idx_t n, const float* x,
                                       b200_idx_t nprobe, float* centroid_dis, b200_idx_t* assign);

/* indexIVF_stats (faiss/IndexIVF.cpp:1764) is per handle here */
void b200_IndexIVFFlat_get_stats(const B200IndexIVFFlat* index, B200IndexIVFStats* stats);
void b200_IndexIVFFlat_reset_stats(B200IndexIVFFlat* index);
/* device time in ms of the stages of the LAST search call: {coarse, plan, scan, select, total} */
void b200_IndexIVFFlat_last_timings(const B200IndexIVFFlat* index, double* ms5);

/* ... and the same five figures summed over every search call since reset_stats (so a timed loop
 * need not stop after each call to read them); returns the number of calls */
unsigned long long b200_IndexIVFFlat_stage_time_sums(const B200IndexIVFFlat* index, double* ms5);
/* Queries that were searched a second time on the exact kernel, since the handle was created. By
 * default ("tc_split" = 1) the list scan uses the tensor cores as a FILTER (one tf32 product, known
 * error bound) and the k-select proves per query that the exactly re-scored candidates contain the
 * exact top k; a query it cannot prove (many rows within the bound of the k-th distance) is redone
 * exactly before the call returns. Results never depend on the option; this counter says how
 * often the second pass ran. A handle whose second pass takes more than 1/64 of a batch keeps more
 * spare candidates from then on, and falls back to the 3xTF32 split if that is not enough.
 * (Searches in this mode read one 4-byte counter back before they return, so they complete
 * synchronously even with device buffers.) No counterpart in the reference. */
unsigned long long b200_IndexIVFFlat_exact_redo_count(const B200IndexIVFFlat* index);

/* Run on a caller-owned CUDA stream (cudaStream_t as void*; NULL = the handle's own stream, a
 * non-blocking one; pass cudaStreamLegacy, (void*)0x1, to be ordered with the legacy default
 * stream). search / search_preassigned / quantize calls whose buffers are ALL device pointers
 * return as soon as their work is enqueued on this stream (no host synchronization): results are
 * ordered with later work on the stream, statistics and stage times are read out when asked for,
 * and a failure only the device can see -- a key >= nlist in a preassigned table,
 * faiss/IndexIVF.cpp:520-524 -- is reported by b200_IndexIVFFlat_synchronize or by the next call
 * on the handle instead of by the call that caused it. */
int b200_IndexIVFFlat_set_stream(B200IndexIVFFlat* index, void* cuda_stream);
/* wait for the asynchronous calls in flight; returns their deferred failure, if any */
int b200_IndexIVFFlat_synchronize(const B200IndexIVFFlat* index);

/* Kernel-selection knobs of the scan (what faiss exposes through ParameterSpace::set_index_parameter,
 * faiss/AutoTune.h:127-170, for its own indexes): "simt_max_avg_q", "tcr_max_avg_q" (queries per
 * probed list below which the exact warp kernel / the rows-major tensor-core kernel takes the
 * scan), "tc_mode" (0 auto, 1 queries-major, 2 rows-major), "use_tc", "use_graph", "tc_split" (1 =
 * one tf32 product per 8 dimensions + checked re-rank, the default; 3 = 3xTF32 with a fixed re-rank
 * margin), "one_margin" (spare candidates kept per partial list in that mode, default 6),
 * "one_max_avg_q" (queries per probed list up to which that kernel is preferred, for 1024-row
 * lists; the cap grows with the cube of the list length), "one_min_items" (work items a search must
 * have for that kernel: 8 waves of CTAs by default), "coarse_one_min_nlist" (centroid tables at
 * least this long are scanned in the same checked form above nprobe 24), "use_pair_plan" (small
 * batches: one plan launch per search instead of three). They change
 * which kernels run, never the results. Also here, for want of a struct accessor in a C API: the
 * quantizer's ClusteringParameters as IndexIVF exposes them (index.cp, faiss/IndexIVF.h:39-41,
 * faiss/Clustering.h:23-64): "cp_niter", "cp_max_points_per_centroid", "cp_seed" -- read by train.
 * Unknown names fail with -2. */
int b200_IndexIVFFlat_set_option(B200IndexIVFFlat* index, const char* name, double value);

/* Number of CUDA kernels this library has launched in this process (all handles). */
unsigned long long b200_kernel_launches(void);
/* ... of which tensor-core (tcgen05) scan launches; tests use it to prove which path ran */
unsigned long long b200_tensor_core_launches(void);

/* Debugging aid: a pipeline wait of the tensor-core scan that makes no progress for ~2 s traps
 * (the call fails with a CUDA launch failure) instead of hanging the GPU. Returns 1 and fills
 * out8 = {1 + which wait, CTA, thread, iteration, parity, 0, 0, 0} if that has happened in this
 * process, else 0. */
int b200_debug_trap_info(unsigned long long* out8);

/* merge_knn_results (faiss/utils/Heap.cpp:158-232) as used by IndexShardsIVF::search
 * (faiss/IndexShardsIVF.cpp:224-242): all_* are [nshard][n][k] per-shard sorted tables, host or
 * device; runs on `device`, on `cuda_stream` (may be NULL). */
int b200_merge_knn_results(B200MetricType metric, int nshard, b200_idx_t n, b200_idx_t k,
                           const float* all_distances, const b200_idx_t* all_labels,
                           float* distances, b200_idx_t* labels, int device, void* cuda_stream);

/* ---- list-sharded index over several GPUs ---------------------------------------------------
 * IndexShardsIVF (faiss/IndexShardsIVF.h:21-44; C API of the family: c_api/IndexShards_c.h) with
 * shards that hold contiguous list RANGES (shard_type 4 / SUBSET_TYPE_INVLIST,
 * faiss/gpu/GpuCloner.cpp:303-314): one process per GPU, each creating ITS rank of the index.
 * The reference fans a search out to its shards with host threads and merges on the CPU
 * (faiss/IndexShardsIVF.cpp:195-242); here the exchange steps are NCCL all-gathers over NVLink on
 * the shard's stream and the merge runs on the device. Every call that takes the sharded handle
 * is COLLECTIVE: all ranks make it, in the same order. */
typedef struct B200IndexShardsIVFFlat_H B200IndexShardsIVFFlat;
#define B200_NCCL_UNIQUE_ID_BYTES 128
/* ncclGetUniqueId: one rank calls it and hands the 128 bytes to the others (MPI, sockets, a
 * torch.distributed store, ...) */
int b200_nccl_get_unique_id(void* out128);
/* rank `rank` of `world`; the rank's shard lives on CUDA device `device` and owns the lists
 * [rank * nlist / world, (rank + 1) * nlist / world). world == 1 needs no id (may be NULL). */
int b200_IndexShardsIVFFlat_new(B200IndexShardsIVFFlat** p_index, size_t d, size_t nlist,
                                B200MetricType metric, int device, int rank, int world,
                                const void* nccl_unique_id128);
void b200_IndexShardsIVFFlat_free(B200IndexShardsIVFFlat* index);
/* faiss_IndexShards_at, c_api/IndexShards_c.h:41: this rank's shard, BORROWED (nprobe, options,
 * statistics, list access, set_stream go through it) */
B200IndexIVFFlat* b200_IndexShardsIVFFlat_local(B200IndexShardsIVFFlat* index);
int b200_IndexShardsIVFFlat_rank(const B200IndexShardsIVFFlat* index);
int b200_IndexShardsIVFFlat_world(const B200IndexShardsIVFFlat* index);
b200_idx_t b200_IndexShardsIVFFlat_ntotal(const B200IndexShardsIVFFlat* index);
/* root >= 0: rank root's table is broadcast (the others may pass NULL); root < 0: every rank
 * passes the same table */
int b200_IndexShardsIVFFlat_set_centroids(B200IndexShardsIVFFlat* index, const float* centroids,
                                          int root);
/* IndexShardsIVF::train, faiss/IndexShardsIVF.cpp:102-135: the common quantizer is trained once,
 * on rank root with its n vectors, and shared */
int b200_IndexShardsIVFFlat_train(B200IndexShardsIVFFlat* index, b200_idx_t n, const float* x,
                                  int root);
/* IndexShardsIVF::add_with_ids, faiss/IndexShardsIVF.cpp:137-154, distributed: each rank brings
 * its OWN n vectors (n may differ per rank and may be 0); they are quantized where they are and
 * sent to the rank that owns their list (NCCL send/recv). xids NULL: ids continue from ntotal in
 * rank order. */
int b200_IndexShardsIVFFlat_add_with_ids(B200IndexShardsIVFFlat* index, b200_idx_t n,
                                         const float* x, const b200_idx_t* xids);
/* the same with the list numbers given (faiss_IndexIVFFlat_add_core, c_api/IndexIVFFlat_c.h:60-65) */
int b200_IndexShardsIVFFlat_add_core(B200IndexShardsIVFFlat* index, b200_idx_t n, const float* x,
                                     const b200_idx_t* xids, const int64_t* precomputed_idx);
/* IndexShardsIVF::search, faiss/IndexShardsIVF.cpp:158-243. root < 0: every rank passes the same
 * n queries; root >= 0: rank root's queries are broadcast (the others pass a device buffer to
 * receive them, or NULL). Every rank receives the merged tables. All-device calls return once
 * their work is enqueued on the shard's stream. */
int b200_IndexShardsIVFFlat_search(const B200IndexShardsIVFFlat* index, b200_idx_t n,
                                   const float* x, b200_idx_t k, float* distances,
                                   b200_idx_t* labels, int root);
/* device ms of the phases of the searches since the last reset, summed: {query broadcast, coarse
 * quantizer (this rank's query slice), probe-table all-gather, local plan + scan + k-select,
 * result all-gather, merge, total}; returns the number of search calls */
unsigned long long b200_IndexShardsIVFFlat_phase_time_sums(const B200IndexShardsIVFFlat* index,
                                                           double* ms7);
void b200_IndexShardsIVFFlat_reset_phase_times(B200IndexShardsIVFFlat* index);
int b200_IndexShardsIVFFlat_synchronize(const B200IndexShardsIVFFlat* index);

/* faiss_write_index_fname / faiss_read_index_fname (c_api/index_io_c.h:31,44) for this index type:
 * the reference's own file format ("IwFl" header, IndexFlat quantizer "IxF2"/"IxFI", "ilar"
 * ArrayInvertedLists; faiss/impl/index_write.cpp:635-638, index_read.cpp:782-787), so an index
 * written by the reference loads onto `device` and an index built here is readable by the
 * reference. io_flags is accepted for signature parity (no mmap: the lists live in HBM). */
int b200_write_index_fname(const B200IndexIVFFlat* index, const char* fname);
int b200_read_index_fname(const char* fname, int io_flags, int device, B200IndexIVFFlat** p_out);

/* ---- training ------------------------------------------------------------------------------
 * faiss_Index_train, c_api/Index_c.h:72 == IndexIVF::train faiss/IndexIVF.cpp:1404 -> k-means of
 * Level1Quantizer::train_q1 (:66-127; niter 10, <= 256 points per centroid, spherical for inner
 * product, :55,:179-182). A no-op when centroids are installed already; the index must be empty. */
int b200_IndexIVFFlat_train(B200IndexIVFFlat* index, b200_idx_t n, const float* x);

/* ---- IDSelector family ---------------------------------------------------------------------
 * faiss/impl/IDSelector.h:19-170, C API c_api/impl/AuxIndexStructures_c.h:51-116. Batch and
 * Bitmap copy their input (host pointers) to `device`. Not/And/Or/XOr borrow their operands,
 * which must outlive them, as in the reference. */
typedef struct B200IDSelector_H B200IDSelector;
int b200_IDSelectorRange_new(B200IDSelector** p_sel, b200_idx_t imin, b200_idx_t imax);
int b200_IDSelectorBatch_new(B200IDSelector** p_sel, size_t n, const b200_idx_t* indices,
                             int device);
int b200_IDSelectorBitmap_new(B200IDSelector** p_sel, size_t n, const uint8_t* bitmap, int device);
int b200_IDSelectorNot_new(B200IDSelector** p_sel, const B200IDSelector* sel);
int b200_IDSelectorAnd_new(B200IDSelector** p_sel, const B200IDSelector* lhs,
                           const B200IDSelector* rhs);
int b200_IDSelectorOr_new(B200IDSelector** p_sel, const B200IDSelector* lhs,
                          const B200IDSelector* rhs);
int b200_IDSelectorXOr_new(B200IDSelector** p_sel, const B200IDSelector* lhs,
                           const B200IDSelector* rhs);
/* faiss_IDSelector_is_member, AuxIndexStructures_c.h:54 */
int b200_IDSelector_is_member(const B200IDSelector* sel, b200_idx_t id);
void b200_IDSelector_free(B200IDSelector* sel);

/* ---- SearchParametersIVF -------------------------------------------------------------------
 * faiss/IndexIVF.h:73-88, C API c_api/IndexIVF_c.h:22-35. sel may be NULL and is borrowed. */
typedef struct B200SearchParametersIVF_H B200SearchParametersIVF;
int b200_SearchParametersIVF_new(B200SearchParametersIVF** p_sp);
int b200_SearchParametersIVF_new_with(B200SearchParametersIVF** p_sp, B200IDSelector* sel,
                                      size_t nprobe, size_t max_codes);
void b200_SearchParametersIVF_free(B200SearchParametersIVF* sp);
const B200IDSelector* b200_SearchParametersIVF_sel(const B200SearchParametersIVF* sp);
size_t b200_SearchParametersIVF_nprobe(const B200SearchParametersIVF* sp);
void b200_SearchParametersIVF_set_nprobe(B200SearchParametersIVF* sp, size_t nprobe);
size_t b200_SearchParametersIVF_max_codes(const B200SearchParametersIVF* sp);
void b200_SearchParametersIVF_set_max_codes(B200SearchParametersIVF* sp, size_t max_codes);

/* IndexIVF::max_codes (faiss/IndexIVF.h:96), used when a search passes no params */
size_t b200_IndexIVFFlat_max_codes(const B200IndexIVFFlat* index);
void b200_IndexIVFFlat_set_max_codes(B200IndexIVFFlat* index, size_t max_codes);

/* faiss_Index_search_with_params, c_api/Index_c.h:128-135: nprobe / max_codes / selector of
 * `params` (faiss/IndexIVF.cpp:416-450,597-631; faiss/IndexIVFFlat.cpp:166) */
int b200_IndexIVFFlat_search_with_params(const B200IndexIVFFlat* index, b200_idx_t n,
                                         const float* x, b200_idx_t k,
                                         const B200SearchParametersIVF* params, float* distances,
                                         b200_idx_t* labels);
/* IndexIVF::search_preassigned with its trailing arguments (faiss/IndexIVF.h:235-246):
 * store_pairs != 0 returns list_no << 32 | offset as labels (faiss/IndexIVFFlat.cpp:172);
 * params may be NULL */
int b200_IndexIVFFlat_search_preassigned_with_params(
        const B200IndexIVFFlat* index, b200_idx_t n, const float* x, b200_idx_t k,
        const b200_idx_t* assign, const float* centroid_dis, float* distances, b200_idx_t* labels,
        int store_pairs, const B200SearchParametersIVF* params);

/* the fork's IndexIVF::search_stats (faiss/IndexIVF.cpp:727-862, faiss/IndexIVF.h:28-32,322-330):
 * search + one QueryLatencyStats per query (HOST array of n). quantization_us is the batch's
 * coarse time / n as in the reference (:767-771); list_scan_us is the batch's plan + scan +
 * select device time shared out by each query's scanned rows (a batched kernel has no per-query
 * clock); total_us is their sum. params may be NULL. */
typedef struct B200QueryLatencyStats {
    double total_us;
    double quantization_us;
    double list_scan_us;
} B200QueryLatencyStats;
int b200_IndexIVFFlat_search_stats(const B200IndexIVFFlat* index, b200_idx_t n, const float* x,
                                   b200_idx_t k, float* distances, b200_idx_t* labels,
                                   const B200SearchParametersIVF* params,
                                   B200QueryLatencyStats* per_query_stats);

/* ---- range search --------------------------------------------------------------------------
 * faiss::RangeSearchResult (faiss/impl/AuxIndexStructures.h:30-58), C API
 * c_api/impl/AuxIndexStructures_c.h:20-49: query i owns entries lims[i]..lims[i+1]. The arrays
 * are host memory owned by the object and valid until the next range search into it / free. */
typedef struct B200RangeSearchResult_H B200RangeSearchResult;
int b200_RangeSearchResult_new(B200RangeSearchResult** p_rsr, b200_idx_t nq);
void b200_RangeSearchResult_free(B200RangeSearchResult* rsr);
size_t b200_RangeSearchResult_nq(const B200RangeSearchResult* rsr);
void b200_RangeSearchResult_lims(B200RangeSearchResult* rsr, size_t** lims);
void b200_RangeSearchResult_labels(B200RangeSearchResult* rsr, b200_idx_t** labels,
                                   float** distances);
/* faiss_Index_range_search, c_api/Index_c.h:148-153 == IndexIVF::range_search
 * faiss/IndexIVF.cpp:1203-1241: L2 keeps dis < radius, inner product dis > radius
 * (faiss/IndexIVFFlat.cpp:196); params may be NULL (nprobe and selector are honoured) */
int b200_IndexIVFFlat_range_search(const B200IndexIVFFlat* index, b200_idx_t n, const float* x,
                                   float radius, B200RangeSearchResult* result,
                                   const B200SearchParametersIVF* params);
/* IndexIVF::range_search_preassigned, faiss/IndexIVF.cpp:1243-1399 */
int b200_IndexIVFFlat_range_search_preassigned(const B200IndexIVFFlat* index, b200_idx_t n,
                                               const float* x, float radius,
                                               const b200_idx_t* assign, const float* centroid_dis,
                                               B200RangeSearchResult* result, int store_pairs,
                                               const B200SearchParametersIVF* params);

#ifdef __cplusplus
}
#endif
#endif /* B200IVF_C_H */
