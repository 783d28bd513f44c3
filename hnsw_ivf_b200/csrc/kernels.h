// kernels.h -- launch interfaces of the hand-written sm_100a kernels (host side).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>

namespace b200ivf {

// number of kernels this library has launched (bench.py reports it as gpu_launches)
extern std::atomic<unsigned long long> g_kernel_launches;
// ... and how many of them were tensor-core (tcgen05) scans
extern std::atomic<unsigned long long> g_tc_launches;
// timing experiments only: bits that switch stages of the tensor-core scans off (scan_tcr.cu);
// from B200IVF_TC_DBG at load, or set_option("tc_dbg"). Results are WRONG while it is non-zero.
extern std::atomic<int> g_tc_dbg;
inline void count_launches(int n) {
    g_kernel_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed);
}

// ---- K2 (and K1 as a scan of the centroid table): ivf_scan ------------------------------------
// One WARP per work item (list, sub-group of <= RQ queries probing it, row split). Replaces
// IVFFlatScanner::scan_codes (faiss/IndexIVFFlat.cpp:155-179) run per (query, probe) by
// IndexIVF::search_preassigned (faiss/IndexIVF.cpp:597-631).
struct ScanArgs {
    const float4* slab;         // interleaved 32-row blocks (common.cuh)
    const uint32_t* list_blk0;  // [nlists] first block of each list
    const int32_t* list_len;    // [nlists] rows in each list
    const int32_t* grp_start;   // [nlists+1] first pair slot of each list's query group
    const int32_t* pair_query;  // [npairs] query number of each pair slot (nullptr: identity)
    const uint32_t* item_off;   // [nlists+1] exclusive prefix of ceil(count/RQ)*nsplit
    int nlists;
    int nsplit;                 // row splits per list
    int k;                      // candidates kept per (pair, split)
    int cap;                    // per-query candidate buffer entries, power of two >= k + 32
    int d;                      // true dimension
    int nf4;                    // float4 columns per row = ceil(d/4)
    const float* queries;       // [n][d] row-major, device
    uint64_t* part;             // [npairs][nsplit][part_mul][k] keys out
    int part_mul = 1;           // 2 when the tensor-core scan shares `part` (it leaves two lists
                                // per (pair, split)): the second one is filled with sentinels
    // SearchParametersIVF (either one non-null selects the filtered kernel, rv must be 4):
    const uint32_t* row_mask = nullptr;  // [slab blocks] selector mask, bit = row in block
    const int32_t* pair_limit = nullptr; // [npairs] rows of the list that count for the pair
};

// rq in {1,4,8}; warps_per_cta chosen by the caller so that shared memory fits.
size_t scan_smem_bytes(int rq, int warps_per_cta, int nf4, int cap);
// rv: 32-row blocks per warp step, 4 (streaming) or 1 (rq == 1 only: one block per step with a
// deeper unroll, for row splits of a single block in the latency-bound small-batch regime)
cudaError_t launch_ivf_scan(int metric, int rq, int warps_per_cta, uint32_t max_items,
                            const ScanArgs& a, cudaStream_t st, int rv = 4);

// ---- K2 for lists probed by many queries: CTA-tiled scan (scan_tile.cu) ------------------------
// One CTA per work item (list, tile of <= kTileTQ queries probing it, row split).
constexpr int kTileTQ = 32;
constexpr int kTensorMapBytes = 128;
struct TileScanArgs {
    const void* tmap;           // HOST pointer to the 128-byte tensor map of the slab
                                // (make_slab_tensor_map); passed to the kernel by value
    const uint32_t* list_blk0;
    const int32_t* list_len;
    const int32_t* grp_start;   // [nlists+1]
    const int32_t* pair_query;  // [npairs] (nullptr: identity)
    const uint32_t* item_off;   // [nlists+1] exclusive prefix of ceil(count/TQ)*nsplit over the
                                // lists given to this kernel (0 items for the others)
    int nlists;
    int nsplit;
    int k;
    int cap;                    // power of two > k
    int d;
    int nf4;
    int nstages;                // bulk-copy ring depth, 2..4
    const float* queries;
    uint64_t* part;             // [npairs][nsplit][k]
};
size_t scan_tile_smem_bytes(int nf4, int cap, int nstages);
cudaError_t make_slab_tensor_map(void* map_out, const void* slab, int nf4, uint64_t nblocks);
cudaError_t launch_ivf_scan_tile(int metric, uint32_t max_items, const TileScanArgs& a,
                                 cudaStream_t st);

// ---- K1 and K2 on the tensor cores (scan_tc.cu): tcgen05 3xTF32, d <= 128 ----------------------
// One CTA per work item (list, tile of <= kTcTQ queries probing it, row split). Scores are
// approximate (~1e-6 of |x||y|): launch_select must be given rerank = true downstream.
constexpr int kTcTQ = 128;
struct TcScanArgs {
    const void* tmap;           // HOST pointer to the tensor map from make_slab_tensor_map_tc
    const uint32_t* list_blk0;
    const int32_t* list_len;
    const int32_t* grp_start;   // [nlists+1]
    const int32_t* pair_query;  // [npairs] (nullptr: identity)
    const uint32_t* item_off;   // [nlists+1] exclusive prefix of ceil(count/kTcTQ)*nsplit
    const int32_t* item_list = nullptr; // [items] list of each item (else: binary search)
    const float* norms;         // [slab rows] |y|^2 (read for METRIC_L2 only)
    int nlists;
    int nsplit;
    int kpart;                  // keys kept per (pair, split): 16 or 32, from scan_tc_kpart
    int d;
    int nf4;                    // <= 32
    int nstages;                // TMA ring depth (kTcStages)
    int tq = 0;                 // rows-major kernel only: queries per item (scan_tcr_tq)
    const float* queries;
    uint64_t* part;             // [npairs][nsplit][2][kpart]: one list per epilogue group
    // [n] per-query pruning bound shared by all the work items of a search: ordered_bits() of the
    // best KP-th score any item has seen for the query so far (the reference keeps ONE heap per
    // query across its probes, faiss/IndexIVF.cpp:597-631; this is that heap's top). Filled with
    // ordered_bits(FLT_MAX) before the launch (launch_fill_u32). nullptr: off.
    uint32_t* qthr = nullptr;
    // SearchParametersIVF (rows-major kernel only): selector row mask ([slab blocks], bit = row in
    // block) and max_codes row limits per pair slot; nullptr = unfiltered
    const uint32_t* row_mask = nullptr;
    const int32_t* pair_limit = nullptr;
    // 1 = ONE tcgen05.mma per 8 dimensions (tf32(x) . trunc_tf32(y)) instead of the 3xTF32 split:
    // scores are then good to |x||y| (2^-10 + 2^-11 + slack) only, and the k-select must be told
    // (SelectArgs::err_coef) so that it re-scores everything inside that window and checks that
    // nothing outside it could matter
    int one_mma = 0;
    unsigned long long* trap = nullptr; // host-mapped note of a wait that gave up (tc_trap_info_*)
    long long* trace = nullptr; // timing experiments only (B200IVF_TC_TRACE=<item>): clock64
    uint32_t trace_item = 0;    // stamps of one CTA, [role][tile]
    int dbg = 0;                // timing experiments only (B200IVF_TC_DBG / "tc_dbg"; results are
                                // wrong while a bit is set): 1 = the epilogue drops every score,
                                // 4 = split warps do not convert, 8 = no MMAs (rows-major one-MMA
                                // form) / stages recycled without loads (queries-major), 16 = no
                                // TMA loads (rows-major), 32 = no query-tile prologue loads
};
constexpr int kTcStages = 9;
// entries per partial list the tensor-core scan keeps for a request of k: 16 for k <= 12, 32 for
// k <= 24 (always a margin over k: the scores are approximate); 0 = not supported (d > 128, k > 24)
int scan_tc_kpart(int nf4, int k);
size_t scan_tc_smem_bytes(int nstages);
cudaError_t make_slab_tensor_map_tc(void* map_out, const void* slab, int nf4, uint64_t nblocks);
cudaError_t launch_ivf_scan_tc(int metric, uint32_t max_items, const TcScanArgs& a,
                               cudaStream_t st);

// ---- rows-major tensor-core scan (scan_tcr.cu): same arguments, <= kTcrTQ queries per item, ONE
// sorted list of kpart keys per (pair, split); nstages is chosen by the launcher. kpart up to 256.
constexpr int kTcrTQ = 64;
// d up to 1024; the queries per item shrink with the dimension (the query tile lives in shared
// memory): scan_tcr_tq(nf4, kpart), 0 = unsupported. scan_tcr_kpart: as scan_tc_kpart.
int scan_tcr_tq(int nf4, int kp, bool one = false);
// entries per partial list under the ONE-MMA filter (TcScanArgs::one_mma): the kept set must hold
// everything inside the error window of the k-th score: margin max(margin_min, k / 4); 0 = none
int scan_tcr_kpart_one(int nf4, int k, int margin_min);
// ... and of the two-CTAs-per-SM shape (<= 32 queries per item, half the shared memory): chosen by
// the launcher whenever TcScanArgs::tq is at most this. 0 = not available for (nf4, kp).
int scan_tcr_tq_small(int nf4, int kp, bool one = false);
int scan_tcr_kpart(int nf4, int k);
cudaError_t launch_ivf_scan_tcr(int metric, uint32_t max_items, const TcScanArgs& a,
                                cudaStream_t st);
// A pipeline wait of the tensor-core scan that makes no progress for ~2 s traps (the launch fails)
// instead of hanging the GPU; before it does it notes {1 + which wait, CTA, thread, iteration,
// parity} in host-mapped memory. nullptr until the first tensor-core launch.
unsigned long long* tc_trap_info_device();
const unsigned long long* tc_trap_info_host();

// ---- planning: invert the probe table into per-list query groups -------------------------------
struct PlanArgs {
    const int32_t* coarse_ids;  // [n][nprobe], -1 = no probe (GLOBAL list numbers)
    int n, nprobe, nlists;      // nlists: lists this plan covers = [list_lo, list_lo + nlists)
    int list_lo = 0;            // first list of the handle's range (a list-range shard plans only
                                // over its own lists: probes of other lists are simply not its)
    int nlist_total = 0;        // lists of the whole index: keys >= this are invalid (0: = nlists)
    const int32_t* list_len;    // [nlists], LOCAL numbering (list list_lo + i at i), as every
                                // per-list array below
    int32_t* count;             // [nlists] scratch (zeroed by launch_plan)
    int32_t* fill;              // [nlists] scratch (zeroed by launch_plan)
    int32_t* grp_start;         // [nlists+1] out
    uint32_t* item_off;         // [nlists+1] out
    int32_t* pair_query;        // [n*nprobe] out
    int32_t* slot_of;           // [n*nprobe] out: pair slot of (q,p) or -1
    int rq, nsplit;
    // lists with more than tile_min queries go to the tiled kernel (tile_tq queries per item) and
    // get 0 items in item_off; the others get 0 items in item_off_tile. tile_tq = 0: no tiling.
    int tile_tq, tile_min;
    uint32_t* item_off_tile;    // [nlists+1] out (may be nullptr when tile_tq == 0)
    int32_t* item_list_tile = nullptr; // [tile items] out: list of each tile item (optional)
    unsigned long long* stats;  // [4] += {nq, nlist visited, ndis, invalid keys}
    // max_codes: rows that count for each (query, probe) from launch_probe_limits, and the same
    // re-ordered by pair slot for the scan (both nullptr: every probed list counts whole)
    const int32_t* limit_qp = nullptr; // [n][nprobe]
    int32_t* limit_slot = nullptr;     // [n*nprobe] out
    // Pair-major plan (small batches): every (query, probe) pair becomes a "list" of its own -- pair
    // i at index i, one query each, nsplit items -- described by vlen / vblk0 (length and first
    // block of the probed list, 0 rows for a skipped probe). One launch, no per-list counters, no
    // prefix scan over nlist entries (38 us of a 120 us batch-1 search at nlist = 16384); the scan
    // is then given nlists = n * nprobe and these two arrays as its list tables. grp_start and
    // item_off must hold n * nprobe + 1 entries. Needs rq == 1, no tiling, no limits.
    int32_t* vlen = nullptr;           // [n*nprobe] out; non-null selects this plan
    uint32_t* vblk0 = nullptr;         // [n*nprobe] out
    const uint32_t* list_blk0 = nullptr; // [nlists] (pair-major plan only)
};
// max_codes walk of IndexIVF::search_preassigned (faiss/IndexIVF.cpp:597-631): writes the rows
// that count per (query, probe) (limit_qp, may be nullptr), turns the probes after the stop into
// -1 in coarse_ids, and leaves each query's nscan in query_ndis (may be nullptr).
// max_codes <= 0 is not accepted: pass LLONG_MAX for "no limit".
cudaError_t launch_probe_limits(int32_t* coarse_ids, int n, int nprobe, int nlists,
                                const int32_t* list_len, long long max_codes, int32_t* limit_qp,
                                unsigned long long* query_ndis, cudaStream_t st);
cudaError_t launch_plan(const PlanArgs& a, cudaStream_t st);
// the one-list plan used when the "list" is the centroid table and every query probes it
cudaError_t launch_plan_single(int32_t* grp_start, uint32_t* item_off, int n, uint32_t items,
                               cudaStream_t st);

// ---- range search (range.cu): IndexIVF::range_search_preassigned + scan_codes_range --------------
// One warp per pair slot of a plan made with rq = 1, nsplit = 1 (item == pair slot). Two passes
// with identical arithmetic: count, then fill at lims[query] + pair_off[slot].
struct RangeArgs {
    const float4* slab;
    const uint32_t* list_blk0;
    const int32_t* list_len;
    const int32_t* grp_start;   // [nlists+1]
    const int32_t* pair_query;  // [npairs]
    int nlists;
    int d, nf4;
    const float* queries;
    float radius;               // L2: dis < radius, inner product: dis > radius
    const uint32_t* row_mask = nullptr; // selector mask or nullptr
    const int64_t* ids = nullptr;       // slab row -> user id (nullptr: the row itself)
    int32_t* pair_cnt = nullptr;        // [npairs] out of the count pass
    const uint32_t* pair_off = nullptr; // [npairs] offset of the pair inside its query's run
    const unsigned long long* lims = nullptr; // [n+1] first result of each query
    float* out_dis = nullptr;
    int64_t* out_ids = nullptr;
};
cudaError_t launch_range_scan(int metric, bool fill, int64_t max_pairs, const RangeArgs& a,
                              cudaStream_t st);
// pair_cnt + slot_of -> pair_off, qcnt[n] (hits per query), lims[n+1] (exclusive prefix)
cudaError_t launch_range_offsets(const int32_t* slot_of, const int32_t* pair_cnt, int n, int nprobe,
                                 uint32_t* pair_off, unsigned long long* qcnt,
                                 unsigned long long* lims, cudaStream_t st);
// store_pairs labels: slab row -> list_no << 32 | offset (faiss/invlists/DirectMap.h:22-24)
cudaError_t launch_rows_to_pairs(int64_t* labels, int64_t n, const uint32_t* list_blk0, int nlists,
                                 cudaStream_t st);

// ---- K3: per-query k-select over the partial candidate lists ------------------------------------
// Replaces heap_reorder (faiss/utils/Heap.h:420-450) / the CPU heap of search_preassigned.
struct SelectArgs {
    const uint64_t* part;       // [slots][nsplit][kpart]
    const int32_t* slot_of;     // [n][nseg] pair slot or -1 (nullptr: slot = q, nseg must be 1)
    int n, nseg, nsplit, kpart; // kpart = entries per partial list
    int k;                      // results per query
    int metric;
    const int64_t* ids;         // slab row -> user id (nullptr: id = row)
    float* out_dis;             // [n][k]
    int64_t* out_ids64;         // [n][k] or nullptr
    int32_t* out_ids32;         // [n][k] or nullptr
    // exact re-rank (needed when any partial list came from the tensor-core scan): the best
    // k + rerank_extra keys are re-scored in plain fp32 (dimension order 0..d-1, the arithmetic of
    // scan.cu) from the slab and the query before the final ordering. slab == nullptr: off.
    const float4* slab = nullptr;
    const float* queries = nullptr;
    int d = 0, nf4 = 0;
    int rerank_extra = 0;
    // Scores that came from the ONE-MMA tensor-core filter (TcScanArgs::one_mma) are only good to
    //   E(q) = err_coef * (|x_q|^2 + max|y|^2)   (L2)     err_coef * |x_q| * max|y|   (inner product)
    // of the exact ones. err_coef > 0 turns the re-rank into a CHECKED one: after the k + rerank_extra
    // best approximate keys have been re-scored exactly, the smallest approximate score B that was
    // NOT re-scored (the next merged key, the gather bound, or the scan's own pruning bound
    // qthr[q]: whatever the scan dropped scored at least that) must satisfy B - E > exact k-th
    // score. Then nothing outside the re-scored set can reach the exact top k (not even as a tie)
    // and the result is the exact one; otherwise the query number is appended to flag_list for
    // the caller to redo on the exact path.
    float err_coef = 0.f;
    const float* ymax2 = nullptr;      // device scalar: max over the slab rows of |y|^2
    const uint32_t* qthr = nullptr;    // [n] ordered bits of the scan's final pruning bound
    int32_t* flag_list = nullptr;      // [n]
    int32_t* flag_count = nullptr;     // device counter (zeroed by the caller)
};
cudaError_t launch_select(const SelectArgs& a, cudaStream_t st);

// ---- multi-shard merge (faiss/utils/Heap.cpp:158-232 semantics, on device) ---------------------
cudaError_t launch_merge_shards(int metric, int nshard, int n, int k, const float* all_dis,
                                const int64_t* all_ids, float* out_dis, int64_t* out_ids,
                                cudaStream_t st);

// ---- layout kernels (build path) ----------------------------------------------------------------
// scatter row-major vectors into the interleaved slab: dst_row[i] < 0 skips vector i
cudaError_t launch_scatter_rows(const float* x, const int64_t* xids, const int64_t* dst_row,
                                int64_t n, int d, int nf4, float* slab, int64_t* ids,
                                int64_t id_base, cudaStream_t st);
// move whole 32-row blocks of every list from an old slab to a new one
cudaError_t launch_move_blocks(const float4* old_slab, const int64_t* old_ids,
                               const uint32_t* old_blk0, const int32_t* old_len,
                               const uint32_t* new_blk0, int nlists, int nf4, float4* new_slab,
                               int64_t* new_ids, cudaStream_t st);
// gather a list back to row-major (get_list / reconstruct)
cudaError_t launch_gather_rows(const float* slab, int64_t row0, int64_t nrows, int d, int nf4,
                               float* out, cudaStream_t st);
// |row|^2 of every slab row (the L2 epilogue of the tensor-core scan): rows [0, nrows)
// max_out (optional, zeroed by the caller): max |row|^2
cudaError_t launch_row_norms(const float4* slab, int64_t nrows, int nf4, float* norms,
                             cudaStream_t st, float* max_out = nullptr);
// flagged queries of a checked re-rank: rows idx[0..cnt) of a table of `words` 4-byte words per row
cudaError_t launch_gather_rows_w(const void* in, const int32_t* idx, int64_t cnt, int words,
                                 void* out, cudaStream_t st);
cudaError_t launch_scatter_rows_w(const void* in, const int32_t* idx, int64_t cnt, int words,
                                  void* out, cudaStream_t st);
cudaError_t launch_scatter_results(const float* Df, const int64_t* If, const int32_t* idx,
                                   int64_t cnt, int k, float* D, int64_t* I, cudaStream_t st);
cudaError_t launch_fill_u32(uint32_t* p, int64_t n, uint32_t v, cudaStream_t st);
// k-means update (Clustering.cpp:150-221 compute_centroids): centroid l = sum of the rows of list l
// in list order (= input order, the reference's summation order) times 1/len; empty lists give 0
cudaError_t launch_list_means(const float* slab, const uint32_t* list_blk0, const int32_t* list_len,
                              int nlists, int d, int nf4, float* centroids, cudaStream_t st);
// out[i] = x[idx[i]] for row-major [.][d] tables
cudaError_t launch_gather_by_index(const float* x, const int64_t* idx, int64_t n, int d, float* out,
                                   cudaStream_t st);
// out[i] = in[idx[i]], or base + idx[i] when in == nullptr (ids of vectors added without ids)
cudaError_t launch_gather_i64(const int64_t* in, const int64_t* idx, int64_t n, int64_t base,
                              int64_t* out, cudaStream_t st);
cudaError_t launch_i64_to_i32(const int64_t* in, int32_t* out, int64_t n, int32_t lo, int32_t hi,
                              unsigned long long* bad, cudaStream_t st);

} // namespace b200ivf
