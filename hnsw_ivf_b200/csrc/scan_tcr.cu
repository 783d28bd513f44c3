// scan_tcr.cu -- the tensor-core scan, rows-major form: K1 (coarse quantizer) and K2 (list scan)
// for d <= 128, any number of queries per list.
//
// Same contract as scan.cu (IVFFlatScanner::scan_codes, faiss/IndexIVFFlat.cpp:155-179, applied to
// every (query, probe) pair; for K1 knn_L2sqr / knn_inner_product over the centroid table,
// faiss/utils/distances.cpp:885-920,836-871): the best KP keys of every (pair, row split) are left
// in `part`, sorted. Scores are 3xTF32 products (fp32 accumulate), good to ~1e-6 of |x||y|: enough
// to SELECT; select.cu re-scores the survivors in plain fp32 before the final ordering.
//
// Work item = (list, tile of <= 64 queries probing it, row split); one CTA each.
//   D[128 rows x N] (TMEM, two buffers) += A[128 rows x 8] (TMEM) * B[N x 8]^T (smem)
// The list rows are the M side of the MMA and the queries the N side, because the time of a
// tcgen05.mma is max(M,128) * N / 256 cycles: N follows the number of queries that really probe the
// list (rounded up to 16), M = 128 costs the same whether it is full or not. What the issuing lane
// executes per MMA matters as much as the pipe time at these small N (round 1
// mistook an issue-bound loop for "an A-in-TMEM MMA costs ~60 cycles whatever N is"), so the MMA
// role issues straight-line blocks and the number of MMAs per 8 dimensions is kept down.
//
// TWO forms, chosen at compile time (template parameter ONE):
//  ONE (the default list scan): a single MMA per 8 dimensions, tf32_rn(x) . trunc_tf32(y) -- A
//  straight from the TMA stage, B = x_hi alone, no split warps, no lo ring. The scores are then a
//  FILTER with a known error bound; select.cu re-scores what could matter exactly and proves the
//  rest away (kernels.h, SelectArgs::err_coef). 4 epilogue warps + TMA + MMA warp, <= 64 queries
//  per item, two CTAs per SM.
//  3xTF32 (coarse quantizer; handles that fell back): x.y = xh.yh + xl.yh + xh.yl in TWO MMAs:
//  * B: the query tile, scaled by -1 (inner product) or -2 (L2), split into tf32 hi / lo parts and
//    STACKED along N: rows 0..n16-1 of the operand are x_hi, rows n16..2*n16-1 are x_lo (K-major
//    no-swizzle canonical layout: 8 rows x 16 bytes = one core matrix, SBO = 128 B, LBO =
//    (2*n16 + 1) * 16 B: odd, so that the coalesced prologue stores are bank-conflict free).
//      MMA 1: D[:, 0 : 2*n16] (+)= y_hi * [x_hi | x_lo]^T   (N = 2*n16)
//      MMA 2: D[:, 0 : n16]    += y_lo * x_hi^T             (N = n16, same buffer, first n16 rows)
//    and the epilogue adds columns j and n16 + j. (Three separate MMAs were 2.3 k cycles per
//    128 x 96-dim tile, exactly the time the tile takes to arrive from HBM at 6.5 TB/s -- the kernel
//    could not be HBM-bound; two are 1.5 k.)
//  * A: list rows arrive by TMA (box {128 floats, 4 blocks, 8 columns} of the slab = 128 rows x 32
//    dims per stage). The hi part needs no work at all: the raw fp32 stage in shared memory IS a
//    K-major no-swizzle tf32 operand (the slab's block layout is the canonical one, and the tensor
//    core ignores the low 13 mantissa bits, tools/tc_probe.cu), so MMA 1 reads A straight from the
//    TMA stage through a shared-memory descriptor. Only lo = y - trunc_tf32(y) is computed: the
//    split warps read a stage once (lane = row) and write lo to tensor memory with tcgen05.st (the
//    tensor core truncates it to tf32 itself: 2^-20 |y|, the size of the dropped lo*lo term). A raw
//    stage goes back to the TMA when both its readers are done: the split warps have it in
//    registers and the MMAs that read it have completed (tcgen05.commit). Half the tcgen05.st
//    traffic and half the conversion work of writing both parts -- the split warps' stage latency
//    was what paced the kernel at ~60 % of the HBM rate.
//  * One elected thread issues the MMAs (2 per 8 dims, see above).
//  * Epilogue (8 warps, thread = row = TMEM lane): per 16 queries two tcgen05.ld (the two column
//    groups) and 16 adds; score (+ |y|^2
//    for L2) is compared with the query's pruning bound held in shared memory; a warp ballot skips
//    the column when no row passes (the steady state). Passing (score, row) keys are appended to
//    the query's candidate buffer (C = max(64, 2 KP) entries: KP kept + free ones) with a shared-memory atomic; at the
//    end of a tile the warps compact the buffers that filled up (warp bitonic sort, keep KP,
//    tighten the bound, publish it in qthr for the other CTAs working on the same query). A tile
//    that overflows a buffer is re-examined for the dropped rows after the compaction (at most 8
//    rounds: only while the bound is still loose, i.e. the first tile of the first lists).
#include "common.cuh"
#include "kernels.h"
#include "tc_ptx.cuh"

#include <cstdio>
#include <cstdlib>

namespace b200ivf {

namespace {

using namespace ptx;

// Two CTA shapes. BIG: 8 epilogue + 8 split warps, all of an SM's shared and tensor memory, one
// CTA per SM (many queries per item: the epilogue and the B operand need the room). SMALL: 4 + 4
// warps, half the shared memory, 256 TMEM columns, <= 32 queries per item -- TWO CTAs per SM, so
// that one CTA's per-item costs (launch, query-tile prologue, pipeline fill and drain: ~7 us per
// item, measured by forcing row splits; 2 ms of a 12 ms scan on the C4 shard) are covered by the
// other CTA's streaming.
template <bool SMALL, bool ONE = false>
struct RShape {
    static constexpr int EpiWarps = SMALL ? 4 : 8;      // warps 0 .. EpiWarps-1
    static constexpr int TmaWarp = EpiWarps;
    static constexpr int MmaWarp = EpiWarps + 1;
    static constexpr int SplitWarp0 = EpiWarps + 2;
    // groups of 4 (one warp per TMEM quarter); none under the one-MMA filter (no lo operand)
    static constexpr int SplitWarps = ONE ? 0 : (SMALL ? 4 : 8);
    static constexpr int SplitGroups = SplitWarps / 4;
    static constexpr int Threads = (SplitWarp0 + SplitWarps) * 32; // 320 / 576; ONE: 192 / 320
    static constexpr int LoaderWarps = EpiWarps + 1 + SplitWarps;  // all but the TMA producer
    static constexpr int CtasPerSM = SMALL ? 2 : 1;
    static constexpr int AStages = SMALL ? 4 : 8;       // lo ring in tensor memory
    static constexpr uint32_t TmemCols = ONE ? 128 : (SMALL ? 256 : 512);
    static constexpr uint32_t DStride = ONE ? 64 : (SMALL ? 64 : 128); // columns per accumulator
    static constexpr uint32_t ColsA = SMALL ? 128 : 256;  // first column of the lo ring
    static constexpr int MaxTQ = (SMALL && !ONE) ? 32 : 64;
};
constexpr size_t kRSmemBig = 227 * 1024, kRSmemSmall = 113 * 1024;
constexpr int kRKC = 8;           // float4 columns per stage (4 MMA k-steps)
constexpr int kRTileBlocks = 4;   // 32-row blocks per tile (M = 128)
constexpr int kRStageF4 = kRKC * kRTileBlocks * kBlockRows; // 1024 float4 = 16 KB
constexpr int kRMaxStages = 12;   // raw ring (as deep as the item's query tile leaves room for)
constexpr int kRAStages = 8;      // barrier slots of the lo ring (RShape::AStages of them in use)
constexpr uint32_t kRColsD = 0;   // two accumulator buffers, RShape::DStride columns apart
constexpr uint32_t kRALBO = kRTileBlocks * 512, kRASBO = 128; // A (hi) straight from a raw stage
constexpr int kRMaxChunks = 256;  // float4 columns of a query (d <= 1024)
constexpr uint32_t kRBSBO = 128;
// Candidate buffer of a query: C entries = the KP kept + free ones. A buffer is compacted
// (sorted, cut to KP, bound tightened) once it holds CompactAt entries, which leaves C -
// CompactAt slots for the rest of the tile before rows start to be dropped and re-examined.
// A compaction stops the whole epilogue at two block barriers, and the epilogue is the stage that
// paces the kernel when few queries meet in a list: with 16 free slots and a compaction every 8
// new candidates there was about one per tile (10 queries x ~1 candidate per tile each); 64 entries
// compacted at 48 make it one tile in four.
// KP up to 256 (the reference's reservoir range, faiss/impl/ResultHandler.h:369-477: k >= 100):
// C = 2 KP entries, compacted when three quarters full.
__host__ __device__ constexpr int tcr_cand(int kp, bool one = false) {
    // (one-MMA kernels, KP = 16: 32 entries, so that 64 queries' buffers fit beside the ring in
    // the half of an SM's shared memory a two-CTAs-per-SM item has)
    return (one && kp == 16) ? 32 : (kp <= 32 ? 64 : 2 * kp);
}
__host__ __device__ constexpr int tcr_compact_at(int kp, bool one = false) {
    return tcr_cand(kp, one) - tcr_cand(kp, one) / 4;
}

__device__ __forceinline__ float tf32_trunc(float x) {
    return __uint_as_float(__float_as_uint(x) & 0xffffe000u);
}
__device__ __forceinline__ float tf32_rn(float x) {
    return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
}
__device__ __forceinline__ float tf32_lo(float y) {
    return tf32_rn(y - tf32_trunc(y));
}
// ... for the list rows: exact residual, truncated to tf32 by the tensor core itself
__device__ __forceinline__ float tf32_lo_raw(float y) {
    return y - tf32_trunc(y);
}

template <int KP, bool ONE = false>
struct RCfg {
    static constexpr int C = tcr_cand(KP, ONE);     // candidate entries per query
    static constexpr int CompactAt = tcr_compact_at(KP, ONE);
    static_assert(KP + 8 <= CompactAt, "compaction must leave room above the kept entries");
};

// ascending bitonic sort of one key per lane
__device__ __forceinline__ uint64_t warp_sort32(uint64_t key, int lane) {
#pragma unroll
    for (int size = 2; size <= 32; size <<= 1) {
#pragma unroll
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            const uint64_t other = __shfl_xor_sync(0xffffffffu, key, stride);
            const bool up = (lane & size) == 0 || size == 32;
            const bool lower = (lane & stride) == 0;
            const bool take_min = up == lower;
            const uint64_t mn = key < other ? key : other, mx = key < other ? other : key;
            key = take_min ? mn : mx;
        }
    }
    return key;
}
// ascending bitonic sort of two keys per lane: positions lane (a) and lane + 32 (b)
__device__ __forceinline__ void warp_sort64(uint64_t& a, uint64_t& b, int lane) {
    // sort a ascending and b descending, then one bitonic merge of 64
    a = warp_sort32(a, lane);
    b = warp_sort32(b, lane);
    b = __shfl_sync(0xffffffffu, b, 31 - lane);
    {
        const uint64_t mn = a < b ? a : b, mx = a < b ? b : a;
        a = mn;
        b = mx;
    }
#pragma unroll
    for (int stride = 16; stride > 0; stride >>= 1) {
        const uint64_t oa = __shfl_xor_sync(0xffffffffu, a, stride);
        const uint64_t ob = __shfl_xor_sync(0xffffffffu, b, stride);
        const bool lower = (lane & stride) == 0;
        a = lower ? (a < oa ? a : oa) : (a < oa ? oa : a);
        b = lower ? (b < ob ? b : ob) : (b < ob ? ob : b);
    }
}

// v[j] for a per-lane j (registers cannot be indexed dynamically: a 4-level select tree)
__device__ __forceinline__ uint32_t sel16(const uint32_t (&v)[16], int j) {
    uint32_t a8[8], a4[4], a2[2];
#pragma unroll
    for (int i = 0; i < 8; ++i)
        a8[i] = (j & 1) ? v[2 * i + 1] : v[2 * i];
#pragma unroll
    for (int i = 0; i < 4; ++i)
        a4[i] = (j & 2) ? a8[2 * i + 1] : a8[2 * i];
#pragma unroll
    for (int i = 0; i < 2; ++i)
        a2[i] = (j & 4) ? a4[2 * i + 1] : a4[2 * i];
    return (j & 8) ? a2[1] : a2[0];
}
// the bound the epilogue's filter uses: 2^-21 looser than the real one, so that the filter (which
// rounds differently from the exact test that follows it) never rejects what the test would accept
__device__ __forceinline__ float loosen(float x) {
    return fmaf(fabsf(x), 4.76837158e-7f, x);
}

struct RSmem {
    float4* raw;        // [nst][kRStageF4]
    unsigned char* b;   // query tile: per 16-byte K chunk, n16 rows of x_hi then n16 rows of x_lo
    uint64_t* cand;     // [kTcrTQ][C]
    uint32_t* thr;      // [kTcrTQ] ordered bits of the pruning bound (full score)
    float* thrp;        // [kTcrTQ] bound minus |x|^2: what the raw accumulator is compared with
    float* thrf;        // [kTcrTQ] loosen(thrp): what the epilogue's filter uses
    float* xn;          // [kTcrTQ] |x|^2 (L2) or 0
    int* cnt;           // [kTcrTQ] entries in the candidate buffer (may run past C: overflow)
    int* qid;           // [kTcrTQ] query number
    int* lim;           // [kTcrTQ] rows of the list that count for the query (max_codes)
    int* flag;          // [4] by barrier epoch: 1 = some buffer wants compaction, 2 = overflowed
    uint64_t* full;     // [kRMaxStages] TMA landed
    uint64_t* empty;    // [kRMaxStages] the split warps have the stage in registers AND the MMAs
                        //               that read it are done
    uint64_t* a_rdy;    // [kRAStages] operand stage written to tensor memory
    uint64_t* a_empty;  // [kRAStages] MMAs that read the operand stage are done
    uint64_t* d_full;   // [2] accumulator buffer complete
    uint64_t* d_empty;  // [2] accumulator buffer drained by the epilogue
    uint32_t* tmem;     // [1]
};

// Shared memory is carved per ITEM: the B operand and the candidate buffers are sized by the item's
// query count (n16) and dimension (nch chunks), and the raw ring gets every 16 KB stage that is
// left -- with few queries the kernel is an HBM stream and lives on bytes in flight.
//   B: nch chunks x LBO bytes, LBO = (2 * n16 + 1) * 16 (odd multiple of 16: the coalesced
//   prologue stores of one query row hit 8 different 16-byte bank groups)
__host__ __device__ inline uint32_t tcr_lbo(int n16, bool one = false) {
    // (one-MMA kernels keep x_hi alone: n16 rows per chunk)
    return (uint32_t)((one ? n16 : 2 * n16) + 1) * 16u;
}
__host__ __device__ inline size_t tcr_rest_bytes(int kp, int n16, int nch, bool one = false) {
    return (size_t)nch * tcr_lbo(n16, one) + (size_t)n16 * tcr_cand(kp, one) * 8 +
            (size_t)kTcrTQ * 7 * 4 + 4 * 4 + (2 * kRMaxStages + 2 * kRAStages + 4) * 8 + 16;
}
__host__ __device__ inline int tcr_stages(size_t smem_total, int kp, int n16, int nch,
                                          bool even = true, bool one = false) {
    const size_t rest = tcr_rest_bytes(kp, n16, nch, one);
    const int n = smem_total > rest ? (int)((smem_total - rest) / (kRStageF4 * sizeof(float4))) : 0;
    // EVEN, always (3xTF32 kernels with two split-warp groups): the groups take alternate
    // iterations, so with an even ring a stage is always consumed by the same group, which
    // therefore saw the stage's previous phase complete before it waits for the next one. With an
    // odd ring the groups swap stages every lap; a group that runs ahead then waits on parity p of
    // a barrier still in the phase BEFORE the one it thinks it follows, the wait passes at once
    // (try_wait.parity is true for the phase preceding the current one), and the ring protocol
    // falls apart (seen as rare launch failures on cold first searches at d = 960, where the ring
    // had 5 stages). One split group, or none, has no such constraint: even = false.
    const int c = n > kRMaxStages ? kRMaxStages : n;
    return even ? (c & ~1) : c;
}

template <int KP, bool ONE>
__device__ __forceinline__ RSmem carve_r(unsigned char* raw, int nst, int n16, int nch) {
    RSmem s;
    const size_t bbytes = (size_t)nch * tcr_lbo(n16, ONE);
    s.raw = reinterpret_cast<float4*>(raw);
    s.b = raw + (size_t)nst * kRStageF4 * sizeof(float4);
    s.cand = reinterpret_cast<uint64_t*>(s.b + bbytes);
    s.thr = reinterpret_cast<uint32_t*>(s.cand + (size_t)n16 * RCfg<KP, ONE>::C);
    s.thrp = reinterpret_cast<float*>(s.thr + kTcrTQ);
    s.thrf = s.thrp + kTcrTQ;
    s.xn = s.thrf + kTcrTQ;
    s.cnt = reinterpret_cast<int*>(s.xn + kTcrTQ);
    s.qid = s.cnt + kTcrTQ;
    s.lim = s.qid + kTcrTQ;
    s.flag = s.lim + kTcrTQ;
    s.full = reinterpret_cast<uint64_t*>(s.flag + 4);
    s.empty = s.full + kRMaxStages;
    s.a_rdy = s.empty + kRMaxStages;
    s.a_empty = s.a_rdy + kRAStages;
    s.d_full = s.a_empty + kRAStages;
    s.d_empty = s.d_full + 2;
    s.tmem = reinterpret_cast<uint32_t*>(s.d_empty + 2);
    return s;
}

// Compaction of the candidate buffers a warp owns (queries = warp mod 8), out of line: sort, keep
// the best KP, tighten the query's bound and publish it for the other CTAs.
template <int KP, int kREpiWarps, bool ONE>
__device__ __noinline__ void tcr_compact(uint64_t* cand, int* cnt, uint32_t* thr, float* thrp,
                                         float* thrf, const float* xn, const int* qid,
                                         uint32_t* qthr, int nqw, int warp, int lane,
                                         bool final_pass) {
    constexpr int C = RCfg<KP, ONE>::C;
    const int qmine = lane * kREpiWarps + warp;
    const bool need = qmine < nqw && (final_pass || cnt[qmine] >= RCfg<KP, ONE>::CompactAt);
    unsigned todo = __ballot_sync(0xffffffffu, need);
    while (todo != 0u) {
        const int qi = (__ffs(todo) - 1) * kREpiWarps + warp;
        todo &= todo - 1u;
        const int n = min(cnt[qi], C);
        uint64_t* buf = cand + (size_t)qi * C;
        uint64_t kth;
        if (C == 32) {
            uint64_t k0 = lane < n ? buf[lane] : kSentinelKey;
            k0 = warp_sort32(k0, lane);
            if (lane < KP)
                buf[lane] = k0;
            kth = __shfl_sync(0xffffffffu, k0, (KP - 1) & 31);
        } else if (C == 64) {
            uint64_t k0 = lane < n ? buf[lane] : kSentinelKey;
            uint64_t k1 = lane + 32 < n ? buf[lane + 32] : kSentinelKey;
            warp_sort64(k0, k1, lane);
            if (lane < KP)
                buf[lane] = k0;
            if (lane + 32 < KP)
                buf[lane + 32] = k1;
            kth = KP <= 32 ? __shfl_sync(0xffffffffu, k0, (KP - 1) & 31)
                           : __shfl_sync(0xffffffffu, k1, (KP - 1) & 31);
        } else {
            // larger buffers: bitonic network in shared memory (rare: one compaction per
            // C / 4 accepted candidates of a query)
            for (int i = n + lane; i < C; i += 32)
                buf[i] = kSentinelKey;
            __syncwarp();
            warp_sort_keys(buf, C, lane);
            kth = buf[KP - 1];
        }
        if (lane == 0) {
            cnt[qi] = min(n, KP);
            if (n >= KP) {
                const uint32_t tb = (uint32_t)(kth >> 32);
                const uint32_t old = atomicMin(&thr[qi], tb);
                if (tb < old) {
                    const float tp = from_ordered_bits(tb) - xn[qi];
                    thrp[qi] = tp;
                    thrf[qi] = loosen(tp);
                    if (qthr)
                        atomicMin(qthr + qid[qi], tb);
                }
            }
        }
        __syncwarp();
    }
}



// One chunk of 16 queries of the epilogue (thread = row): filter the 16 accumulators against the
// queries' loosened bounds (one vote per chunk: in the steady state no row passes and the warp
// moves on), then the lanes that hold candidates walk their own columns, test them exactly and
// append them to the queries' buffers. Returns the columns this lane pushed past a full buffer
// (re-examined after the compaction); fl collects what the block must do after the tile
// (1 = compact, 2 = a buffer overflowed).
template <int KP, bool ONE>
__device__ __forceinline__ uint32_t tcr_filter_chunk(const uint32_t (&v)[16], int ch, bool first,
                                                     uint32_t dbits, float nrm, float nrmf,
                                                     uint32_t rowid, int row_in_list, bool limited,
                                                     const RSmem& s, int& fl, int& n_events) {
    constexpr int C = RCfg<KP, ONE>::C;
    float dd[16];
#pragma unroll
    for (int j4 = 0; j4 < 4; ++j4) {
        const float4 th = *reinterpret_cast<const float4*>(s.thrf + ch * 16 + j4 * 4);
        dd[4 * j4 + 0] = __uint_as_float(v[4 * j4 + 0]) - th.x;
        dd[4 * j4 + 1] = __uint_as_float(v[4 * j4 + 1]) - th.y;
        dd[4 * j4 + 2] = __uint_as_float(v[4 * j4 + 2]) - th.z;
        dd[4 * j4 + 3] = __uint_as_float(v[4 * j4 + 3]) - th.w;
    }
    float mn = dd[0];
#pragma unroll
    for (int j = 1; j < 16; ++j)
        mn = fminf(mn, dd[j]);
    const bool hit = first ? (mn + nrmf <= 0.f) : dbits != 0u;
    if (!__any_sync(0xffffffffu, hit))
        return first ? 0u : dbits;
    uint32_t pm = 0u;
    if (first) {
#pragma unroll
        for (int j = 0; j < 16; ++j)
            pm |= (dd[j] + nrmf <= 0.f) ? (1u << j) : 0u;
    } else {
        pm = dbits;
    }
    uint32_t ndrop = 0u;
    while (pm != 0u) {
        const int j = __ffs(pm) - 1;
        pm &= pm - 1u;
        ++n_events;
        const int qi = ch * 16 + j;
        const float sc = __uint_as_float(sel16(v, j)) + nrm;
        if (sc <= s.thrp[qi] && (!limited || row_in_list < s.lim[qi])) {
            const int slot = atomicAdd(&s.cnt[qi], 1);
            if (slot < C) {
                s.cand[(size_t)qi * C + slot] = make_key(sc + s.xn[qi], rowid);
            } else {
                ndrop |= 1u << j;
                fl |= 2;
            }
            if (slot + 1 >= RCfg<KP, ONE>::CompactAt)
                fl |= 1;
        }
    }
    return ndrop;
}

// wait with the waiting time added to a trace counter (timing experiments only)
#define TCR_WAIT(bar, par, acc, code, aux)                               \
    do {                                                                 \
        if (tracing) {                                                   \
            const long long w0_ = clock64();                             \
            mbar_wait_code(bar, par, a.trap, code, (unsigned)(aux));     \
            acc += clock64() - w0_;                                      \
        } else {                                                         \
            mbar_wait_code(bar, par, a.trap, code, (unsigned)(aux));     \
        }                                                                \
    } while (0)

#define TCR_TRACE(role, idx)                                             \
    do {                                                                 \
        if (tracing && (idx) < 64)                                       \
            a.trace[(role) * 64 + (idx)] = clock64();                    \
    } while (0)

template <int METRIC, int KP, bool SMALL, bool ONE>
__global__ void __launch_bounds__(RShape<SMALL, ONE>::Threads, RShape<SMALL, ONE>::CtasPerSM)
        ivf_scan_tcr_kernel(const TcScanArgs a, const __grid_constant__ CUtensorMap tmap) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    using SH = RShape<SMALL, ONE>;
    constexpr int kREpiWarps = SH::EpiWarps, kRTmaWarp = SH::TmaWarp, kRMmaWarp = SH::MmaWarp;
    constexpr int kRSplitWarp0 = SH::SplitWarp0;
    constexpr int kRLoaderWarps = SH::LoaderWarps, kRSplitGroups = SH::SplitGroups;
    constexpr int kAStages = SH::AStages;
    constexpr uint32_t kRColsA = SH::ColsA, kDStride = SH::DStride;
    constexpr int C = RCfg<KP, ONE>::C;
    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;

    // ---- which item --------------------------------------------------------------------------
    const uint32_t item = blockIdx.x;
    if (item >= a.item_off[a.nlists])
        return;
    int lo_l = 0, hi_l = a.nlists;
    if (a.item_list) {
        lo_l = a.item_list[item]; // written by the plan: no dependent-load chain
    } else {
        while (hi_l - lo_l > 1) {
            const int mid = (lo_l + hi_l) >> 1;
            if (a.item_off[mid] <= item)
                lo_l = mid;
            else
                hi_l = mid;
        }
    }
    const int l = lo_l;
    const int g0 = a.grp_start[l];
    const int cnt_l = a.grp_start[l + 1] - g0;
    const int nsub = (cnt_l + a.tq - 1) / a.tq;
    const uint32_t local = item - a.item_off[l];
    const int split = (int)(local / (uint32_t)nsub);
    const int sub = (int)(local % (uint32_t)nsub);
    // the queries of the list are spread evenly over its nsub tiles
    const int qa = (int)(((long long)cnt_l * sub) / nsub);
    const int qb = (int)(((long long)cnt_l * (sub + 1)) / nsub);
    const int q0 = g0 + qa;
    const int nqw = qb - qa;                  // 1 .. kTcrTQ
    const int n16 = (nqw + 15) & ~15;         // MMA N

    const int len = a.list_len[l];
    const int nb = (len + kBlockRows - 1) / kBlockRows;
    const int b0 = (int)(((long long)nb * split) / a.nsplit);
    const int b1 = (int)(((long long)nb * (split + 1)) / a.nsplit);
    const uint32_t blk0 = a.list_blk0[l];
    const int nf4 = a.nf4;
    const int ntiles = (b1 - b0 + kRTileBlocks - 1) / kRTileBlocks;
    const int nsg = (nf4 + kRKC - 1) / kRKC;  // stages per tile
    const int nks = (nf4 + 1) >> 1;           // MMA k-steps (8 dims each) per tile
    const int niter = ntiles * nsg;
    const int nst = tcr_stages((size_t)a.nstages * 1024, KP, n16, 2 * nks, !SMALL && !ONE, ONE);
    const uint32_t lbo = tcr_lbo(n16, ONE);

    const RSmem s = carve_r<KP, ONE>(smem_raw, nst, n16, 2 * nks);
    const bool tracing = a.trace != nullptr && item == a.trace_item;
    // one MMA per k-step (y_hi . x_hi only): the tensor core is a FILTER with a known error bound,
    // see TcScanArgs::one_mma
    constexpr bool one = ONE;
    if (tid == 0) {
        TCR_TRACE(0, 0);
        if (tracing) {
            a.trace[0 * 64 + 2] = ntiles;
            a.trace[0 * 64 + 3] = nqw;
        }
    }

    if (tid == 0) {
        for (int i = 0; i < kRMaxStages; ++i) {
            mbar_init(&s.full[i], 1);
            mbar_init(&s.empty[i], one ? 1 : 4 + 1); // one split group (4 warps) + the MMA commit
        }
        for (int i = 0; i < kRAStages; ++i) {
            mbar_init(&s.a_rdy[i], 4);
            mbar_init(&s.a_empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&s.d_full[i], 1);
            mbar_init(&s.d_empty[i], kREpiWarps);
        }
        for (int i = 0; i < 4; ++i)
            s.flag[i] = 0;
        fence_barrier_init();
    }
    if (warp == kRMmaWarp) {
        tmem_alloc(s.tmem, SH::TmemCols);
        tmem_relinquish();
    }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem = *s.tmem;
    if (tid == 0)
        TCR_TRACE(0, 1);

    if (warp == kRTmaWarp) {
        // ======================= TMA producer ===================================================
        if (lane == 0 && ntiles > 0) {
            tma_prefetch_desc(&tmap);
            int st = 0;
            uint32_t ph = 0;
            long long w_empty = 0;
            for (int t = 0; t < ntiles; ++t) {
                const int blk = (int)blk0 + b0 + t * kRTileBlocks;
                for (int sg = 0; sg < nsg; ++sg) {
                    if (t * nsg + sg >= nst)
                        TCR_WAIT(&s.empty[st], ph ^ 1u, w_empty, 1, t * nsg + sg);
                    if (a.dbg & 16) {
                        mbar_arrive(&s.full[st]);
                    } else {
                        mbar_expect_tx(&s.full[st], (uint32_t)(kRStageF4 * sizeof(float4)));
                        tma_load_3d(s.raw + (size_t)st * kRStageF4, &tmap, 0, blk, sg * kRKC,
                                    &s.full[st]);
                    }
                    if (++st == nst) {
                        st = 0;
                        ph ^= 1u;
                    }
                }
                TCR_TRACE(1, t);
            }
            if (tracing)
                a.trace[9 * 64 + 4] = w_empty;
        }
        return; // takes no part in the named barriers below
    }

    // ---- prologue: the query tile -> shared memory (B operand), 17 warps -----------------------
    {
        const int wl = warp < kRTmaWarp ? warp : warp - 1;
        const float scale = METRIC == kMetricL2 ? -2.f : -1.f;
        const bool vec = (a.d & 3) == 0 && (reinterpret_cast<uintptr_t>(a.queries) & 15) == 0;
        const int nch = 2 * nks; // chunks the MMAs read (the last one may be padding)
        constexpr int kPer = (SH::MaxTQ + kRLoaderWarps - 1) / kRLoaderWarps;
        int qn[kPer];
        float nn[kPer];
#pragma unroll
        for (int u = 0; u < kPer; ++u) {
            const int n = wl + u * kRLoaderWarps;
            qn[u] = n < nqw ? (a.pair_query ? a.pair_query[q0 + n] : q0 + n) : -1;
            nn[u] = 0.f;
        }
        // lane = chunk (float4 column) of the query row: one coalesced 512-byte read per 32 chunks
        for (int cb = 0; cb < nch && !(a.dbg & 32); cb += 32) {
            const int c = cb + lane;
            float4 xv[kPer];
#pragma unroll
            for (int u = 0; u < kPer; ++u) {
                xv[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (qn[u] >= 0 && c < nf4) {
                    const float* src = a.queries + (size_t)qn[u] * a.d;
                    const int dd = 4 * c;
                    if (vec) {
                        xv[u] = __ldg(reinterpret_cast<const float4*>(src + dd));
                    } else {
                        xv[u].x = __ldg(src + dd);
                        xv[u].y = dd + 1 < a.d ? __ldg(src + dd + 1) : 0.f;
                        xv[u].z = dd + 2 < a.d ? __ldg(src + dd + 2) : 0.f;
                        xv[u].w = dd + 3 < a.d ? __ldg(src + dd + 3) : 0.f;
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < kPer; ++u) {
                const int n = wl + u * kRLoaderWarps;
                if (n < n16 && c < nch) {
                    const float4 x = xv[u];
                    nn[u] += x.x * x.x + x.y * x.y + x.z * x.z + x.w * x.w;
                    float4 h, lo;
                    h.x = x.x * scale;
                    h.y = x.y * scale;
                    h.z = x.z * scale;
                    h.w = x.w * scale;
                    if (one) { // no lo part follows: round to nearest instead of the MMA's truncation
                        h.x = tf32_rn(h.x);
                        h.y = tf32_rn(h.y);
                        h.z = tf32_rn(h.z);
                        h.w = tf32_rn(h.w);
                    }
                    lo.x = tf32_lo(h.x);
                    lo.y = tf32_lo(h.y);
                    lo.z = tf32_lo(h.z);
                    lo.w = tf32_lo(h.w);
                    const size_t off = (size_t)c * lbo + (size_t)n * 16;
                    *reinterpret_cast<float4*>(s.b + off) = h;                     // row n
                    if (!ONE)
                        *reinterpret_cast<float4*>(s.b + off + (size_t)n16 * 16) = lo; // row n16 + n
                }
            }
        }
#pragma unroll
        for (int u = 0; u < kPer; ++u) {
            const int n = wl + u * kRLoaderWarps;
            if (n < n16) { // (warp-uniform)
                float v = nn[u];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1)
                    v += __shfl_xor_sync(0xffffffffu, v, o);
                if (lane == 0) {
                    const int q = qn[u];
                    const float xnq = (METRIC == kMetricL2 && q >= 0) ? v : 0.f;
                    uint32_t tb = ordered_bits(kNeutralScore);
                    if (q >= 0 && a.qthr)
                        tb = min(tb, a.qthr[q]);
                    s.qid[n] = q;
                    s.lim[n] = (a.pair_limit != nullptr && q >= 0) ? a.pair_limit[q0 + n]
                                                                    : 0x7fffffff;
                    s.xn[n] = xnq;
                    s.cnt[n] = 0;
                    s.thr[n] = tb;
                    // a padding column (q < 0) must never accept a row
                    const float tp = q >= 0 ? from_ordered_bits(tb) - xnq : -kNeutralScore;
                    s.thrp[n] = tp;
                    s.thrf[n] = loosen(tp);
                }
            }
        }
        fence_proxy_async_smem(); // B is read by the tensor core through the async proxy
        named_bar_sync(1, kRLoaderWarps * 32);
    }
    if (tid == 0)
        TCR_TRACE(0, 4);

    if (warp == kRMmaWarp) {
        // ======================= MMA issuer =====================================================
        // The whole warp runs the loop (converged control flow); one elected lane issues. What the
        // issuing lane executes between two tcgen05.mma is what paces this role (an MMA with N <= 64
        // is 16-32 cycles of tensor pipe, tools/tc_probe.cu): the block under the elect is straight
        // line, with every operand prepared before it and nothing to decide inside.
        // MMA 1: N = 2 * n16 (x_hi | x_lo stacked), MMA 2: N = n16 (x_hi: the first n16 rows of the
        // same chunks); one_mma: MMA 1 alone with N = n16
        const uint32_t idesc2 = make_idesc_tf32(kRTileBlocks * kBlockRows, 2 * n16);
        const uint32_t idesc1 = make_idesc_tf32(kRTileBlocks * kBlockRows, n16);
        const uint64_t desc0 = make_smem_desc_kmajor_noswz(0, lbo, kRBSBO);
        const uint64_t adesc0 = make_smem_desc_kmajor_noswz(0, kRALBO, kRASBO);
        const uint32_t b0a = smem_u32(s.b), raw0 = smem_u32(s.raw);
        const uint64_t kb_step = (uint64_t)((2u * lbo) >> 4);  // 2 chunks of B per k-step
        constexpr uint64_t ka_step = (2 * kRALBO) >> 4;        // 2 chunks of the raw stage
        tc_fence_after_sync();
        int as = 0, rst = 0;
        uint32_t aph = 0, rph = 0;
        long long w_ardy = 0, w_dempty = 0, w_issue = 0;
        for (int t = 0; t < ntiles; ++t) {
            const int b = t & 1;
            if (t >= 2) {
                TCR_WAIT(&s.d_empty[b], (uint32_t)(((t >> 1) & 1) ^ 1), w_dempty, 2, t);
                tc_fence_after_sync();
            }
            const uint32_t tD = tmem + kRColsD + (uint32_t)b * kDStride;
            for (int sg = 0; sg < nsg; ++sg) {
                if (one)
                    TCR_WAIT(&s.full[rst], rph, w_ardy, 3, t * nsg + sg);
                else
                    TCR_WAIT(&s.a_rdy[as], aph, w_ardy, 3, t * nsg + sg);
                tc_fence_after_sync();
                const long long i0_ = tracing ? clock64() : 0;
                const int ks0 = sg * (kRKC / 2);
                const int ksn = nks - ks0; // >= 1 k-steps of this stage
                const uint32_t alo = tmem + kRColsA + (uint32_t)as * 32u;
                // hi: the raw stage itself (2 chunks = 2 * kRALBO bytes per k-step)
                const uint64_t dah =
                        adesc0 + (uint64_t)((raw0 + (uint32_t)rst * kRStageF4 * 16u) >> 4);
                // the B descriptor advances by 2 chunks (2 * LBO bytes, in 16-byte units) per k-step
                const uint64_t db = desc0 + (uint64_t)((b0a + (uint32_t)(ks0 * 2) * lbo) >> 4);
                const uint32_t acc0 = sg != 0 ? 1u : 0u; // the first MMA of a tile overwrites
                const bool last = sg == nsg - 1;
                if (elect_one_sync()) {
                    if (one) {
                        if (a.dbg & 8) { // timing experiment: no MMAs
                        } else if (ksn >= kRKC / 2) {
#pragma unroll
                            for (int ks = 0; ks < kRKC / 2; ++ks)
                                mma_tf32_ss(tD, dah + ks * ka_step, db + ks * kb_step, idesc1,
                                            ks ? 1u : acc0);
                        } else {
                            for (int ks = 0; ks < ksn; ++ks)
                                mma_tf32_ss(tD, dah + ks * ka_step, db + ks * kb_step, idesc1,
                                            ks ? 1u : acc0);
                        }
                        mma_commit(&s.empty[rst]); // the raw stage may be refilled
                    } else {
                        if (ksn >= kRKC / 2) {
#pragma unroll
                            for (int ks = 0; ks < kRKC / 2; ++ks) {
                                mma_tf32_ss(tD, dah + ks * ka_step, db + ks * kb_step, idesc2,
                                            ks ? 1u : acc0);
                                mma_tf32_ts(tD, alo + ks * 8, db + ks * kb_step, idesc1, 1u);
                            }
                        } else {
                            for (int ks = 0; ks < ksn; ++ks) {
                                mma_tf32_ss(tD, dah + ks * ka_step, db + ks * kb_step, idesc2,
                                            ks ? 1u : acc0);
                                mma_tf32_ts(tD, alo + ks * 8, db + ks * kb_step, idesc1, 1u);
                            }
                        }
                        mma_commit(&s.empty[rst]);
                        mma_commit(&s.a_empty[as]);
                    }
                    if (last)
                        mma_commit(&s.d_full[b]);
                }
                __syncwarp();
                if (++rst == nst) {
                    rst = 0;
                    rph ^= 1u;
                }
                if (++as == kAStages) {
                    as = 0;
                    aph ^= 1u;
                }
                if (tracing)
                    w_issue += clock64() - i0_;
            }
            if (lane == 0)
                TCR_TRACE(2, t);
        }
        if (tracing && lane == 0) {
            a.trace[9 * 64 + 2] = w_ardy;
            a.trace[9 * 64 + 3] = w_dempty;
            a.trace[9 * 64 + 6] = w_issue;
        }
    } else if (!ONE && warp >= kRSplitWarp0) {
        // ======================= split warps: raw stage -> (hi, lo) in tensor memory ============
        // Two groups of 4 warps (one per TMEM lane quarter) take alternate stages, so that the
        // load -> convert -> tcgen05.st -> wait chain of one stage overlaps the next stage's.
        const int sw = warp - kRSplitWarp0;
        const int quarter = warp & 3;             // the TMEM lanes this warp may touch
        const int grp = sw >> 2;
        const int row = quarter * 32 + lane;
        const uint32_t lane_addr = (uint32_t)(quarter * 32) << 16;
        long long w_full = 0, w_aempty = 0;
        for (int it = grp; it < niter; it += (kRSplitGroups > 0 ? kRSplitGroups : 1)) {
            const int st = it % nst;
            const uint32_t ph = (uint32_t)((it / nst) & 1);
            const int as = it % kAStages;
            const uint32_t aph = (uint32_t)((it / kAStages) & 1);
            TCR_WAIT(&s.full[st], ph, w_full, 4, it);
            const float4* src = s.raw + (size_t)st * kRStageF4 + row;
            if (it >= kAStages) {
                TCR_WAIT(&s.a_empty[as], aph ^ 1u, w_aempty, 5, it);
                tc_fence_after_sync();
            }
            const uint32_t ta = tmem + lane_addr + kRColsA + (uint32_t)as * 32u;
#pragma unroll
            for (int h = 0; h < 2 && !(a.dbg & 4); ++h) {
                uint32_t lo[16];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const float4 y = src[(h * 4 + u) * 128];
                    lo[4 * u + 0] = __float_as_uint(tf32_lo_raw(y.x));
                    lo[4 * u + 1] = __float_as_uint(tf32_lo_raw(y.y));
                    lo[4 * u + 2] = __float_as_uint(tf32_lo_raw(y.z));
                    lo[4 * u + 3] = __float_as_uint(tf32_lo_raw(y.w));
                }
                tmem_st_32x32b_x16(ta + (uint32_t)h * 16u, lo);
            }
            __syncwarp();
            if (lane == 0)
                mbar_arrive(&s.empty[st]); // the stage is in registers: the TMA may refill it
            tmem_st_wait();
            tc_fence_before_sync();
            __syncwarp();
            if (lane == 0)
                mbar_arrive(&s.a_rdy[as]);
            if (lane == 0 && quarter == 0 && (it + 1) % nsg == 0)
                TCR_TRACE(3, it / nsg);
        }
        if (tracing && lane == 0 && quarter == 0 && grp == 0) {
            a.trace[9 * 64 + 0] = w_full;
            a.trace[9 * 64 + 1] = w_aempty;
        }
    } else {
        // ======================= epilogue: 8 warps, thread = row of the tile ====================
        // Warps e and e + 4 share a TMEM lane quarter and take alternate 16-query chunks.
        const int quarter = warp & 3, half = warp >> 2;
        const uint32_t lane_addr = (uint32_t)(quarter * 32) << 16;
        const int nch16 = n16 >> 4;
        const int rows_split = (b1 - b0) * kBlockRows;       // rows of this split (padded blocks)
        const int row_lim = min(rows_split, len - b0 * kBlockRows);
        // per-tile refresh of the bounds from the other CTAs: thread j looks after query j
        const int myq = tid < nqw ? s.qid[tid] : -1;
        long long w_dfull = 0;
        int ep = 0; // barrier epochs: flag[ep & 3] collects what the pushes of the epoch want

        for (int t = 0; t < ntiles; ++t) {
            const int b = t & 1;
            const int r_in_split = t * (kRTileBlocks * kBlockRows) + quarter * 32 + lane;
            bool valid = r_in_split < row_lim;
            const uint32_t rowid = (blk0 + (uint32_t)b0) * kBlockRows + (uint32_t)r_in_split;
            // SearchParametersIVF::sel (IVFFlatScanner's use_sel arm, faiss/IndexIVFFlat.cpp:166):
            // the selector was evaluated into one mask word per 32-row block, bit = lane
            if (a.row_mask != nullptr && valid)
                valid = (__ldg(a.row_mask + (rowid >> 5)) >> lane) & 1u;
            // max_codes: only the first pair_limit[pair] rows of the list count for a query
            // (scan_one_list's list_size_max, faiss/IndexIVF.cpp:545-548)
            const int row_in_list = b0 * kBlockRows + r_in_split;
            // a row that does not exist scores +inf
            const float nrm = !valid ? __int_as_float(0x7f800000)
                                     : (METRIC == kMetricL2 ? __ldg(a.norms + rowid) : 0.f);
            const float nrmf = nrm * (1.f - 4.76837158e-7f); // the filter errs on the accept side
            const uint32_t gthr = (myq >= 0 && a.qthr) ? __ldcg(a.qthr + myq) : 0xffffffffu;
            // The warp's chunks of 16 queries: ch = half + u * (epilogue warps per quarter), u < 2
            constexpr int kChStride = kREpiWarps / 4;
            constexpr int kChPerWarp = (SH::MaxTQ / 16 + kChStride - 1) / kChStride;
            static_assert(ONE || kChPerWarp == 2, "two chunks per epilogue warp (16 drop bits each)");
            uint64_t dropw = 0u; // 16 bits per chunk: columns pushed past a full buffer
            const bool limited = a.pair_limit != nullptr;
            TCR_WAIT(&s.d_full[b], (uint32_t)((t >> 1) & 1), w_dfull, 6, t);
            tc_fence_after_sync();
            const uint32_t tD = tmem + lane_addr + kRColsD + (uint32_t)b * kDStride;
            // one-MMA filter: ALL of the tile's accumulator columns this warp owns are fetched at
            // once (one tcgen05.ld round trip per tile, not one per chunk: that round trip is
            // what an epilogue warp spends its time on) and the buffer goes back to the MMA warp
            // before anything is filtered. (The 3xTF32 form has twice the columns to fetch and
            // eight more warps to share the register file with: it goes chunk by chunk.)
            uint32_t vv[ONE ? kChPerWarp : 1][16];
            if (ONE) {
                if (kChStride == 1) {
                    // (columns 0 .. n16-1 are this warp's: ONE load instruction for all of them)
                    constexpr int u1 = ONE && kChPerWarp > 1 ? 1 : 0, u2 = ONE && kChPerWarp > 2 ? 2 : 0,
                                  u3 = ONE && kChPerWarp > 3 ? 3 : 0;
                    if (nch16 == 4)
                        tmem_ld_32x32b_x64_split(tD, vv[0], vv[u1], vv[u2], vv[u3]);
                    else if (nch16 == 3) {
                        tmem_ld_32x32b_x32_split(tD, vv[0], vv[u1]);
                        tmem_ld_32x32b_x16(tD + 32u, vv[u2]);
                    } else if (nch16 == 2)
                        tmem_ld_32x32b_x32_split(tD, vv[0], vv[u1]);
                    else
                        tmem_ld_32x32b_x16(tD, vv[0]);
                } else {
#pragma unroll
                    for (int u = 0; u < kChPerWarp; ++u) {
                        const int ch = half + u * kChStride;
                        if (ch < nch16)
                            tmem_ld_32x32b_x16(tD + (uint32_t)ch * 16u, vv[ONE ? u : 0]);
                    }
                }
                tmem_ld_wait();
                tc_fence_before_sync();
                __syncwarp();
                if (lane == 0)
                    mbar_arrive(&s.d_empty[b]); // accumulator buffer drained
            }
            int n_events = 0, n_rounds = 0;
            long long c_scan = 0, c_comp = 0;
            for (int round = 0;; ++round) {
                long long tt0 = tracing ? clock64() : 0;
                ++n_rounds;
                int fl = 0;
                if (ONE) {
#pragma unroll
                    for (int u = 0; u < kChPerWarp; ++u) {
                        const int ch = half + u * kChStride;
                        if (ch >= nch16 || (a.dbg & 1))
                            continue;
                        const uint32_t nd = tcr_filter_chunk<KP, ONE>(
                                vv[ONE ? u : 0], ch, round == 0,
                                (uint32_t)(dropw >> (u * 16)) & 0xffffu, nrm, nrmf, rowid,
                                row_in_list, limited, s, fl, n_events);
                        dropw = (dropw & ~(0xffffull << (u * 16))) | ((uint64_t)nd << (u * 16));
                    }
                } else {
#pragma unroll 1
                    for (int u = 0; u < kChPerWarp; ++u) {
                        const int ch = half + u * kChStride;
                        if (ch >= nch16)
                            break;
                        // y_hi.x_hi + y_lo.x_hi sits in column j, y_hi.x_lo in column n16 + j
                        uint32_t w[16];
                        tmem_ld_32x32b_x16(tD + (uint32_t)ch * 16u, vv[0]);
                        tmem_ld_32x32b_x16(tD + (uint32_t)(n16 + ch * 16), w);
                        tmem_ld_wait();
#pragma unroll
                        for (int j = 0; j < 16; ++j)
                            vv[0][j] = __float_as_uint(__uint_as_float(vv[0][j]) +
                                                       __uint_as_float(w[j]));
                        if (a.dbg & 1)
                            continue;
                        const uint32_t nd = tcr_filter_chunk<KP, ONE>(
                                vv[0], ch, round == 0, (uint32_t)(dropw >> (u * 16)) & 0xffffu, nrm,
                                nrmf, rowid, row_in_list, limited, s, fl, n_events);
                        dropw = (dropw & ~(0xffffull << (u * 16))) | ((uint64_t)nd << (u * 16));
                    }
                }
                if (fl)
                    atomicOr(&s.flag[ep & 3], fl);
                long long tt1 = tracing ? clock64() : 0;
                c_scan += tt1 - tt0;
                named_bar_sync(2, kREpiWarps * 32);
                const int f = s.flag[ep & 3];
                if (tid == 0)
                    s.flag[(ep + 2) & 3] = 0;
                ++ep;
                if (round == 0 && myq >= 0 && gthr < s.thr[tid]) {
                    // another CTA has a better bound for this query
                    atomicMin(&s.thr[tid], gthr);
                    const float tp = from_ordered_bits(gthr) - s.xn[tid];
                    s.thrp[tid] = tp;
                    s.thrf[tid] = loosen(tp);
                }
                if (f == 0) {
                    c_comp += (tracing ? clock64() : 0) - tt1;
                    break; // the common case: one barrier per tile
                }
                tcr_compact<KP, kREpiWarps, ONE>(s.cand, s.cnt, s.thr, s.thrp, s.thrf, s.xn, s.qid, a.qthr, nqw, warp, lane,
                                false);
                named_bar_sync(2, kREpiWarps * 32);
                c_comp += (tracing ? clock64() : 0) - tt1;
                if (!(f & 2) || round >= 16) // (8 rounds drain any tile; the cap is a safety net)
                    break;
            }
            if (!ONE) { // accumulator buffer drained (re-examination rounds read it again)
                tc_fence_before_sync();
                __syncwarp();
                if (lane == 0)
                    mbar_arrive(&s.d_empty[b]);
            }
            if (tid == 0) {
                TCR_TRACE(4, t);
                if (tracing && t < 64) {
                    a.trace[5 * 64 + t] = n_rounds;
                    a.trace[6 * 64 + t] = n_events;
                    a.trace[7 * 64 + t] = c_scan;
                    a.trace[8 * 64 + t] = c_comp;
                }
            }
        }
        if (tracing && tid == 0)
            a.trace[9 * 64 + 5] = w_dfull;
        // ---- item done: best KP of every query, sorted, to `part` --------------------------------
        named_bar_sync(2, kREpiWarps * 32);
        tcr_compact<KP, kREpiWarps, ONE>(s.cand, s.cnt, s.thr, s.thrp, s.thrf, s.xn, s.qid, a.qthr, nqw, warp, lane, true);
        __syncwarp();
        for (int qi = warp; qi < nqw; qi += kREpiWarps) {
            const int n = s.cnt[qi];
            uint64_t* out = a.part + ((size_t)(q0 + qi) * a.nsplit + split) * KP;
            const uint64_t* buf = s.cand + (size_t)qi * C;
            for (int i = lane; i < KP; i += 32)
                out[i] = i < n ? buf[i] : kSentinelKey;
        }
    }

    // ---- teardown -----------------------------------------------------------------------------------
    tc_fence_before_sync();
    named_bar_sync(1, kRLoaderWarps * 32);
    if (warp == kRMmaWarp) {
        tc_fence_after_sync();
        tmem_dealloc(tmem, SH::TmemCols);
    }
    if (tid == 0)
        TCR_TRACE(0, 5);
}

} // namespace

// ---- host side ----------------------------------------------------------------------------------
// Host-mapped note a trapping wait leaves behind (mbar_wait_code): {1 + which wait, CTA, thread,
// iteration, parity}. Pinned host memory, so it survives the context the trap kills.
static unsigned long long* g_trap_host = nullptr;
static unsigned long long* g_trap_dev = nullptr;
unsigned long long* tc_trap_info_device() {
    if (!g_trap_host) {
        void* h = nullptr;
        if (cudaHostAlloc(&h, 8 * sizeof(unsigned long long), cudaHostAllocMapped) != cudaSuccess) {
            cudaGetLastError();
            return nullptr;
        }
        g_trap_host = static_cast<unsigned long long*>(h);
        for (int i = 0; i < 8; ++i)
            g_trap_host[i] = 0;
        void* d = nullptr;
        if (cudaHostGetDevicePointer(&d, h, 0) != cudaSuccess) {
            cudaGetLastError();
            return nullptr;
        }
        g_trap_dev = static_cast<unsigned long long*>(d);
    }
    return g_trap_dev;
}
const unsigned long long* tc_trap_info_host() {
    return g_trap_host;
}

template <int METRIC, int KP, bool SMALL, bool ONE>
static cudaError_t launch_tcr_shape(uint32_t max_items, const TcScanArgs& a_in, cudaStream_t st) {
    TcScanArgs a = a_in;
    const size_t smem = SMALL ? kRSmemSmall : kRSmemBig;
    a.nstages = (int)(smem / 1024); // the kernel carves per item (tcr_stages)
    if (a.tq < 16 || a.tq > RShape<SMALL, ONE>::MaxTQ || (a.tq & 15) ||
        tcr_stages(smem, KP, a.tq, 2 * ((a.nf4 + 1) >> 1), !SMALL && !ONE, ONE) < 3)
        return cudaErrorInvalidValue;
    const int dbg = g_tc_dbg.load(std::memory_order_relaxed); // B200IVF_TC_DBG / "tc_dbg"
    a.dbg = a.nlists > 1 ? dbg : 0; // timing experiments; never for the coarse quantizer
    auto kern = ivf_scan_tcr_kernel<METRIC, KP, SMALL, ONE>;
    cudaError_t e =
            cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess)
        return e;
    if (max_items == 0)
        return cudaSuccess;
    a.trap = tc_trap_info_device();
    static const char* tr = getenv("B200IVF_TC_TRACE");
    static const char* trc = getenv("B200IVF_TC_TRACE_COARSE");
    if (tr && (a.nlists > 1) == (trc == nullptr)) {
        static long long* buf = nullptr;
        if (!buf)
            cudaMalloc(&buf, 10 * 64 * sizeof(long long));
        cudaMemsetAsync(buf, 0, 10 * 64 * sizeof(long long), st);
        a.trace = buf;
        a.trace_item = (uint32_t)atoi(tr);
    }
    kern<<<max_items, RShape<SMALL, ONE>::Threads, smem, st>>>(
            a, *reinterpret_cast<const CUtensorMap*>(a.tmap));
    count_launches(1);
    g_tc_launches.fetch_add(1, std::memory_order_relaxed);
    if (a.trace) {
        long long h[10 * 64];
        cudaStreamSynchronize(st);
        cudaMemcpy(h, a.trace, sizeof(h), cudaMemcpyDeviceToHost);
        const long long t0 = h[0];
        fprintf(stderr, "TCRTRACE item %u ntiles %lld nq %lld: sync %lld b_rdy %lld end %lld\n",
                a.trace_item, h[2], h[3], h[1] - t0, h[4] - t0, h[5] - t0);
        const char* names[5] = {"", "tma", "mma", "split", "epi"};
        for (int r = 1; r < 5; ++r) {
            fprintf(stderr, "TCRTRACE %-6s", names[r]);
            for (int i = 0; i < 24 && i < (int)h[2]; ++i)
                fprintf(stderr, " %lld", h[r * 64 + i] ? h[r * 64 + i] - t0 : -1);
            fprintf(stderr, "\n");
        }
        fprintf(stderr,
                "TCRTRACE waits: split full %lld a_empty %lld | mma a_rdy %lld d_empty %lld | tma empty "
                "%lld | epi d_full %lld | mma issue %lld\n",
                h[9 * 64 + 0], h[9 * 64 + 1], h[9 * 64 + 2], h[9 * 64 + 3], h[9 * 64 + 4],
                h[9 * 64 + 5], h[9 * 64 + 6]);
        const char* cn[4] = {"rounds", "events", "c_scan", "c_comp"};
        for (int r = 5; r < 9; ++r) {
            fprintf(stderr, "TCRTRACE %-6s", cn[r - 5]);
            for (int i = 0; i < 24 && i < (int)h[2]; ++i)
                fprintf(stderr, " %lld", h[r * 64 + i]);
            fprintf(stderr, "\n");
        }
    }
    return cudaGetLastError();
}

// tq <= 32: the two-CTAs-per-SM shape, else the whole-SM one
template <int METRIC, int KP>
static cudaError_t launch_tcr_one(uint32_t max_items, const TcScanArgs& a, cudaStream_t st) {
    const int nch = 2 * ((a.nf4 + 1) >> 1);
    if (a.one_mma) {
        const bool small = a.tq <= RShape<true, true>::MaxTQ &&
                tcr_stages(kRSmemSmall, KP, a.tq, nch, false, true) >= 3;
        return small ? launch_tcr_shape<METRIC, KP, true, true>(max_items, a, st)
                     : launch_tcr_shape<METRIC, KP, false, true>(max_items, a, st);
    }
    const bool small = a.tq <= RShape<true, false>::MaxTQ &&
            tcr_stages(kRSmemSmall, KP, a.tq, nch, false) >= 3;
    return small ? launch_tcr_shape<METRIC, KP, true, false>(max_items, a, st)
                 : launch_tcr_shape<METRIC, KP, false, false>(max_items, a, st);
}

// queries per item for dimension nf4*4 and kp kept keys: the largest multiple of 16 (<= kTcrTQ) that
// leaves the raw ring at least 4 stages; 0 = the kernel cannot take this shape
int scan_tcr_tq(int nf4, int kp, bool one) {
    if (nf4 < 1 || nf4 > kRMaxChunks ||
        (kp != 16 && kp != 32 && kp != 64 && kp != 128 && kp != 256))
        return 0;
    const int nch = 2 * ((nf4 + 1) >> 1);
    for (int tq = kTcrTQ; tq >= 16; tq -= 16)
        if (tcr_stages(227 * 1024, kp, tq, nch, !one, one) >= 4)
            return tq;
    return 0;
}
// queries per item of the two-CTAs-per-SM shape (<= 32; <= 64 for the one-MMA kernels, whose
// accumulators are half as wide), 0 = that shape cannot take (nf4, kp)
int scan_tcr_tq_small(int nf4, int kp, bool one) {
    if (!scan_tcr_tq(nf4, kp, one))
        return 0;
    const int nch = 2 * ((nf4 + 1) >> 1);
    for (int tq = one ? RShape<true, true>::MaxTQ : RShape<true, false>::MaxTQ; tq >= 16; tq -= 16)
        if (tcr_stages(kRSmemSmall, kp, tq, nch, false, one) >= 3)
            return tq;
    return 0;
}
// keys kept per partial list for a request of k: always a few more than k, because the scores
// that select them are 3xTF32 ones and the exact fp32 re-rank (select.cu) can only re-order what
// was kept: margin >= 4 for k <= 12, >= 8 above
int scan_tcr_kpart(int nf4, int k) {
    const int kp = k + 4 <= 16 ? 16
            : k + 8 <= 32      ? 32
            : k + 8 <= 64      ? 64
            : k + 8 <= 128     ? 128
            : k + 8 <= 256     ? 256
                               : 0;
    return kp && scan_tcr_tq(nf4, kp, false) ? kp : 0;
}

int scan_tcr_kpart_one(int nf4, int k, int margin_min) {
    const int need = k + (k / 4 > margin_min ? k / 4 : margin_min);
    const int kp = need <= 16  ? 16
            : need <= 32       ? 32
            : need <= 64       ? 64
            : need <= 128      ? 128
            : need <= 256      ? 256
                               : 0;
    return kp && scan_tcr_tq(nf4, kp, true) ? kp : 0;
}

cudaError_t launch_ivf_scan_tcr(int metric, uint32_t max_items, const TcScanArgs& a,
                                cudaStream_t st) {
    if (a.nf4 > kRMaxChunks || a.nf4 < 1)
        return cudaErrorInvalidValue;
    if (a.kpart == 16)
        return metric == kMetricL2 ? launch_tcr_one<kMetricL2, 16>(max_items, a, st)
                                   : launch_tcr_one<kMetricInnerProduct, 16>(max_items, a, st);
    if (a.kpart == 32)
        return metric == kMetricL2 ? launch_tcr_one<kMetricL2, 32>(max_items, a, st)
                                   : launch_tcr_one<kMetricInnerProduct, 32>(max_items, a, st);
    if (a.kpart == 64)
        return metric == kMetricL2 ? launch_tcr_one<kMetricL2, 64>(max_items, a, st)
                                   : launch_tcr_one<kMetricInnerProduct, 64>(max_items, a, st);
    if (a.kpart == 128)
        return metric == kMetricL2 ? launch_tcr_one<kMetricL2, 128>(max_items, a, st)
                                   : launch_tcr_one<kMetricInnerProduct, 128>(max_items, a, st);
    if (a.kpart == 256)
        return metric == kMetricL2 ? launch_tcr_one<kMetricL2, 256>(max_items, a, st)
                                   : launch_tcr_one<kMetricInnerProduct, 256>(max_items, a, st);
    return cudaErrorInvalidValue;
}

} // namespace b200ivf
