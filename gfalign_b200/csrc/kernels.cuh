// gfalign_b200/csrc/kernels.cuh — sm_100a kernels of the scoring engine (declarations + layout).
//
// Data layout in HBM (all little-endian, dense):
//   off   u32[A+1]        CSR offsets of the alignment table            (as handed in)
//   steps u32[S]          oriented steps, node<<1|minus, global dense node ids   (edge support)
//   -- scoring view, rebuilt by gfe_set_universe (length-bucketed, universe-local, 16-bit) --
//   bstep u16[..]         for every length m = 1..kMaxBucket a dense [count_m][row_stride(m)] matrix:
//                         m local steps (local<<1|minus; every node outside the universe is
//                         local id L), then m-1 PAIR IDS (id of the consecutive oriented pair
//                         (B_j, B_j+1) in the pair dictionary, id^1 = its reverse complement;
//                         2C = pair with a node outside the universe), zero padding to a multiple
//                         of 4; then the steps of the "long" alignments (m > kMaxBucket) back to back
//   pkeys/pids u32[H]     pair dictionary: open-addressing hash of canonical pair keys
//                         min(x<<16|y, (y^1)<<16|(x^1)) -> canonical id c; pair id = 2c + flipped
//   masks u32[chunks][E][W]  per batch: bitsets over the chunk's 32*W candidate paths, one entry per
//                         node (N: on the path, T: visited twice, F: node of path[0]; 4 words
//                         interleaved per path word), per length
//                         (LM[m]: n < m) and per pair id (M: the path takes this oriented pair)
//   ldesc uint2[n_long]   long alignments: (element offset into bstep, m)
//   perm  u32[A]          sorted row -> original alignment index (rows of length 0 first, then
//                         m = 1, 2, ... kMaxBucket, then long)
//   ptab  u16[P][stride]  per candidate path: node -> first-position table, entry
//                         0 = node not on the path, else twice<<15 | (pos+1)<<1 | minus
//   pstep u16[..]         candidate paths in local ids (CSR by poff u32[P+1])
//   pnext u16[..]         same CSR: next position visiting the same node (0xFFFF = none)
//
// Reference semantics implemented (file:line in /root/reference):
//   filter            src/eval.cpp:81-91
//   verdict           src/eval.cpp:92-98   (good <=> T_fw == 0 or T_rc == 0)
//   T (traceback)     src/alignments.cpp:499-554
//   edge support      src/alignments.cpp:353-403, src/eval.cpp:42-60
#pragma once
#include <cuda_runtime.h>
#include <cstdlib>
#include <stdint.h>

namespace gfe {

constexpr int kBlock = 256;           // threads per CTA for the streaming kernels
constexpr int kMaxBucket = 16;        // alignment lengths with their own dense matrix
constexpr int kNumBuckets = kMaxBucket + 2;   // 0, 1..16, long
constexpr int kLongBucket = kMaxBucket + 1;
constexpr int kExactRegSteps = 32;    // widest alignment the register DP handles
constexpr int kMaxChunk = 64;         // candidate paths per CTA pass (tables in shared memory)
constexpr uint32_t kTabAbsent = 0;
constexpr int kBitThreads = 512;      // threads per CTA of the bitset kernel (one CTA per SM)
constexpr uint32_t kSlowHole = 0xFFFFFFFFu;   // unused slot of the second-look list (bitset.cu push_slow8)
constexpr int kWorkCounters = 8;      // [0] exact entries [1] exact overflow [2] second-look slots [3] overflow [4] holes among [2]
constexpr int kPlanes = 12;           // bit-sliced counter planes per lane (flush before 2^12 rows)
constexpr uint32_t kPairHashCap = 1u << 16;   // slots of the pair dictionary
constexpr uint32_t kPairEmpty = 0xFFFFFFFFu;
constexpr uint32_t kBloom = 128;      // hashed "broken triple" entries per orientation (bitset kernel)
constexpr size_t kBitsetSmemBudget = 208 * 1024;   // mask table + counters per CTA

// Hash of the consecutive-pair pair (qa, qb) = the triple (b_j-1, b_j, b_j+1), symmetric under
// reverse complement up to the low bit: triple_slot(qb^1, qa^1) == triple_slot(qa, qb) ^ 1.
__host__ __device__ inline uint32_t triple_slot(uint32_t qa, uint32_t qb) {
    const uint32_t key = (qa << 16) | qb, twin = ((qb ^ 1u) << 16) | (qa ^ 1u);
    const uint32_t canon = key < twin ? key : twin;
    return (((canon * 0x9E3779B1u) >> 25) << 1) | (key != canon ? 1u : 0u);   // 2*kBloom slots
}

// (row = m steps + (m-1) pair ids as u16, then (m-2) one-byte triple slots — triple_slot of the two
// pairs around each interior step — padded to 8 bytes)
__host__ __device__ constexpr uint32_t row_stride(int m) {
    return m <= 0 ? 0u : (uint32_t)((2 * m - 1 + (m > 2 ? (m - 1) / 2 : 0) + 3) & ~3);
}

struct Buckets {
    uint32_t count[kNumBuckets];          // rows per bucket
    uint32_t row_start[kNumBuckets + 1];  // first sorted row of each bucket
    uint32_t tile_start[kNumBuckets + 1]; // first kBlock-row tile of each bucket (bucket 0: none)
    uint32_t step_base[kNumBuckets];      // element offset of the bucket's matrix in bstep
    const uint16_t *bstep;
    const uint2 *ldesc;
    const uint32_t *perm;
};

struct ScoreArgs {
    Buckets bk;
    const uint32_t *poff;     // P+1
    const uint16_t *pstep;
    const uint16_t *ptab;     // P rows
    const uint16_t *pnext;
    uint32_t tab_stride;      // entries per row (multiple of 8)
    uint32_t P;
    uint32_t chunk;           // paths per CTA (<= kMaxChunk)
    int filter;
    uint32_t *counts;         // P x {bad, good}
    uint2 *work;              // exact-DP worklist (path, sorted row)
    uint32_t *work_count;     // [0] entries, [1] overflow flag
    uint32_t work_cap;
    // count record of path p, accumulator set s (list entries carry s in bit 30 of the path field):
    // counts[count_stride * p + 2 * s + {0: bad, 1: good}]; plain batches: stride 2, one set
    uint32_t count_stride = 2;
    const uint8_t *rshift = nullptr;   // deduplicated view: sorted row -> k, the row counts 2^k alignments
    uint32_t *err = nullptr;  // bit 2: a row of more than GFE_MAX_STEPS steps passed the id filter (reference: UB)
    uint32_t path_tag = 0;    // OR-ed into the path field of the entries score_paths_kernel queues (set << 30)
};
constexpr uint32_t kPathMask = 0x3FFFFFFFu;   // path index inside a list entry's x (bit 31 guess, bit 30 set)
__host__ __device__ inline uint32_t entry_path(uint32_t x) { return x & kPathMask; }
__host__ __device__ inline size_t entry_count_base(uint32_t stride, uint32_t x) {
    return (size_t)stride * (x & kPathMask) + (((x >> 30) & 1u) << 1);
}

struct PairDict {
    uint32_t *keys;           // kPairHashCap slots, kPairEmpty = free
    uint32_t *ids;            // canonical id of the slot's key
    uint32_t *counter;        // [0] distinct keys inserted, [1] ids handed out, [2] overflow flag
};

// mask-table geometry of one batch (entries are W words each)
struct MaskGeom {
    uint32_t W;               // words per entry = chunk paths / 32 (power of two <= 32)
    uint32_t L;               // universe size; node v, word w: {N, F+|F-} at (v*W + w)*2, {F+, F-} in BitsetArgs::fplane (global memory)
    uint32_t node_base;       // node plane starts at entry node_base (= 2*kBloom: BT comes first, at entry 0)
    uint32_t lm_base;         // LM[m] at lm_base + m, m = 0..kMaxBucket
    uint32_t pair_base;       // M[pid] at pair_base + pid, pid <= 2C+1 (2C, 2C+1: always zero)
    uint32_t bt_base;         // BT[h] at bt_base + h, h < 2*kBloom: paths with a broken triple hashing to h
    uint32_t entries;         // E, multiple of 4
    uint32_t n_pairs;         // C
};

// Units [unit_start, next group's unit_start) score rows [row0, row0 + nrows) of the matrix of bucket m;
// every row of the group counts 2^shift alignments (shift > 0 only in the deduplicated view).
constexpr int kMaxGroups = 512;
struct __align__(16) UnitGroup {
    uint32_t unit_start, row0, nrows;
    uint16_t m, shift;
};

struct BitsetArgs {
    Buckets bk;
    MaskGeom mg;
    const uint32_t *masks;    // [chunks][E][W]
    const uint32_t *fplane;   // [chunks][L+1][W] x {F+, F-}: path[0] is this node in that orientation (second look)
    uint32_t P;
    int filter;
    uint32_t *counts;         // P x {bad, good}
    uint32_t unit_rows;       // rows per scheduling unit (multiple of 8 * 32/W, set per launch)
    uint32_t unit_start[kMaxBucket + 2];   // plain view: first unit of buckets 1..kMaxBucket; [kMaxBucket+1] = total
    const UnitGroup *groups;  // class-sorted view: groups (length, class) in ascending unit order; else nullptr
    uint32_t n_groups, n_units;
    uint32_t *unit_ctr;       // per chunk: next unit (dynamic scheduling), zeroed per launch
    uint2 *slow;              // (path, sorted row) pairs the bitsets cannot settle
    uint32_t *work_count;     // [0..1] exact list count/overflow, [2..3] slow list count/overflow
    uint32_t slow_cap;
};

// Inverted index of the scoring view (incremental search scoring): for every universe
// node v and row length m the bucket-local ids of the rows of length m that contain v — except the rows
// of F = {rows containing the root node (path[0] of every candidate: the only rows that can overhang
// the path start, SURVEY.md A.3 class iv)}, which form the groups of pseudo-node L.  Rows naming a
// node outside the universe appear nowhere (the id filter drops them for every path).  Long rows
// (m > kMaxBucket) are not indexed: they are scored afresh for every path like F.
constexpr int kIdxLens = kMaxBucket + 1;      // group (v, m) at v * kIdxLens + m, m = 1..kMaxBucket
struct InvIndex {
    const uint32_t *off;      // (L + 1) * kIdxLens + 1
    const uint32_t *rows;     // bucket-local row ids
};

// Incremental batch: chunks of <= 32 paths (one word, W = 1) that END IN THE SAME NODE v.  Pass 0 scores
// the rows of list(v) (accumulator set 0), pass 1 the rows of F (set 1).
struct IncrArgs {
    Buckets bk;
    MaskGeom mg;              // W = 1
    InvIndex inv;
    const uint32_t *masks;    // [chunks][entries]
    const uint32_t *fplane;   // [chunks][L+1] x {F+, F-}
    const uint32_t *chunk_node;   // per chunk: local node of pass 0, 0xFFFFFFFF = no pass 0
    uint32_t chunks;
    uint32_t *counts;         // [chunks * 32][4]: {bad, good} of set 0, {bad, good} of set 1
    uint32_t unit_rows;
    uint32_t *unit_ctr;       // per chunk, zeroed per launch
    uint2 *slow;
    uint32_t *work_count;
    uint32_t slow_cap;
};

struct PairArgs {             // evalPath: one path against every alignment
    Buckets bk;
    const uint16_t *pstep;    // the path, local ids
    uint32_t n;
    const uint16_t *ptab;     // one row
    const uint16_t *pnext;    // next visit of the same node, per path position (0xFFFF = none)
    uint32_t tab_entries;
    int filter;
    int32_t *t_fw, *t_rc;     // may be null; indexed by ORIGINAL alignment index
    uint32_t *counts;         // {bad, good}
    uint32_t *work;           // sorted rows needing the DP
    uint32_t *work_count;
    uint32_t work_cap;
};

struct __align__(16) EdgeSlot {   // evalGFA: one directed arc of the graph (edges.cu)
    unsigned long long key;       // from_step << 32 | to_step; ~0 = empty
    uint32_t count;               // index of the arc's counter
    uint32_t rep;                 // lowest slot (2i = record i, 2i+1 = its reverse complement) carrying it
};

struct EdgeBuild {                // evalGFA: what launch_edge_build fills from the link records
    EdgeSlot *table;
    uint32_t cap;
    uint32_t *deg;                // arcs per oriented step (2 * n_nodes)
    uint4 *adj;                   // dense adjacency records {to_0, to_1, to_2, first counter}
    uint32_t steps2;              // 2 * n_nodes
    uint32_t *cursor;             // one word: counters handed out
};

struct EdgeStreamArgs {
    const uint32_t *steps;
    const uint32_t *rowstart;     // bitmap over steps: bit s = step s starts an alignment
    uint64_t S;                   // total steps
    const uint4 *adj;
    uint32_t *cnt;                // one counter per directed arc (zeroed by the caller)
    uint32_t steps2;
    uint32_t n_arcs;              // counters (upper bound 2 E)
    const EdgeSlot *table;        // arcs of crowded steps (more than three per step)
    uint32_t cap;
};

struct TraceArgs {            // evalPath detail: traceback columns of one path vs a batch of alignments
    const uint32_t *off;      // original CSR (global node ids)
    const uint32_t *steps;
    const uint32_t *path;     // n oriented steps, global ids
    uint32_t n;
    uint64_t first;           // first alignment of the batch
    uint32_t count;
    const uint8_t *use_rc;    // per pair: align against the reverse complement (may be null)
    uint32_t W;               // decision words per DP row = ceil(max m of the batch / 16)
    uint32_t *dirs;           // n * count * W words
    uint32_t *long_b;         // count * long_stride (only when the batch has m > 32)
    int16_t *long_rows;       // count * (long_stride + 1)
    uint32_t long_stride;
    const uint64_t *ops_start_base;   // per pair: start of its (n + m)-byte region in ops
    uint8_t *ops;
    uint64_t *ops_first;      // out: index of the pair's first column in ops
    uint32_t *ops_len;        // out: columns
    int32_t *score;           // out (may be null): traceback score
};

struct FilterArgs {           // filter mode: rows whose nodes all lie in a set, with at least min_nodes steps
    const uint32_t *off;
    const uint32_t *steps;
    uint64_t A;
    const uint32_t *node_bits;   // bitmap over n_nodes
    uint32_t n_nodes;
    int32_t min_nodes;
    uint8_t *keep;
    uint64_t S;                  // total steps
    uint32_t *bad_bits;          // scratch, (S + 31) / 32 words: step names a node outside the set
};

// hist: kNumBuckets counters, [kNumBuckets] = max length, [kNumBuckets+1] = steps of long rows
void launch_length_histogram(const uint32_t *off, uint64_t A, uint32_t *hist, cudaStream_t st);
// cursors: kNumBuckets row cursors + [kNumBuckets] long-step cursor (zeroed by the caller)
void launch_pair_collect(const uint32_t *off, const uint32_t *steps, const uint32_t *node2local,
                         uint64_t A, uint32_t L, uint32_t max_pairs, const PairDict &pd,
                         cudaStream_t st);
void launch_pair_assign(const PairDict &pd, cudaStream_t st);
// n_pairs = 0xFFFFFFFF: no dictionary (pair ids are not written)
// pair_bias is added to every stored pair id, so a row holds ABSOLUTE table entries (MaskGeom::pair_base + id)
void launch_bucket_scatter(const uint32_t *off, const uint32_t *steps, const uint32_t *node2local,
                           uint64_t A, const Buckets &bk, uint16_t *bstep, uint2 *ldesc,
                           uint32_t *perm, uint32_t *cursors, const PairDict &pd, uint32_t n_pairs,
                           uint32_t pair_bias, cudaStream_t st, bool only_long = false);
// multiplicity compression (kernels.cu): distinct rows up to reverse complement, class k = bit k of the
// multiplicity; gcount has kDedupStats + 2 words (zeroed by the caller)
struct DedupSlot {
    unsigned long long key;   // fingerprint << 32 | representative alignment; ~0 = empty
    uint32_t w, pad;          // multiplicity
};
constexpr uint32_t kDedupStats = (kMaxBucket + 1) * 32;
void launch_dedup_insert(const uint32_t *off, const uint32_t *steps, const uint32_t *node2local, uint64_t A,
                         DedupSlot *slots, uint32_t cap /* power of two */, cudaStream_t st);
void launch_dedup_count(const DedupSlot *slots, uint32_t cap, const uint32_t *off, uint32_t *gcount, cudaStream_t st);
void launch_dedup_scatter(const DedupSlot *slots, uint32_t cap, const uint32_t *off, const uint32_t *steps,
                          const uint32_t *node2local, const Buckets &bk, uint16_t *bstep, const uint32_t *class_base,
                          uint32_t *cursors, uint8_t *rshift, const PairDict &pd, uint32_t n_pairs, uint32_t pair_bias,
                          cudaStream_t st);

// mask-table geometry as a function of the universe size L (entries of W words): BT first (so its
// byte offsets need no base), then the node plane, LM, and the pair masks on an even entry
__host__ __device__ constexpr uint32_t geom_node_base() { return 2u * kBloom; }
__host__ __device__ constexpr uint32_t geom_lm_base(uint32_t L) { return geom_node_base() + 2u * (L + 1u); }
__host__ __device__ constexpr uint32_t geom_pair_base(uint32_t L) { return (geom_lm_base(L) + kMaxBucket + 2u) & ~1u; }
// Programmatic dependent launch (the chain prologue -> tables -> bitset scoring -> second look ->
// exact DP of one batch): a kernel launched with launch_pdl may be scheduled while its predecessor
// in the stream drains; it must call pdl_wait() before it touches anything a predecessor wrote
// (the wait returns when every earlier grid has completed and flushed).  pdl_trigger(), placed
// AFTER the kernel's own wait, lets its successor be scheduled early -- so at most two kernels of
// the chain overlap and a waiting kernel can only be early with respect to its direct predecessor.
// The first value a kernel reads from that predecessor goes through a volatile load: a load on the
// read-only path (const __restrict__ / __ldg) counts as invariant and may be hoisted above the
// wait.  Without the launch attribute both calls are no-ops (plain <<<>>> launches stay correct).
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;"); }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(int edge, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                              cudaStream_t st, Args &&...args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    static const int mask = getenv("GFE_PDL_MASK") ? atoi(getenv("GFE_PDL_MASK")) : 15;   // debugging aid
    cfg.attrs = attr;
    cfg.numAttrs = (mask >> edge) & 1;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif

// One launch in front of every scoring batch: clears the batch's accumulators (count records, work
// counters, mask tables, unit counters) and, for paths given as raw device steps, maps them to
// universe-local 16-bit steps (psteps == nullptr: the host already did).
struct PrologueArgs {
    uint32_t *zero[4];
    uint32_t zero_words[4];
    const uint32_t *psteps;
    const uint32_t *map;
    uint32_t n_nodes, L;
    uint16_t *out;
    uint32_t total;
    uint32_t *err;
    uint32_t *sticky;         // may be null: the work counters' overflow flags (zero[1][1], [3]) of the PREVIOUS
                              // batch are OR-ed in here (bit1) before they are cleared
};
void launch_prologue(const PrologueArgs &a, int sm_count, cudaStream_t st);
// *out = bit0 a path step outside the universe, bit1 a worklist overflowed (gfe_device_status, on the
// device); err[0] (bad step) and err[1] (sticky overflow of earlier batches) are cleared once reported
void launch_status_flag(uint32_t *err, const uint32_t *wc, uint32_t *out, cudaStream_t st);
// the same word without clearing anything (the host API reads the header itself afterwards)
void launch_flag_word(const uint32_t *err, const uint32_t *wc, uint32_t *out, cudaStream_t st);
// *out |= (*sticky ? 2 : 0)
void launch_flag_or(const uint32_t *sticky, uint32_t *out, cudaStream_t st);
// masks == nullptr: tables only (legacy kernel); else also ORs path p into the batch's bitsets
void launch_build_tables(const uint32_t *poff, const uint16_t *pstep, uint32_t P, uint32_t L,
                         uint32_t tab_stride, uint16_t *ptab, uint16_t *pnext, uint32_t *masks,
                         uint32_t *fplane, const MaskGeom &mg, const PairDict &pd, cudaStream_t st);
// tiles [tile_lo, tile_hi) of the bucket geometry; add_empty: count the step-less rows as good
void launch_score_paths(const ScoreArgs &a, int sm_count, uint32_t tile_lo, uint32_t tile_hi,
                        bool add_empty, cudaStream_t st);
void launch_score_bitset(const BitsetArgs &a, int sm_count, cudaStream_t st);
void launch_score_incr(const IncrArgs &a, int sm_count, cudaStream_t st);
// inverted index: count -> scan -> fill.  cnt/off have (L + 1) * kIdxLens + 1 words
void launch_index_count(const Buckets &bk, uint32_t L, uint32_t root_local, uint32_t *cnt, cudaStream_t st);
void launch_index_scan(const uint32_t *cnt, uint32_t groups, uint32_t *off, uint32_t *cursor, cudaStream_t st);
void launch_index_fill(const Buckets &bk, uint32_t L, uint32_t root_local, uint32_t *cursor, uint32_t *rows,
                       cudaStream_t st);
// candidate paths given as (base path, suffix) pairs -> local 16-bit steps at out[poff[p]..]; desc[p] =
// {base id or 0xFFFFFFFF, first suffix step, suffix steps, steps to drop at the end}
struct AssembleArgs {
    const uint4 *desc;
    uint32_t P;
    const uint32_t *base_off;     // raw base paths (global ids)
    const uint32_t *base_steps;
    const uint32_t *suffix_steps;
    const uint32_t *poff;         // P + 1, lengths as the host computed them
    const uint32_t *map;          // node -> local
    uint32_t n_nodes, L;
    uint16_t *out;
    uint32_t *err;
};
void launch_assemble_paths(const AssembleArgs &a, cudaStream_t st);
// candidate-batch sharding: raw slot records -> per-item records {d bad, d good, fresh bad, fresh good}; slot_item[k] =
// item | kind << 30 (0 child, 1 parent copy, 2 full scan; 0xFFFFFFFF padding), slot_aux[k] = index into `full` (kind 2)
void launch_compose_items(const uint32_t *slot_item, const uint32_t *slot_aux, uint32_t P_dev, const uint32_t *inc,
                          const uint32_t *full, int32_t *out, cudaStream_t st);
void launch_slow_worklist(const ScoreArgs &a, const uint2 *slow, const uint32_t *slow_count,
                          uint32_t slow_cap, cudaStream_t st);
void launch_exact_worklist(const ScoreArgs &a, cudaStream_t st);
void launch_exact_worklist_long(const ScoreArgs &a, uint32_t max_m, cudaStream_t st);
void launch_pair_classify(const PairArgs &a, int sm_count, cudaStream_t st);
void launch_pair_exact(const PairArgs &a, cudaStream_t st);
void launch_pair_exact_long(const PairArgs &a, uint32_t max_m, cudaStream_t st);
uint32_t edge_table_capacity(uint32_t E);
void launch_rowstart(const uint32_t *off, uint64_t A, uint32_t *rowstart, cudaStream_t st);
void launch_edge_build(const uint32_t *d_ef, const uint32_t *d_et, uint32_t E, const EdgeBuild &b, cudaStream_t st);
void launch_edge_stream(const EdgeStreamArgs &a, int sm_count, cudaStream_t st);
void launch_edge_scatter(const EdgeSlot *table, uint32_t cap, const uint32_t *cnt, uint32_t *counts, cudaStream_t st);
void launch_edge_weights(const uint32_t *d_ef, const uint32_t *d_et, uint32_t E, const EdgeSlot *table,
                         uint32_t cap, const uint32_t *cnt, uint32_t *weights, cudaStream_t st);
int configure_edge_kernels();
int configure_bitset_kernels();
void launch_max_len(const uint32_t *off, uint64_t first, uint32_t count, uint32_t *out,
                    cudaStream_t st);
void launch_trace_bases(const uint32_t *off, uint64_t first, uint32_t count, uint32_t n,
                        uint64_t *base, cudaStream_t st);
void launch_pair_trace(const TraceArgs &a, cudaStream_t st);
void launch_filter_rows(const FilterArgs &a, int sm_count, cudaStream_t st);
int configure_kernels();   // cudaFuncSetAttribute for the large-shared-memory kernels

}  // namespace gfe
