// gfalign_b200/csrc/kernels.cuh — sm_100a kernels of the scoring engine (declarations + layout).
//
// Data layout in HBM (all little-endian, dense):
//   off   u32[A+1]        CSR offsets of the alignment table            (as handed in)
//   steps u32[S]          oriented steps, node<<1|minus, global dense node ids   (edge support)
//   -- scoring view, rebuilt by gfe_set_universe (length-bucketed, universe-local, 16-bit) --
//   bstep u16[..]         for every length m = 1..kMaxBucket a dense [count_m][m] matrix of local
//                         steps (local<<1|minus; every node outside the universe is local id L),
//                         then the steps of the "long" alignments (m > kMaxBucket) back to back
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
#include <stdint.h>

namespace gfe {

constexpr int kBlock = 256;           // threads per CTA for the streaming kernels
constexpr int kMaxBucket = 16;        // alignment lengths with their own dense matrix
constexpr int kNumBuckets = kMaxBucket + 2;   // 0, 1..16, long
constexpr int kLongBucket = kMaxBucket + 1;
constexpr int kExactRegSteps = 32;    // widest alignment the register DP handles
constexpr int kMaxChunk = 64;         // candidate paths per CTA pass (tables in shared memory)
constexpr uint32_t kTabAbsent = 0;

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
};

struct PairArgs {             // evalPath: one path against every alignment
    Buckets bk;
    const uint16_t *pstep;    // the path, local ids
    uint32_t n;
    const uint16_t *ptab;     // one row
    uint32_t tab_entries;
    int filter;
    int32_t *t_fw, *t_rc;     // may be null; indexed by ORIGINAL alignment index
    uint32_t *counts;         // {bad, good}
    uint32_t *work;           // sorted rows needing the DP
    uint32_t *work_count;
    uint32_t work_cap;
};

struct EdgeArgs {
    const uint32_t *off;
    const uint32_t *steps;
    uint64_t A;
    const uint32_t *row_off;  // 2*n_nodes+1, adjacency rows by oriented from-step
    const uint2 *row_ent;     // (to step, counter slot)
    uint32_t n_steps_space;   // 2*n_nodes
    uint32_t *counts;         // n_slots
    uint32_t n_slots;
};

// hist: kNumBuckets counters, [kNumBuckets] = max length, [kNumBuckets+1] = steps of long rows
void launch_length_histogram(const uint32_t *off, uint64_t A, uint32_t *hist, cudaStream_t st);
// cursors: kNumBuckets row cursors + [kNumBuckets] long-step cursor (zeroed by the caller)
void launch_bucket_scatter(const uint32_t *off, const uint32_t *steps, const uint32_t *node2local,
                           uint64_t A, const Buckets &bk, uint16_t *bstep, uint2 *ldesc,
                           uint32_t *perm, uint32_t *cursors, cudaStream_t st);
void launch_localize_paths(const uint32_t *psteps, const uint32_t *node2local, uint32_t n_nodes,
                           uint32_t L, uint16_t *out, uint32_t total, uint32_t *err,
                           cudaStream_t st);
void launch_build_tables(const uint32_t *poff, const uint16_t *pstep, uint32_t P, uint32_t L,
                         uint32_t tab_stride, uint16_t *ptab, uint16_t *pnext, cudaStream_t st);
void launch_score_paths(const ScoreArgs &a, int sm_count, cudaStream_t st);
void launch_exact_worklist(const ScoreArgs &a, cudaStream_t st);
void launch_exact_worklist_long(const ScoreArgs &a, uint32_t max_m, cudaStream_t st);
void launch_pair_classify(const PairArgs &a, int sm_count, cudaStream_t st);
void launch_pair_exact(const PairArgs &a, cudaStream_t st);
void launch_pair_exact_long(const PairArgs &a, uint32_t max_m, cudaStream_t st);
void launch_edge_support(const EdgeArgs &a, int sm_count, cudaStream_t st);
int configure_kernels();   // cudaFuncSetAttribute for the large-shared-memory kernels

}  // namespace gfe
