// gfalign_b200/csrc/kernels.cuh — sm_100a kernels of the scoring engine (declarations + layout).
//
// Data layout in HBM (all little-endian, dense):
//   off   u32[A+1]        CSR offsets of the alignment table
//   steps u32[S]          oriented steps, node<<1|minus, global dense node ids   (edge support)
//   lstep u16[S]          the same steps in universe-local ids: local<<1|minus, every node outside
//                         the universe is local id L (= n_universe)               (scoring)
//   ptab  u16[P][stride]  per candidate path: node -> first position table, entry
//                         0 = node not on the path, else (pos+1)<<2 | minus<<1 | visited_twice
//   pstep u16[..]         candidate paths in local ids (CSR by poff u32[P+1])
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
constexpr int kRegSteps = 16;         // alignment steps kept in registers by score_paths
constexpr int kExactRegSteps = 32;    // widest alignment the register DP handles
constexpr uint32_t kTabAbsent = 0;

struct ScoreArgs {
    const uint32_t *off;      // A+1
    const uint16_t *lstep;    // S
    uint64_t A;
    const uint32_t *poff;     // P+1
    const uint16_t *pstep;
    const uint16_t *ptab;     // P rows
    uint32_t tab_stride;      // entries per row (multiple of 8)
    uint32_t P;
    uint32_t chunk;           // paths per CTA (tables resident in shared memory)
    int filter;
    uint32_t *counts;         // P x {bad, good}
    uint2 *work;              // exact-DP worklist (path, alignment)
    uint32_t *work_count;     // [0] short entries, [1] overflow flag
    uint32_t work_cap;
};

struct PairArgs {             // evalPath: one path against every alignment
    const uint32_t *off;
    const uint16_t *lstep;
    uint64_t A;
    const uint16_t *pstep;    // the path, local ids
    uint32_t n;
    const uint16_t *ptab;     // one row
    uint32_t tab_entries;
    int filter;
    int32_t *t_fw, *t_rc;     // may be null
    uint32_t *counts;         // {bad, good}
    uint32_t *work;           // alignment ids needing the DP
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

void launch_localize_steps(const uint32_t *steps, const uint32_t *node2local, uint16_t *lstep,
                           uint64_t S, cudaStream_t st);
void launch_localize_paths(const uint32_t *psteps, const uint32_t *node2local, uint32_t n_nodes,
                           uint32_t L, uint16_t *out, uint32_t total, uint32_t *err,
                           cudaStream_t st);
void launch_build_tables(const uint32_t *poff, const uint16_t *pstep, uint32_t P, uint32_t L,
                         uint32_t tab_stride, uint16_t *ptab, cudaStream_t st);
void launch_score_paths(const ScoreArgs &a, int sm_count, cudaStream_t st);
void launch_exact_worklist(const ScoreArgs &a, cudaStream_t st);
void launch_exact_worklist_long(const ScoreArgs &a, uint32_t max_m, cudaStream_t st);
void launch_pair_classify(const PairArgs &a, int sm_count, cudaStream_t st);
void launch_pair_exact(const PairArgs &a, cudaStream_t st);
void launch_pair_exact_long(const PairArgs &a, uint32_t max_m, cudaStream_t st);
void launch_edge_support(const EdgeArgs &a, int sm_count, cudaStream_t st);
int configure_kernels();   // cudaFuncSetAttribute for the large-shared-memory kernels

}  // namespace gfe
