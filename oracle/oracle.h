/*
 * oracle/oracle.h — CPU restatement of gfalign's scoring hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library; the product (libgfalign_b200.so, the gfalign-b200 CLI) never links it.
 *
 * Parity status: PINNED.  Checked against (a) the compiled, unmodified reference
 * (oracle/_ref/gfalign and oracle/_ref/libgfalign_ref.so, built by oracle/Makefile from
 * /root/reference) pair by pair on randomized inputs, (b) the reference's only golden for the path
 * (validateFiles/test.6.tst) and (c) the binary-generated goldens in tests/golden/ — see
 * tests/test_oracle.py.
 *
 * Encoding shared with the product: an oriented step is u32 `node << 1 | (orientation == '-')`
 * over dense node ids; alignments are CSR (offsets[A+1], steps[offsets[A]]).
 */
#ifndef GFALIGN_ORACLE_H
#define GFALIGN_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define ORC_FILTERED INT32_MIN /* per-alignment score slot of an alignment removed by the filter */

/* Traceback score T of alignPaths(0,-1,-1,A,B,dp) — reference src/alignments.cpp:499-561. */
int32_t orc_pair_score(const uint32_t *A, uint32_t n, const uint32_t *B, uint32_t m);

/* Same, also emitting the gapped alignment (reversed back to forward order) as two arrays of
 * oriented steps with 0xFFFFFFFF for a gap; returns the length.  src/alignments.cpp:511-554. */
uint32_t orc_pair_alignment(const uint32_t *A, uint32_t n, const uint32_t *B, uint32_t m,
                            uint32_t *SA, uint32_t *SB, int32_t *score);

/* evaluatePath — reference src/eval.cpp:67-108.  best/t_fw/t_rc may be NULL (A entries else). */
void orc_evaluate_path(const uint32_t *path, uint32_t n, const uint32_t *offsets,
                       const uint32_t *steps, uint64_t A, int filter, uint32_t *bad,
                       uint32_t *good, uint32_t *unaligned, int32_t *t_fw, int32_t *t_rc);

/* P paths (CSR) against all alignments; `threads` pthreads, each over a contiguous alignment range.
 * counts = P × {bad, good}.  Used as the multi-core CPU baseline ("port"). */
void orc_evaluate_paths(const uint32_t *path_offsets, const uint32_t *path_steps, uint32_t P,
                        const uint32_t *offsets, const uint32_t *steps, uint64_t A, int filter,
                        int threads, uint32_t *counts);

/* InAlignments::buildEdgeGraph + the lookup loop of evalGFA — src/alignments.cpp:353-403,
 * src/eval.cpp:34-61.  Edges are (from_step, to_step) oriented-step pairs, E of them, in GFA
 * L-line order; weights[E] receives the RC:i value of each. */
void orc_edge_support(const uint32_t *offsets, const uint32_t *steps, uint64_t A, uint32_t n_nodes,
                      const uint32_t *edge_from, const uint32_t *edge_to, uint32_t E,
                      uint32_t *weights);

#ifdef __cplusplus
}
#endif
#endif
