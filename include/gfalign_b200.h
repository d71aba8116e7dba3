/*
 * gfalign_b200.h — C ABI of the B200-native scoring engine for gfalign's hot path.
 *
 * The reference (vgl-hub/gfalign) has no plugin/FFI layer: its seam is three free functions,
 *     evalGFA / evalPath / dijkstra                         (reference include/eval.h:4,6,8)
 * called from Input::read()                                  (reference src/input-gfalign.cpp:94,110,121)
 * plus the two pure inner functions they spend their time in,
 *     evaluatePath(path, graph, alignmentPaths, filter, print)   (reference src/eval.cpp:67-108)
 *     InAlignments::buildEdgeGraph(hash1, hash2, uId)            (reference src/alignments.cpp:353-403)
 * Each entry point below names the reference interface it replaces.  INTEGRATION.md shows the
 * few lines a gfalign maintainer adds to call them.
 *
 * Conventions
 *   - plain C, plain pointers and sizes; the caller owns every host buffer, the engine owns all
 *     device memory; one engine = one CUDA device; a handle is not re-entrant.
 *   - every function returns GFE_OK (0) or a negative gfe_status; gfe_last_error() gives the text.
 *     There is NO CPU fallback: without a usable CUDA device gfe_create fails with GFE_ERR_CUDA.
 *   - oriented step (uint32): node << 1 | (orientation == '-'), node = dense id in [0, n_nodes).
 *     This replaces reference `struct Step {int32 id; char orientation;}` (include/alignments.h:11-21).
 *   - alignments are CSR: offsets[A+1] (offsets[0] == 0) into steps[offsets[A]]; this replaces
 *     `std::vector<Path> alignmentPaths` (reference src/eval.cpp:121, src/alignments.cpp:451-457).
 *   - reference limits are kept: a candidate path and an alignment have at most GFE_MAX_STEPS steps
 *     (reference MAX_N = 1001, include/alignments.h:246: longer inputs are undefined there; here
 *     they are rejected with GFE_ERR_LIMIT).  A longer ALIGNMENT is accepted into the table: under the
 *     id filter of search scoring the reference drops it before its DP unless all its nodes are on the
 *     candidate path (src/eval.cpp:81-91); GFE_ERR_LIMIT is raised only when one does pass the filter,
 *     and by every unfiltered call (evalPath, filter = 0).
 */
#ifndef GFALIGN_B200_H
#define GFALIGN_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GFE_MAX_STEPS 1000u        /* MAX_N - 1, reference include/alignments.h:246            */
#define GFE_MAX_UNIVERSE 32766u    /* nodes a candidate path may draw from (u16 local steps)   */
#define GFE_SCORE_FILTERED INT32_MIN

typedef enum gfe_status {
    GFE_OK = 0,
    GFE_ERR_CUDA = -1,     /* no device / CUDA runtime error (text in gfe_last_error)          */
    GFE_ERR_ARG = -2,      /* NULL pointer, malformed CSR, step out of range                    */
    GFE_ERR_LIMIT = -3,    /* path or alignment longer than GFE_MAX_STEPS, universe too large   */
    GFE_ERR_STATE = -4,    /* call order: alignments / universe not set                         */
    GFE_ERR_NOMEM = -5
} gfe_status;

typedef struct gfe_engine gfe_engine;

/* ---- lifetime --------------------------------------------------------------------------------- */
int gfe_create(int cuda_device, gfe_engine **out);
void gfe_destroy(gfe_engine *e);
const char *gfe_last_error(const gfe_engine *e);   /* e may be NULL: error of a failed gfe_create */
const char *gfe_version(void);

/* ---- alignment table ---------------------------------------------------------------------------
 * Replaces InAlignments::getPaths() (reference src/alignments.cpp:451-457) as the table every
 * scoring call reads.  Uploads the CSR to HBM (host pointers), or adopts device pointers the
 * caller keeps alive (gfe_set_alignments_device; e.g. tensors of a torch.distributed rank).
 * A == 0 is legal.  n_nodes is the size of the dense node-id space. */
int gfe_set_alignments(gfe_engine *e, const uint32_t *offsets, const uint32_t *steps, uint64_t A,
                       uint32_t n_nodes);
int gfe_set_alignments_device(gfe_engine *e, const uint32_t *d_offsets, const uint32_t *d_steps,
                              uint64_t A, uint64_t total_steps, uint32_t n_nodes);

/* The universe = the nodes candidate paths may contain (search: nodelist + source + destination,
 * reference src/eval.cpp:125-128; evalPath: the nodes of the path).  The engine keeps a second,
 * 16-bit "local" copy of the steps in which every node outside the universe is one shared id that
 * no path contains, so the scoring kernels index a small shared-memory table.  Must be called
 * after gfe_set_alignments and before the scoring calls; may be called again. */
int gfe_set_universe(gfe_engine *e, const uint32_t *nodes, uint32_t n_universe);

/* ---- search scoring: evaluatePath x P -----------------------------------------------------------
 * Replaces `evaluatePath(newPath, inSequences, alignmentPaths, true)` (reference
 * src/eval.cpp:162, body :67-108) for P candidate paths at once.  paths are CSR over oriented
 * steps (global dense ids, all inside the universe).  filter != 0 applies the id-set filter of
 * src/eval.cpp:81-91.  counts[2p] = badAlignments, counts[2p+1] = goodAlignments of path p —
 * bit-exact with the reference's Needleman-Wunsch + traceback verdict. */
int gfe_score_paths(gfe_engine *e, const uint32_t *path_offsets, const uint32_t *path_steps,
                    uint32_t P, int filter, uint32_t *counts);

/* ---- search scoring, incremental ------------------------------------------------------------------
 * The search expands a path by ONE step at a time (reference src/eval.cpp:146-163: newPath = u + v,
 * then evaluatePath(newPath)).  Relative to its parent, a child can change the verdict only of the
 * alignments that contain the appended node (a new occurrence must end at the new step; a newly
 * passing alignment must name the new node) and of the alignments that contain the root node
 * (path[0]; the start-overhang class of the traceback).  The engine keeps an inverted
 * node -> alignment index of the scoring view and scores exactly those.
 *
 * Candidate paths are given as (base path, suffix) pairs so that a batch of siblings / a speculated
 * subtree shares its long common prefixes: path i = bases[item_base[i]] ++ suffix i.  All paths start
 * at the same node (the root).  Per path, counts[4i..4i+3] =
 *   mode 0 (full):   bad, good of evaluatePath(path) over ALL alignments; then bad, good over the
 *                    alignments of the fresh set F (root-node and > 16-step alignments, empty ones)
 *   mode 1 (delta):  bad(path) - bad(parent), good(path) - good(parent) over the alignments OUTSIDE
 *                    F, parent = path minus its last step; then bad, good of the path over F.
 *                    Requires the parent to have at least 16 steps.
 * so  counts(child) = [counts(parent) - F(parent)] + delta(child) + F(child), bit-exact.
 * Only with the id filter (search); returns GFE_ERR_STATE when the universe has no pair dictionary
 * (the caller then uses gfe_score_paths). */
typedef struct gfe_path_batch {
    uint32_t n_bases;
    const uint32_t *base_offsets;     /* n_bases + 1 */
    const uint32_t *base_steps;       /* oriented steps, global dense ids */
    uint32_t P;
    const uint32_t *item_base;        /* P: index of the item's base path */
    const uint32_t *suffix_offsets;   /* P + 1 */
    const uint32_t *suffix_steps;
    const uint8_t *mode;              /* P: 0 full, 1 delta */
} gfe_path_batch;
int gfe_score_paths_incremental(gfe_engine *e, const gfe_path_batch *batch, int32_t *counts /* P x 4 */);

/* Same with every buffer already on the device and no host synchronisation: work is enqueued on
 * `cuda_stream` (a cudaStream_t; NULL = the CUDA default stream).  d_counts (P x 2, uint32) is zeroed
 * and filled by the enqueued work, so a collective (NCCL all-reduce over alignment shards) can
 * be enqueued right behind it on the same stream. */
int gfe_score_paths_device(gfe_engine *e, const uint32_t *d_path_offsets,
                           const uint32_t *d_path_steps, uint32_t P, uint32_t total_path_steps,
                           int filter, uint32_t *d_counts, void *cuda_stream);

/* ---- evalPath scoring: per-alignment traceback scores -------------------------------------------
 * Replaces the unfiltered `evaluatePath(path, ..., false, true)` of evalPath (reference
 * src/eval.cpp:238): for ONE path, T_fw and T_rc (alignPaths(0,-1,-1,...).alignmentScore,
 * src/eval.cpp:92-93) of every alignment, and the bad/good totals of src/eval.cpp:94-98.
 * The universe is set internally to the path's nodes.  t_fw / t_rc (A entries each) may be NULL
 * when only the totals are wanted.  With filter != 0 filtered alignments get GFE_SCORE_FILTERED. */
int gfe_score_pairs(gfe_engine *e, const uint32_t *path_steps, uint32_t n, int filter,
                    int32_t *t_fw, int32_t *t_rc, uint32_t *bad, uint32_t *good);

/* ---- evalPath detail: the gapped alignment strings ------------------------------------------------
 * Replaces get_optimal_alignment's SA/SB (reference src/alignments.cpp:511-554) as printed by
 * PairwisePathAlignment::getAlignment (include/alignments.h:98-122) from evaluatePath
 * (src/eval.cpp:100-105): for ONE path and the alignments [first, first + count), the traceback of
 * alignPaths(path, B) where B is the alignment (use_rc[k] == 0 or use_rc == NULL) or its reverse
 * complement (use_rc[k] != 0; the reference prints the rc alignment unless T_fw > T_rc).
 * Output: one op byte per alignment column, columns in path order:
 *     0  path step == alignment step          1  both present, different
 *     2  path step against a gap              3  alignment step against a gap in the path
 * Pair k's columns are ops[ops_first[k] .. ops_first[k] + ops_len[k]); the caller sizes ops to
 * (offsets[first+count] - offsets[first]) + count * n bytes.  score (may be NULL) receives the
 * traceback score of the chosen direction (== t_fw / t_rc of gfe_score_pairs). */
int gfe_trace_pairs(gfe_engine *e, const uint32_t *path_steps, uint32_t n, uint64_t first,
                    uint32_t count, const uint8_t *use_rc, uint8_t *ops, uint64_t ops_capacity,
                    uint64_t *ops_first, uint32_t *ops_len, int32_t *score);

/* ---- filter mode --------------------------------------------------------------------------------
 * Replaces InAlignments::filterAlignmentByNodelist (reference src/alignments.cpp:459-472 with
 * isContained :96-114 and pathNodesCount :116-135): keep[a] = 1 iff every step of alignment a names
 * a node of `nodes` and the alignment has at least min_nodes steps. */
int gfe_filter_alignments(gfe_engine *e, const uint32_t *nodes, uint32_t n, int32_t min_nodes,
                          uint8_t *keep /* A */);

/* ---- filter mode without device allocation ------------------------------------------------------------
 * No counterpart in the reference (it has no device memory to manage); same result as
 * gfe_filter_alignments.  For hosts where this library must not call the CUDA allocator: processes in
 * which a framework's caching allocator or a memory manager has already claimed nearly all of HBM (a
 * cudaMalloc from here would fail or break their accounting), and services with other streams in
 * flight on the GPU, which cudaMalloc / cudaFree would stall.  Such a caller holds the table in device
 * memory of its own and passes everything as DEVICE pointers on the CURRENT CUDA device:
 *   d_offsets[A + 1], d_steps[total_steps]  the alignment CSR (total_steps == d_offsets[A]);
 *   d_nodes[n]   the node set (ids >= n_nodes are ignored; gfe_filter_alignments rejects them);
 *   d_keep[A]    output, 1 = every step names a node of the set and the row has >= min_nodes steps.
 * The work is enqueued on `cuda_stream` (a cudaStream_t; NULL = default stream) and NOT waited for.
 * No cudaMalloc / cudaFree / allocating routine is called: the filter's two bitmaps live in a fixed
 * 8.125 MiB array inside the library image, resident on every GPU that loads the library until the
 * process ends.  Hence the limits below (GFE_ERR_LIMIT beyond them, text in err).
 * One array per GPU, shared by all calls and never cleared between them except where a call needs
 * zeros: calls that overlap on the same GPU (two streams) race and return garbage — order them.
 * The first call on each device looks up the array's device address and the SM count (it may load the
 * library's kernels onto that GPU); make it at start-up, before any stream capture.
 * err (may be NULL) receives the message of a failed call. */
#define GFE_STATIC_FILTER_MAX_STEPS (1u << 26)   /* 67 108 864 steps in the table */
#define GFE_STATIC_FILTER_MAX_NODES (1u << 20)   /* size of the dense node-id space */
int gfe_filter_alignments_static(const uint32_t *d_offsets, const uint32_t *d_steps, uint64_t A,
                                 uint64_t total_steps, uint32_t n_nodes, const uint32_t *d_nodes,
                                 uint32_t n, int32_t min_nodes, uint8_t *d_keep, void *cuda_stream,
                                 char *err, size_t err_len);

/* ---- evalGFA: per-edge read support -------------------------------------------------------------
 * Replaces InAlignments::buildEdgeGraph (reference src/alignments.cpp:353-403) + the per-edge
 * lookup of evalGFA (src/eval.cpp:42-60).  edges are the graph's L lines in file order as
 * (from, to) oriented steps; weights[i] is the value the reference writes as `RC:i:<w>`
 * (including its 2k-1 rule for self-reverse-complement edges). */
int gfe_edge_support(gfe_engine *e, const uint32_t *edge_from, const uint32_t *edge_to, uint32_t E,
                     uint32_t *weights);

/* Device-side half of the above for sharded alignments: raw traversal counters, to be summed across
 * shards (all-reduce) before gfe_edge_weights_from_counts.  d_counts has 2*E uint32 slots (slot 2i =
 * record i as written, 2i+1 = its reverse complement).  Every directed arc is counted once, at the
 * LOWEST slot that carries it (duplicate records and records that are each other's reverse
 * complement share arcs); the other slots stay 0.  gfe_edge_weights_from_counts resolves that. */
int gfe_edge_traversals_device(gfe_engine *e, const uint32_t *edge_from, const uint32_t *edge_to,
                               uint32_t E, uint32_t *d_counts, void *cuda_stream);
int gfe_edge_weights_from_counts(const uint32_t *edge_from, const uint32_t *edge_to, uint32_t E,
                                 const uint32_t *counts /* 2*E, host */, uint32_t *weights);

/* ---- sharded alignments: the one exchange step ----------------------------------------------------
 * The reference has no multi-process mode; this is the exchange SURVEY.md 8(e) adds for alignment
 * shards (the sum the single evaluatePath loop of src/eval.cpp:14-28 forms in one address space).
 * Sums the n uint32 count slots of `world` ranks (one process per GPU of a box) by reading every
 * peer's record over NVLink peer memory, one small kernel per rank, stream-ordered.
 * peer_bases[q] = rank q's exchange buffer as mapped into THIS process (e.g. torch symmetric memory
 * buffer_ptrs), laid out as  u32 data[2][n_cap]; u32 flags[GFE_PEER_MAX]  and zero-filled once
 * before the first call.  Epoch e = 1, 2, ... (the same on every rank): the caller has this rank's
 * records written to data[e & 1] by earlier work on `cuda_stream`; d_out receives the sums.
 * The wait for the peers is bounded (20 s, GFE_PEER_TIMEOUT_MS): a rank whose peer never reaches the
 * epoch gets every output slot set to GFE_PEER_TIMEOUT instead of hanging. */
#define GFE_PEER_MAX 16
#define GFE_PEER_TIMEOUT 0xFFFFFFFFu
int gfe_peer_sum_u32(const uint64_t *peer_bases, int world, int rank, uint32_t n_cap, uint32_t n,
                     uint32_t epoch, uint32_t *d_out, void *cuda_stream);

/* The sharded form of gfe_score_paths as ONE call (host paths in, summed host counts out): this rank's
 * kernels accumulate into its half of the exchange buffer, the rank's status word rides along as one
 * more slot, gfe_peer_sum_u32 sums all shards, one device->host copy returns the result.  Every rank
 * of the box calls it with the same batch, in the same order (the search is replayed in lock step on
 * every rank, reference src/eval.cpp:134-189); if any rank overflowed a worklist, all ranks repeat
 * the batch together.  gfe_exchange wraps the peer-mapped buffer (layout above; n_cap >= 2P + 1). */
typedef struct gfe_exchange gfe_exchange;
int gfe_exchange_create(const uint64_t *peer_bases, int world, int rank, uint32_t n_cap,
                        gfe_exchange **out);
void gfe_exchange_destroy(gfe_exchange *x);
/* Candidate-batch sharding for tables too small to fill N GPUs (SURVEY.md 8e "alternative when A is
 * small"; north_star: "sharding DFS subtrees"): every rank holds the WHOLE alignment table, the
 * speculation subtrees of a batch (its base paths) are dealt round-robin over the ranks, each rank scores
 * its share with gfe_score_paths_incremental_sharded, and the sum of the per-candidate records — zeros
 * from the ranks that did not score a candidate — is the all-gather.  gfe_score_paths_sharded then
 * needs no exchange (every rank scores the batch itself). */
int gfe_exchange_set_batch_sharding(gfe_exchange *x, int on);
int gfe_score_paths_sharded(gfe_engine *e, gfe_exchange *x, const uint32_t *path_offsets,
                            const uint32_t *path_steps, uint32_t P, int filter, uint32_t *counts);

/* ---- best-first search driver -------------------------------------------------------------------
 * Replaces dijkstra() (reference src/eval.cpp:110-193) on packed inputs: the bidirected adjacency
 * of InSequences::buildEdgeGraph (reference gfalibs/src/gfa.cpp:1076-1097) as CSR over nodes
 * (adj_offsets[n_nodes+1]; per entry orientation0 / target node / orientation1, in L-line order),
 * the NodeTable budgets (reference include/nodetable.h), source, destination, and the
 * -m / --min-nodes / --return-all-paths options.  Rows are delivered in emission order through
 * `emit`; the search itself (Fibonacci-heap order, `alt` keys, hamiltonian flag) follows the
 * reference exactly, and every evaluatePath call is answered by the CUDA kernels. */
typedef struct gfe_search_row {
    uint64_t path_counter;
    uint32_t bad, good;
    int32_t alt;
    uint32_t size, uniques;
    int hamiltonian;
    const uint32_t *steps;      /* `size` oriented steps, valid during the callback only */
} gfe_search_row;
typedef void (*gfe_row_fn)(const gfe_search_row *row, void *user);

typedef struct gfe_graph {
    uint32_t n_nodes;
    const uint32_t *adj_offsets;   /* n_nodes + 1 */
    const uint32_t *adj_target;    /* target node of each adjacency entry */
    const uint8_t *adj_or0;        /* 0 '+', 1 '-' : orientation the source node is left in */
    const uint8_t *adj_or1;        /* orientation the target node is entered in */
} gfe_graph;

typedef struct gfe_search_params {
    uint32_t source, destination;      /* dense node ids */
    const uint32_t *budget_nodes;      /* NodeTable records: node, count (count >= 1) */
    const uint32_t *budget_counts;
    uint32_t n_budget;
    uint32_t node_count;               /* NodeTable::nodeCount incl. the two add() calls */
    uint32_t max_steps, min_nodes;
    int return_all_paths;
    uint32_t speculate;                /* max candidate paths per kernel batch (0 = default) */
} gfe_search_params;

typedef struct gfe_search_stats {
    uint64_t extractions, evaluations, kernel_batches, paths_scored, rows;
    int reached_max_steps;
    double seconds_scoring, seconds_total;
} gfe_search_stats;

int gfe_search(gfe_engine *e, const gfe_graph *g, const gfe_search_params *p, gfe_row_fn emit,
               void *user, gfe_search_stats *stats);

/* gfe_search over alignment shards: every rank calls it with its own engine (holding its shard) and the
 * box's exchange buffer; evaluatePath is answered by gfe_score_paths_sharded, so all ranks emit the
 * same rows. */
/* gfe_score_paths_incremental over alignment shards: deltas and fresh-set counts are sums over
 * alignments, so every rank scores the same batch against its shard (its own inverted index) and the
 * raw records are summed by the peer kernel before the per-path composition.  n_cap of the exchange
 * buffer must cover 4 slots per candidate plus padding (2^18 slots serve batches of 4096 paths). */
int gfe_score_paths_incremental_sharded(gfe_engine *e, gfe_exchange *x, const gfe_path_batch *batch,
                                        int32_t *counts /* P x 4 */);

int gfe_search_sharded(gfe_engine *e, gfe_exchange *x, const gfe_graph *g, const gfe_search_params *p,
                       gfe_row_fn emit, void *user, gfe_search_stats *stats);

/* The same driver with the evaluatePath calls answered by a caller-supplied backend: `score`
 * receives a batch of P candidate paths (CSR, oriented steps) and must fill counts[2p] = bad,
 * counts[2p+1] = good exactly as gfe_score_paths(filter = 1) would; non-zero return aborts the
 * search with that status.  This is the seam for alignment tables sharded over several engines /
 * GPUs (score every shard, sum the counts: one all-reduce) — the emission order stays that of
 * the single canonical heap replay (reference src/eval.cpp:134-189). */
typedef int (*gfe_score_fn)(void *score_user, const uint32_t *path_offsets,
                            const uint32_t *path_steps, uint32_t P, uint32_t *counts);
int gfe_search_with_scorer(gfe_score_fn score, void *score_user, const gfe_graph *g,
                           const gfe_search_params *p, gfe_row_fn emit, void *user,
                           gfe_search_stats *stats);

/* ---- text packers (host only, no GPU needed) ------------------------------------------------------
 * GFA1 / GAF text -> the packed arrays the entry points above take.  Replaces, for this path,
 * gfalibs' readGFA (reference gfalibs/src/input-gfa.cpp:497-641: S and L records, name -> id),
 * InSequences::buildEdgeGraph (gfalibs/src/gfa.cpp:1076-1097), InAlignments::load +
 * InAlignment::GAFpathToPath (reference src/alignments.cpp:143-235, 75-94).  Nodes get dense ids in
 * order of first appearance (S records, then names first seen on L records); a walk name the
 * graph does not know becomes node 0, as the reference's headersToIds[...] default-inserts 0
 * (src/alignments.cpp:86).  The GAF is parsed by `threads` host threads (<= 0: all cores). */
typedef struct gfe_text gfe_text;
typedef struct gfe_text_view {
    uint32_t n_nodes, n_links;
    const uint32_t *edge_from, *edge_to;   /* L records in file order, oriented steps */
    gfe_graph graph;                       /* adjacency for gfe_search */
    uint64_t A;                            /* alignments (0 when no GAF was given) */
    const uint32_t *offsets, *steps;       /* CSR for gfe_set_alignments */
} gfe_text_view;
int gfe_text_open(const char *gfa_path, const char *gaf_path /* may be NULL */, int threads,
                  gfe_text **out, char *err, size_t err_len);
void gfe_text_close(gfe_text *t);
int gfe_text_get(const gfe_text *t, gfe_text_view *out);
const char *gfe_text_node_name(const gfe_text *t, uint32_t node);
int64_t gfe_text_node_id(const gfe_text *t, const char *name);   /* -1: unknown */
const char *gfe_text_qname(const gfe_text *t, uint64_t a, uint32_t *len);   /* not NUL-terminated */

/* ---- introspection (bench / tests) --------------------------------------------------------------*/
typedef struct gfe_counters {
    uint64_t kernel_launches;     /* kernels of this library launched since gfe_create */
    uint64_t exact_pairs;         /* pairs that needed the exact DP (ambiguous class) */
    uint64_t h2d_bytes, d2h_bytes;
    uint64_t slow_pairs;          // pairs the bitset kernel handed to the second-look kernel
} gfe_counters;
int gfe_get_counters(const gfe_engine *e, gfe_counters *out);
/* Multiplicity compression of the scoring view (on by default; GFE_DEDUP=0 or enabled = 0 turns it off;
 * takes effect at the next gfe_set_universe).  The verdict of evaluatePath's inner loop
 * (reference src/eval.cpp:80-99) depends only on the alignment's oriented node list and is the same
 * for an alignment and its reverse complement, so alignments with equal walks are scored once and
 * counted with their multiplicity — counts stay bit-exact.  stats: rows of 1..16 steps, distinct
 * walks among them, rows of the compressed view (one per set bit of each multiplicity), and whether
 * full scans (gfe_score_paths*) currently read it (kept only when it saves a quarter of the rows). */
int gfe_set_dedup(gfe_engine *e, int enabled);
int gfe_get_dedup_stats(const gfe_engine *e, uint64_t *rows, uint64_t *distinct, uint64_t *view_rows, int *in_use);
/* After the *_device entry points (which do not synchronise): waits for the device and reports the
 * flags of ALL batches enqueued since the previous report (they are sticky across batches and are
 * cleared by the report) — bit0: a candidate path named a node outside the universe, bit1: a
 * worklist overflowed (results of that batch are incomplete; re-run it through gfe_score_paths,
 * which grows the list and retries by itself), bit2: an alignment of more than GFE_MAX_STEPS steps
 * passed the id filter of a path (the reference cannot score it; gfe_score_paths: GFE_ERR_LIMIT). */
int gfe_device_status(gfe_engine *e, uint32_t *flags);
/* The same flags written to a device word, stream-ordered behind the *_device calls on `cuda_stream`
 * (no synchronisation): sharded callers add the word to the exchanged record so that every rank
 * learns of any rank's overflow and repeats the batch through the host call together. */
int gfe_device_status_to(gfe_engine *e, uint32_t *d_flags, void *cuda_stream);
/* Milliseconds the last scoring / edge-support call spent in its dominant kernel (CUDA events on
 * the engine's stream); 0 if timing was not enabled with gfe_set_timing(e, 1). */
int gfe_set_timing(gfe_engine *e, int enabled);
float gfe_last_kernel_ms(const gfe_engine *e);
/* gfe_score_pairs: milliseconds of the exact-DP kernels behind the classify pass (0 without timing). */
float gfe_last_aux_kernel_ms(const gfe_engine *e);
/* Synchronises the device, sums the CUDA-event durations of every dominant-kernel launch timed
 * since the previous call (any stream the launches were enqueued on) and resets the record. */
int gfe_kernel_time_stats(gfe_engine *e, double *total_ms, uint32_t *launches);

#ifdef __cplusplus
}
#endif
#endif /* GFALIGN_B200_H */
