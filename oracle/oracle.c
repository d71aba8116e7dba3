/*
 * oracle/oracle.c — CPU restatement of gfalign's scoring hot path.  TEST INFRASTRUCTURE ONLY
 * (see oracle.h for who may load it and for the parity-pinning status).
 *
 * Written from the behaviour of the reference, not from its text: flat u32 arrays, one DP matrix
 * shared by all pairs of one evaluatePath call exactly as the reference shares `dp`
 * (src/eval.cpp:79), same loop bounds, same tie-break order.
 */
#include "oracle.h"
#include <stdlib.h>
#include <string.h>
#include <pthread.h>

#define GAP_STEP 0xFFFFFFFFu

typedef struct {
    int32_t *cell;   /* (rows) x (width) ints, zero-initialised once per evaluate call */
    uint32_t width;
} dp_t;

static inline int32_t *dp_at(const dp_t *d, uint32_t i, uint32_t j) {
    return d->cell + (size_t)i * d->width + j;
}

static inline int32_t max3(int32_t a, int32_t b, int32_t c) {
    int32_t t = b > c ? b : c;
    return a > t ? a : t;
}

/* needleman_wunsch(n, m, dp, 0, -1, -1, A, B) — src/alignments.cpp:499-509.
 * Row 0 is initialised for columns 0..n (loop bound n, not m); column 0 is never written. */
static void fill_dp(const dp_t *d, const uint32_t *A, uint32_t n, const uint32_t *B, uint32_t m) {
    for (uint32_t i = 0; i <= n; ++i) *dp_at(d, 0, i) = -(int32_t)i;
    for (uint32_t i = 1; i <= n; ++i)
        for (uint32_t j = 1; j <= m; ++j) {
            int32_t S = (A[i - 1] == B[j - 1]) ? 0 : -1;
            *dp_at(d, i, j) = max3(*dp_at(d, i - 1, j - 1) + S,
                                   *dp_at(d, i - 1, j) + (j < m ? -1 : 0),
                                   *dp_at(d, i, j - 1) - 1);
        }
}

/* get_optimal_alignment — src/alignments.cpp:511-554.  Returns the traceback score; when SA/SB
 * are given also records the gapped sequences (in traceback order, caller reverses). */
static int32_t traceback(const dp_t *d, const uint32_t *A, uint32_t n, const uint32_t *B,
                         uint32_t m, uint32_t *SA, uint32_t *SB, uint32_t *len) {
    int32_t score = 0, sb_len = 0;
    uint32_t ii = n, jj = m, k = 0;
    while (ii != 0 || jj != 0) {
        uint32_t a = GAP_STEP, b = GAP_STEP;
        if (ii == 0) {
            b = B[--jj];
        } else if (jj == 0) {
            a = A[--ii];
        } else {
            int32_t S = (A[ii - 1] == B[jj - 1]) ? 0 : -1;
            if (*dp_at(d, ii, jj) == *dp_at(d, ii - 1, jj - 1) + S) {
                a = A[--ii];
                b = B[--jj];
                ++sb_len;
                score += S;
            } else if (*dp_at(d, ii - 1, jj) >= *dp_at(d, ii, jj - 1)) {
                a = A[--ii];
                if (sb_len > 0) score -= 1;
            } else {
                b = B[--jj];
                ++sb_len;
                score -= 1;
            }
        }
        if (SA) { SA[k] = a; SB[k] = b; }
        ++k;
    }
    if (len) *len = k;
    return score;
}

static int dp_alloc(dp_t *d, uint32_t n, uint32_t max_m) {
    uint32_t w = (max_m > n ? max_m : n) + 1;
    d->width = w;
    d->cell = (int32_t *)calloc((size_t)(n + 1) * w, sizeof(int32_t));
    return d->cell != NULL;
}

int32_t orc_pair_score(const uint32_t *A, uint32_t n, const uint32_t *B, uint32_t m) {
    dp_t d;
    if (!dp_alloc(&d, n, m)) return INT32_MIN;
    fill_dp(&d, A, n, B, m);
    int32_t t = traceback(&d, A, n, B, m, NULL, NULL, NULL);
    free(d.cell);
    return t;
}

uint32_t orc_pair_alignment(const uint32_t *A, uint32_t n, const uint32_t *B, uint32_t m,
                            uint32_t *SA, uint32_t *SB, int32_t *score) {
    dp_t d;
    uint32_t len = 0;
    if (!dp_alloc(&d, n, m)) return 0;
    fill_dp(&d, A, n, B, m);
    int32_t t = traceback(&d, A, n, B, m, SA, SB, &len);
    free(d.cell);
    for (uint32_t i = 0; i < len / 2; ++i) { /* SA.reverse(); SB.reverse(); */
        uint32_t x = SA[i]; SA[i] = SA[len - 1 - i]; SA[len - 1 - i] = x;
        x = SB[i]; SB[i] = SB[len - 1 - i]; SB[len - 1 - i] = x;
    }
    if (score) *score = t;
    return len;
}

/* Path::reverseComplement — include/alignments.h:64-70 */
static void revcomp(const uint32_t *B, uint32_t m, uint32_t *out) {
    for (uint32_t j = 0; j < m; ++j) out[j] = B[m - 1 - j] ^ 1u;
}

static uint32_t max_len(const uint32_t *offsets, uint64_t lo, uint64_t hi) {
    uint32_t mx = 0;
    for (uint64_t a = lo; a < hi; ++a) {
        uint32_t m = offsets[a + 1] - offsets[a];
        if (m > mx) mx = m;
    }
    return mx;
}

/* One evaluatePath over alignments [lo, hi) — src/eval.cpp:76-99. */
static void evaluate_range(const uint32_t *path, uint32_t n, const uint32_t *offsets,
                           const uint32_t *steps, uint64_t lo, uint64_t hi, int filter,
                           uint32_t n_nodes_hint, uint32_t *bad, uint32_t *good,
                           uint32_t *unaligned, int32_t *t_fw, int32_t *t_rc) {
    (void)n_nodes_hint;
    uint32_t mx = max_len(offsets, lo, hi);
    dp_t d;
    uint32_t *rc = (uint32_t *)malloc(sizeof(uint32_t) * (mx ? mx : 1));
    /* uIds set of the path (ids only, orientation ignored) as a sorted array + bsearch */
    uint32_t *ids = (uint32_t *)malloc(sizeof(uint32_t) * (n ? n : 1));
    uint32_t max_id = 0;
    for (uint32_t i = 0; i < n; ++i) { ids[i] = path[i] >> 1; if (ids[i] > max_id) max_id = ids[i]; }
    uint8_t *present = (uint8_t *)calloc((size_t)max_id + 1, 1);
    for (uint32_t i = 0; i < n; ++i) present[ids[i]] = 1;
    if (!dp_alloc(&d, n, mx)) abort();
    uint32_t b = 0, g = 0, u = 0;
    for (uint64_t a = lo; a < hi; ++a) {
        const uint32_t *B = steps + offsets[a];
        uint32_t m = offsets[a + 1] - offsets[a];
        if (filter) {
            int skip = 0;
            for (uint32_t j = 0; j < m; ++j) {
                uint32_t id = B[j] >> 1;
                if (id > max_id || !present[id]) { skip = 1; ++u; }
            }
            if (skip) {
                if (t_fw) t_fw[a] = ORC_FILTERED;
                if (t_rc) t_rc[a] = ORC_FILTERED;
                continue;
            }
        }
        fill_dp(&d, path, n, B, m);
        int32_t fw = traceback(&d, path, n, B, m, NULL, NULL, NULL);
        revcomp(B, m, rc);
        fill_dp(&d, path, n, rc, m);
        int32_t rv = traceback(&d, path, n, rc, m, NULL, NULL, NULL);
        int32_t best = fw > rv ? fw : rv;
        if (best < 0) ++b; else ++g;
        if (t_fw) t_fw[a] = fw;
        if (t_rc) t_rc[a] = rv;
    }
    *bad = b; *good = g; *unaligned = u;
    free(d.cell); free(rc); free(ids); free(present);
}

void orc_evaluate_path(const uint32_t *path, uint32_t n, const uint32_t *offsets,
                       const uint32_t *steps, uint64_t A, int filter, uint32_t *bad,
                       uint32_t *good, uint32_t *unaligned, int32_t *t_fw, int32_t *t_rc) {
    evaluate_range(path, n, offsets, steps, 0, A, filter, 0, bad, good, unaligned, t_fw, t_rc);
}

typedef struct {
    const uint32_t *path_offsets, *path_steps, *offsets, *steps;
    uint32_t P; uint64_t lo, hi; int filter;
    uint32_t *counts;   /* private P x 2 */
} eval_job_t;

static void *eval_worker(void *arg) {
    eval_job_t *j = (eval_job_t *)arg;
    for (uint32_t p = 0; p < j->P; ++p) {
        uint32_t u;
        evaluate_range(j->path_steps + j->path_offsets[p],
                       j->path_offsets[p + 1] - j->path_offsets[p], j->offsets, j->steps, j->lo,
                       j->hi, j->filter, 0, &j->counts[2 * p], &j->counts[2 * p + 1], &u, NULL,
                       NULL);
    }
    return NULL;
}

/* `threads` pthreads, each owning a contiguous alignment range and scoring all P paths on it. */
void orc_evaluate_paths(const uint32_t *path_offsets, const uint32_t *path_steps, uint32_t P,
                        const uint32_t *offsets, const uint32_t *steps, uint64_t A, int filter,
                        int threads, uint32_t *counts) {
    if (threads < 1) threads = 1;
    memset(counts, 0, sizeof(uint32_t) * 2 * (size_t)P);
    eval_job_t *jobs = (eval_job_t *)calloc((size_t)threads, sizeof(eval_job_t));
    pthread_t *tid = (pthread_t *)calloc((size_t)threads, sizeof(pthread_t));
    for (int c = 0; c < threads; ++c) {
        eval_job_t j = {path_offsets, path_steps, offsets, steps, P, A * (uint64_t)c / (uint64_t)threads,
                        A * (uint64_t)(c + 1) / (uint64_t)threads, filter,
                        (uint32_t *)calloc((size_t)P * 2 + 1, sizeof(uint32_t))};
        jobs[c] = j;
        if (threads == 1) eval_worker(&jobs[c]);
        else pthread_create(&tid[c], NULL, eval_worker, &jobs[c]);
    }
    for (int c = 0; c < threads; ++c) {
        if (threads > 1) pthread_join(tid[c], NULL);
        for (uint32_t k = 0; k < 2 * P; ++k) counts[k] += jobs[c].counts[k];
        free(jobs[c].counts);
    }
    free(jobs); free(tid);
}

/* ---- edge support: InAlignments::buildEdgeGraph, src/alignments.cpp:353-403 ------------------ */

typedef struct { uint8_t or0, or1; uint32_t id, weight; } adj_edge_t;     /* Edge, struct.h:83-92 */
typedef struct { adj_edge_t *e; uint32_t n, cap; } adj_list_t;

static adj_edge_t *adj_find(adj_list_t *l, uint8_t or0, uint32_t id, uint8_t or1) {
    for (uint32_t i = 0; i < l->n; ++i)   /* Edge::operator== ignores weight, struct.cpp:63-65 */
        if (l->e[i].or0 == or0 && l->e[i].or1 == or1 && l->e[i].id == id) return &l->e[i];
    return NULL;
}

static void adj_push(adj_list_t *l, uint8_t or0, uint32_t id, uint8_t or1, uint32_t w) {
    if (l->n == l->cap) {
        l->cap = l->cap ? l->cap * 2 : 4;
        l->e = (adj_edge_t *)realloc(l->e, sizeof(adj_edge_t) * l->cap);
    }
    adj_edge_t x = {or0, or1, id, w};
    l->e[l->n++] = x;
}

void orc_edge_support(const uint32_t *offsets, const uint32_t *steps, uint64_t A, uint32_t n_nodes,
                      const uint32_t *edge_from, const uint32_t *edge_to, uint32_t E,
                      uint32_t *weights) {
    adj_list_t *adj = (adj_list_t *)calloc(n_nodes ? n_nodes : 1, sizeof(adj_list_t));
    for (uint64_t a = 0; a < A; ++a) {
        uint32_t m = offsets[a + 1] - offsets[a];
        const uint32_t *B = steps + offsets[a];
        for (uint32_t j = 0; j + 1 < m; ++j) {           /* GAFpathToEdges, :409-445 */
            uint32_t id1 = B[j] >> 1, id2 = B[j + 1] >> 1;
            uint8_t or1 = B[j] & 1, or2 = B[j + 1] & 1;  /* 0 = '+', 1 = '-' */
            adj_edge_t *it = adj_find(&adj[id1], or1, id2, or2);
            if (!it) {
                adj_push(&adj[id1], or1, id2, or2, 1);
                if (!adj_find(&adj[id2], or2 ^ 1, id1, or1 ^ 1))
                    adj_push(&adj[id2], or2 ^ 1, id1, or1 ^ 1, 1);
            } else {
                it->weight++;
                adj_edge_t *it2 = adj_find(&adj[id2], or2 ^ 1, id1, or1 ^ 1);
                it2->weight++;
            }
        }
    }
    for (uint32_t e = 0; e < E; ++e) {                    /* evalGFA, src/eval.cpp:42-60 */
        adj_edge_t *it = adj_find(&adj[edge_from[e] >> 1], edge_from[e] & 1, edge_to[e] >> 1,
                                  edge_to[e] & 1);
        weights[e] = it ? it->weight : 0;
    }
    for (uint32_t i = 0; i < n_nodes; ++i) free(adj[i].e);
    free(adj);
}
