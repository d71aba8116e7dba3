// oracle/ref_shim.cpp — TEST INFRASTRUCTURE ONLY.
// Our own thin C-ABI wrapper around the UNMODIFIED reference objects (compiled by oracle/Makefile
// from the sources where they lie under /root/reference; nothing is copied).  It lets the tests
// call the reference's own pair arithmetic — alignPaths(), src/alignments.cpp:556-561, and
// Path::reverseComplement(), include/alignments.h:64-70 — in-process, pair by pair.
// The globals below are the ones the reference's main.cpp defines (src/main.cpp:49-61) and the
// other translation units reference as extern.
#include <stdlib.h>
#include <unistd.h>
#include <string>
#include <thread>
#include <mutex>
#include <vector>
#include <queue>
#include <stack>
#include <functional>
#include <chrono>
#include <iostream>
#include <fstream>
#include <sstream>

#include <parallel-hashmap/phmap.h>

#include "log.h"
#include "global.h"
#include "uid-generator.h"
#include "bed.h"
#include "struct.h"
#include "functions.h"
#include "threadpool.h"
#include "gfa-lines.h"
#include "gfa.h"
#include "sak.h"
#include "stream-obj.h"
#include "input-agp.h"
#include "input-filters.h"
#include "input-gfa.h"
#include "nodetable.h"
#include "alignments.h"
#include "input-gfalign.h"
#include "eval.h"

std::string version = "ref-shim";
std::chrono::high_resolution_clock::time_point start = std::chrono::high_resolution_clock::now();
int verbose_flag;
Log lg;
std::vector<Log> logs;
int tabular_flag;
int maxThreads = 0;
std::mutex mtx;
ThreadPool<std::function<bool()>> threadPool;
UserInputGfalign userInput;

static Path toPath(const uint32_t *s, uint32_t n) {
    Path p;
    for (uint32_t i = 0; i < n; ++i) p.push_back((int32_t)(s[i] >> 1), (s[i] & 1) ? '-' : '+');
    return p;
}

typedef int dp_matrix_t[MAX_N][MAX_N];

extern "C" {

// T_fw and T_rc of one (path, alignment) pair exactly as evaluatePath computes them
// (src/eval.cpp:92-93): the dp matrix is zeroed once and shared by both calls.
int ref_pair_scores(const uint32_t *A, uint32_t n, const uint32_t *B, uint32_t m, int32_t *t_fw,
                    int32_t *t_rc) {
    if (n >= MAX_N || m >= MAX_N) return 1;
    dp_matrix_t *dp = (dp_matrix_t *)calloc(1, sizeof(dp_matrix_t));
    Path a = toPath(A, n), b = toPath(B, m);
    PairwisePathAlignment fw = alignPaths(0, -1, -1, a, b, *dp);
    PairwisePathAlignment rc = alignPaths(0, -1, -1, a, b.reverseComplement(), *dp);
    *t_fw = fw.alignmentScore;
    *t_rc = rc.alignmentScore;
    free(dp);
    return 0;
}

// Many alignments (CSR) against one path with ONE shared dp matrix, as in evaluatePath's loop.
int ref_path_scores(const uint32_t *A, uint32_t n, const uint32_t *offsets, const uint32_t *steps,
                    uint64_t nAln, int32_t *t_fw, int32_t *t_rc) {
    if (n >= MAX_N) return 1;
    dp_matrix_t *dp = (dp_matrix_t *)calloc(1, sizeof(dp_matrix_t));
    Path a = toPath(A, n);
    for (uint64_t k = 0; k < nAln; ++k) {
        uint32_t m = offsets[k + 1] - offsets[k];
        if (m >= MAX_N) { free(dp); return 1; }
        Path b = toPath(steps + offsets[k], m);
        t_fw[k] = alignPaths(0, -1, -1, a, b, *dp).alignmentScore;
        t_rc[k] = alignPaths(0, -1, -1, a, b.reverseComplement(), *dp).alignmentScore;
    }
    free(dp);
    return 0;
}

}  // extern "C"
