// tests/native/engine_stub.cpp — TEST INFRASTRUCTURE.  The engine entry points the search driver
// (gfalign_b200/csrc/search.cpp, compiled unchanged next to this file) calls, answered by the CPU oracle
// (oracle/oracle.c) instead of the CUDA kernels, so that the driver's host logic — incremental
// composition main/fresh/delta, effective-key speculation, incremental path materialisation, heap
// replay — runs in the CPU test suite against the reference's golden rows.  Never part of the product.
#include <cstdint>
#include <cstring>
#include <map>
#include <vector>

#include "../../include/gfalign_b200.h"
#include "../../oracle/oracle.h"

struct gfe_engine { int unused; };
struct gfe_exchange { int unused; };

namespace {
std::vector<uint32_t> g_off, g_steps;              // the alignment table
struct Split { std::vector<uint32_t> f_off, f_steps, m_off, m_steps; };
std::map<uint32_t, Split> g_split;                 // root node -> fresh set F / the rest
uint64_t g_delta_items = 0, g_full_items = 0;

const Split &split_for(uint32_t root) {
    auto it = g_split.find(root);
    if (it != g_split.end()) return it->second;
    Split s;
    s.f_off.push_back(0);
    s.m_off.push_back(0);
    const size_t A = g_off.size() - 1;
    for (size_t a = 0; a < A; ++a) {
        const uint32_t o0 = g_off[a], m = g_off[a + 1] - o0;
        bool fresh = m == 0 || m > 16;             // the engine's F: empty rows, rows of more than 16 steps,
        for (uint32_t j = 0; j < m && !fresh; ++j) fresh = (g_steps[o0 + j] >> 1) == root;   // rows with the root node
        std::vector<uint32_t> &off = fresh ? s.f_off : s.m_off, &st = fresh ? s.f_steps : s.m_steps;
        st.insert(st.end(), g_steps.begin() + o0, g_steps.begin() + o0 + m);
        off.push_back((uint32_t)st.size());
    }
    return g_split.emplace(root, std::move(s)).first->second;
}

void eval(const std::vector<uint32_t> &off, const std::vector<uint32_t> &st, const uint32_t *path, uint32_t n,
          uint32_t *bad, uint32_t *good) {
    uint32_t un = 0;
    static const uint32_t none = 0;
    orc_evaluate_path(path, n, off.data(), st.empty() ? &none : st.data(), off.size() - 1, 1, bad, good, &un, nullptr, nullptr);
}
}  // namespace

extern "C" {

void stub_set_alignments(const uint32_t *off, const uint32_t *steps, uint64_t A) {
    g_off.assign(off, off + A + 1);
    g_steps.assign(steps, steps + off[A]);
    g_split.clear();
    g_delta_items = g_full_items = 0;
}
void stub_counters(uint64_t *delta_items, uint64_t *full_items) {
    *delta_items = g_delta_items;
    *full_items = g_full_items;
}

int gfe_set_universe(gfe_engine *, const uint32_t *, uint32_t) { return GFE_OK; }

int gfe_score_paths(gfe_engine *, const uint32_t *po, const uint32_t *ps, uint32_t P, int filter, uint32_t *counts) {
    orc_evaluate_paths(po, ps, P, g_off.data(), g_steps.data(), g_off.size() - 1, filter, 1, counts);
    return GFE_OK;
}
int gfe_score_paths_sharded(gfe_engine *e, gfe_exchange *, const uint32_t *po, const uint32_t *ps, uint32_t P, int filter,
                            uint32_t *counts) {
    return gfe_score_paths(e, po, ps, P, filter, counts);
}

int gfe_score_paths_incremental(gfe_engine *, const gfe_path_batch *b, int32_t *c) {
    std::vector<uint32_t> path;
    for (uint32_t i = 0; i < b->P; ++i) {
        const uint32_t kb = b->item_base[i];
        path.assign(b->base_steps + b->base_offsets[kb], b->base_steps + b->base_offsets[kb + 1]);
        path.insert(path.end(), b->suffix_steps + b->suffix_offsets[i], b->suffix_steps + b->suffix_offsets[i + 1]);
        if ((b->base_steps[0] >> 1) != (path[0] >> 1)) return GFE_ERR_ARG;     // every candidate starts at the root
        const Split &s = split_for(path[0] >> 1);
        const uint32_t n = (uint32_t)path.size();
        uint32_t fb, fg, mb, mg;
        eval(s.f_off, s.f_steps, path.data(), n, &fb, &fg);
        eval(s.m_off, s.m_steps, path.data(), n, &mb, &mg);
        if (b->mode[i]) {
            if (n < 17) return GFE_ERR_ARG;        // the ABI's contract: delta needs a parent of >= 16 steps
            uint32_t pb, pg;
            eval(s.m_off, s.m_steps, path.data(), n - 1, &pb, &pg);
            c[4 * i] = (int32_t)mb - (int32_t)pb;
            c[4 * i + 1] = (int32_t)mg - (int32_t)pg;
            ++g_delta_items;
        } else {
            c[4 * i] = (int32_t)(mb + fb);
            c[4 * i + 1] = (int32_t)(mg + fg);
            ++g_full_items;
        }
        c[4 * i + 2] = (int32_t)fb;
        c[4 * i + 3] = (int32_t)fg;
    }
    return GFE_OK;
}
int gfe_score_paths_incremental_sharded(gfe_engine *e, gfe_exchange *, const gfe_path_batch *b, int32_t *c) {
    return gfe_score_paths_incremental(e, b, c);
}

}  // extern "C"

// library-internal vote of gfe_search_sharded (engine.cu): the stub is a single rank that has the path
int gfe_internal_incremental_consensus(gfe_engine *, gfe_exchange *, int *all_have) {
    *all_have = 1;
    return GFE_OK;
}
