// gfalign_b200/csrc/search.cpp — gfe_search: the best-first path search of `gfalign search`
// (reference dijkstra(), src/eval.cpp:110-193) replayed on the host with every evaluatePath call
// (src/eval.cpp:162) answered by the CUDA kernels.
//
// What must match the reference exactly, and how it is kept:
//   * extraction order — the reference's priority queue is a Fibonacci heap whose tie-breaks are
//     structural (gfalibs/include/fibonacci-heap.h:55-126, :219-271).  OrderHeap below performs
//     the same sequence of structural operations on index-linked nodes.
//   * successor admission — orientation continuity (src/eval.cpp:137) and the remaining-visit
//     budget of the EXTRACTED path's own NodeTable (:142-143); budgets drop only for
//     non-destination pushes (:165-167); the source placed at start is not a push.
//   * key — alt = bad - good - uniques (:163); "better" rule and row emission (:176-183);
//     hamiltonian flag (include/nodetable.h:56-67).
// What is free: evaluatePath(path) is a pure function of the path, so the driver may score paths
// it has not been asked about yet.  On a miss it sends the children of the extracted path plus a
// breadth-first subtree below them in ONE kernel batch and caches the counts on the path tree.
#include <chrono>
#include <climits>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <deque>
#include <queue>
#include <utility>
#include <vector>

#include "../../include/gfalign_b200.h"

namespace {

// ---- Fibonacci heap with the reference's structural behaviour ----------------------------------
class OrderHeap {
    struct Node {
        int key, degree;
        int parent, child, left, right;
        uint32_t value;
    };
    std::vector<Node> N;
    std::vector<int> freelist, table;
    int min_ = -1;
    uint32_t num_ = 0;

    int alloc() {
        if (!freelist.empty()) {
            int i = freelist.back();
            freelist.pop_back();
            return i;
        }
        N.push_back(Node());
        return (int)N.size() - 1;
    }
    // fibonacci-heap.h:145-165 (_existingToRoot): left of the current minimum, strict-less update
    void to_root(int x) {
        N[x].parent = -1;
        if (min_ != -1) {
            const int ml = N[min_].left;
            N[min_].left = x;
            N[x].right = min_;
            N[x].left = ml;
            N[ml].right = x;
            if (N[min_].key > N[x].key) min_ = x;
        } else {
            min_ = x;
            N[x].left = N[x].right = x;
        }
    }
    // :166-180 (_removeNodeFromRoot)
    void unlink(int x) {
        if (N[x].right != x) {
            N[N[x].right].left = N[x].left;
            N[N[x].left].right = N[x].right;
        }
        const int p = N[x].parent;
        if (p != -1) {
            N[p].child = (N[p].degree == 1) ? -1 : N[x].right;
            N[p].degree--;
        }
    }
    // :184-201 (_addChild): appended left of the first child
    void add_child(int parent, int c) {
        if (N[parent].degree == 0) {
            N[parent].child = c;
            N[c].left = N[c].right = c;
        } else {
            const int first = N[parent].child, last = N[first].left;
            N[first].left = c;
            N[c].right = first;
            N[c].left = last;
            N[last].right = c;
        }
        N[c].parent = parent;
        N[parent].degree++;
    }
    // :219-271 (_consolidate)
    void consolidate() {
        if (num_ <= 1) return;
        table.clear();
        int roots = 0, it = min_;
        do {
            ++roots;
            it = N[it].right;
        } while (it != min_);
        int curr = min_;
        for (int c = 0; c < roots; ++c) {
            int x = curr;
            curr = N[curr].right;
            int deg = N[x].degree;
            for (;;) {
                while (deg >= (int)table.size()) table.push_back(-1);
                if (table[deg] == -1) {
                    table[deg] = x;
                    break;
                }
                int y = table[deg];
                if (N[x].key > N[y].key) std::swap(x, y);   // ties: the root in hand stays parent
                if (y == x) break;
                unlink(y);
                add_child(x, y);
                table[deg] = -1;
                ++deg;
            }
        }
        min_ = -1;
        for (size_t i = 0; i < table.size(); ++i)
            if (table[i] != -1) to_root(table[i]);
    }

public:
    uint32_t size() const { return num_; }
    // :55-86
    void insert(uint32_t value, int key) {
        const int x = alloc();
        N[x].key = key;
        N[x].degree = 0;
        N[x].parent = N[x].child = -1;
        N[x].left = N[x].right = x;
        N[x].value = value;
        if (min_ != -1) {
            const int ml = N[min_].left;
            N[min_].left = x;
            N[x].right = min_;
            N[x].left = ml;
            N[ml].right = x;
        }
        if (min_ == -1 || N[min_].key > key) min_ = x;
        ++num_;
    }
    // :87-126
    uint32_t extract_min() {
        const int m = min_;
        const int deg = N[m].degree;
        int c = N[m].child;
        for (int i = 0; i < deg; ++i) {
            const int r = c;
            c = N[c].right;
            to_root(r);
        }
        unlink(m);
        --num_;
        if (num_ == 0) {
            min_ = -1;
        } else {
            min_ = N[m].right;
            const int ml = N[m].left;
            N[min_].left = ml;
            N[ml].right = min_;
            consolidate();
        }
        const uint32_t v = N[m].value;
        freelist.push_back(m);
        return v;
    }
};

constexpr uint32_t kNoParent = 0xFFFFFFFFu;
constexpr uint32_t kHeapCap = 10000000u;   // fibonacci-heap.h:27; past it the reference misbehaves

struct PathNode {
    uint32_t parent;
    uint32_t step;         // oriented step (root: node<<1, orientation decided by the first edge)
    uint32_t depth;        // number of steps of the path ending here
    uint32_t uniques;
    uint32_t bad = 0, good = 0;
    uint32_t child_begin = 0, child_count = 0;
    uint8_t root_or = 0;   // depth-2 nodes: orientation the source is left in (src/eval.cpp:148-149)
    bool scored = false, expanded = false;
    bool kids_scored = false;        // every child carries its counts (or is in flight in the current batch)
    // counts over the alignments outside the fresh set F (incremental scoring, gfe_score_paths_incremental)
    int32_t main_bad = 0, main_good = 0;
    int32_t key = 0;                 // alt = bad - good - uniques, the path's heap key (src/eval.cpp:163)
    // speculation priority: the replay reaches a cached path only after every ancestor was extracted,
    // i.e. not before the largest key on the way down from the heap: eff = max(key, eff(parent))
    int32_t eff = INT32_MIN;
    bool extracted = false;
};

struct Driver {
    gfe_score_fn score;
    void *score_user;
    const gfe_graph *g;
    const gfe_search_params *prm;
    std::vector<PathNode> T;
    std::vector<uint32_t> budget;     // initial NodeTable counts by node (0 = no record)
    std::vector<uint32_t> occ;        // scratch: visits of each node along the current path
    std::vector<uint32_t> base_path;  // materialised path of the node being expanded
    std::vector<uint32_t> suffix;
    // batch
    std::vector<uint32_t> b_nodes, b_off, b_steps, b_counts;
    gfe_search_stats st{};

    // path of node x into `out` (root orientation resolved)
    void materialise(uint32_t x, std::vector<uint32_t> &out) {
        out.resize(T[x].depth);
        uint32_t i = T[x].depth, cur = x, second = kNoParent;
        while (cur != kNoParent) {
            out[--i] = T[cur].step;
            if (T[cur].depth == 2) second = cur;
            cur = T[cur].parent;
        }
        if (second != kNoParent) out[0] = (out[0] & ~1u) | T[second].root_or;
    }

    // children of x per src/eval.cpp:136-151; `path` is x's materialised path and occ[] its visits
    void expand(uint32_t x, const std::vector<uint32_t> &path) { expand(x, path.back(), path[0] >> 1); }
    // the same from the path's ends: `last` = its last step (== T[x].step), root_node = node of path[0]
    void expand(uint32_t x, uint32_t last, uint32_t root_node) {
        PathNode &nx = T[x];
        nx.expanded = true;
        const uint32_t begin = (uint32_t)T.size();
        const uint32_t node = last >> 1;
        const bool is_root = T[x].depth == 1;
        uint32_t count = 0;
        for (uint32_t k = g->adj_offsets[node]; k < g->adj_offsets[node + 1]; ++k) {
            if (!is_root && (last & 1u) != g->adj_or0[k]) continue;
            const uint32_t v = g->adj_target[k];
            const uint32_t pushes = occ[v] - (v == root_node ? 1u : 0u);
            if (budget[v] <= pushes) continue;   // no record, or no visits left
            PathNode c;
            c.parent = x;
            c.step = (v << 1) | g->adj_or1[k];
            c.depth = T[x].depth + 1;
            c.uniques = T[x].uniques + (occ[v] == 0 ? 1u : 0u);
            c.root_or = g->adj_or0[k];
            T.push_back(c);
            ++count;
        }
        T[x].child_begin = begin;
        T[x].child_count = count;
    }

    void add_to_batch(uint32_t node_id, const std::vector<uint32_t> &prefix,
                      const std::vector<uint32_t> &suf) {
        b_nodes.push_back(node_id);
        b_steps.insert(b_steps.end(), prefix.begin(), prefix.end());
        b_steps.insert(b_steps.end(), suf.begin(), suf.end());
        b_off.push_back((uint32_t)b_steps.size());
    }

    // Make sure every child of x carries its (bad, good); speculate below them while at it.
    int score_children(uint32_t x) {
        materialise(x, base_path);
        for (uint32_t s : base_path) occ[s >> 1]++;
        if (!T[x].expanded) expand(x, base_path);
        bool missing = false;
        for (uint32_t c = 0; c < T[x].child_count; ++c)
            if (!T[T[x].child_begin + c].scored) missing = true;
        int rc = GFE_OK;
        if (missing) {
            b_nodes.clear(); b_steps.clear(); b_off.assign(1, 0);
            const uint32_t limit = prm->speculate ? prm->speculate : 256u;
            // breadth-first below x: (tree node, suffix of steps below x)
            std::deque<uint32_t> frontier;
            frontier.push_back(x);
            const uint32_t base_depth = T[x].depth;
            bool base_fixed = base_path.size() > 1;
            while (!frontier.empty() && b_nodes.size() < limit) {
                const uint32_t y = frontier.front();
                frontier.pop_front();
                // suffix of y below x
                suffix.resize(T[y].depth - base_depth);
                for (uint32_t cur = y, i = (uint32_t)suffix.size(); cur != x; cur = T[cur].parent)
                    suffix[--i] = T[cur].step;
                if (y != x) {
                    if (T[y].depth > GFE_MAX_STEPS - 1) continue;   // children would exceed MAX_N
                    if (!T[y].expanded) {
                        for (uint32_t s : suffix) occ[s >> 1]++;
                        // expand() wants the full path only for its ends and length
                        std::vector<uint32_t> full(base_path);
                        full.insert(full.end(), suffix.begin(), suffix.end());
                        if (!base_fixed) full[0] = (full[0] & ~1u) | T[first_below(y, x)].root_or;
                        expand(y, full);
                        for (uint32_t s : suffix) occ[s >> 1]--;
                    }
                }
                for (uint32_t c = 0; c < T[y].child_count; ++c) {
                    // the children of x itself are required; everything deeper is speculation
                    if (y != x && b_nodes.size() >= limit) break;
                    const uint32_t ci = T[y].child_begin + c;
                    if (!T[ci].scored) {
                        suffix.push_back(T[ci].step);
                        if (!base_fixed) {
                            // x is the root: its orientation comes from the first edge of this child
                            std::vector<uint32_t> pre(base_path);
                            const uint32_t fb = first_below(ci, x);
                            pre[0] = (pre[0] & ~1u) | T[fb].root_or;
                            add_to_batch(ci, pre, suffix);
                        } else {
                            add_to_batch(ci, base_path, suffix);
                        }
                        suffix.pop_back();
                    }
                    if ((T[ci].step >> 1) != prm->destination) frontier.push_back(ci);
                }
            }
            const uint32_t P = (uint32_t)b_nodes.size();
            b_counts.resize((size_t)P * 2);
            auto t0 = std::chrono::steady_clock::now();
            rc = score(score_user, b_off.data(), b_steps.data(), P, b_counts.data());
            st.seconds_scoring +=
                std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            st.kernel_batches++;
            st.paths_scored += P;
            if (rc == GFE_OK)
                for (uint32_t i = 0; i < P; ++i) {
                    PathNode &n = T[b_nodes[i]];
                    n.bad = b_counts[2 * i];
                    n.good = b_counts[2 * i + 1];
                    n.scored = true;
                }
        }
        for (uint32_t s : base_path) occ[s >> 1]--;
        return rc;
    }

    // ---- engine-backed scoring: incremental, several speculation roots ------------------------------
    // evaluatePath(child) differs from evaluatePath(parent) only on the alignments that contain the
    // appended node or the root node (gfe_score_paths_incremental); the batch therefore carries, for
    // every candidate, (base path of its speculation root, suffix below it) and asks for the delta.
    // Roots: the extracted path x (its children are REQUIRED) and the best other frontier paths of the
    // heap (lowest key first) whose children are not scored yet — the paths the replay will extract next.
    gfe_engine *eng = nullptr;
    gfe_exchange *xch = nullptr;      // alignment shards: the same batches on every rank, records summed over the box
    // (key, node) of every scored, non-final path whose children are not scored yet — in the heap or
    // still below it in the cached tree: where the replay can miss next, best key first
    // min-heap with lazy deletion: an entry is live while its node still lacks scored children and its
    // priority is the node's current eff
    struct FrontCmp {
        bool operator()(const std::pair<int32_t, uint32_t> &a, const std::pair<int32_t, uint32_t> &b) const {
            return a > b;
        }
    };
    std::priority_queue<std::pair<int32_t, uint32_t>, std::vector<std::pair<int32_t, uint32_t>>, FrontCmp> front;
    std::vector<std::pair<int32_t, uint32_t>> front_keep;
    // candidates per batch / speculation roots per batch / live front entries examined per batch: measured
    // optimum on cfg3 (profiles/r02_search_sweep_effkey.txt): many roots with a small subtree each
    uint32_t spec_width = 4096, spec_roots = 256, spec_scan = 4096;
    std::vector<uint32_t> t_base_off, t_base_steps, t_item_base, t_suf_off, t_suf_steps, t_items, t_parents;
    std::vector<uint8_t> t_mode;
    std::vector<int32_t> t_counts;
    std::vector<uint32_t> root_path, t_kids;

    // The materialised path of ONE tree node at a time (`cur`), with occ[] = its visit counts; moving to
    // another node costs the tree distance, not the depth: speculation roots are mostly close relatives.
    uint32_t cur = kNoParent;
    std::vector<uint32_t> cur_path, chain, climb;   // chain[d-1] = tree node of depth d on the way to cur
    void move_to(uint32_t r) {
        climb.clear();
        uint32_t a = r;
        while (a != kNoParent && (cur == kNoParent || T[a].depth > cur_path.size() || chain[T[a].depth - 1] != a)) {
            climb.push_back(a);
            a = T[a].parent;
        }
        const size_t keep = a == kNoParent ? 0 : T[a].depth;
        while (cur_path.size() > keep) {
            occ[cur_path.back() >> 1]--;
            cur_path.pop_back();
            chain.pop_back();
        }
        for (size_t i = climb.size(); i-- > 0;) {
            const uint32_t y = climb[i];
            cur_path.push_back(T[y].step);
            chain.push_back(y);
            occ[T[y].step >> 1]++;
        }
        // the source's orientation is the one its first edge left it in (src/eval.cpp:148-149)
        if (cur_path.size() >= 2) cur_path[0] = (cur_path[0] & ~1u) | T[chain[1]].root_or;
        cur = r;
    }
    void leave_path() {   // back to the neutral state the full-scan driver expects (occ[] all zero)
        for (uint32_t s : cur_path) occ[s >> 1]--;
        cur_path.clear();
        chain.clear();
        cur = kNoParent;
    }

    // BFS below root r (path `rp`, occ[] holds its visits): one item per unscored child.  A node's
    // children are taken all or none; the walk stops when the next node's children no longer fit
    // `budget` (the children of r itself are taken regardless when `required`).
    void speculate_below(uint32_t r, const std::vector<uint32_t> &rp, uint32_t budget, bool required) {
        const uint32_t base_depth = T[r].depth;
        const bool base_fixed = rp.size() > 1;     // else r is the search root: orientation per first edge
        uint32_t kb[2] = {kNoParent, kNoParent};   // base id per root orientation (base_fixed: kb[0] only)
        auto base_for = [&](uint32_t orient) -> uint32_t {
            const uint32_t k = base_fixed ? 0u : orient;
            if (kb[k] == kNoParent) {
                kb[k] = (uint32_t)t_base_off.size() - 1;
                const size_t at = t_base_steps.size();
                t_base_steps.insert(t_base_steps.end(), rp.begin(), rp.end());
                if (!base_fixed) t_base_steps[at] = (rp[0] & ~1u) | orient;
                t_base_off.push_back((uint32_t)t_base_steps.size());
            }
            return kb[k];
        };
        std::deque<uint32_t> frontier;
        frontier.push_back(r);
        uint32_t taken = 0, visited = 0;
        while (!frontier.empty() && visited < 4 * budget + 16) {
            const uint32_t y = frontier.front();
            frontier.pop_front();
            ++visited;
            if (T[y].depth > GFE_MAX_STEPS - 1) continue;   // children would exceed MAX_N
            suffix.resize(T[y].depth - base_depth);
            for (uint32_t cur = y, i = (uint32_t)suffix.size(); cur != r; cur = T[cur].parent)
                suffix[--i] = T[cur].step;
            if (!T[y].expanded) {
                for (uint32_t s : suffix) occ[s >> 1]++;
                expand(y, T[y].step, rp[0] >> 1);
                for (uint32_t s : suffix) occ[s >> 1]--;
            }
            uint32_t open = 0;
            for (uint32_t c = 0; c < T[y].child_count; ++c) open += T[T[y].child_begin + c].scored ? 0u : 1u;
            if (open && !(required && y == r) && taken + open > budget) break;
            for (uint32_t c = 0; c < T[y].child_count; ++c) {
                const uint32_t ci = T[y].child_begin + c;
                if (!T[ci].scored) {
                    const uint32_t orient = base_fixed ? 0u : T[y == r ? ci : first_below(y, r)].root_or;
                    t_items.push_back(ci);
                    t_parents.push_back(y);
                    t_item_base.push_back(base_for(orient));
                    t_suf_steps.insert(t_suf_steps.end(), suffix.begin(), suffix.end());
                    t_suf_steps.push_back(T[ci].step);
                    t_suf_off.push_back((uint32_t)t_suf_steps.size());
                    t_mode.push_back(T[y].depth >= 16 ? 1 : 0);   // delta needs a parent of >= 16 steps
                    ++taken;
                }
                if ((T[ci].step >> 1) != prm->destination) frontier.push_back(ci);
            }
            t_kids.push_back(y);
        }
    }

    int score_children_tree(uint32_t x) {
        if (T[x].kids_scored) return GFE_OK;     // set only once every child carries its counts
        t_base_off.assign(1, 0);
        t_base_steps.clear(); t_item_base.clear(); t_suf_off.assign(1, 0); t_suf_steps.clear();
        t_items.clear(); t_parents.clear(); t_mode.clear(); t_kids.clear();
        const uint32_t width = prm->speculate ? prm->speculate : spec_width;
        // root 0: the extracted path
        move_to(x);
        speculate_below(x, cur_path, std::max<uint32_t>(width / 2, 1), true);
        // further roots: the best frontier paths of the heap whose children are not scored
        if (width > 1 && spec_roots > 1) {
            uint32_t picked = 0, scanned = 0;
            const uint32_t per_root = std::max<uint32_t>((width - width / 2) / (spec_roots - 1), 1);
            front_keep.clear();
            while (!front.empty() && picked + 1 < spec_roots && scanned < spec_scan) {
                const std::pair<int32_t, uint32_t> top = front.top();
                front.pop();
                const uint32_t r = top.second;
                if (T[r].kids_scored || T[r].eff != top.first) continue;          // stale entry
                ++scanned;
                if (r == x || T[r].depth >= GFE_MAX_STEPS) {
                    front_keep.push_back(top);
                    continue;
                }
                move_to(r);
                speculate_below(r, cur_path, per_root, false);
                front_keep.push_back(top);   // dropped below once its children are in the batch
                ++picked;
            }
        }
        const uint32_t P = (uint32_t)t_items.size();
        if (P == 0) {
            T[x].kids_scored = true;
            for (const auto &k : front_keep)
                if (!T[k.second].kids_scored) front.push(k);
            front_keep.clear();
            return GFE_OK;
        }
        t_counts.resize((size_t)P * 4);
        gfe_path_batch b;
        b.n_bases = (uint32_t)t_base_off.size() - 1;
        b.base_offsets = t_base_off.data();
        b.base_steps = t_base_steps.data();
        b.P = P;
        b.item_base = t_item_base.data();
        b.suffix_offsets = t_suf_off.data();
        b.suffix_steps = t_suf_steps.data();
        b.mode = t_mode.data();
        auto t0 = std::chrono::steady_clock::now();
        const int rc = xch ? gfe_score_paths_incremental_sharded(eng, xch, &b, t_counts.data())
                           : gfe_score_paths_incremental(eng, &b, t_counts.data());
        st.seconds_scoring += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        st.kernel_batches++;
        st.paths_scored += P;
        if (rc != GFE_OK) return rc;
        for (uint32_t i = 0; i < P; ++i) {
            PathNode &n = T[t_items[i]];
            const int32_t *c = &t_counts[4 * (size_t)i];
            if (t_mode[i]) {
                const PathNode &par = T[t_parents[i]];
                n.main_bad = par.main_bad + c[0];
                n.main_good = par.main_good + c[1];
            } else {
                n.main_bad = c[0] - c[2];
                n.main_good = c[1] - c[3];
            }
            n.bad = (uint32_t)(n.main_bad + c[2]);
            n.good = (uint32_t)(n.main_good + c[3]);
            n.scored = true;
            n.key = (int32_t)n.bad - (int32_t)n.good - (int32_t)n.uniques;   // src/eval.cpp:163
            const PathNode &pp = T[t_parents[i]];
            n.eff = pp.extracted ? n.key : std::max(n.key, pp.eff);
            if ((n.step >> 1) != prm->destination && n.depth < GFE_MAX_STEPS) front.push({n.eff, t_items[i]});
        }
        for (uint32_t y : t_kids) T[y].kids_scored = true;   // their front entries are stale from now on
        for (const auto &k : front_keep)
            if (!T[k.second].kids_scored) front.push(k);
        front_keep.clear();
        return GFE_OK;
    }

    // the ancestor of y (or y itself) that is a direct child of x
    uint32_t first_below(uint32_t y, uint32_t x) const {
        uint32_t cur = y;
        while (T[cur].parent != x) cur = T[cur].parent;
        return cur;
    }
};

}  // namespace

static int run_search(gfe_score_fn score, void *score_user, const gfe_graph *g,
                      const gfe_search_params *p, gfe_row_fn emit, void *user,
                      gfe_search_stats *stats, gfe_engine *eng = nullptr, gfe_exchange *xch = nullptr) {
    const auto t_begin = std::chrono::steady_clock::now();
    Driver d;
    d.score = score;
    d.score_user = score_user;
    d.g = g;
    d.prm = p;
    d.budget.assign(g->n_nodes, 0);
    d.occ.assign(g->n_nodes, 0);
    for (uint32_t i = 0; i < p->n_budget; ++i) {
        if (p->budget_nodes[i] >= g->n_nodes) return GFE_ERR_ARG;
        d.budget[p->budget_nodes[i]] = p->budget_counts[i];
    }
    if (eng) {
        d.eng = eng;
        d.xch = xch;
        if (const char *v = getenv("GFE_SPEC_WIDTH")) d.spec_width = (uint32_t)std::max(1, atoi(v));
        if (const char *v = getenv("GFE_SPEC_ROOTS")) d.spec_roots = (uint32_t)std::max(1, atoi(v));
        if (const char *v = getenv("GFE_SPEC_SCAN")) d.spec_scan = (uint32_t)std::max(1, atoi(v));
    }
    OrderHeap Q;
    PathNode root;
    root.parent = kNoParent;
    root.step = p->source << 1;
    root.depth = 1;
    root.uniques = 1;
    root.scored = true;
    d.T.reserve(1u << 20);
    d.T.push_back(root);
    Q.insert(0, 0);

    uint64_t path_counter = 0;
    uint32_t steps = 0, best_uniques = 0;
    int32_t best_alt = INT32_MAX;
    std::vector<uint32_t> row_path;
    std::vector<uint32_t> visits(g->n_nodes, 0);
    int rc = GFE_OK;

    const bool dbg = getenv("GFE_DEBUG_TIMING") != nullptr;
    double t_heap = 0, t_score = 0, t_kids = 0;
    auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    while (Q.size() > 0 && steps < p->max_steps) {
        const double t0 = dbg ? now() : 0;
        const uint32_t x = Q.extract_min();
        const double t1 = dbg ? now() : 0;
        t_heap += t1 - t0;
        d.T[x].extracted = true;
        if (d.T[x].depth < GFE_MAX_STEPS) {
            rc = eng ? d.score_children_tree(x) : d.score_children(x);
            if (rc == GFE_ERR_STATE && eng && !xch) {
                // no pair dictionary for this universe: the incremental path is off, use full scans
                // (single engine only: over shards the choice is made for all ranks together, up front)
                eng = nullptr;
                d.eng = nullptr;
                d.leave_path();
                rc = d.score_children(x);
            }
            if (rc != GFE_OK) break;
        } else {
            // the reference's dp[MAX_N][MAX_N] cannot hold a longer path (include/alignments.h:246)
            d.T[x].expanded = true;
            d.T[x].child_count = 0;
        }
        const double t2 = dbg ? now() : 0;
        t_score += t2 - t1;
        for (uint32_t c = 0; c < d.T[x].child_count; ++c) {
            const uint32_t ci = d.T[x].child_begin + c;
            const PathNode &n = d.T[ci];
            const int32_t alt = (int32_t)n.bad - (int32_t)n.good - (int32_t)n.uniques;
            d.st.evaluations++;
            if ((n.step >> 1) != p->destination) {
                if (Q.size() >= kHeapCap) {
                    rc = GFE_ERR_LIMIT;   // reference: "reached limit of the Fibonacci heap"
                    break;
                }
                Q.insert(ci, alt);
                if (eng && d.T[ci].eff != alt) {      // now in the heap: reachable at its own key
                    d.T[ci].eff = alt;
                    if (!d.T[ci].kids_scored) d.front.push({alt, ci});
                }

            } else {
                ++path_counter;
                d.materialise(ci, row_path);
                // NodeTable::checkHamiltonian, include/nodetable.h:56-67
                bool ham = (uint64_t)row_path.size() + 2 == p->node_count;
                if (ham) {
                    for (uint32_t s : row_path) visits[s >> 1]++;
                    for (uint32_t i = 0; i < p->n_budget && ham; ++i)
                        if (visits[p->budget_nodes[i]] != p->budget_counts[i]) ham = false;
                    for (uint32_t s : row_path) visits[s >> 1]--;
                }
                bool print = false;
                if (n.uniques >= p->min_nodes &&
                    (best_uniques < n.uniques || (best_uniques == n.uniques && best_alt > alt))) {
                    best_alt = alt;
                    best_uniques = n.uniques;
                    print = true;
                }
                if ((p->return_all_paths || print) && emit) {
                    gfe_search_row row;
                    row.path_counter = path_counter;
                    row.bad = n.bad;
                    row.good = n.good;
                    row.alt = alt;
                    row.size = (uint32_t)row_path.size();
                    row.uniques = n.uniques;
                    row.hamiltonian = ham ? 1 : 0;
                    row.steps = row_path.data();
                    emit(&row, user);
                    d.st.rows++;
                }
            }
        }
        if (dbg) t_kids += now() - t2;
        if (rc != GFE_OK) break;
        ++steps;
    }
    if (dbg)
        fprintf(stderr, "[gfe search] extract_min %.3f s, score_children (incl. GPU %.3f s) %.3f s, insert/emit %.3f s, "
                        "tree nodes %zu\n", t_heap, d.st.seconds_scoring, t_score, t_kids, d.T.size());
    d.st.extractions = steps;
    d.st.reached_max_steps = steps >= p->max_steps ? 1 : 0;
    d.st.seconds_total =
        std::chrono::duration<double>(std::chrono::steady_clock::now() - t_begin).count();
    if (stats) *stats = d.st;
    return rc;
}

static int engine_scorer(void *user, const uint32_t *path_offsets, const uint32_t *path_steps,
                         uint32_t P, uint32_t *counts) {
    return gfe_score_paths(static_cast<gfe_engine *>(user), path_offsets, path_steps, P, 1, counts);
}

// universe = every node that owns a NodeTable record (nodelist + source + destination)
static void search_universe(const gfe_graph *g, const gfe_search_params *p, std::vector<uint32_t> &uni) {
    std::vector<uint8_t> in(g->n_nodes, 0);
    for (uint32_t i = 0; i < p->n_budget; ++i)
        if (p->budget_counts[i]) in[p->budget_nodes[i]] = 1;
    in[p->source] = in[p->destination] = 1;
    for (uint32_t v = 0; v < g->n_nodes; ++v)
        if (in[v]) uni.push_back(v);
}

static int check_search_args(const gfe_graph *g, const gfe_search_params *p) {
    if (!g || !p || !g->adj_offsets) return GFE_ERR_ARG;
    if (p->source >= g->n_nodes || p->destination >= g->n_nodes) return GFE_ERR_ARG;
    for (uint32_t i = 0; i < p->n_budget; ++i)
        if (p->budget_nodes[i] >= g->n_nodes) return GFE_ERR_ARG;
    return GFE_OK;
}

// engine.cu (C++ linkage, not part of the C ABI): one-word vote over the exchange buffer
int gfe_internal_incremental_consensus(gfe_engine *e, gfe_exchange *x, int *all_have);

extern "C" int gfe_search(gfe_engine *e, const gfe_graph *g, const gfe_search_params *p,
                          gfe_row_fn emit, void *user, gfe_search_stats *stats) {
    if (!e) return GFE_ERR_ARG;
    int rc = check_search_args(g, p);
    if (rc != GFE_OK) return rc;
    std::vector<uint32_t> uni;
    search_universe(g, p, uni);
    rc = gfe_set_universe(e, uni.data(), (uint32_t)uni.size());
    if (rc != GFE_OK) return rc;
    // GFE_SEARCH_FULLSCAN=1: every candidate through gfe_score_paths (the checker of the incremental path)
    const char *fs = getenv("GFE_SEARCH_FULLSCAN");
    return run_search(engine_scorer, e, g, p, emit, user, stats, (fs && fs[0] == '1') ? nullptr : e);
}

namespace {
struct ShardedUser {
    gfe_engine *e;
    gfe_exchange *x;
};
int sharded_scorer(void *user, const uint32_t *path_offsets, const uint32_t *path_steps, uint32_t P,
                   uint32_t *counts) {
    ShardedUser *u = static_cast<ShardedUser *>(user);
    return gfe_score_paths_sharded(u->e, u->x, path_offsets, path_steps, P, 1, counts);
}
}  // namespace

extern "C" int gfe_search_sharded(gfe_engine *e, gfe_exchange *x, const gfe_graph *g,
                                  const gfe_search_params *p, gfe_row_fn emit, void *user,
                                  gfe_search_stats *stats) {
    if (!e || !x) return GFE_ERR_ARG;
    int rc = check_search_args(g, p);
    if (rc != GFE_OK) return rc;
    std::vector<uint32_t> uni;
    search_universe(g, p, uni);
    rc = gfe_set_universe(e, uni.data(), (uint32_t)uni.size());
    if (rc != GFE_OK) return rc;
    ShardedUser u{e, x};
    const char *fs = getenv("GFE_SEARCH_FULLSCAN");
    bool full = fs && fs[0] == '1';
    if (!full) {
        // a rank whose shard has no pair dictionary can only run full scans, and the two scoring entries
        // exchange records of different shapes: all ranks take the incremental path, or none does
        int all_have = 0;
        rc = gfe_internal_incremental_consensus(e, x, &all_have);
        if (rc != GFE_OK) return rc;
        full = !all_have;
    }
    return run_search(sharded_scorer, &u, g, p, emit, user, stats, full ? nullptr : e, full ? nullptr : x);
}

extern "C" int gfe_search_with_scorer(gfe_score_fn score, void *score_user, const gfe_graph *g,
                                      const gfe_search_params *p, gfe_row_fn emit, void *user,
                                      gfe_search_stats *stats) {
    if (!score) return GFE_ERR_ARG;
    int rc = check_search_args(g, p);
    if (rc != GFE_OK) return rc;
    return run_search(score, score_user, g, p, emit, user, stats);
}
