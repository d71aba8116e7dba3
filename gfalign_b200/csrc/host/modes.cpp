// gfalign_b200/csrc/host/modes.cpp — the four tools of `gfalign-b200`.  Host code parses the text,
// keeps the reference's stdout formats and error texts; every score, count and weight comes from
// the CUDA engine through the C ABI (there is no CPU scoring here).
#include "cli.h"

#include <zlib.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "host.h"

namespace gfh {

namespace {

// CUDA start-up (driver initialisation + context: ~1 s on a B200 box) runs on its own thread while
// the text inputs are parsed; the first engine call joins it.
struct Engine;
Engine *g_starting = nullptr;
void finish_start();

[[noreturn]] void die(const std::string &msg) {
    finish_start();   // never exit() with the CUDA start-up still in flight
    fprintf(stderr, "%s\n", msg.c_str());
    exit(EXIT_FAILURE);
}

struct Engine {
    gfe_engine *e = nullptr;
    int rc = GFE_OK;
    std::string err;
    std::thread th;
    explicit Engine(int device) {
        g_starting = this;
        th = std::thread([this, device] {
            rc = gfe_create(device, &e);
            if (rc != GFE_OK) err = gfe_last_error(nullptr);   // thread-local message: read it here
        });
    }
    void wait() {
        if (th.joinable()) th.join();
        if (g_starting == this) g_starting = nullptr;
    }
    // joins the start-up; dies the way the synchronous constructor did when there is no device
    gfe_engine *get() {
        wait();
        if (rc != GFE_OK) die(std::string("gfalign-b200: ") + err);
        return e;
    }
    ~Engine() {
        wait();
        if (e) gfe_destroy(e);
    }
    void check(int r) const {
        if (r != GFE_OK) die(std::string("gfalign-b200: ") + gfe_last_error(e));
    }
};

void finish_start() {
    if (g_starting) g_starting->wait();
}

double now_s() {
    using namespace std::chrono;
    return duration<double>(steady_clock::now().time_since_epoch()).count();
}

// buffered stdout / file writer (evalPath prints one long line per alignment)
struct Out {
    FILE *f;
    gzFile gz = nullptr;
    std::string buf;
    explicit Out(FILE *f_) : f(f_) { buf.reserve(1 << 22); }
    void flush() {
        if (buf.empty()) return;
        if (gz) gzwrite(gz, buf.data(), (unsigned)buf.size());
        else fwrite(buf.data(), 1, buf.size(), f);
        buf.clear();
    }
    void put(std::string_view s) {
        buf.append(s.data(), s.size());
        if (buf.size() > (1u << 22) - 65536) flush();
    }
    void put(char c) { buf.push_back(c); }
    void num(long long v) {
        char t[32];
        int n = snprintf(t, sizeof t, "%lld", v);
        buf.append(t, (size_t)n);
    }
    ~Out() { flush(); }
};

void load_graph(const Options &o, Graph &g) {
    std::string err;
    if (o.gfa.empty()) die("Error: no input sequence (-f) given.");
    if (!g.load(o.gfa, err)) die("Error: " + err);
}

void load_gaf(const Options &o, const NameTable &names, uint32_t unknown, bool records, Gaf &gaf) {
    std::string err;
    const int threads = o.threads > 0 ? o.threads : hardware_threads();
    if (!gaf.load(o.gaf, names, unknown, threads, records, err)) die("Error: " + err);
}

// std::cout << double with the stream's default formatting (%g, precision 6), or fixed with two
// decimals once an OutputStream to a file exists (gfalibs/src/output.cpp:26-27)
std::string fmt_double(double d, bool fixed2) {
    char t[64];
    if (fixed2) snprintf(t, sizeof t, "%.2f", d);
    else snprintf(t, sizeof t, "%g", d);
    return t;
}

double gfa_round(double d) {   // gfalibs/include/functions.h:215-220
    if (std::isnan(d)) return NAN;
    return std::round(d * 100) / 100;
}

struct AlignmentSummary {     // InAlignments totals, include/alignments.h:178-190
    unsigned long long n = 0, totQLen = 0, totAlgSeq = 0, totPlus = 0, totMinus = 0, totPLen = 0,
                       totMapq = 0, totMatches = 0, totBlockLen = 0, primary = 0, secondary = 0,
                       supplementary = 0, terminalSupplementary = 0;
};

void add_totals(const Gaf &gaf, AlignmentSummary &s) {   // AlignmentStats::add, src/alignments.cpp:237-246
    for (const GafRecord &r : gaf.records) {
        s.totQLen += r.qLen;
        s.totAlgSeq += (unsigned long long)(r.qEnd - r.qStart);   // unsigned arithmetic as the reference
        r.strand == '+' ? ++s.totPlus : ++s.totMinus;
        s.totPLen += r.pLen;
        s.totMatches += r.matches;
        s.totBlockLen += r.blockLen;
        s.totMapq += r.mapq;
    }
}

// InAlignments::printStats, src/alignments.cpp:260-277
void print_stats(const AlignmentSummary &s, bool fixed2) {
    auto avg = [&](unsigned long long v) { return fmt_double(gfa_round((double)v / (double)s.n), fixed2); };
    const double tot = (double)(s.totPlus + s.totMinus);
    printf("+++Alignment summary+++: \n");
    printf("# alignments: %llu\n", s.n);
    printf("Average read length: %s\n", avg(s.totQLen).c_str());
    printf("Average aligned sequence: %s\n", avg(s.totAlgSeq).c_str());
    printf("Alignment orientation (+/-): %llu(%s%%):%llu(%s%%)\n", s.totPlus,
           fmt_double(gfa_round((double)s.totPlus / tot * 100), fixed2).c_str(), s.totMinus,
           fmt_double(gfa_round((double)s.totMinus / tot * 100), fixed2).c_str());
    printf("Average path length: %s\n", avg(s.totPLen).c_str());
    printf("Average alignment quality: %s\n", avg(s.totMapq).c_str());
    printf("Average matches #: %s\n", avg(s.totMatches).c_str());
    printf("Average block length: %s\n", avg(s.totBlockLen).c_str());
    printf("Primary alignments: %llu\n", s.primary);
    printf("Secondary alignments: %llu\n", s.secondary);
    printf("Supplementary alignments: %llu\n", s.supplementary);
    printf("Terminal supplementary alignments: %llu\n", s.terminalSupplementary);
}

// InAlignment::print (src/alignments.cpp:52-73): the record re-serialised from its parsed columns —
// numbers through stoi/to_string, tags as label(2):type(1):third ':'-field
void print_record(const Gaf &gaf, uint64_t a, Out &out) {
    std::string_view line = gaf.line(a);
    std::vector<std::string_view> cols;
    size_t p = 0;
    for (;;) {
        size_t t = line.find('\t', p);
        if (t == std::string_view::npos) {
            cols.push_back(line.substr(p));
            break;
        }
        cols.push_back(line.substr(p, t - p));
        p = t + 1;
    }
    const GafRecord &r = gaf.records[a];
    out.put(cols[0]); out.put('\t');
    out.num(r.qLen); out.put('\t'); out.num(r.qStart); out.put('\t'); out.num(r.qEnd); out.put('\t');
    out.put(r.strand); out.put('\t');
    out.put(cols[5]); out.put('\t');
    out.num(r.pLen); out.put('\t'); out.num(r.pStart); out.put('\t'); out.num(r.pEnd); out.put('\t');
    out.num(r.matches); out.put('\t'); out.num(r.blockLen); out.put('\t'); out.num(r.mapq);
    for (size_t i = 12; i < cols.size(); ++i) {
        std::string_view c = cols[i];
        const size_t c1 = c.find(':');
        const size_t c2 = c1 == std::string_view::npos ? c1 : c.find(':', c1 + 1);
        if (c2 == std::string_view::npos || c1 < 2 || c2 == c1 + 1) continue;   // malformed tag (UB in the reference)
        const size_t c3 = c.find(':', c2 + 1);
        out.put('\t');
        out.put(c.substr(0, 2)); out.put(':');
        out.put(c[c1 + 1]); out.put(':');
        out.put(c.substr(c2 + 1, c3 == std::string_view::npos ? c3 : c3 - c2 - 1));
    }
    out.put('\n');
}

}  // namespace

// ---- search (dijkstra, src/eval.cpp:110-193) -------------------------------------------------------

namespace {
struct RowSink {
    const Graph *g;
    Out *out;
};
void emit_row(const gfe_search_row *r, void *user) {   // the row of src/eval.cpp:183
    RowSink *s = static_cast<RowSink *>(user);
    Out &o = *s->out;
    o.num((long long)r->path_counter); o.put('\t');
    o.num(r->bad); o.put('\t');
    o.num(r->good); o.put('\t');
    o.num(r->alt); o.put('\t');
    o.num(r->size); o.put('\t');
    o.num(r->uniques); o.put('\t');
    o.put(r->hamiltonian ? 'T' : 'F'); o.put('\t');
    for (uint32_t i = 0; i < r->size; ++i) {
        o.put(s->g->names.name(r->steps[i] >> 1));
        o.put((r->steps[i] & 1) ? '-' : '+');
        if (i + 1 < r->size) o.put(',');
    }
    o.put('\n');
    o.flush();   // rows appear as they are found, like the reference's std::endl
}
}  // namespace

int run_search(const Options &o) {
    Engine eng(o.device);   // CUDA start-up overlaps the parsing below
    const double t0 = now_s();
    Graph g;
    load_graph(o, g);
    if (g.names.size() == 0) die("Error: the graph has no segments.");
    Gaf gaf;
    if (!o.gaf.empty()) load_gaf(o, g.names, 0, false, gaf);
    else gaf.offsets.assign(1, 0);
    // headersToIds[source] / [destination] default-insert 0 for unknown names (src/eval.cpp:126-127)
    const int64_t s = g.names.find(o.source), d = g.names.find(o.destination);
    const uint32_t source = s < 0 ? 0 : (uint32_t)s, destination = d < 0 ? 0 : (uint32_t)d;
    NodeBudget nb;
    std::string err;
    if (o.node_file.empty()) {   // NodeTable of a file that cannot be opened: no records
        nb.nodes = {source};
        nb.counts = {1};
        if (destination != source) {
            nb.nodes.push_back(destination);
            nb.counts.push_back(1);
        }
        nb.node_count = 2;
    } else if (!load_nodelist(o.node_file, g, source, destination, nb, err)) {
        die(err);
    }
    const double t1 = now_s();
    eng.get();
    eng.check(gfe_set_alignments(eng.e, gaf.offsets.data(), gaf.steps.data(), gaf.A, g.names.size()));
    gfe_graph gv = g.view();
    gfe_search_params p;
    memset(&p, 0, sizeof p);
    p.source = source;
    p.destination = destination;
    p.budget_nodes = nb.nodes.data();
    p.budget_counts = nb.counts.data();
    p.n_budget = (uint32_t)nb.nodes.size();
    p.node_count = nb.node_count;
    p.max_steps = o.max_steps;
    p.min_nodes = o.min_nodes;
    p.return_all_paths = o.return_all_paths ? 1 : 0;
    p.speculate = o.speculate;
    Out out(stdout);
    RowSink sink{&g, &out};
    gfe_search_stats st;
    memset(&st, 0, sizeof st);
    const double t2 = now_s();
    eng.check(gfe_search(eng.e, &gv, &p, emit_row, &sink, &st));
    if (st.reached_max_steps) {
        out.put("Reached maximum number of steps (");
        out.num((long long)st.extractions);
        out.put(")\n");
    }
    out.flush();
    if (o.timing)
        fprintf(stderr,
                "[gfalign-b200] parse %.3f s, engine setup %.3f s, search %.3f s (scoring %.3f s): "
                "%llu extractions, %llu evaluatePath calls, %llu kernel batches, %llu paths scored\n",
                t1 - t0, t2 - t1, st.seconds_total, st.seconds_scoring,
                (unsigned long long)st.extractions, (unsigned long long)st.evaluations,
                (unsigned long long)st.kernel_batches, (unsigned long long)st.paths_scored);
    return EXIT_SUCCESS;
}

// ---- evalPath (src/eval.cpp:196-242) ------------------------------------------------------------------

int run_evalpath(const Options &o) {
    Engine eng(o.device);   // CUDA start-up overlaps the parsing below
    Graph g;
    load_graph(o, g);
    // path string: components split at ';' / ',' (the separator stays on the component and is
    // dropped again), last character = orientation (src/eval.cpp:203-226)
    std::vector<uint32_t> path;
    std::string shown;
    {
        const std::string &ps = o.path;
        if (ps.empty()) die("Error: cannot handle starting gap. Terminating.");
        size_t p = 0;
        while (p <= ps.size()) {
            size_t q = ps.find_first_of(";,", p);
            std::string comp = ps.substr(p, q == std::string::npos ? q : q - p);
            if (comp.empty()) die("Error: empty path component. Terminating.");
            const char orient = comp.back();
            comp.pop_back();
            const int64_t id = g.names.find(comp);
            if (id < 0) die("Error: cannot find node (" + comp + "). Terminating.");
            path.push_back(((uint32_t)id << 1) | (orient == '-'));
            if (!shown.empty()) shown += ',';
            shown += g.names.name((uint32_t)id);
            shown += orient;
            if (q == std::string::npos) break;
            p = q + 1;
        }
    }
    const uint32_t n = (uint32_t)path.size();
    std::vector<uint32_t> ids(n);
    for (uint32_t i = 0; i < n; ++i) ids[i] = path[i] >> 1;
    std::sort(ids.begin(), ids.end());
    const uint32_t uniques = (uint32_t)(std::unique(ids.begin(), ids.end()) - ids.begin());
    Gaf gaf;
    if (!o.gaf.empty()) load_gaf(o, g.names, 0, false, gaf);
    else gaf.offsets.assign(1, 0);
    eng.get();
    eng.check(gfe_set_alignments(eng.e, gaf.offsets.data(), gaf.steps.data(), gaf.A, g.names.size()));
    std::vector<int32_t> tfw(gaf.A), trc(gaf.A);
    uint32_t bad = 0, good = 0;
    eng.check(gfe_score_pairs(eng.e, path.data(), n, 0, tfw.data(), trc.data(), &bad, &good));
    Out out(stdout);
    out.put(shown);
    out.put('\n');
    // traceback columns in batches; the reference prints the forward alignment only when it
    // scores strictly higher (src/eval.cpp:101)
    const uint64_t batch_bytes = 1ull << 28;
    std::vector<uint8_t> use_rc, ops;
    std::vector<uint64_t> first;
    std::vector<uint32_t> len;
    uint64_t a0 = 0;
    while (a0 < gaf.A) {
        uint64_t a1 = a0, bytes = 0;
        while (a1 < gaf.A && a1 - a0 < (1u << 22)) {
            const uint64_t add = (uint64_t)(gaf.offsets[a1 + 1] - gaf.offsets[a1]) + n;
            if (a1 > a0 && bytes + add > batch_bytes) break;
            bytes += add;
            ++a1;
        }
        const uint32_t count = (uint32_t)(a1 - a0);
        use_rc.resize(count);
        for (uint32_t k = 0; k < count; ++k) use_rc[k] = !(tfw[a0 + k] > trc[a0 + k]);
        ops.resize(bytes);
        first.resize(count);
        len.resize(count);
        eng.check(gfe_trace_pairs(eng.e, path.data(), n, a0, count, use_rc.data(), ops.data(), bytes,
                                  first.data(), len.data(), nullptr));
        for (uint32_t k = 0; k < count; ++k) {
            const uint64_t a = a0 + k;
            const uint32_t *bs = gaf.steps.data() + gaf.offsets[a];
            const uint32_t m = gaf.offsets[a + 1] - gaf.offsets[a];
            const uint8_t *op = ops.data() + first[k];
            uint32_t i = 0, j = 0;   // next path / alignment step
            // PairwisePathAlignment::getAlignment(idsToHeaders, true), include/alignments.h:111-120
            for (uint32_t c = 0; c < len[k]; ++c) {
                uint32_t b = 0;
                if (op[c] != 2) {
                    b = use_rc[k] ? (bs[m - 1 - j] ^ 1u) : bs[j];
                    ++j;
                }
                if (op[c] == 2) {
                    out.buf.append(g.names.name(path[i] >> 1).size() + 1, '-');
                } else if (op[c] == 0) {
                    out.buf.append(g.names.name(b >> 1).size() + 1, '.');
                } else {
                    out.put(g.names.name(b >> 1));
                    out.put((b & 1) ? '-' : '+');
                }
                out.put(',');
                if (op[c] != 3) ++i;
            }
            out.put('\t');
            out.put(gaf.qname(a));
            out.put('\t');
            out.num(std::max(tfw[a], trc[a]));
            out.put('\n');
            if (out.buf.size() > (1u << 21)) out.flush();
        }
        a0 = a1;
    }
    const int32_t alt = (int32_t)bad - (int32_t)good - (int32_t)uniques;
    out.num(bad); out.put('\t'); out.num(good); out.put('\t'); out.num(alt); out.put('\t');
    out.num(n); out.put('\t'); out.num(uniques); out.put('\n');
    out.flush();
    return EXIT_SUCCESS;
}

// ---- evalGFA (Input::read case 1, src/input-gfalign.cpp:80-96; evalGFA, src/eval.cpp:34-61) ---------

namespace {

// InAlignments::countSupplementary, src/alignments.cpp:329-351
void count_supplementary(const Gaf &gaf, std::vector<uint64_t> group, AlignmentSummary &s,
                         bool print_terminal, Out &out) {
    std::stable_sort(group.begin(), group.end(), [&](uint64_t x, uint64_t y) {
        return gaf.records[x].qStart < gaf.records[y].qStart;
    });
    unsigned int pos = 0, count = 0;
    for (uint64_t a : group) {
        const GafRecord &r = gaf.records[a];
        if (pos != 0 && r.qStart > pos) {
            ++s.supplementary;
            ++count;
        }
        pos = r.qEnd;
    }
    if (group.size() == 2 && count == 1) {
        const GafRecord &l = gaf.records[group[0]], &r = gaf.records[group[1]];
        if (l.pEnd >= l.pLen - 500 && r.pStart <= 500) {   // unsigned arithmetic, as the reference
            ++s.terminalSupplementary;
            if (print_terminal) {
                print_record(gaf, group[0], out);
                print_record(gaf, group[1], out);
            }
        }
    }
}

// InAlignments::markDuplicates, src/alignments.cpp:305-327 (the running `alignments` vector also
// holds the primaries seen since the last flush — kept as is)
void mark_duplicates(const Gaf &gaf, const std::vector<uint64_t> &order, AlignmentSummary &s,
                     bool print_terminal, Out &out) {
    std::string_view prev;
    bool have_prev = false;
    std::vector<uint64_t> group;
    for (size_t i = 0; i < order.size(); ++i) {
        const uint64_t a = order[i];
        group.push_back(a);
        // prevQname starts as the empty string: a first record with an empty name is "secondary"
        const bool same = have_prev ? gaf.qname(a) == prev : gaf.qname(a).empty();
        if (same) {
            ++s.secondary;
            if (i + 1 == order.size() || gaf.qname(order[i + 1]) != gaf.qname(a)) {
                count_supplementary(gaf, group, s, print_terminal, out);
                group.clear();
            }
        } else {
            ++s.primary;
            prev = gaf.qname(a);
            have_prev = true;
        }
    }
}

}  // namespace

int run_evalgfa(const Options &o) {
    // CUDA start-up overlaps the parsing and the alignment summary; not needed at all without -o
    std::unique_ptr<Engine> engp;
    if (!o.out_file.empty()) engp.reset(new Engine(o.device));
    Graph g;
    load_graph(o, g);
    if (o.gaf.empty()) return EXIT_SUCCESS;   // nothing to do without alignments (src/input-gfalign.cpp:83,92)
    Gaf gaf;
    load_gaf(o, g.names, 0, true, gaf);
    Out out(stdout);
    {
        // sortAlignmentsByNameAscending + markDuplicates + printStats (src/input-gfalign.cpp:84-90)
        std::vector<uint64_t> order(gaf.A);
        std::iota(order.begin(), order.end(), 0);
        std::stable_sort(order.begin(), order.end(),
                         [&](uint64_t x, uint64_t y) { return gaf.qname(x) < gaf.qname(y); });
        AlignmentSummary s;
        s.n = gaf.A;
        add_totals(gaf, s);
        mark_duplicates(gaf, order, s, o.terminal_alignments, out);
        out.flush();
        print_stats(s, false);
        fflush(stdout);
    }
    if (o.out_file.empty()) return EXIT_SUCCESS;
    Engine &eng = *engp;
    eng.get();
    eng.check(gfe_set_alignments(eng.e, gaf.offsets.data(), gaf.steps.data(), gaf.A, g.names.size()));
    const uint32_t E = (uint32_t)g.links.size();
    std::vector<uint32_t> w(E);
    eng.check(gfe_edge_support(eng.e, g.edge_from.data(), g.edge_to.data(), E, w.data()));
    // decorated GFA: header, S records, L records + RC:i:<w> (Tag appended at src/eval.cpp:58-59),
    // then every other record.  The reference re-serialises the whole graph through gfalibs'
    // writer; here S / L / other records are passed through as read (P overlaps become '*' as
    // gfalibs prints them) — the L records and their RC tags are the part this engine owns.
    FILE *f = fopen(o.out_file.c_str(), "w");
    if (!f) die("Error: cannot write " + o.out_file);
    {
        Out fo(f);
        const bool gz = o.out_file.size() > 3 && o.out_file.compare(o.out_file.size() - 3, 3, ".gz") == 0;
        if (gz) {
            fclose(f);
            fo.gz = gzopen(o.out_file.c_str(), "wb");
            if (!fo.gz) die("Error: cannot write " + o.out_file);
        }
        fo.put("H\tVN:Z:1.2\n");
        const char *base = g.bytes.data, *end = g.bytes.data + g.bytes.size;
        auto each_line = [&](auto &&fn) {
            for (const char *p = base; p < end;) {
                const void *q = memchr(p, '\n', (size_t)(end - p));
                const char *eol = q ? static_cast<const char *>(q) : end;
                if (eol > p) fn(std::string_view(p, (size_t)(eol - p)));
                p = eol + 1;
            }
        };
        each_line([&](std::string_view l) {
            if (l[0] == 'S') { fo.put(l); fo.put('\n'); }
        });
        for (uint32_t i = 0; i < E; ++i) {
            fo.put(std::string_view(base + g.links[i].line_off, g.links[i].line_len));
            fo.put("\tRC:i:");
            fo.num(w[i]);
            fo.put('\n');
        }
        each_line([&](std::string_view l) {
            if (l[0] == 'S' || l[0] == 'L' || l[0] == 'H') return;
            if (l[0] == 'P') {   // name, steps, '*' for the overlaps, tags
                size_t t1 = l.find('\t'), t2 = t1 == l.npos ? t1 : l.find('\t', t1 + 1),
                       t3 = t2 == l.npos ? t2 : l.find('\t', t2 + 1);
                if (t3 != l.npos) {
                    size_t t4 = l.find('\t', t3 + 1);
                    fo.put(l.substr(0, t3));
                    fo.put("\t*");
                    if (t4 != l.npos) fo.put(l.substr(t4));
                    fo.put('\n');
                    return;
                }
            }
            fo.put(l);
            fo.put('\n');
        });
        fo.flush();
        if (fo.gz) gzclose(fo.gz);
        else fclose(f);
    }
    return EXIT_SUCCESS;
}

// ---- filter (Input::read case 4, src/input-gfalign.cpp:113-119) -----------------------------------

int run_filter(const Options &o) {
    Engine eng(o.device);   // CUDA start-up overlaps the parsing below
    if (o.gaf.empty()) return EXIT_SUCCESS;
    // readNodelist: every line of the node file is a name (no column split, src/input-gfalign.cpp:62-72)
    NameTable set;
    if (!o.node_file.empty()) {
        FileBytes nf;
        std::string err;
        if (!nf.open(o.node_file, err)) die("Error: " + err);
        for (const char *p = nf.data, *end = nf.data + nf.size; p < end;) {
            const void *q = memchr(p, '\n', (size_t)(end - p));
            const char *eol = q ? static_cast<const char *>(q) : end;
            set.intern(std::string_view(p, (size_t)(eol - p)));
            p = eol + 1;
        }
    }
    set.freeze();
    const uint32_t N = set.size();
    Gaf gaf;
    load_gaf(o, set, N, true, gaf);   // names outside the list share the id N
    eng.get();
    eng.check(gfe_set_alignments(eng.e, gaf.offsets.data(), gaf.steps.data(), gaf.A, N + 1));
    std::vector<uint32_t> nodes(N);
    std::iota(nodes.begin(), nodes.end(), 0);
    std::vector<uint8_t> keep(gaf.A);
    eng.check(gfe_filter_alignments(eng.e, nodes.data(), N, (int32_t)o.min_nodes, keep.data()));
    if (o.out_file.empty()) return EXIT_SUCCESS;
    // outputAlignments (src/alignments.cpp:283-303): a file name is always a file here; the stats
    // are printed first (their totals still cover every record that was read), in fixed notation
    AlignmentSummary s;
    add_totals(gaf, s);
    for (uint64_t a = 0; a < gaf.A; ++a) s.n += keep[a];
    FILE *f = fopen(o.out_file.c_str(), "w");
    if (!f) die("Error: cannot write " + o.out_file);
    print_stats(s, true);
    fflush(stdout);
    {
        Out fo(f);
        for (uint64_t a = 0; a < gaf.A; ++a)
            if (keep[a]) print_record(gaf, a, fo);
    }
    fclose(f);
    return EXIT_SUCCESS;
}

}  // namespace gfh
