// gfalign_b200/csrc/host/pack.cpp — GFA / GAF / nodelist text -> packed arrays (see host.h).
#include "host.h"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>

#include <algorithm>
#include <cstring>
#include <thread>

namespace gfh {

int hardware_threads() {
    unsigned n = std::thread::hardware_concurrency();
    return n ? (int)n : 1;
}

// ---- FileBytes ----------------------------------------------------------------------------------

bool FileBytes::open(const std::string &path, std::string &err) {
    close();
    int fd = ::open(path.c_str(), O_RDONLY);
    if (fd < 0) {
        err = "cannot open " + path;
        return false;
    }
    unsigned char magic[2] = {0, 0};
    ssize_t got = ::pread(fd, magic, 2, 0);
    struct stat st;
    const bool regular = fstat(fd, &st) == 0 && S_ISREG(st.st_mode);
    if (got == 2 && magic[0] == 0x1f && magic[1] == 0x8b) {   // gzip: inflate to memory
        ::close(fd);
        gzFile gz = gzopen(path.c_str(), "rb");
        if (!gz) {
            err = "cannot open " + path;
            return false;
        }
        std::vector<char> buf(1 << 20);
        int n;
        while ((n = gzread(gz, buf.data(), (unsigned)buf.size())) > 0)
            owned_.insert(owned_.end(), buf.data(), buf.data() + n);
        gzclose(gz);
        data = owned_.data();
        size = owned_.size();
        return true;
    }
    if (regular && st.st_size > 0) {
        void *m = mmap(nullptr, (size_t)st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
        ::close(fd);
        if (m == MAP_FAILED) {
            err = "cannot map " + path;
            return false;
        }
        madvise(m, (size_t)st.st_size, MADV_SEQUENTIAL);
        map_ = m;
        map_len_ = (size_t)st.st_size;
        data = static_cast<const char *>(m);
        size = map_len_;
        return true;
    }
    // empty file, pipe or other non-regular input: read it
    char buf[1 << 16];
    ssize_t n;
    while ((n = ::read(fd, buf, sizeof buf)) > 0) owned_.insert(owned_.end(), buf, buf + n);
    ::close(fd);
    data = owned_.data();
    size = owned_.size();
    return true;
}

void FileBytes::close() {
    if (map_) munmap(map_, map_len_);
    map_ = nullptr;
    map_len_ = 0;
    owned_.clear();
    data = nullptr;
    size = 0;
}

// ---- NameTable ----------------------------------------------------------------------------------

uint64_t NameTable::hash(std::string_view s) {
    uint64_t h = 0x9E3779B97F4A7C15ull ^ (s.size() * 0xff51afd7ed558ccdull);
    const char *p = s.data();
    size_t n = s.size();
    while (n >= 8) {
        uint64_t w;
        memcpy(&w, p, 8);
        h = (h ^ w) * 0xff51afd7ed558ccdull;
        h ^= h >> 32;
        p += 8;
        n -= 8;
    }
    uint64_t w = 0;
    memcpy(&w, p, n);
    h = (h ^ w) * 0xc4ceb9fe1a85ec53ull;
    h ^= h >> 29;
    return h;
}

void NameTable::grow() {
    size_t cap = slots_.empty() ? 1024 : slots_.size() * 2;
    while (cap < names_.size() * 2 + 16) cap *= 2;
    slots_.assign(cap, 0);
    mask_ = cap - 1;
    for (uint32_t id = 0; id < names_.size(); ++id) {
        uint64_t i = hashes_[id] & mask_;
        while (slots_[i]) i = (i + 1) & mask_;
        slots_[i] = id + 1;
    }
}

void NameTable::freeze() { grow(); }

int64_t NameTable::find(std::string_view name) const {
    if (slots_.empty()) return -1;
    const uint64_t h = hash(name);
    uint64_t i = h & mask_;
    while (uint32_t s = slots_[i]) {
        const uint32_t id = s - 1;
        if (hashes_[id] == h && names_[id].size() == name.size() &&
            memcmp(names_[id].data(), name.data(), name.size()) == 0)
            return id;
        i = (i + 1) & mask_;
    }
    return -1;
}

uint32_t NameTable::intern(std::string_view name) {
    int64_t f = find(name);
    if (f >= 0) return (uint32_t)f;
    if ((names_.size() + 1) * 2 > slots_.size()) {
        names_.emplace_back(name);
        hashes_.push_back(hash(name));
        grow();
        return (uint32_t)names_.size() - 1;
    }
    const uint64_t h = hash(name);
    names_.emplace_back(name);
    hashes_.push_back(h);
    uint64_t i = h & mask_;
    while (slots_[i]) i = (i + 1) & mask_;
    slots_[i] = (uint32_t)names_.size();
    return (uint32_t)names_.size() - 1;
}

// ---- line / column helpers ----------------------------------------------------------------------

namespace {

inline const char *find_eol(const char *p, const char *end) {
    const void *q = memchr(p, '\n', (size_t)(end - p));
    return q ? static_cast<const char *>(q) : end;
}

// next tab-delimited column of [p, eol); advances p past the tab (readDelimited semantics:
// the last column runs to the end of the line, empty columns are kept)
inline std::string_view next_col(const char *&p, const char *eol, bool &more) {
    const void *q = memchr(p, '\t', (size_t)(eol - p));
    if (!q) {
        std::string_view v(p, (size_t)(eol - p));
        p = eol;
        more = false;
        return v;
    }
    const char *t = static_cast<const char *>(q);
    std::string_view v(p, (size_t)(t - p));
    p = t + 1;
    more = true;
    return v;
}

// std::stoi on a column: optional blanks, sign, digits; anything after the digits is ignored.
inline bool parse_int(std::string_view v, int64_t &out) {
    size_t i = 0;
    while (i < v.size() && (v[i] == ' ' || (v[i] >= '\t' && v[i] <= '\r'))) ++i;
    bool neg = false;
    if (i < v.size() && (v[i] == '+' || v[i] == '-')) neg = v[i++] == '-';
    if (i >= v.size() || v[i] < '0' || v[i] > '9') return false;
    int64_t x = 0;
    while (i < v.size() && v[i] >= '0' && v[i] <= '9') {
        x = x * 10 + (v[i++] - '0');
        if (x > 0x7fffffffll + 1) return false;   // stoi would throw out_of_range
    }
    out = neg ? -x : x;
    return !(out > 0x7fffffffll);
}

}  // namespace

// ---- GFA ----------------------------------------------------------------------------------------


bool Graph::load(const std::string &path, std::string &err) {
    if (!bytes.open(path, err)) return false;
    const char *p = bytes.data, *end = bytes.data + bytes.size;
    // GFA2 detection as the reference does it: version from the H record, else from the first
    // record type (gfalibs/src/input-gfa.cpp:165-205)
    if (p < end) {
        const char *eol = find_eol(p, end);
        std::string_view first(p, (size_t)(eol - p));
        bool gfa2 = false;
        if (first.size() >= 1 && first[0] == 'H') {
            size_t v = first.find("VN:Z:");
            if (v != std::string_view::npos && v + 5 < first.size() && first[v + 5] == '2') gfa2 = true;
        } else if (!first.empty() && (first[0] == 'G' || first[0] == 'O' || first[0] == 'E')) {
            gfa2 = true;
        }
        if (gfa2) {
            err = "GFA2 input is not supported by gfalign-b200 (convert to GFA1)";
            return false;
        }
    }
    // pass 1: S records name the segments in file order
    for (const char *q = p; q < end;) {
        const char *eol = find_eol(q, end);
        if (eol - q >= 2 && q[0] == 'S' && q[1] == '\t') {
            const char *c = q + 2;
            bool more;
            std::string_view nm = next_col(c, eol, more);
            names.intern(nm);
        }
        q = eol + 1;
    }
    has_segment.assign(names.size(), 1);
    // pass 2: L records; a name first seen here gets the next id (input-gfa.cpp:593-617)
    for (const char *q = p; q < end;) {
        const char *eol = find_eol(q, end);
        if (eol - q >= 2 && q[0] == 'L' && q[1] == '\t') {
            const char *c = q + 2;
            bool more;
            std::string_view n1 = next_col(c, eol, more);
            std::string_view o1 = more ? next_col(c, eol, more) : std::string_view();
            std::string_view n2 = more ? next_col(c, eol, more) : std::string_view();
            std::string_view o2 = more ? next_col(c, eol, more) : std::string_view();
            if (o1.empty() || o2.empty()) {
                err = "malformed L record in " + path;
                return false;
            }
            LinkLine l;
            l.from = names.intern(n1) << 1 | (o1[0] == '-');
            l.to = names.intern(n2) << 1 | (o2[0] == '-');
            l.line_off = (uint64_t)(q - bytes.data);
            l.line_len = (uint32_t)(eol - q);
            links.push_back(l);
        }
        q = eol + 1;
    }
    has_segment.resize(names.size(), 0);
    names.freeze();
    edge_from.resize(links.size());
    edge_to.resize(links.size());
    for (size_t i = 0; i < links.size(); ++i) {
        edge_from[i] = links[i].from;
        edge_to[i] = links[i].to;
    }
    build_adjacency();
    return true;
}

// InSequences::buildEdgeGraph (gfalibs/src/gfa.cpp:1076-1097): per L record the forward entry
// {or1, id2, or2} at id1 unless already there (then the record is skipped), and the reverse entry
// {!or2, id1, !or1} at id2 with no presence check; rows keep insertion order.
void Graph::build_adjacency() {
    struct Ent { uint32_t target; uint8_t or0, or1; };
    const uint32_t n = names.size();
    std::vector<std::vector<Ent>> rows(n);
    for (const LinkLine &l : links) {
        const uint32_t a = l.from >> 1, b = l.to >> 1;
        const uint8_t oa = l.from & 1, ob = l.to & 1;
        bool present = false;
        for (const Ent &e : rows[a])
            if (e.target == b && e.or0 == oa && e.or1 == ob) { present = true; break; }
        if (present) continue;
        rows[a].push_back({b, oa, ob});
        rows[b].push_back({a, (uint8_t)(ob ^ 1), (uint8_t)(oa ^ 1)});
    }
    adj_offsets.assign(n + 1, 0);
    for (uint32_t i = 0; i < n; ++i) adj_offsets[i + 1] = adj_offsets[i] + (uint32_t)rows[i].size();
    adj_target.resize(adj_offsets[n]);
    adj_or0.resize(adj_offsets[n]);
    adj_or1.resize(adj_offsets[n]);
    for (uint32_t i = 0; i < n; ++i) {
        uint32_t k = adj_offsets[i];
        for (const Ent &e : rows[i]) {
            adj_target[k] = e.target;
            adj_or0[k] = e.or0;
            adj_or1[k] = e.or1;
            ++k;
        }
    }
}

gfe_graph Graph::view() const {
    gfe_graph g;
    g.n_nodes = names.size();
    g.adj_offsets = adj_offsets.data();
    g.adj_target = adj_target.data();
    g.adj_or0 = adj_or0.data();
    g.adj_or1 = adj_or1.data();
    return g;
}


// ---- GAF ----------------------------------------------------------------------------------------


namespace {

struct GafChunk {
    std::vector<uint32_t> steps, lens, line_len, qname_len;
    std::vector<uint64_t> line_off;
    std::vector<GafRecord> records;
    std::string err;
};

// `>a<b...` -> oriented steps, one per piece that starts at a '>' / '<' (or at the string start);
// the piece's first character is the orientation mark whatever it is: '>' -> '+', else '-'
// (InAlignment::GAFpathToPath, src/alignments.cpp:75-94)
inline void pack_walk(std::string_view walk, const NameTable &names, uint32_t unknown_id,
                      std::vector<uint32_t> &out) {
    size_t i = 0;
    const size_t n = walk.size();
    while (i < n) {
        size_t j = i + 1;
        while (j < n && walk[j] != '>' && walk[j] != '<') ++j;
        const int64_t id = names.find(walk.substr(i + 1, j - i - 1));
        out.push_back(((id < 0 ? unknown_id : (uint32_t)id) << 1) | (walk[i] != '>'));
        i = j;
    }
}

void parse_gaf_chunk(const char *base, const char *p, const char *end, const NameTable &names,
                     uint32_t unknown_id, bool keep_records, GafChunk &c) {
    while (p < end) {
        const char *eol = find_eol(p, end);
        if (eol == p) {   // empty line
            p = eol + 1;
            continue;
        }
        const char *q = p;
        bool more;
        std::string_view col[12];
        int nc = 0;
        do {
            col[nc++] = next_col(q, eol, more);
        } while (more && nc < 12);
        if (nc < 12) {
            c.err = "GAF record with fewer than 12 columns: " + std::string(p, (size_t)(eol - p));
            return;
        }
        const size_t before = c.steps.size();
        pack_walk(col[5], names, unknown_id, c.steps);
        c.lens.push_back((uint32_t)(c.steps.size() - before));
        c.line_off.push_back((uint64_t)(p - base));
        c.line_len.push_back((uint32_t)(eol - p));
        c.qname_len.push_back((uint32_t)col[0].size());
        if (keep_records) {
            static const int idx[9] = {1, 2, 3, 6, 7, 8, 9, 10, 11};
            int64_t v[9];
            for (int k = 0; k < 9; ++k)
                if (!parse_int(col[idx[k]], v[k])) {
                    c.err = "GAF record with a non-numeric column: " + std::string(p, (size_t)(eol - p));
                    return;
                }
            GafRecord r;
            r.qLen = (uint32_t)v[0]; r.qStart = (uint32_t)v[1]; r.qEnd = (uint32_t)v[2];
            r.pLen = (uint32_t)v[3]; r.pStart = (uint32_t)v[4]; r.pEnd = (uint32_t)v[5];
            r.matches = (uint32_t)v[6]; r.blockLen = (uint32_t)v[7]; r.mapq = (uint32_t)v[8];
            r.strand = col[4].empty() ? '\0' : col[4][0];
            c.records.push_back(r);
        }
        p = eol + 1;
    }
}

}  // namespace

bool Gaf::load(const std::string &path, const NameTable &names, uint32_t unknown_id, int threads,
               bool keep_records, std::string &err) {
    if (!bytes.open(path, err)) return false;
    const char *base = bytes.data, *end = bytes.data + bytes.size;
    if (threads < 1) threads = 1;
    const size_t min_chunk = 1u << 20;
    int T = (int)std::min<size_t>((size_t)threads, bytes.size / min_chunk + 1);
    // chunk boundaries on record starts
    std::vector<const char *> cut(T + 1);
    cut[0] = base;
    cut[T] = end;
    for (int t = 1; t < T; ++t) {
        const char *p = base + bytes.size / T * t;
        if (p < cut[t - 1]) p = cut[t - 1];
        const char *e = find_eol(p, end);
        cut[t] = e < end ? e + 1 : end;
    }
    std::vector<GafChunk> chunks(T);
    {
        std::vector<std::thread> pool;
        for (int t = 1; t < T; ++t)
            pool.emplace_back(parse_gaf_chunk, base, cut[t], cut[t + 1], std::cref(names),
                              unknown_id, keep_records, std::ref(chunks[t]));
        parse_gaf_chunk(base, cut[0], cut[1], names, unknown_id, keep_records, chunks[0]);
        for (auto &th : pool) th.join();
    }
    uint64_t tot_a = 0, tot_s = 0;
    std::vector<uint64_t> a0(T), s0(T);
    for (int t = 0; t < T; ++t) {
        if (!chunks[t].err.empty()) {
            err = chunks[t].err;
            return false;
        }
        a0[t] = tot_a;
        s0[t] = tot_s;
        tot_a += chunks[t].lens.size();
        tot_s += chunks[t].steps.size();
    }
    if (tot_s > 0xFFFFFFFFull || tot_a >= 0xFFFFFFFFull) {
        err = "GAF too large for 32-bit step offsets";
        return false;
    }
    A = tot_a;
    offsets.resize(A + 1);
    steps.resize(tot_s);
    line_off.resize(A);
    line_len.resize(A);
    qname_len.resize(A);
    if (keep_records) records.resize(A);
    offsets[0] = 0;
    auto merge = [&](int t) {
        const GafChunk &c = chunks[t];
        uint64_t s = s0[t];
        for (size_t i = 0; i < c.lens.size(); ++i) {
            s += c.lens[i];
            offsets[a0[t] + i + 1] = (uint32_t)s;
        }
        if (!c.steps.empty()) memcpy(&steps[s0[t]], c.steps.data(), c.steps.size() * 4);
        if (!c.lens.empty()) {
            memcpy(&line_off[a0[t]], c.line_off.data(), c.lens.size() * 8);
            memcpy(&line_len[a0[t]], c.line_len.data(), c.lens.size() * 4);
            memcpy(&qname_len[a0[t]], c.qname_len.data(), c.lens.size() * 4);
            if (keep_records)
                memcpy(&records[a0[t]], c.records.data(), c.lens.size() * sizeof(GafRecord));
        }
    };
    {
        std::vector<std::thread> pool;
        for (int t = 1; t < T; ++t) pool.emplace_back(merge, t);
        merge(0);
        for (auto &th : pool) th.join();
    }
    return true;
}

// ---- nodelist -----------------------------------------------------------------------------------

// NodeTable(nodeFile, headersToIds) then add(source), add(destination) (src/eval.cpp:125-128,
// include/nodetable.h:16-54): count < 1 drops the line, nodeCount grows by every kept line's count
// (and by the two add() calls) whether or not the record was new; insert() keeps the first record.
bool load_nodelist(const std::string &path, const Graph &g, uint32_t source, uint32_t destination,
                   NodeBudget &out, std::string &err) {
    FileBytes f;
    if (!f.open(path, err)) return false;
    std::vector<int64_t> slot(g.names.size(), -1);
    auto insert = [&](uint32_t node, uint32_t count) {
        if (slot[node] >= 0) return;
        slot[node] = (int64_t)out.nodes.size();
        out.nodes.push_back(node);
        out.counts.push_back(count);
    };
    const char *p = f.data, *end = f.data + f.size;
    while (p < end) {
        const char *eol = find_eol(p, end);
        const char *q = p;
        bool more;
        std::string_view name = next_col(q, eol, more);
        uint32_t count = 1;
        if (more) {
            std::string_view c = next_col(q, eol, more);
            int64_t v;
            if (!parse_int(c, v)) {
                err = "Error: invalid count in node file (" + std::string(name) + ")";
                return false;
            }
            count = (uint32_t)(int32_t)v;
            if (count < 1) {
                p = eol + 1;
                continue;
            }
        }
        out.node_count += count;
        const int64_t id = g.names.find(name);
        if (id < 0) {
            err = "Error: node not in graph (pIUd: " + std::string(name) + ")";
            return false;
        }
        insert((uint32_t)id, count);
        p = eol + 1;
    }
    insert(source, 1);
    out.node_count += 1;
    insert(destination, 1);
    out.node_count += 1;
    return true;
}

}  // namespace gfh
