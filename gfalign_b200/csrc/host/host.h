// gfalign_b200/csrc/host/host.h — host side of the drop-in CLI: text -> packed arrays.
//
// Replaces, for the hot path only, what the reference does with gfalibs' readGFA
// (gfalibs/src/input-gfa.cpp:497-641: S and L records, name -> id tables), InAlignments::load
// (src/alignments.cpp:143-235: GAF records), InAlignment::GAFpathToPath / GAFpathToEdges
// (src/alignments.cpp:75-94, 409-445: walk string -> oriented steps) and NodeTable's constructor
// (include/nodetable.h:16-42).  Everything lands in flat arrays: oriented step =
// node << 1 | (orientation == '-') over dense node ids, alignments as CSR.
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <string_view>
#include <vector>

#include "../../../include/gfalign_b200.h"

namespace gfh {

// Read-only view of a whole file (mmap for regular files; gzip'd files are inflated to memory).
struct FileBytes {
    const char *data = nullptr;
    size_t size = 0;
    bool open(const std::string &path, std::string &err);
    void close();
    ~FileBytes() { close(); }
    FileBytes() = default;
    FileBytes(const FileBytes &) = delete;
    FileBytes &operator=(const FileBytes &) = delete;

private:
    void *map_ = nullptr;
    size_t map_len_ = 0;
    std::vector<char> owned_;
};

// name -> dense id; open addressing over string_views that point into `names`.
class NameTable {
public:
    uint32_t intern(std::string_view name);          // id of name, appended if new
    int64_t find(std::string_view name) const;       // -1 if unknown
    uint32_t size() const { return (uint32_t)names_.size(); }
    const std::string &name(uint32_t id) const { return names_[id]; }
    void freeze();                                    // rebuilds slots; find() is then thread-safe

private:
    static uint64_t hash(std::string_view s);
    void grow();
    std::vector<std::string> names_;
    std::vector<uint32_t> slots_;                     // id + 1, 0 = empty
    std::vector<uint64_t> hashes_;
    uint64_t mask_ = 0;
};

struct LinkLine {            // one L record, file order
    uint32_t from, to;       // oriented steps
    uint64_t line_off;       // where the record's text starts in the GFA bytes
    uint32_t line_len;       // without the newline
};

struct Graph {
    FileBytes bytes;                      // kept: evalGFA re-emits the records
    NameTable names;                      // segments in order of first appearance (S, then L)
    std::vector<uint8_t> has_segment;     // node has an S record
    std::vector<LinkLine> links;
    std::vector<uint32_t> edge_from, edge_to;
    // InSequences::buildEdgeGraph (gfalibs/src/gfa.cpp:1076-1097) as CSR over nodes
    std::vector<uint32_t> adj_offsets, adj_target;
    std::vector<uint8_t> adj_or0, adj_or1;
    bool load(const std::string &path, std::string &err);
    void build_adjacency();
    gfe_graph view() const;
    std::string step_name(uint32_t step) const {   // "name+" / "name-"
        return names.name(step >> 1) + ((step & 1) ? '-' : '+');
    }
};

struct GafRecord {           // numeric columns 2..12 of a GAF line (InAlignment, include/alignments.h:124-141)
    uint32_t qLen, qStart, qEnd, pLen, pStart, pEnd, matches, blockLen, mapq;
    char strand;
};

struct Gaf {
    FileBytes bytes;
    uint64_t A = 0;
    std::vector<uint32_t> offsets, steps;       // CSR, file order
    std::vector<uint64_t> line_off;             // A: start of each record
    std::vector<uint32_t> line_len, qname_len;  // A
    std::vector<GafRecord> records;             // A when keep_records
    // Walk names are looked up in `names` (frozen); a name it does not hold becomes `unknown_id`:
    // 0 for search / evalPath / evalGFA, as headersToIds[...] default-inserts 0
    // (src/alignments.cpp:86); names.size() for filter mode, where the table is the nodelist.
    bool load(const std::string &path, const NameTable &names, uint32_t unknown_id, int threads,
              bool keep_records, std::string &err);
    std::string_view qname(uint64_t a) const { return {bytes.data + line_off[a], qname_len[a]}; }
    std::string_view line(uint64_t a) const { return {bytes.data + line_off[a], line_len[a]}; }
};

struct NodeBudget {          // NodeTable (include/nodetable.h) after the two add() calls of dijkstra
    std::vector<uint32_t> nodes, counts;
    uint32_t node_count = 0;
};
// returns false with err set to the reference's message for an unknown node
bool load_nodelist(const std::string &path, const Graph &g, uint32_t source, uint32_t destination,
                   NodeBudget &out, std::string &err);

int hardware_threads();

}  // namespace gfh
