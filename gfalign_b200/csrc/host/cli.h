// gfalign_b200/csrc/host/cli.h — options and modes of the drop-in CLI `gfalign-b200`
// (same tools, options and stdout as the reference's src/main.cpp for search / evalPath / evalGFA
// / filter; the scoring runs on the GPU through the C ABI of include/gfalign_b200.h).
#pragma once
#include <cstdint>
#include <string>

namespace gfh {

struct Options {                     // UserInputGfalign, reference include/input-gfalign.h:4-14
    int mode = -1;                   // 1 evalGFA, 3 search, 4 filter, 5 evalPath
    std::string gfa, gaf, node_file, out_file, source, destination, path;
    uint32_t max_steps = 100000;     // dijkstraSteps
    uint32_t min_nodes = 0;
    bool return_all_paths = false;
    bool terminal_alignments = false, sort_alignment = false, align_stats = false;
    int threads = 0;                 // -j: host threads of the text packers (0 = all cores)
    int device = 0;                  // --device (extension): CUDA device
    uint32_t speculate = 0;          // --speculate (extension): candidate paths per kernel batch
    bool verbose = false, cmd = false, timing = false;
};

int run_search(const Options &o);
int run_evalpath(const Options &o);
int run_evalgfa(const Options &o);
int run_filter(const Options &o);

}  // namespace gfh
