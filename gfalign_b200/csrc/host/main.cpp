// gfalign_b200/csrc/host/main.cpp — `gfalign-b200 <tool> [options]`: drop-in for the reference CLI
// (src/main.cpp) on the tools whose hot path runs on the GPU: search, evalPath, evalGFA, filter.
// `align` (external GraphAligner) and `subgraph` are outside this engine's scope and say so.
#include <getopt.h>
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "cli.h"

using gfh::Options;

static void print_tools() {
    printf("gfalign-b200 [options] [tool] [arguments]\n-h for additional help.\n");
    printf("\nTools:\n");
    printf("evalGFA\n");
    printf("evalPath\n");
    printf("search\n");
    printf("filter\n");
    exit(0);
}

static void need_file(const char *p) {   // ifFileExists, gfalibs/include/functions.h:264-276
    if (access(p, F_OK)) {
        printf("Error - file does not exist: %s\n", p);
        exit(1);
    }
}

static void version_and_exit() {
    printf("gfalign-b200 v0.1 (gfalign v0.0.1 semantics, sm_100a engine)\n");
    exit(0);
}

static void help_and_exit(int mode) {
    switch (mode) {
        case 1:
            printf("gfalign-b200 evalGFA [options]\n\nOptions:\n");
            printf("-f --input-sequence sequence input file (GFA1).\n");
            printf("-g --input-alignment alignment input file (currently supports: GAF).\n");
            printf("-o --out-format ouput to file (currently supports: GFA).\n");
            printf("--sort-alignment output alignment sorted by query name.\n");
            printf("--output-terminal-alignments output terminal alignments.\n");
            break;
        case 3:
            printf("gfalign-b200 search [options]\nOptions:\n");
            printf("-d --destination <string> destination node.\n");
            printf("-f --input-sequence <filename> sequence input file (GFA1).\n");
            printf("-g --input-alignment alignment input file (currently supports: GAF).\n");
            printf("-m --max-steps <int> limit graph exploration.\n");
            printf("-n --node-file <filename> list of nodes available to the search.\n");
            printf("-s --source <string> source node.\n");
            printf("--return-all-paths return all viable paths as they are discovered, not only better ones (default: false).\n");
            printf("--min-nodes <int> do not report paths with less than int nodes (default: 0).\n");
            break;
        case 4:
            printf("gfalign-b200 filter [options]\nOptions:\n");
            printf("-g --input-alignment alignment input file (currently supports: GAF).\n");
            printf("-n --node-file <filename> list of nodes available to the search.\n");
            printf("-o --out-format ouput to file or stdout (currently supports: GAF).\n");
            printf("--min-nodes <int> retain alignments mapping to at least int nodes.\n");
            break;
        case 5:
            printf("gfalign-b200 evalPath [options]\nOptions:\n");
            printf("-p --path in GFA format.\n");
            printf("-f --input-sequence <filename> sequence input file (GFA1).\n");
            printf("-g --input-alignment alignment input file (currently supports: GAF).\n");
            break;
    }
    printf("-j --threads <int> host threads for parsing (default: all).\n");
    printf("--device <int> CUDA device (default: 0).\n");
    exit(0);
}

int main(int argc, char **argv) {
    if (argc == 1) print_tools();
    Options o;
    const std::string tool = argv[1];
    if (tool == "evalGFA") o.mode = 1;
    else if (tool == "search") o.mode = 3;
    else if (tool == "filter") o.mode = 4;
    else if (tool == "evalPath") o.mode = 5;
    else if (tool == "align" || tool == "subgraph") {
        fprintf(stderr, "mode '%s' is not part of gfalign-b200 (scoring engine only); use gfalign.\n",
                argv[1]);
        return EXIT_FAILURE;
    } else {
        fprintf(stderr, "mode '%s' does not exist. Terminating.\n", argv[1]);
        return EXIT_FAILURE;
    }
    int flag_all = 0, flag_sort = 0, flag_term = 0, flag_verbose = 0, flag_cmd = 0, flag_stats = 0,
        flag_timing = 0;
    static struct option long_options[] = {
        {"destination", required_argument, 0, 'd'},
        {"input-sequence", required_argument, 0, 'f'},
        {"input-alignment", required_argument, 0, 'g'},
        {"max-steps", required_argument, 0, 'm'},
        {"node-file", required_argument, 0, 'n'},
        {"out-format", required_argument, 0, 'o'},
        {"path", required_argument, 0, 'p'},
        {"source", required_argument, 0, 's'},
        {"threads", required_argument, 0, 'j'},
        {"min-nodes", required_argument, 0, 1},
        {"device", required_argument, 0, 2},
        {"speculate", required_argument, 0, 3},
        {"return-all-paths", no_argument, &flag_all, 1},
        {"graph-statistics", no_argument, &flag_stats, 1},
        {"sort-alignment", no_argument, &flag_sort, 1},
        {"output-terminal-alignments", no_argument, &flag_term, 1},
        {"timing", no_argument, &flag_timing, 1},
        {"cmd", no_argument, &flag_cmd, 1},
        {"verbose", no_argument, &flag_verbose, 1},
        {"version", no_argument, 0, 'v'},
        {"help", no_argument, 0, 'h'},
        {0, 0, 0, 0}};
    // each tool accepts the reference's option letters for that tool (src/main.cpp:254-640)
    const char *shorts = o.mode == 1 ? "-:f:g:j:o:vh"
                       : o.mode == 3 ? "-:d:f:g:j:m:n:o:s:vh"
                       : o.mode == 4 ? "-:g:j:n:o:vh"
                                     : "-:p:f:g:j:vh";
    for (;;) {
        int idx = 0;
        const int c = getopt_long(argc, argv, shorts, long_options, &idx);
        if (c == -1) break;
        switch (c) {
            case 'd': o.destination = optarg; break;
            case 'f': need_file(optarg); o.gfa = optarg; break;
            case 'g': need_file(optarg); o.gaf = optarg; o.align_stats = true; break;
            case 'm': o.max_steps = (uint32_t)atoi(optarg); break;
            case 'n': need_file(optarg); o.node_file = optarg; break;
            case 'o': o.out_file = optarg; break;
            case 'p': o.path = optarg; break;
            case 's': o.source = optarg; break;
            case 'j': o.threads = atoi(optarg); break;
            case 1: o.min_nodes = (uint32_t)atoi(optarg); break;
            case 2: o.device = atoi(optarg); break;
            case 3: o.speculate = (uint32_t)atoi(optarg); break;
            case 'v': version_and_exit(); break;
            case 'h': help_and_exit(o.mode); break;
            default: break;   // ':' missing argument, 0 flag options, positional arguments
        }
    }
    o.return_all_paths = flag_all;
    o.sort_alignment = flag_sort;
    o.terminal_alignments = flag_term;
    o.verbose = flag_verbose;
    o.cmd = flag_cmd;
    o.timing = flag_timing;
    if (o.sort_alignment || o.terminal_alignments) o.align_stats = false;   // as the align tool does
    if (flag_stats)
        fprintf(stderr, "--graph-statistics is not part of gfalign-b200 (ignored)\n");
    if (o.cmd) {
        for (int i = 0; i < argc; ++i) printf("%s ", argv[i]);
        printf("\n");
    }
    switch (o.mode) {
        case 1: return gfh::run_evalgfa(o);
        case 3: return gfh::run_search(o);
        case 4: return gfh::run_filter(o);
        case 5: return gfh::run_evalpath(o);
    }
    return EXIT_FAILURE;
}
