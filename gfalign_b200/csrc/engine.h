// gfalign_b200/csrc/engine.h — engine state shared by engine.cu (C ABI) and search.cpp (driver).
#pragma once
#include <cuda_runtime.h>

#include <cstddef>
#include <cstdint>
#include <vector>

#include "../../include/gfalign_b200.h"
#include "kernels.cuh"

template <typename T>
struct DevBuf {
    T *ptr = nullptr;
    size_t cap = 0;   // elements
};
template <typename T>
struct PinBuf {
    T *ptr = nullptr;
    size_t cap = 0;
};

struct gfe_engine {
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;   // last timed launch (points into ev_ring)
    std::vector<cudaEvent_t> ev_ring;           // pairs recorded since the last gfe_kernel_time_stats
    size_t ev_used = 0;
    bool timing = false;
    float last_kernel_ms = 0.f, last_aux_ms = 0.f;
    cudaEvent_t ev_aux = nullptr;
    char error[512] = "";

    // alignment table (owned copies, or adopted device pointers)
    DevBuf<uint32_t> own_off, own_steps;
    const uint32_t *d_off = nullptr, *d_steps = nullptr;
    uint64_t A = 0, S = 0;
    uint32_t n_nodes = 0, max_m = 0;
    bool have_alignments = false;

    // universe-local view
    std::vector<uint32_t> node2local;
    DevBuf<uint32_t> d_node2local;
    DevBuf<uint16_t> d_bstep;        // length-bucketed local steps
    DevBuf<uint2> d_ldesc;
    DevBuf<uint32_t> d_perm, d_hist, d_cursors;
    gfe::Buckets bk{};
    uint64_t bstep_elems = 0;
    uint32_t L = 0, tab_stride = 8;
    bool have_universe = false;

    // scheduling groups of the bitset kernel, one table per W in {1,2,4,8,16,32} (index log2 W), for
    // the plain view [0] and the deduplicated view [1]; built with the universe
    DevBuf<gfe::UnitGroup> d_groups[2];
    uint32_t n_groups[2][6] = {}, n_units[2][6] = {}, unit_rows[2][6] = {};
    // multiplicity-compressed view (K1 full scans): distinct rows up to reverse complement, class k =
    // bit k of the multiplicity (kernels.cu dedup_*); built with the universe when it pays
    DevBuf<gfe::DedupSlot> d_dedup;
    DevBuf<uint16_t> d_bstep2;
    DevBuf<uint2> d_ldesc2;
    DevBuf<uint8_t> d_rshift2;
    DevBuf<uint32_t> d_dedup_misc;       // class histogram / class bases / cursors
    gfe::Buckets bk2{};
    bool have_dedup = false, dedup_enabled = true;
    uint64_t dedup_rows = 0, dedup_view_rows = 0, dedup_distinct = 0;

    // pair dictionary of the universe-local view + per-batch bitsets (bitset kernel)
    DevBuf<uint32_t> d_pair_keys, d_pair_ids, d_pair_ctr, d_masks, d_unit_ctr;
    DevBuf<uint2> d_slow;
    uint32_t n_pairs = 0;
    bool have_dict = false;
    size_t slow_cap = 1u << 20;
    int force_legacy = 0;            // GFE_FORCE_LEGACY=1: A/B runs of the pre-bitset kernel

    // inverted index of the scoring view (incremental search scoring), built lazily per (universe, root)
    DevBuf<uint32_t> d_inv_off, d_inv_cursor, d_inv_rows, d_chunk_node;
    DevBuf<uint4> d_desc;
    uint32_t inv_root = 0xFFFFFFFFu;
    bool have_index = false;
    uint64_t inv_entries = 0;
    int incr_retries = 0;
    double incr_t[3] = {0, 0, 0};     // GFE_DEBUG_TIMING: host prep / enqueue / wait seconds of the incremental batches
    uint64_t incr_n[4] = {0, 0, 0, 0};
    // host scratch of gfe_score_paths_incremental (kept to avoid per-batch allocation)
    struct IncrSlot { uint32_t key, item, parent; };
    std::vector<uint32_t> h_present, h_chunk_node, h_slot_dev, h_len_dev, h_full_of_item;
    std::vector<IncrSlot> h_slots, h_slots2;
    std::vector<uint32_t> h_key_count;

    // candidate-path staging and results
    PinBuf<unsigned char> h_stage;
    DevBuf<unsigned char> d_stage;
    DevBuf<uint16_t> d_ptab, d_plocal, d_pnext;
    DevBuf<uint32_t> d_counts, d_err;
    PinBuf<uint32_t> h_counts;
    uint32_t h_wc_storage[4] = {0, 0, 0, 0};
    uint32_t *h_wc = h_wc_storage;

    // exact-DP worklists
    DevBuf<uint2> d_work;
    DevBuf<uint32_t> d_work_count, d_pair_work;
    size_t work_cap = 1u << 20;
    int shard_retries = 0;           // collective batch repeats of gfe_score_paths_sharded

    // evalPath outputs
    DevBuf<int32_t> d_tfw, d_trc;

    // evalPath traceback / filter scratch
    DevBuf<uint32_t> d_tr_path, d_tr_misc, d_tr_dirs, d_tr_len, d_tr_longb;
    DevBuf<uint64_t> d_tr_base, d_tr_first;
    DevBuf<int32_t> d_tr_score;
    DevBuf<int16_t> d_tr_longrows;
    DevBuf<uint8_t> d_tr_ops, d_tr_rc;

    // edge support
    DevBuf<uint32_t> d_edge_ft, d_edge_counts;     // link records (from | to), per-record weights
    DevBuf<gfe::EdgeSlot> d_edge_table;
    DevBuf<uint4> d_edge_adj;                      // dense adjacency records, one per oriented step
    DevBuf<uint32_t> d_edge_deg, d_edge_cnt;       // arcs per step (+ the counter cursor), one counter per arc
    uint32_t edge_cap = 0;
    DevBuf<uint32_t> d_rowstart;                   // bit s: step s starts an alignment
    bool have_rowstart = false;

    // counters
    uint64_t kernel_launches = 0, exact_pairs = 0, slow_pairs = 0, h2d_bytes = 0, d2h_bytes = 0;

    void free_all() {
        auto fd = [](void *p) { if (p) cudaFree(p); };
        auto fh = [](void *p) { if (p) cudaFreeHost(p); };
        fd(own_off.ptr); fd(own_steps.ptr); fd(d_node2local.ptr); fd(d_bstep.ptr);
        fd(d_ldesc.ptr); fd(d_perm.ptr); fd(d_hist.ptr); fd(d_cursors.ptr);
        fd(d_stage.ptr); fd(d_ptab.ptr); fd(d_plocal.ptr); fd(d_pnext.ptr); fd(d_counts.ptr); fd(d_err.ptr);
        fd(d_work.ptr); fd(d_work_count.ptr); fd(d_pair_work.ptr); fd(d_tfw.ptr); fd(d_trc.ptr);
        fd(d_edge_ft.ptr); fd(d_edge_counts.ptr); fd(d_edge_table.ptr); fd(d_rowstart.ptr);
        fd(d_edge_adj.ptr); fd(d_edge_deg.ptr); fd(d_edge_cnt.ptr);
        fd(d_pair_keys.ptr); fd(d_pair_ids.ptr); fd(d_pair_ctr.ptr); fd(d_masks.ptr);
        fd(d_unit_ctr.ptr); fd(d_slow.ptr);
        fd(d_groups[0].ptr); fd(d_groups[1].ptr); fd(d_dedup.ptr); fd(d_bstep2.ptr); fd(d_ldesc2.ptr);
        fd(d_rshift2.ptr); fd(d_dedup_misc.ptr);
        fd(d_inv_off.ptr); fd(d_inv_cursor.ptr); fd(d_inv_rows.ptr); fd(d_chunk_node.ptr); fd(d_desc.ptr);
        fd(d_tr_path.ptr); fd(d_tr_misc.ptr); fd(d_tr_dirs.ptr); fd(d_tr_len.ptr); fd(d_tr_longb.ptr);
        fd(d_tr_base.ptr); fd(d_tr_first.ptr); fd(d_tr_score.ptr); fd(d_tr_longrows.ptr);
        fd(d_tr_ops.ptr); fd(d_tr_rc.ptr);
        fh(h_stage.ptr); fh(h_counts.ptr);
    }
};
