// gfalign_b200/csrc/edges.cu — evalGFA: per-edge read support (K3).
//
// Reference semantics (InAlignments::buildEdgeGraph, src/alignments.cpp:353-403 + evalGFA lookup,
// src/eval.cpp:42-60): every consecutive oriented step pair (x, y) of every alignment is one
// traversal event of the directed arc (x, y) and of its reverse complement (y^1, x^1); an L record
// carries the events of both; a self-reverse-complement ("hairpin") record is inserted with
// weight 1 and then incremented twice per event => 2k - 1.
//
// Device data, rebuilt per call from the E link records (all on the device, no host hashing):
//   arcs   EdgeSlot[cap]   open addressing on the DIRECTED arc key x<<32|y, one entry per distinct arc
//                          of the records and of their reverse complements: {key, counter index, lowest
//                          record slot carrying it}; cap = 4 E.  Used to build the adjacency, for
//                          the few arcs that overflow it, and by the per-record weight lookup.
//   adj    uint4[2N]       DENSE adjacency, one 16-byte record per oriented step x:
//                          {to_0, to_1, to_2, base}: the arcs leaving x and the index of the first of
//                          its counters.  An event (x, y) costs ONE 16-byte load (no hashing, no
//                          canonical key, no probe loop) and one RED on cnt[base + j].  A step with more
//                          than three arcs keeps two in the record, to_2 = kMore sends the rest to `arcs`.
//   cnt    u32[#arcs]      one counter per directed arc, the arcs of one step consecutive
// Kernels:
//   arc_insert_kernel      one thread per record slot (2i = record i, 2i+1 = its reverse complement): CAS
//                          insert, the winner takes the arc's position among the arcs of its step
//   adj_base_kernel        one thread per oriented step: base = atomicAdd(cursor, degree)
//   adj_fill_kernel        one thread per table entry: counter index, adjacency record
//   rowstart_kernel        once per alignment table: bitmap over the step array, bit s = step s
//                          starts a row (S/8 bytes; lets the stream ignore the offsets array)
//   edge_stream_kernel     the HBM stream, flat over the step array: a thread owns 4 consecutive
//                          steps (one 128-bit load) + the next one; steps s, s+1 form an event
//                          unless s+1 starts a row.  Graphs whose adjacency + counters fit shared
//                          memory are counted there, flushed once per CTA (persistent grid).
//   edge_scatter_kernel    counter -> counts[lowest record slot of the arc]        (sharded flow)
//   edge_weights_kernel    one thread per record: k = events of the arc + of its reverse complement,
//                          hairpin 2k-1
// Algorithmic bytes: 4(m+1) per alignment; the stream itself reads 4m + m/8.
// Bound: every event is one uncoalesced 16-byte load and one RED, i.e. two LSU lane-operations on
// random addresses; the measured per-SM rate for those (~1.3 cycles per lane, B300_MICROARCH.md
// "Atomics") puts the floor of cfg4's 6.1e7 events at ~0.55 ms — the kernel is LSU/L2-atomic bound,
// not HBM bound; walks of a random graph give a CTA no two events of the same arc to combine.
#include "kernels.cuh"

namespace gfe {

namespace {

constexpr uint64_t kEmptyKey = ~0ull;
constexpr uint32_t kNone = 0xFFFFFFFFu;      // no arc in this position of an adjacency record
constexpr uint32_t kMore = 0xFFFFFFFEu;      // to_2: the step has further arcs, look in the arc table
constexpr size_t kEdgeSmemBytes = 160 * 1024;   // adjacency + counters of the shared-memory variant

__device__ __forceinline__ uint32_t arc_hash(uint64_t key, uint32_t cap) {
    // the HIGH half of the product depends on every key bit; a second multiply breaks up the runs a
    // chain graph's consecutive ids would leave
    uint64_t h = key * 0x9E3779B97F4A7C15ull;
    h ^= h >> 32;
    h *= 0xFF51AFD7ED558CCDull;
    return (uint32_t)(((h >> 32) * cap) >> 32);
}

__device__ __forceinline__ uint64_t arc_key(uint32_t x, uint32_t y) { return ((uint64_t)x << 32) | y; }

__global__ void arc_clear_kernel(EdgeSlot *tab, uint32_t cap, uint32_t *deg, uint4 *adj, uint32_t steps2,
                                 uint32_t *cursor) {
    const uint32_t i0 = blockIdx.x * blockDim.x + threadIdx.x, n = gridDim.x * blockDim.x;
    for (uint32_t i = i0; i < cap; i += n) {
        EdgeSlot s;
        s.key = kEmptyKey;
        s.count = 0;            // counter index of the arc (position among its step's arcs until adj_fill)
        s.rep = 0xFFFFFFFFu;
        tab[i] = s;
    }
    for (uint32_t i = i0; i < steps2; i += n) {
        deg[i] = 0;
        adj[i] = make_uint4(kNone, kNone, kNone, 0u);
    }
    if (i0 == 0) *cursor = 0;
}

// record slot -> directed arc (2i: record i as written, 2i+1: its reverse complement)
__device__ __forceinline__ uint64_t slot_key(const uint32_t *ef, const uint32_t *et, uint32_t slot) {
    const uint32_t i = slot >> 1;
    const uint32_t x = ef[i], y = et[i];
    return (slot & 1u) ? arc_key(y ^ 1u, x ^ 1u) : arc_key(x, y);
}

__global__ void arc_insert_kernel(const uint32_t *ef, const uint32_t *et, uint32_t E, EdgeSlot *tab,
                                  uint32_t cap, uint32_t *deg) {
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= 2 * E) return;
    const uint64_t key = slot_key(ef, et, slot);
    uint32_t p = arc_hash(key, cap);
    for (;;) {
        const unsigned long long prev =
            atomicCAS(reinterpret_cast<unsigned long long *>(&tab[p].key), kEmptyKey, key);
        if (prev == kEmptyKey) tab[p].count = atomicAdd(&deg[(uint32_t)(key >> 32)], 1u);   // the one winner
        if (prev == kEmptyKey || prev == key) {
            atomicMin(&tab[p].rep, slot);
            return;
        }
        p = p + 1 == cap ? 0 : p + 1;
    }
}

__global__ void adj_base_kernel(const uint32_t *deg, uint32_t steps2, uint4 *adj, uint32_t *cursor) {
    const uint32_t x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= steps2) return;
    const uint32_t d = deg[x];
    if (d) adj[x].w = atomicAdd(cursor, d);
}

__global__ void adj_fill_kernel(EdgeSlot *tab, uint32_t cap, const uint32_t *deg, uint4 *adj) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= cap) return;
    const uint64_t key = tab[i].key;
    if (key == kEmptyKey) return;
    const uint32_t x = (uint32_t)(key >> 32), y = (uint32_t)key, j = tab[i].count, d = deg[x];
    uint32_t *rec = reinterpret_cast<uint32_t *>(adj + x);
    tab[i].count = rec[3] + j;                      // counter index: the arcs of x are consecutive
    if (d <= 3u || j < 2u) rec[j] = y;
    else if (j == 2u) rec[2] = kMore;               // positions >= 2 of a crowded step: arc table only
}

__device__ __forceinline__ uint32_t arc_counter(const EdgeSlot *tab, uint32_t cap, uint64_t key) {
    uint32_t p = arc_hash(key, cap);
    for (;;) {
        const uint64_t k = tab[p].key;
        if (k == key) return tab[p].count;
        if (k == kEmptyKey) return kNone;
        p = p + 1 == cap ? 0 : p + 1;
    }
}

// bit s of rowstart <=> step s is the first step of an alignment
__global__ void rowstart_kernel(const uint32_t *off, uint64_t A, uint32_t *rowstart) {
    for (uint64_t r = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; r < A;
         r += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t o0 = off[r], o1 = off[r + 1];
        if (o1 > o0) atomicOr(&rowstart[o0 >> 5], 1u << (o0 & 31));
    }
}

// The HBM stream, flat over the step array: a thread owns 4 consecutive steps (one 128-bit load)
// plus the step after them; steps s and s+1 form an event unless s+1 starts a row.
template <bool kSmem, bool kVec>
__global__ void __launch_bounds__(256) edge_stream_kernel(EdgeStreamArgs a) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    const uint4 *adj = a.adj;
    uint32_t *cnt = a.cnt;
    if (kSmem) {
        uint4 *s_adj = reinterpret_cast<uint4 *>(s_raw);
        uint32_t *s_cnt = reinterpret_cast<uint32_t *>(s_adj + a.steps2);
        for (uint32_t i = threadIdx.x; i < a.steps2; i += blockDim.x) s_adj[i] = a.adj[i];
        for (uint32_t i = threadIdx.x; i < a.n_arcs; i += blockDim.x) s_cnt[i] = 0;
        adj = s_adj;
        cnt = s_cnt;
        __syncthreads();
    }
    const uint64_t n_quads = (a.S + 3) / 4;
    for (uint64_t q = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; q < n_quads;
         q += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t s0 = q * 4;
        uint32_t v[5];
        if (kVec && s0 + 4 <= a.S) {
            const uint4 w = __ldg(reinterpret_cast<const uint4 *>(a.steps + s0));
            v[0] = w.x; v[1] = w.y; v[2] = w.z; v[3] = w.w;
        } else {
#pragma unroll
            for (int u = 0; u < 4; ++u) v[u] = s0 + u < a.S ? __ldg(a.steps + s0 + u) : 0u;
        }
        v[4] = s0 + 4 < a.S ? __ldg(a.steps + s0 + 4) : 0u;
        // start bits of steps s0+1 .. s0+4
        const uint64_t b0 = s0 + 1;
        uint32_t bits = __ldg(a.rowstart + (b0 >> 5)) >> (b0 & 31);
        if ((b0 & 31) > 28) bits |= __ldg(a.rowstart + (b0 >> 5) + 1) << (32 - (b0 & 31));
        uint4 rec[4];
        bool live[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            // a step outside the graph's id space has no record (the engine validates alignment steps
            // against n_nodes, and the records cover exactly 2 * n_nodes oriented steps)
            live[u] = s0 + u + 1 < a.S && !((bits >> u) & 1u) && v[u] < a.steps2;
            if (live[u]) rec[u] = kSmem ? adj[v[u]] : __ldg(adj + v[u]);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (!live[u]) continue;
            const uint32_t y = v[u + 1];
            uint32_t c = kNone;
            if (rec[u].x == y) c = rec[u].w;
            else if (rec[u].y == y) c = rec[u].w + 1u;
            else if (rec[u].z == y) c = rec[u].w + 2u;
            else if (rec[u].z == kMore) c = arc_counter(a.table, a.cap, arc_key(v[u], y));
            if (c != kNone) atomicAdd(cnt + c, 1u);       // else: the pair is not a link of the graph
        }
    }
    if (kSmem) {
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < a.n_arcs; i += blockDim.x)
            if (cnt[i]) atomicAdd(a.cnt + i, cnt[i]);
    }
}

__global__ void edge_scatter_kernel(const EdgeSlot *tab, uint32_t cap, const uint32_t *cnt, uint32_t *counts) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= cap) return;
    const EdgeSlot s = tab[i];
    if (s.key != kEmptyKey) {
        const uint32_t c = cnt[s.count];
        if (c) counts[s.rep] = c;
    }
}

__global__ void edge_weights_kernel(const uint32_t *ef, const uint32_t *et, uint32_t E,
                                    const EdgeSlot *tab, uint32_t cap, const uint32_t *cnt, uint32_t *weights) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= E) return;
    const uint32_t x = ef[i], y = et[i];
    const bool hairpin = y == (x ^ 1u);             // its reverse complement is itself: one arc
    uint64_t k = cnt[arc_counter(tab, cap, arc_key(x, y))];
    if (!hairpin) k += cnt[arc_counter(tab, cap, arc_key(y ^ 1u, x ^ 1u))];
    weights[i] = (uint32_t)(hairpin ? (k ? 2 * k - 1 : 0) : k);
}

}  // namespace

uint32_t edge_table_capacity(uint32_t E) {
    uint64_t cap = (uint64_t)E * 4 + 64;        // <= 2 E directed arcs, load <= 1/2
    if (cap > 0xFFFFFFF0ull) cap = 0xFFFFFFF0ull;
    return (uint32_t)cap;
}

int configure_edge_kernels() {
    cudaError_t e = cudaFuncSetAttribute(edge_stream_kernel<true, true>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kEdgeSmemBytes);
    if (e != cudaSuccess) return (int)e;
    return (int)cudaFuncSetAttribute(edge_stream_kernel<true, false>,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kEdgeSmemBytes);
}

void launch_rowstart(const uint32_t *off, uint64_t A, uint32_t *rowstart, cudaStream_t st) {
    if (A == 0) return;
    uint64_t blocks = (A + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    rowstart_kernel<<<(unsigned)blocks, 256, 0, st>>>(off, A, rowstart);
}

void launch_edge_build(const uint32_t *d_ef, const uint32_t *d_et, uint32_t E, const EdgeBuild &b, cudaStream_t st) {
    const uint32_t most = b.cap > b.steps2 ? b.cap : b.steps2;
    arc_clear_kernel<<<(most + 255) / 256 < 2368 ? (most + 255) / 256 : 2368, 256, 0, st>>>(b.table, b.cap, b.deg, b.adj,
                                                                                          b.steps2, b.cursor);
    if (!E) return;
    arc_insert_kernel<<<(2 * E + 255) / 256, 256, 0, st>>>(d_ef, d_et, E, b.table, b.cap, b.deg);
    adj_base_kernel<<<(b.steps2 + 255) / 256, 256, 0, st>>>(b.deg, b.steps2, b.adj, b.cursor);
    adj_fill_kernel<<<(b.cap + 255) / 256, 256, 0, st>>>(b.table, b.cap, b.deg, b.adj);
}

void launch_edge_stream(const EdgeStreamArgs &a, int sm_count, cudaStream_t st) {
    if (a.S < 2) return;
    const uint64_t n_blk = ((a.S + 3) / 4 + 255) / 256;
    const bool vec = (reinterpret_cast<uintptr_t>(a.steps) & 15) == 0;
    const size_t smem = (size_t)a.steps2 * 16 + (size_t)a.n_arcs * 4;
    if (smem <= kEdgeSmemBytes) {
        const uint64_t grid = n_blk < (uint64_t)sm_count ? n_blk : (uint64_t)sm_count;
        if (vec) edge_stream_kernel<true, true><<<(unsigned)grid, 256, smem, st>>>(a);
        else edge_stream_kernel<true, false><<<(unsigned)grid, 256, smem, st>>>(a);
    } else {
        const uint64_t grid = n_blk < (uint64_t)sm_count * 64 ? n_blk : (uint64_t)sm_count * 64;
        if (vec) edge_stream_kernel<false, true><<<(unsigned)grid, 256, 0, st>>>(a);
        else edge_stream_kernel<false, false><<<(unsigned)grid, 256, 0, st>>>(a);
    }
}

void launch_edge_scatter(const EdgeSlot *table, uint32_t cap, const uint32_t *cnt, uint32_t *counts, cudaStream_t st) {
    edge_scatter_kernel<<<(cap + 255) / 256, 256, 0, st>>>(table, cap, cnt, counts);
}

void launch_edge_weights(const uint32_t *d_ef, const uint32_t *d_et, uint32_t E, const EdgeSlot *table,
                         uint32_t cap, const uint32_t *cnt, uint32_t *weights, cudaStream_t st) {
    if (E) edge_weights_kernel<<<(E + 255) / 256, 256, 0, st>>>(d_ef, d_et, E, table, cap, cnt, weights);
}

}  // namespace gfe
