// gfalign_b200/csrc/edges.cu — evalGFA: per-edge read support (K3).
//
// Reference semantics (InAlignments::buildEdgeGraph, src/alignments.cpp:353-403 + evalGFA lookup,
// src/eval.cpp:42-60): every consecutive oriented step pair (x, y) of every alignment is one
// traversal event of the directed pair (x, y) and of its reverse complement (y^1, x^1); an L record
// carries the events of both; a self-reverse-complement ("hairpin") record is inserted with
// weight 1 and then incremented twice per event => 2k - 1.
//
// Device data, rebuilt per call from the E link records (all on the device, no host hashing):
//   table  EdgeSlot[cap]   open addressing on the CANONICAL pair key min(x<<32|y, (y^1)<<32|(x^1)) —
//                          a pair and its reverse complement are one entry, since only their sum
//                          is ever used; 16-byte entries {key, count, representative slot}: the
//                          probe and the atomic of one event touch the same 32-byte sector;
//                          cap = 3 E (34 MB for a 700 k-link graph, L2-resident)
//   counts u32[2E]         slot 2i = record i as written, 2i+1 = its reverse complement; the events
//                          of both orientations are reported at the lowest slot that carries the
//                          canonical orientation, every other slot stays 0
// Kernels:
//   edge_insert_kernel     one thread per slot: CAS insert of canonical keys, atomicMin of the slot id
//   rowstart_kernel        once per alignment table: bitmap over the step array, bit s = step s
//                          starts a row (S/8 bytes; lets the stream ignore the offsets array)
//   edge_stream_kernel     the HBM stream, flat over the step array: a thread owns 4 consecutive
//                          steps (one 128-bit load) + the next one; steps s, s+1 form an event
//                          unless s+1 starts a row.  The four events are probed together.
//                          Small tables (<= 4096 entries) are copied to shared memory and counted
//                          there, flushed once per CTA (persistent grid).
//   edge_scatter_kernel    table entry -> counts[representative]
//   edge_weights_kernel    one thread per record: k = events of either orientation, hairpin 2k-1
// Algorithmic bytes: 4(m+1) per alignment; the stream itself reads 4m + m/8.
#include "kernels.cuh"

namespace gfe {

namespace {

constexpr uint64_t kEmptyKey = ~0ull;
constexpr uint32_t kEdgeSmemCap = 4096;   // table entries that fit the shared-memory variant (64 KB)

__device__ __forceinline__ uint32_t edge_hash(uint64_t key, uint32_t cap) {
    // the HIGH half of the product depends on every key bit (the low half sees only the to-step);
    // a second multiply breaks up the runs a chain graph's consecutive ids would leave
    uint64_t h = key * 0x9E3779B97F4A7C15ull;
    h ^= h >> 32;
    h *= 0xFF51AFD7ED558CCDull;
    return (uint32_t)(((h >> 32) * cap) >> 32);
}

// a directed pair and its reverse complement (y^1, x^1) share one table entry: the smaller key
__device__ __forceinline__ uint64_t canon_key(uint32_t x, uint32_t y) {
    const uint64_t k = ((uint64_t)x << 32) | y, t = ((uint64_t)(y ^ 1u) << 32) | (x ^ 1u);
    return k < t ? k : t;
}

__global__ void edge_clear_kernel(EdgeSlot *tab, uint32_t cap) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < cap; i += gridDim.x * blockDim.x) {
        EdgeSlot s;
        s.key = kEmptyKey;
        s.count = 0;
        s.rep = 0xFFFFFFFFu;
        tab[i] = s;
    }
}

__device__ __forceinline__ uint64_t directed_key(const uint32_t *ef, const uint32_t *et, uint32_t slot) {
    const uint32_t i = slot >> 1;
    const uint32_t x = ef[i], y = et[i];
    return (slot & 1u) ? (((uint64_t)(y ^ 1u) << 32) | (x ^ 1u)) : (((uint64_t)x << 32) | y);
}

__global__ void edge_insert_kernel(const uint32_t *ef, const uint32_t *et, uint32_t E, EdgeSlot *tab,
                                   uint32_t cap) {
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= 2 * E) return;
    const uint64_t key = directed_key(ef, et, slot);
    if (key != canon_key((uint32_t)(key >> 32), (uint32_t)key)) return;   // the twin slot carries it
    uint32_t p = edge_hash(key, cap);
    for (;;) {
        const unsigned long long prev =
            atomicCAS(reinterpret_cast<unsigned long long *>(&tab[p].key), kEmptyKey, key);
        if (prev == kEmptyKey || prev == key) {
            atomicMin(&tab[p].rep, slot);
            return;
        }
        p = p + 1 == cap ? 0 : p + 1;
    }
}

template <bool kSmem>
__device__ __forceinline__ void count_event(EdgeSlot *tab, uint32_t cap, uint64_t key, uint32_t p,
                                            uint64_t first_key) {
    uint64_t k = first_key;
    while (k != key) {
        if (k == kEmptyKey) return;           // the pair is not a link of the graph
        p = p + 1 == cap ? 0 : p + 1;
        k = tab[p].key;
    }
    atomicAdd(&tab[p].count, 1u);
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
    EdgeSlot *tab = a.table;
    if (kSmem) {
        EdgeSlot *st = reinterpret_cast<EdgeSlot *>(s_raw);
        for (uint32_t i = threadIdx.x; i < a.cap; i += blockDim.x) {
            EdgeSlot s = a.table[i];
            s.count = 0;
            st[i] = s;
        }
        tab = st;
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
        uint64_t key[4], fk[4];
        uint32_t p[4];
        bool live[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            live[u] = s0 + u + 1 < a.S && !((bits >> u) & 1u);
            key[u] = canon_key(v[u], v[u + 1]);
            p[u] = edge_hash(key[u], a.cap);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (live[u]) fk[u] = tab[p[u]].key;
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (live[u]) count_event<kSmem>(tab, a.cap, key[u], p[u], fk[u]);
    }
    if (kSmem) {
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < a.cap; i += blockDim.x)
            if (tab[i].count) atomicAdd(&a.table[i].count, tab[i].count);
    }
}

__global__ void edge_scatter_kernel(const EdgeSlot *tab, uint32_t cap, uint32_t *counts) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= cap) return;
    const EdgeSlot s = tab[i];
    if (s.key != kEmptyKey && s.count) counts[s.rep] = s.count;
}

__device__ __forceinline__ uint32_t table_count(const EdgeSlot *tab, uint32_t cap, uint64_t key) {
    uint32_t p = edge_hash(key, cap);
    for (;;) {
        const uint64_t k = tab[p].key;
        if (k == key) return tab[p].count;
        if (k == kEmptyKey) return 0;
        p = p + 1 == cap ? 0 : p + 1;
    }
}

__global__ void edge_weights_kernel(const uint32_t *ef, const uint32_t *et, uint32_t E,
                                    const EdgeSlot *tab, uint32_t cap, uint32_t *weights) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= E) return;
    const uint32_t x = ef[i], y = et[i];
    const bool hairpin = y == (x ^ 1u);
    const uint64_t k = table_count(tab, cap, canon_key(x, y));   // both orientations
    weights[i] = (uint32_t)(hairpin ? (k ? 2 * k - 1 : 0) : k);
}

}  // namespace

uint32_t edge_table_capacity(uint32_t E) {
    uint64_t cap = (uint64_t)E * 3 + 64;        // <= E canonical keys, load <= 1/3: short probe runs
    if (cap > 0xFFFFFFF0ull) cap = 0xFFFFFFF0ull;
    return (uint32_t)cap;
}

int configure_edge_kernels() {
    cudaError_t e = cudaFuncSetAttribute(edge_stream_kernel<true, true>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)(kEdgeSmemCap * sizeof(EdgeSlot)));
    if (e != cudaSuccess) return (int)e;
    return (int)cudaFuncSetAttribute(edge_stream_kernel<true, false>,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)(kEdgeSmemCap * sizeof(EdgeSlot)));
}

void launch_rowstart(const uint32_t *off, uint64_t A, uint32_t *rowstart, cudaStream_t st) {
    if (A == 0) return;
    uint64_t blocks = (A + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    rowstart_kernel<<<(unsigned)blocks, 256, 0, st>>>(off, A, rowstart);
}

void launch_edge_build(const uint32_t *d_ef, const uint32_t *d_et, uint32_t E, EdgeSlot *table,
                       uint32_t cap, cudaStream_t st) {
    edge_clear_kernel<<<(cap + 255) / 256 < 1184 ? (cap + 255) / 256 : 1184, 256, 0, st>>>(table, cap);
    if (E) edge_insert_kernel<<<(2 * E + 255) / 256, 256, 0, st>>>(d_ef, d_et, E, table, cap);
}

void launch_edge_stream(const EdgeStreamArgs &a, int sm_count, cudaStream_t st) {
    if (a.S < 2) return;
    const uint64_t n_blk = ((a.S + 3) / 4 + 255) / 256;
    const bool vec = (reinterpret_cast<uintptr_t>(a.steps) & 15) == 0;
    if (a.cap <= kEdgeSmemCap) {
        const uint64_t grid = n_blk < (uint64_t)sm_count * 3 ? n_blk : (uint64_t)sm_count * 3;
        const size_t smem = a.cap * sizeof(EdgeSlot);
        if (vec) edge_stream_kernel<true, true><<<(unsigned)grid, 256, smem, st>>>(a);
        else edge_stream_kernel<true, false><<<(unsigned)grid, 256, smem, st>>>(a);
    } else {
        const uint64_t grid = n_blk < (uint64_t)sm_count * 64 ? n_blk : (uint64_t)sm_count * 64;
        if (vec) edge_stream_kernel<false, true><<<(unsigned)grid, 256, 0, st>>>(a);
        else edge_stream_kernel<false, false><<<(unsigned)grid, 256, 0, st>>>(a);
    }
}

void launch_edge_scatter(const EdgeSlot *table, uint32_t cap, uint32_t *counts, cudaStream_t st) {
    edge_scatter_kernel<<<(cap + 255) / 256, 256, 0, st>>>(table, cap, counts);
}

void launch_edge_weights(const uint32_t *d_ef, const uint32_t *d_et, uint32_t E, const EdgeSlot *table,
                         uint32_t cap, uint32_t *weights, cudaStream_t st) {
    if (E) edge_weights_kernel<<<(E + 255) / 256, 256, 0, st>>>(d_ef, d_et, E, table, cap, weights);
}

}  // namespace gfe
