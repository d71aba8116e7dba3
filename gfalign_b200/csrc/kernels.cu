// gfalign_b200/csrc/kernels.cu — hand-written sm_100a kernels (integer, HBM/L2/shared-memory bound;
// no tensor cores: the path has no dense contraction).  See kernels.cuh for the data layout and
// DESIGN.md for the per-kernel roofline.
#include "kernels.cuh"

#include <algorithm>

namespace gfe {

namespace {

constexpr unsigned kFull = 0xFFFFFFFFu;
constexpr int GFE_MAX_STEPS_DEV = 1000;   // GFE_MAX_STEPS (include/gfalign_b200.h): longest path / alignment
enum Verdict : int { kNone = 0, kBad = 1, kGood = 2, kExact = 3, kSlow = 4 };

__device__ __forceinline__ int imax(int a, int b) { return a > b ? a : b; }

// ------------------------------------------------------------------------------------------------
// K0a: histogram of alignment lengths (buckets 0, 1..kMaxBucket, long), max length, long steps.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock)
length_histogram_kernel(const uint32_t *__restrict__ off, uint64_t A, uint32_t *__restrict__ hist) {
    __shared__ uint32_t s_h[kNumBuckets + 2];
    if (threadIdx.x < kNumBuckets + 2) s_h[threadIdx.x] = 0;
    __syncthreads();
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t a = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; a < A; a += stride) {
        const uint32_t m = __ldg(off + a + 1) - __ldg(off + a);
        const uint32_t b = m <= (uint32_t)kMaxBucket ? m : (uint32_t)kLongBucket;
        atomicAdd(&s_h[b], 1u);
        atomicMax(&s_h[kNumBuckets], m);
        if (b == (uint32_t)kLongBucket) atomicAdd(&s_h[kNumBuckets + 1], m);
    }
    __syncthreads();
    if (threadIdx.x < kNumBuckets + 2 && s_h[threadIdx.x]) {
        if (threadIdx.x == kNumBuckets) atomicMax(&hist[threadIdx.x], s_h[threadIdx.x]);
        else atomicAdd(&hist[threadIdx.x], s_h[threadIdx.x]);
    }
}

// Pair dictionary.  A consecutive oriented pair (x, y) of local steps and its reverse complement
// (y^1, x^1) share one canonical key; pair id = 2*c + (key != canonical), so id^1 names the
// reverse complement.  A self-reverse-complement pair (y == x^1) has flip 0.
__device__ __forceinline__ uint32_t pair_canonical(uint32_t x, uint32_t y, uint32_t *flip) {
    const uint32_t key = (x << 16) | y, twin = ((y ^ 1u) << 16) | (x ^ 1u);
    *flip = key > twin ? 1u : 0u;
    return key > twin ? twin : key;
}
__device__ __forceinline__ uint32_t pair_hash(uint32_t key) {
    return (key * 0x9E3779B1u) >> 16;   // kPairHashCap = 2^16 slots
}
// slot of `key`, or kPairEmpty when absent
__device__ __forceinline__ uint32_t pair_find(const PairDict &pd, uint32_t key) {
    uint32_t h = pair_hash(key);
    for (uint32_t probe = 0; probe < kPairHashCap; ++probe) {
        const uint32_t k = __ldg(pd.keys + h);
        if (k == key) return h;
        if (k == kPairEmpty) return kPairEmpty;
        h = (h + 1) & (kPairHashCap - 1);
    }
    return kPairEmpty;
}

// K0p: insert the canonical key of every in-universe consecutive pair of every alignment.
__global__ void __launch_bounds__(kBlock)
pair_collect_kernel(const uint32_t *__restrict__ off, const uint32_t *__restrict__ steps,
                    const uint32_t *__restrict__ map, uint64_t A, uint32_t L, uint32_t max_pairs,
                    const PairDict pd) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t a = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; a < A; a += stride) {
        const uint32_t o0 = __ldg(off + a), o1 = __ldg(off + a + 1);
        if (o1 - o0 < 2 || o1 - o0 > (uint32_t)kMaxBucket) continue;   // long rows: legacy kernel
        uint32_t s = __ldg(steps + o0);
        uint32_t x = (__ldg(map + (s >> 1)) << 1) | (s & 1u);
        for (uint32_t k = o0 + 1; k < o1; ++k) {
            s = __ldg(steps + k);
            const uint32_t y = (__ldg(map + (s >> 1)) << 1) | (s & 1u);
            if ((x >> 1) < L && (y >> 1) < L) {
                uint32_t flip;
                const uint32_t key = pair_canonical(x, y, &flip);
                uint32_t h = pair_hash(key);
                bool done = false;
                for (uint32_t probe = 0; probe < kPairHashCap && !done; ++probe) {
                    uint32_t cur = pd.keys[h];
                    if (cur == kPairEmpty) {
                        if (pd.counter[0] >= max_pairs) break;   // too many pairs: no dictionary
                        cur = atomicCAS(pd.keys + h, kPairEmpty, key);
                        if (cur == kPairEmpty) atomicAdd(pd.counter, 1u);
                        done = cur == kPairEmpty || cur == key;
                    } else {
                        done = cur == key;
                    }
                    h = (h + 1) & (kPairHashCap - 1);
                }
                if (!done) pd.counter[2] = 1u;
            }
            x = y;
        }
    }
}

__global__ void pair_assign_kernel(const PairDict pd) {
    const uint32_t h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h < kPairHashCap && pd.keys[h] != kPairEmpty) pd.ids[h] = atomicAdd(pd.counter + 1, 1u);
}

// One row of the scoring view at element offset dst: m local steps, then (bucketed rows only) m-1 pair
// entries, m-2 triple-slot bytes and zero padding (layout: kernels.cuh row_stride).
__device__ __forceinline__ void write_view_row(uint16_t *__restrict__ bstep, uint32_t dst,
                                               const uint32_t *__restrict__ src, uint32_t m, bool is_long,
                                               const uint32_t *__restrict__ map, const PairDict &pd,
                                               uint32_t n_pairs, uint32_t pair_bias) {
    uint32_t prev = 0, prev_pid = 0;
    unsigned char *slots = reinterpret_cast<unsigned char *>(bstep + dst + 2 * m - 1);
    for (uint32_t j = 0; j < m; ++j) {
        const uint32_t s = __ldg(src + j);
        const uint32_t l = (__ldg(map + (s >> 1)) << 1) | (s & 1u);
        bstep[dst + j] = (uint16_t)l;
        if (!is_long && j > 0) {
            uint32_t pid = 2u * n_pairs;   // pair with a node outside the universe
            if (n_pairs != 0xFFFFFFFFu) {
                uint32_t flip;
                const uint32_t key = pair_canonical(prev, l, &flip);
                const uint32_t h = pair_find(pd, key);
                if (h != kPairEmpty) pid = 2u * pd.ids[h] + flip;
            }
            bstep[dst + m + j - 1] = (uint16_t)(pid + pair_bias);   // absolute table entry
            if (j >= 2) slots[j - 2] = (unsigned char)triple_slot(prev_pid, pid);
            prev_pid = pid;
        }
        prev = l;
    }
    if (!is_long)   // zero the padding behind the slot bytes
        for (uint32_t j = 2 * (2 * m - 1) + (m > 2 ? m - 2 : 0); j < 2 * row_stride((int)m); ++j)
            reinterpret_cast<unsigned char *>(bstep + dst)[j] = 0;
}

// K0b: scatter every alignment into its length bucket, translating steps to universe-local 16-bit
// ids on the way (node2local gathered through L2).  Row order inside a bucket is arbitrary; perm
// remembers the original index.  Slots are handed out per CTA (shared-memory ranks + one global
// atomic per bucket per CTA).
__global__ void __launch_bounds__(kBlock)
bucket_scatter_kernel(const uint32_t *__restrict__ off, const uint32_t *__restrict__ steps,
                      const uint32_t *__restrict__ map, uint64_t A, const Buckets bk,
                      uint16_t *__restrict__ bstep, uint2 *__restrict__ ldesc,
                      uint32_t *__restrict__ perm, uint32_t *__restrict__ cursors,
                      const PairDict pd, uint32_t n_pairs, uint32_t pair_bias, const bool only_long) {
    __shared__ uint32_t s_cnt[kNumBuckets + 1], s_base[kNumBuckets + 1];
    const uint64_t n_tiles = (A + kBlock - 1) / kBlock;
    for (uint64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        if (threadIdx.x <= kNumBuckets) s_cnt[threadIdx.x] = 0;
        __syncthreads();
        const uint64_t a = tile * kBlock + threadIdx.x;
        uint32_t o0 = 0, m = 0, b = 0, rank = 0, lrank = 0;
        bool valid = a < A;
        if (valid) {
            o0 = __ldg(off + a);
            m = __ldg(off + a + 1) - o0;
            b = m <= (uint32_t)kMaxBucket ? m : (uint32_t)kLongBucket;
            if (only_long && b != (uint32_t)kLongBucket) valid = false;   // the deduplicated view: short rows come from the table
        }
        if (valid) {
            rank = atomicAdd(&s_cnt[b], 1u);
            if (b == (uint32_t)kLongBucket) lrank = atomicAdd(&s_cnt[kNumBuckets], m);
        }
        __syncthreads();
        if (threadIdx.x <= kNumBuckets && s_cnt[threadIdx.x])
            s_base[threadIdx.x] = atomicAdd(&cursors[threadIdx.x], s_cnt[threadIdx.x]);
        __syncthreads();
        if (valid) {
            const uint32_t slot = s_base[b] + rank;
            if (perm) perm[bk.row_start[b] + slot] = (uint32_t)a;
            uint32_t dst;
            if (b == (uint32_t)kLongBucket) {
                dst = bk.step_base[b] + s_base[kNumBuckets] + lrank;
                ldesc[slot] = make_uint2(dst, m);
            } else {
                dst = bk.step_base[b] + slot * row_stride((int)m);
            }
            write_view_row(bstep, dst, steps + o0, m, b == (uint32_t)kLongBucket, map, pd, n_pairs, pair_bias);
        }
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------------
// Multiplicity compression of the scoring view.  The verdict of (path, alignment) depends only on the
// alignment's universe-local content, and it is the same for an alignment and its reverse complement
// (good <=> T_fw == 0 or T_rc == 0, src/eval.cpp:92-98).  Rows of 1..kMaxBucket steps are therefore
// merged up to reverse complement in a hash table; a distinct row with multiplicity w is written once
// per set bit k of w into class k of its length bucket, and the rows of class k count 2^k alignments
// each (the kernels flush their counters shifted by k).  Real read sets repeat walks at coverage x.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool load_local_row(const uint32_t *__restrict__ off, const uint32_t *__restrict__ steps,
                                               const uint32_t *__restrict__ map, uint32_t a, uint32_t (&loc)[kMaxBucket],
                                               uint32_t *m_out) {
    const uint32_t o0 = __ldg(off + a), m = __ldg(off + a + 1) - o0;
    *m_out = m;
    if (m < 1 || m > (uint32_t)kMaxBucket) return false;
#pragma unroll
    for (int j = 0; j < kMaxBucket; ++j) {
        uint32_t l = 0xFFFFFFFFu;
        if ((uint32_t)j < m) {
            const uint32_t s = __ldg(steps + o0 + j);
            l = (__ldg(map + (s >> 1)) << 1) | (s & 1u);
        }
        loc[j] = l;
    }
    return true;
}
// element j of the reverse complement of a row held in registers (static indexing only)
__device__ __forceinline__ uint32_t rc_elem(const uint32_t (&loc)[kMaxBucket], uint32_t m, int j) {
    uint32_t v = 0;
#pragma unroll
    for (int i = 0; i < kMaxBucket; ++i)
        if ((uint32_t)i == m - 1u - (uint32_t)j) v = loc[i] ^ 1u;
    return v;
}

__global__ void __launch_bounds__(kBlock)
dedup_insert_kernel(const uint32_t *__restrict__ off, const uint32_t *__restrict__ steps,
                    const uint32_t *__restrict__ map, uint64_t A, DedupSlot *__restrict__ slots, uint32_t mask) {
    for (uint64_t a64 = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; a64 < A; a64 += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t a = (uint32_t)a64;
        uint32_t loc[kMaxBucket], m;
        if (!load_local_row(off, steps, map, a, loc, &m)) continue;
        // canonical orientation: the lexicographically smaller of the row and its reverse complement
        uint32_t rc[kMaxBucket];
        int cmp = 0;
#pragma unroll
        for (int j = 0; j < kMaxBucket; ++j) {
            rc[j] = (uint32_t)j < m ? rc_elem(loc, m, j) : 0xFFFFFFFFu;
            if (cmp == 0 && (uint32_t)j < m) cmp = loc[j] < rc[j] ? -1 : (loc[j] > rc[j] ? 1 : 0);
        }
        uint32_t h = 0x811C9DC5u ^ m;
#pragma unroll
        for (int j = 0; j < kMaxBucket; ++j)
            if ((uint32_t)j < m) h = (h ^ (cmp <= 0 ? loc[j] : rc[j])) * 0x01000193u;
        h ^= h >> 15;
        h *= 0x2C1B3C6Du;
        h ^= h >> 12;
        const uint32_t fp = h * 0x9E3779B1u;
        uint32_t p = h & mask;
        for (;;) {
            unsigned long long cur = slots[p].key;
            if (cur == ~0ull) {
                const unsigned long long mine = ((unsigned long long)fp << 32) | a;
                cur = atomicCAS(&slots[p].key, ~0ull, mine);
                if (cur == ~0ull) cur = mine;
            }
            if ((uint32_t)(cur >> 32) == fp) {
                const uint32_t rep = (uint32_t)cur;
                bool same = rep == a;
                if (!same) {
                    uint32_t other[kMaxBucket], m2;
                    if (load_local_row(off, steps, map, rep, other, &m2) && m2 == m) {
                        bool eq_fw = true, eq_rc = true;   // the two rows are equal up to reverse complement
#pragma unroll
                        for (int j = 0; j < kMaxBucket; ++j)
                            if ((uint32_t)j < m) {
                                eq_fw = eq_fw && other[j] == loc[j];
                                eq_rc = eq_rc && other[j] == rc[j];
                            }
                        same = eq_fw || eq_rc;
                    }
                }
                if (same) {
                    atomicAdd(&slots[p].w, 1u);
                    break;
                }
            }
            p = (p + 1) & mask;
        }
    }
}

// class histogram: gcount[m * 32 + k] = distinct rows of length m whose multiplicity has bit k set;
// gcount[kDedupStats] = distinct rows, [kDedupStats + 1] = rows merged
__global__ void __launch_bounds__(kBlock)
dedup_count_kernel(const DedupSlot *__restrict__ slots, uint32_t cap, const uint32_t *__restrict__ off,
                   uint32_t *__restrict__ gcount) {
    __shared__ uint32_t s_h[kDedupStats + 2];
    for (uint32_t i = threadIdx.x; i < kDedupStats + 2; i += blockDim.x) s_h[i] = 0;
    __syncthreads();
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < cap; i += gridDim.x * blockDim.x) {
        const unsigned long long key = slots[i].key;
        if (key == ~0ull) continue;
        const uint32_t rep = (uint32_t)key, m = __ldg(off + rep + 1) - __ldg(off + rep);
        uint32_t w = slots[i].w + 1u;          // the weight word starts at ~0 (memset) and counts inserts
        atomicAdd(&s_h[kDedupStats], 1u);
        atomicAdd(&s_h[kDedupStats + 1], w);
        while (w) {
            const uint32_t k = (uint32_t)__ffs(w) - 1u;
            w &= w - 1u;
            atomicAdd(&s_h[m * 32u + k], 1u);
        }
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < kDedupStats + 2; i += blockDim.x)
        if (s_h[i]) atomicAdd(&gcount[i], s_h[i]);
}

// one view row per (distinct row, set bit k of its multiplicity), class k of bucket m at rows
// [class_base[m * 32 + k], ...); rshift[sorted row] = k
__global__ void __launch_bounds__(kBlock)
dedup_scatter_kernel(const DedupSlot *__restrict__ slots, uint32_t cap, const uint32_t *__restrict__ off,
                     const uint32_t *__restrict__ steps, const uint32_t *__restrict__ map, const Buckets bk,
                     uint16_t *__restrict__ bstep, const uint32_t *__restrict__ class_base,
                     uint32_t *__restrict__ cursors, uint8_t *__restrict__ rshift, const PairDict pd,
                     uint32_t n_pairs, uint32_t pair_bias) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < cap; i += gridDim.x * blockDim.x) {
        const unsigned long long key = slots[i].key;
        if (key == ~0ull) continue;
        const uint32_t rep = (uint32_t)key, o0 = __ldg(off + rep), m = __ldg(off + rep + 1) - o0;
        uint32_t w = slots[i].w + 1u;
        while (w) {
            const uint32_t k = (uint32_t)__ffs(w) - 1u;
            w &= w - 1u;
            const uint32_t row = class_base[m * 32u + k] + atomicAdd(&cursors[m * 32u + k], 1u);
            write_view_row(bstep, bk.step_base[m] + row * row_stride((int)m), steps + o0, m, false, map, pd, n_pairs,
                           pair_bias);
            rshift[bk.row_start[m] + row] = (uint8_t)k;
        }
    }
}

// K0b': batch prologue: clears the accumulators; candidate paths (global ids) -> local ids, a node
// outside the universe raises *err.
__global__ void __launch_bounds__(256) prologue_kernel(PrologueArgs a) {
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nthr = gridDim.x * blockDim.x;
    if (a.sticky && blockIdx.x == 0) {
        // the overflow flags of the previous batch survive its counters (words 1 and 3 are cleared
        // by threads of block 0 below)
        if (threadIdx.x == 0 && a.zero[1] && a.zero_words[1] > 3 && (a.zero[1][1] | a.zero[1][3]))
            atomicOr(a.sticky, 2u);
        __syncthreads();
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        uint32_t *z = a.zero[r];
        const uint32_t n = a.zero_words[r];
        if ((n & 3u) == 0 && (reinterpret_cast<uintptr_t>(z) & 15u) == 0) {
            uint4 *z4 = reinterpret_cast<uint4 *>(z);
            for (uint32_t i = tid; i < n / 4; i += nthr) z4[i] = make_uint4(0, 0, 0, 0);
        } else {
            for (uint32_t i = tid; i < n; i += nthr) z[i] = 0;
        }
    }
    if (a.psteps) {
        for (uint32_t i = tid; i < a.total; i += nthr) {
            const uint32_t s = a.psteps[i], node = s >> 1;
            uint32_t l = node < a.n_nodes ? a.map[node] : a.L;
            if (l >= a.L) { atomicOr(a.err, 1u); l = a.L; }
            a.out[i] = (uint16_t)((l << 1) | (s & 1u));
        }
    }
}

// Which bucket a kBlock-row tile belongs to, and this thread's row in it.
struct RowRef {
    const uint16_t *ptr;
    uint32_t m, sorted;
    bool valid;
};
__device__ __forceinline__ int tile_bucket(const Buckets &bk, uint32_t tile) {
    int b = 1;
#pragma unroll
    for (int k = 2; k < kNumBuckets; ++k) b += (tile >= bk.tile_start[k]) ? 1 : 0;
    return b;
}
__device__ __forceinline__ RowRef tile_row(const Buckets &bk, int b, uint32_t tile) {
    RowRef r;
    const uint32_t row = (tile - bk.tile_start[b]) * kBlock + threadIdx.x;
    r.valid = row < bk.count[b];
    r.sorted = bk.row_start[b] + row;
    if (b == kLongBucket) {
        const uint2 d = r.valid ? __ldg(bk.ldesc + row) : make_uint2(bk.step_base[b], 0u);
        r.ptr = bk.bstep + d.x;
        r.m = d.y;
    } else {
        r.ptr = bk.bstep + bk.step_base[b] + (r.valid ? (size_t)row * row_stride(b) : 0);
        r.m = (uint32_t)b;
    }
    return r;
}
// sorted row -> (steps, m); used by the exact-DP worklist kernels
__device__ __forceinline__ RowRef locate_row(const Buckets &bk, uint32_t sorted) {
    int b = 0;
#pragma unroll
    for (int k = 1; k < kNumBuckets; ++k) b += (sorted >= bk.row_start[k]) ? 1 : 0;
    RowRef r;
    const uint32_t row = sorted - bk.row_start[b];
    r.valid = true;
    r.sorted = sorted;
    if (b == kLongBucket) {
        const uint2 d = __ldg(bk.ldesc + row);
        r.ptr = bk.bstep + d.x;
        r.m = d.y;
    } else {
        r.ptr = bk.bstep + bk.step_base[b] + (size_t)row * row_stride(b);
        r.m = (uint32_t)b;
    }
    return r;
}

// ------------------------------------------------------------------------------------------------
// K0c: node -> first-position table of each candidate path.  One CTA per path; the 32-bit working
// row lives in shared memory ((L+1)*4 B <= 128 KB).  atomicMin on (pos<<1|minus) keeps the first
// visit, a second pass flags nodes visited more than once (NodeTable counts > 1, reference
// include/nodetable.h:4-7).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock)
build_tables_kernel(const uint32_t *__restrict__ poff, const uint16_t *__restrict__ pstep,
                    uint32_t L, uint32_t stride, uint16_t *__restrict__ ptab,
                    uint16_t *__restrict__ pnext, uint32_t *__restrict__ masks,
                    uint32_t *__restrict__ fplane, const MaskGeom mg, const PairDict pd) {
    extern __shared__ uint32_t row[];
    pdl_trigger();   // the scoring kernel may be scheduled now; it waits for this grid before its first read
    const uint32_t p = blockIdx.x;
    const uint32_t o0 = poff[p], n = poff[p + 1] - o0;
    if (n == 0) {            // padding slot of an incremental chunk: no bits, an all-absent table row
        for (uint32_t i = threadIdx.x; i < stride; i += blockDim.x) ptab[(size_t)p * stride + i] = 0;
        return;
    }
    const uint16_t *s = pstep + o0;
    for (uint32_t i = threadIdx.x; i <= L; i += blockDim.x) row[i] = 0xFFFFFFFFu;
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) {
        uint32_t st = s[i], node = st >> 1;
        if (node < L) atomicMin(&row[node], (i << 1) | (st & 1u));
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) {
        uint32_t node = s[i] >> 1;
        if (node < L && ((row[node] & 0x7FFFFFFFu) >> 1) != i) atomicOr(&row[node], 0x80000000u);
    }
    __syncthreads();
    uint16_t *dst = ptab + (size_t)p * stride;
    for (uint32_t i = threadIdx.x; i < stride; i += blockDim.x) {
        uint32_t v = i <= L ? row[i] : 0xFFFFFFFFu;
        uint32_t e = 0;
        if (v != 0xFFFFFFFFu)
            e = ((((v & 0x7FFFFFFFu) >> 1) + 1u) << 2) | ((v & 1u) << 1) | (v >> 31);   // see TabEntry
        dst[i] = (uint16_t)e;
    }
    // next visit of the same node (0xFFFF = none): only positions of twice-visited nodes scan
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) {
        const uint32_t node = s[i] >> 1;
        uint32_t nx = 0xFFFFu;
        if (node < L && (row[node] >> 31) && row[node] != 0xFFFFFFFFu) {
            for (uint32_t k = i + 1; k < n; ++k)
                if ((uint32_t)(s[k] >> 1) == node) { nx = k; break; }
        }
        pnext[o0 + i] = (uint16_t)nx;
    }
    if (masks == nullptr) return;
    // ---- this path's bit in the batch's bitsets (chunk = 32*W paths, entry = W words) ----------
    const uint32_t per_chunk = 32u * mg.W;
    const uint32_t q = p % per_chunk;
    uint32_t *tab0 = masks + (size_t)(p / per_chunk) * mg.entries * mg.W;
    uint32_t *tabw = tab0 + (q >> 5);
    const uint32_t bit = 1u << (q & 31u);
    // node block: {N, F+|F-} per (node i, word w) in the table, {F+, F-} in the global fplane
    for (uint32_t i = threadIdx.x; i < L; i += blockDim.x) {
        const uint32_t v = row[i];
        if (v == 0xFFFFFFFFu) continue;
        uint32_t *e = tab0 + (size_t)mg.node_base * mg.W + ((size_t)i * mg.W + (q >> 5)) * 2;
        atomicOr(e, bit);                                                  // N: on the path
        if (((v & 0x7FFFFFFFu) >> 1) == 0) {                               // path[0] is this node
            atomicOr(e + 1, bit);
            atomicOr(fplane + ((size_t)(p / per_chunk) * (L + 1) * mg.W + (size_t)i * mg.W + (q >> 5)) * 2 +
                         (s[0] & 1u), bit);                                // F+ / F-
        }
    }
    for (uint32_t m = threadIdx.x; m <= (uint32_t)kMaxBucket; m += blockDim.x)
        if (n < m) atomicOr(tabw + (size_t)(mg.lm_base + m) * mg.W, bit);           // LM: n < m
    for (uint32_t i = threadIdx.x; i + 1 < n; i += blockDim.x) {
        uint32_t flip;
        const uint32_t x = s[i], y = s[i + 1];
        const uint32_t key = pair_canonical(x, y, &flip);
        const uint32_t h = pair_find(pd, key);
        if (h == kPairEmpty) continue;       // no alignment carries this pair
        const uint32_t c2 = 2u * __ldg(pd.ids + h);
        atomicOr(tabw + (size_t)(mg.pair_base + c2 + flip) * mg.W, bit);
        if (y == (x ^ 1u)) atomicOr(tabw + (size_t)(mg.pair_base + c2 + 1u) * mg.W, bit);   // its own twin
    }
    // Broken triples.  With a node visited more than once, "every pair of B is taken by the path"
    // no longer implies "B occurs": (y, x) may sit at one visit of x and (x, z) at another.  Such a
    // triple (y, x, z) — both pairs taken, never contiguous — is hashed into BT; so is every real
    // triple around a pair the path takes twice.  B occurs  <=>  all its pairs are taken and none of
    // its triples is broken (bitset.cu checks BT only for rows whose interior touches T).
    __syncthreads();   // pnext of the whole path is written
    auto pid_of = [&](uint32_t x, uint32_t y) -> uint32_t {
        uint32_t flip;
        const uint32_t key = pair_canonical(x, y, &flip);
        const uint32_t h = pair_find(pd, key);
        return h == kPairEmpty ? 0xFFFFFFFFu : 2u * __ldg(pd.ids + h) + flip;
    };
    auto mark = [&](uint32_t y, uint32_t x, uint32_t z) {
        const uint32_t qa = pid_of(y, x), qb = pid_of(x, z);
        if (qa == 0xFFFFFFFFu || qb == 0xFFFFFFFFu) return;   // no alignment carries it
        // a self-reverse-complement pair answers to both of its ids (2c and 2c+1)
        const uint32_t na = (x == (y ^ 1u)) ? 2u : 1u, nb = (z == (x ^ 1u)) ? 2u : 1u;
        for (uint32_t ia = 0; ia < na; ++ia)
            for (uint32_t ib = 0; ib < nb; ++ib) {
                const uint32_t a2 = qa ^ ia, b2 = qb ^ ib;
                const uint32_t slot = triple_slot(a2, b2);
                atomicOr(tabw + (size_t)(mg.bt_base + slot) * mg.W, bit);
                if (a2 == (b2 ^ 1u)) atomicOr(tabw + (size_t)(mg.bt_base + (slot ^ 1u)) * mg.W, bit);
            }
    };
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) {
        const uint32_t node = s[i] >> 1;
        if (node >= L || !(row[node] >> 31) || row[node] == 0xFFFFFFFFu) continue;
        const uint32_t first = (row[node] & 0x7FFFFFFFu) >> 1;
        for (uint32_t v2 = first; v2 != 0xFFFFu; v2 = pnext[o0 + v2]) {
            if (v2 == i || s[v2] != s[i]) continue;
            if (i >= 1 && v2 + 1 < n) {
                // cross triple (pred of visit i, x, succ of visit v2): real at some visit?
                bool real = false;
                for (uint32_t v = first; v != 0xFFFFu && !real; v = pnext[o0 + v])
                    real = v >= 1 && v + 1 < n && s[v] == s[i] && s[v - 1] == s[i - 1] &&
                           s[v + 1] == s[v2 + 1];
                if (!real) mark(s[i - 1], s[i], s[v2 + 1]);
            }
            if (i + 1 < n && v2 + 1 < n && s[v2 + 1] == s[i + 1]) {
                // the pair (x, succ) is taken twice: its surrounding real triples are ambiguous
                if (i >= 1) mark(s[i - 1], s[i], s[i + 1]);
                if (i + 2 < n) mark(s[i], s[i + 1], s[i + 2]);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Exact traceback score T of one (path A, alignment B) direction, forward-only.
//
// The reference fills dp[n+1][m+1] (src/alignments.cpp:499-509) and then walks back from (n,m)
// (:511-554).  The walk is deterministic given dp, so its accumulated score can be carried
// FORWARD with the DP row:  F[i][j]  = score of the walk from (i,j) once a B step was consumed
// (SBlen > 0), F0[i][j] = score of the walk from (i,j) while SBlen == 0 (up-moves free).
//   diag (dp[i][j] == dp[i-1][j-1]+S):   F = F0 = F[i-1][j-1] + S
//   up   (dp[i-1][j] >= dp[i][j-1]):     F = F[i-1][j] - 1,  F0 = F0[i-1][j]
//   left:                                F = F0 = F[i][j-1] - 1
// borders F = F0 = 0 (ii == 0 / jj == 0 moves are free, :517-524); T = F0[n][m].
// Row 0 of dp is -j for j <= n and 0 beyond (loop bound n at :500), column 0 is 0.
// One row of (dp, F, F0) lives in registers: MCAP+1 values each, fully unrolled.
// ------------------------------------------------------------------------------------------------
// Sparse rows.  A row i of the matrix whose path step A[i-1] visits no node of B is all-mismatch,
// and all-mismatch rows drive the DP row into a fixed point within at most m rows:
//     dp[j] = F[j] = F0[j] = -j   for j < m        (diagonal move, checked first, reproduces -j)
//     dp[m] = best so far;  best > -m: "up" move, F[m] -= 1 per row, F0[m] unchanged
//                           best == -m: diagonal, F[m] = F0[m] = -m
// (column 0 is 0, the last column takes up-moves for free, src/alignments.cpp:503-504).  So only the
// rows that touch B, plus the few rows after each until the fixed point is reached again, are
// computed cell by cell; a run of k further all-mismatch rows is applied in closed form.  For an
// evalPath path of 1000 steps against a 5-step alignment that is ~30 rows instead of 1000.
// `touch` = bitmap over path positions (bit p: node of A[p] is a node of B), either orientation.
constexpr int kTouchWords = (GFE_MAX_STEPS_DEV + 31) / 32;

template <int MCAP>
__device__ __forceinline__ int exact_T_regs(const uint16_t *__restrict__ A, int n,
                                            const uint32_t (&B)[MCAP], int m,
                                            const uint32_t *__restrict__ touch) {
    if (m == 0) return 0;   // the walk only consumes A, free
    int dp[MCAP + 1], F[MCAP + 1], F0[MCAP + 1];
#pragma unroll
    for (int j = 0; j <= MCAP; ++j) {
        dp[j] = (j <= n) ? -j : 0;
        F[j] = 0;
        F0[j] = 0;
    }
    const int nw = (n + 31) >> 5;
    int i = 1;
    while (i <= n) {
        const uint32_t a = A[i - 1];
        int diag_dp = 0, diag_F = 0;     // column 0 of row i-1
        int left_dp = 0, left_F = 0;     // column 0 of row i
        bool steady = true;              // columns 1..m-1 of row i at the fixed point
#pragma unroll
        for (int j = 1; j <= MCAP; ++j) {
            if (j > m) break;
            const int S = (a == B[j - 1]) ? 0 : -1;
            const int up_dp = dp[j], up_F = F[j], up_F0 = F0[j];
            const int d = diag_dp + S;
            const int u = up_dp + (j < m ? -1 : 0);
            const int l = left_dp - 1;
            const int v = imax(d, imax(u, l));
            int nF, nF0;
            if (v == d) {
                nF = diag_F + S;
                nF0 = nF;
            } else if (up_dp >= left_dp) {
                nF = up_F - 1;
                nF0 = up_F0;
            } else {
                nF = left_F - 1;
                nF0 = nF;
            }
            diag_dp = up_dp;
            diag_F = up_F;
            dp[j] = v;
            F[j] = nF;
            F0[j] = nF0;
            left_dp = v;
            left_F = nF;
            if (j < m) steady = steady && v == -j && nF == -j && nF0 == -j;
        }
        const int p = i - 1;
        const bool touched = (touch[p >> 5] >> (p & 31)) & 1u;
        if (touched || !steady) {
            ++i;
            continue;
        }
        // next row that touches B (n + 1: none); rows i+1 .. nt-1 in closed form
        int w = i >> 5;
        uint32_t bits = w < nw ? touch[w] & (0xFFFFFFFFu << (i & 31)) : 0u;
        while (!bits && ++w < nw) bits = touch[w];
        const int nt = bits ? w * 32 + __ffs(bits) : n + 1;
        const int k = nt - 1 - i;
        if (k > 0) {
#pragma unroll
            for (int j = 1; j <= MCAP; ++j)
                if (j == m) {
                    if (dp[j] > -m) F[j] -= k;
                    else F[j] = F0[j] = -m;
                }
        }
        i = nt;
    }
    int res = 0;
#pragma unroll
    for (int j = 1; j <= MCAP; ++j)
        if (j == m) res = F0[j];
    return res;
}

// bitmap of the path positions that visit a node of B (tab: node -> first visit, nxt: visit chain)
__device__ __forceinline__ void touch_map(const uint16_t *__restrict__ tab, const uint16_t *__restrict__ nxt,
                                          int n, const uint16_t *__restrict__ b, int m, uint32_t *touch) {
    const int nw = (n + 31) >> 5;
    for (int w = 0; w < nw; ++w) touch[w] = 0;
    for (int j = 0; j < m; ++j) {
        const uint32_t e = tab[b[j] >> 1];
        if (e == kTabAbsent) continue;
        for (int q = (int)(e >> 2) - 1; q != 0xFFFF; q = nxt[q]) touch[q >> 5] |= 1u << (q & 31);
    }
}

// Load an alignment (local ids) into registers, forward and reverse-complemented
// (Path::reverseComplement, reference include/alignments.h:64-70).
template <int MCAP>
__device__ __forceinline__ void load_fw_rc(const uint16_t *__restrict__ b, int m,
                                           uint32_t (&Bf)[MCAP], uint32_t (&Br)[MCAP]) {
#pragma unroll
    for (int j = 0; j < MCAP; ++j) {
        Bf[j] = (j < m) ? (uint32_t)b[j] : 0xFFFFFFFFu;
        Br[j] = (j < m) ? ((uint32_t)b[m - 1 - j] ^ 1u) : 0xFFFFFFFFu;
    }
}

template <int MCAP>
__device__ __noinline__ void exact_both_regs(const uint16_t *__restrict__ A, int n,
                                             const uint16_t *__restrict__ b, int m, bool need_both,
                                             const uint32_t *__restrict__ touch, int *tf, int *tr) {
    uint32_t Bf[MCAP], Br[MCAP];
    load_fw_rc<MCAP>(b, m, Bf, Br);
    int f = exact_T_regs<MCAP>(A, n, Bf, m, touch);
    int r = -1;
    if (need_both || f != 0) r = exact_T_regs<MCAP>(A, n, Br, m, touch);
    *tf = f;
    *tr = r;
}

// tab / nxt: the path's node -> first-visit table row and visit chain (build_tables_kernel)
__device__ __forceinline__ void exact_both(const uint16_t *__restrict__ A, int n,
                                           const uint16_t *__restrict__ tab,
                                           const uint16_t *__restrict__ nxt,
                                           const uint16_t *__restrict__ b, int m, bool need_both,
                                           int *tf, int *tr) {
    uint32_t touch[kTouchWords];
    touch_map(tab, nxt, n, b, m, touch);
    if (m <= 8) exact_both_regs<8>(A, n, b, m, need_both, touch, tf, tr);
    else if (m <= 16) exact_both_regs<16>(A, n, b, m, need_both, touch, tf, tr);
    else exact_both_regs<kExactRegSteps>(A, n, b, m, need_both, touch, tf, tr);
}

// Exact T for alignments longer than kExactRegSteps: one warp per direction, anti-diagonal
// wavefront.  Diagonal d holds the cells (i = d - j, j); three diagonals of dp and F and two of
// F0 live in shared memory (8 arrays of m+1 ints, indexed by j).  Border cells (i == 0 or
// j == 0) are produced on the fly, so the loop starts at d = 0 with nothing to initialise.
__device__ int warp_exact_T(const uint16_t *__restrict__ A, int n, const uint16_t *__restrict__ b,
                            int m, bool rc, int *scratch) {
    const int lane = threadIdx.x & 31;
    const int w = m + 1;
    int *dpA = scratch, *dpB = scratch + w, *dpC = scratch + 2 * w;          // d-2, d-1, d
    int *fA = scratch + 3 * w, *fB = scratch + 4 * w, *fC = scratch + 5 * w;  // F: d-2, d-1, d
    int *zB = scratch + 6 * w, *zC = scratch + 7 * w;                         // F0: d-1, d
    int result = 0;
    for (int d = 0; d <= n + m; ++d) {
        const int jlo = d > n ? d - n : 0;
        const int jhi = d < m ? d : m;
        for (int j = jlo + lane; j <= jhi; j += 32) {
            const int i = d - j;
            int v, f, z;
            if (i == 0) {
                v = (j <= n) ? -j : 0;   // row 0 is written for columns 0..n only (:500)
                f = 0;
                z = 0;
            } else if (j == 0) {
                v = 0;
                f = 0;
                z = 0;
            } else {
                const uint32_t bj = rc ? ((uint32_t)b[m - j] ^ 1u) : (uint32_t)b[j - 1];
                const int S = ((uint32_t)A[i - 1] == bj) ? 0 : -1;
                const int diag_dp = dpA[j - 1], diag_F = fA[j - 1];
                const int up_dp = dpB[j], up_F = fB[j], up_Z = zB[j];
                const int left_dp = dpB[j - 1], left_F = fB[j - 1];
                const int dd = diag_dp + S, u = up_dp + (j < m ? -1 : 0), l = left_dp - 1;
                v = imax(dd, imax(u, l));
                if (v == dd) { f = diag_F + S; z = f; }
                else if (up_dp >= left_dp) { f = up_F - 1; z = up_Z; }
                else { f = left_F - 1; z = f; }
            }
            dpC[j] = v; fC[j] = f; zC[j] = z;
            if (i == n && j == m) result = z;   // last diagonal: its single cell is lane 0's
        }
        __syncwarp();
        int *t = dpA; dpA = dpB; dpB = dpC; dpC = t;
        t = fA; fA = fB; fB = fC; fC = t;
        t = zB; zB = zC; zC = t;
    }
    return __shfl_sync(kFull, result, 0);
}

// ------------------------------------------------------------------------------------------------
// K1: score_paths — evaluatePath(filter) for a chunk of candidate paths x a stream of alignments.
// grid = (alignment-tile workers, path chunks).  The chunk's node->position tables sit in shared
// memory; one thread owns one alignment row (its M steps in registers — rows are bucketed by
// length, so the step loop is fully unrolled and branch-free) and tests it against every table of
// the chunk: one shared-memory lookup per step answers the id filter, the forward occurrence and
// the reverse-complement occurrence at once.
//
// Table entry e = (pos+1)<<2 | minus<<1 | twice, 0 = absent.  With x_j = e_j ^ (minus(B_j)<<1):
//   B occurs forward at pos(B_0)      <=>  all (x_j - 4j) == x_0 above bit 0, bit 1 of x_0 clear
//   rc(B) occurs ending at pos(B_0)   <=>  all (x_j + 4j) == x_0 above bit 0, bit 1 of x_0 set
//   every node on the path            <=>  min_j e_j != 0;   touches path[0]  <=>  min_j e_j < 8
// Verdict rule (proved + fuzzed against the reference, SURVEY.md Appendix A.3, tests/):
//   filtered out -> nothing;  m > n -> good;  B or rc(B) occurs contiguously -> good;
//   otherwise bad, unless B touches a node visited twice (search every visit) or the node of
//   path[0] with a suffix/prefix overhang (exact DP worklist).
// Per path the warp's verdicts are two ballots; lane (p mod 32) keeps the popcounts in registers,
// so the inner loop has no atomics.  The CTA flushes chunk x {bad, good} once at the end.
// ------------------------------------------------------------------------------------------------

// Rare second look at an alignment the table test could not settle (m <= n, no occurrence found
// through first positions).  A = the path (local ids), nxt = next visit of the same node.
__device__ __noinline__ int slow_verdict(const uint16_t *__restrict__ A,
                                         const uint16_t *__restrict__ nxt, int n,
                                         const uint16_t *__restrict__ tab,
                                         const uint16_t *__restrict__ b, int m, bool twice) {
    if (twice) {
        // try every visit of B[0]'s node as the start of a forward occurrence and as the end of a
        // reverse-complement occurrence, comparing steps directly
        const uint32_t e = tab[b[0] >> 1];
        if (e != kTabAbsent) {
            for (int q = (int)(e >> 2) - 1; q != 0xFFFF; q = nxt[q]) {
                if (q + m <= n) {
                    bool ok = true;
                    for (int j = 0; j < m && ok; ++j) ok = (A[q + j] == b[j]);
                    if (ok) return kGood;
                }
                if (q - (m - 1) >= 0) {
                    bool ok = true;
                    for (int j = 0; j < m && ok; ++j) ok = (A[q - j] == (uint16_t)(b[j] ^ 1u));
                    if (ok) return kGood;
                }
            }
        }
    }
    // only if a proper suffix of B (or of rc(B)) equals a prefix of A can the traceback still
    // score 0 (SURVEY.md Appendix A.3 class iv) -> exact DP
    for (int k = 1; k < m; ++k) {
        bool fw = true, rc = true;
        for (int t = 0; t < k && (fw || rc); ++t) {
            const uint16_t a = A[t];
            fw = fw && (b[m - k + t] == a);
            rc = rc && ((uint16_t)(b[k - 1 - t] ^ 1u) == a);
        }
        if (fw || rc) return kExact;
    }
    return kBad;
}

struct Accum {
    uint32_t bad[kMaxChunk / 32], good[kMaxChunk / 32];
};

// verdict bookkeeping shared by the fixed-length and the long-row paths
template <bool kFilter>
__device__ __forceinline__ void settle(const ScoreArgs &a, uint32_t p_glob, const uint16_t *tab,
                                       int n, const uint16_t *row_ptr, int m, uint32_t sorted,
                                       bool valid, uint32_t mn, uint32_t mn_present,
                                       uint32_t orall, uint32_t accf, uint32_t accr, uint32_t x0,
                                       bool &good, bool &bad) {
    const bool pass = mn != 0;
    const bool occ = pass && ((((accf & 0xFFFEu) == 0) && !(x0 & 2u)) ||
                              (((accr & 0xFFFEu) == 0) && (x0 & 2u)));
    const bool gt = m > n;
    const bool twice = orall & 1u;
    const bool special = twice || ((kFilter ? mn : mn_present) >> 2) == 1u;
    const bool in = valid && (kFilter ? pass : true);
    if (in && m > GFE_MAX_STEPS_DEV && a.err) atomicOr(a.err, 4u);   // the reference's dp[][] cannot hold this row
    good = in && (gt || occ);
    bad = in && !gt && !occ && !special;
    if (in && !gt && !occ && special) {
        const uint32_t po = a.poff[p_glob];
        const int v = slow_verdict(a.pstep + po, a.pnext + po, n, tab, row_ptr, m, twice);
        good = v == kGood;
        bad = v == kBad;
        if (v == kExact) {
            const uint32_t idx = atomicAdd(a.work_count, 1u);
            if (idx < a.work_cap) a.work[idx] = make_uint2(p_glob | a.path_tag, sorted);
            else atomicOr(a.work_count + 1, 1u);
        }
    }
}

template <int M, bool kFilter>
__device__ __forceinline__ void score_tile_fixed(const ScoreArgs &a, const uint16_t *s_tab,
                                                 const uint32_t *s_n, uint32_t pc, uint32_t p_base,
                                                 const RowRef &r, Accum &acc, int lane) {
    uint32_t boff[M], bo2[M];   // byte offset of the node's table entry; minus(B_j) << 1
#pragma unroll
    for (int j = 0; j < M; ++j) {
        const uint32_t st = __ldg(r.ptr + j);
        boff[j] = st & 0xFFFEu;
        bo2[j] = (st & 1u) << 1;
    }
    const unsigned char *tabb = reinterpret_cast<const unsigned char *>(s_tab);
    const uint32_t stride_b = a.tab_stride * 2;
#pragma unroll
    for (int h = 0; h < kMaxChunk / 32; ++h) {
        const uint32_t p_lo = h * 32;
        if (p_lo >= pc) break;
        const uint32_t p_hi = min(pc, p_lo + 32);
        for (uint32_t p = p_lo; p < p_hi; ++p) {
            const unsigned char *row = tabb + (size_t)p * stride_b;
            uint32_t e = *reinterpret_cast<const uint16_t *>(row + boff[0]);
            const uint32_t x0 = e ^ bo2[0];
            uint32_t mn = e, mnp = e ? e : 0xFFFFu, orall = e, accf = 0, accr = 0;
#pragma unroll
            for (int j = 1; j < M; ++j) {
                e = *reinterpret_cast<const uint16_t *>(row + boff[j]);
                const uint32_t x = e ^ bo2[j];
                accf |= (x - 4u * j) ^ x0;
                accr |= (x + 4u * j) ^ x0;
                mn = min(mn, e);
                if (!kFilter) mnp = min(mnp, e ? e : 0xFFFFu);
                orall |= e;
            }
            bool good, bad;
            settle<kFilter>(a, p_base + p, reinterpret_cast<const uint16_t *>(row), (int)s_n[p],
                            r.ptr, M, r.sorted, r.valid, mn, mnp, orall, accf, accr, x0, good, bad);
            const unsigned gm = __ballot_sync(kFull, good);
            const unsigned bm = __ballot_sync(kFull, bad);
            if ((uint32_t)lane == (p & 31u)) {
                acc.good[h] += (uint32_t)__popc(gm);
                acc.bad[h] += (uint32_t)__popc(bm);
            }
        }
    }
}

template <bool kFilter>
__device__ __forceinline__ void score_tile_long(const ScoreArgs &a, const uint16_t *s_tab,
                                                const uint32_t *s_n, uint32_t pc, uint32_t p_base,
                                                const RowRef &r, Accum &acc, int lane) {
    const int m = (int)r.m;
#pragma unroll
    for (int h = 0; h < kMaxChunk / 32; ++h) {
        const uint32_t p_lo = h * 32;
        if (p_lo >= pc) break;
        const uint32_t p_hi = min(pc, p_lo + 32);
        for (uint32_t p = p_lo; p < p_hi; ++p) {
            const uint16_t *tab = s_tab + (size_t)p * a.tab_stride;
            uint32_t mn = 0xFFFFu, mnp = 0xFFFFu, orall = 0, accf = 0, accr = 0, x0 = 0;
            if (r.valid) {
                for (int j = 0; j < m; ++j) {
                    const uint32_t st = __ldg(r.ptr + j);
                    const uint32_t e = tab[st >> 1];
                    const uint32_t x = e ^ ((st & 1u) << 1);
                    if (j == 0) x0 = x;
                    accf |= (x - 4u * j) ^ x0;
                    accr |= (x + 4u * j) ^ x0;
                    mn = min(mn, e);
                    mnp = min(mnp, e ? e : 0xFFFFu);
                    orall |= e;
                    if (kFilter && e == 0) break;
                }
            }
            bool good, bad;
            settle<kFilter>(a, p_base + p, tab, (int)s_n[p], r.ptr, m, r.sorted, r.valid, mn, mnp,
                            orall, accf, accr, x0, good, bad);
            const unsigned gm = __ballot_sync(kFull, good);
            const unsigned bm = __ballot_sync(kFull, bad);
            if ((uint32_t)lane == (p & 31u)) {
                acc.good[h] += (uint32_t)__popc(gm);
                acc.bad[h] += (uint32_t)__popc(bm);
            }
        }
    }
}

template <bool kFilter>
__global__ void __launch_bounds__(kBlock)
score_paths_kernel(const ScoreArgs a, const uint32_t tile_lo, const uint32_t tile_hi,
                   const bool add_empty) {
    extern __shared__ __align__(16) unsigned char smem[];
    uint16_t *s_tab = reinterpret_cast<uint16_t *>(smem);
    uint32_t *s_cnt = reinterpret_cast<uint32_t *>(smem + (size_t)a.chunk * a.tab_stride * 2);
    uint32_t *s_n = s_cnt + 2 * a.chunk;

    const uint32_t p_base = blockIdx.y * a.chunk;
    const uint32_t pc = min(a.chunk, a.P - p_base);
    {
        const uint4 *src = reinterpret_cast<const uint4 *>(a.ptab + (size_t)p_base * a.tab_stride);
        uint4 *dst = reinterpret_cast<uint4 *>(s_tab);
        const uint32_t nvec = pc * (a.tab_stride >> 3);
        for (uint32_t i = threadIdx.x; i < nvec; i += blockDim.x) dst[i] = __ldg(src + i);
        for (uint32_t i = threadIdx.x; i < 2 * pc; i += blockDim.x) s_cnt[i] = 0;
        for (uint32_t i = threadIdx.x; i < pc; i += blockDim.x)
            s_n[i] = a.poff[p_base + i + 1] - a.poff[p_base + i];
    }
    __syncthreads();

    const int lane = threadIdx.x & 31;
    Accum acc;
#pragma unroll
    for (int h = 0; h < kMaxChunk / 32; ++h) acc.bad[h] = acc.good[h] = 0;

    for (uint32_t tile = tile_lo + blockIdx.x; tile < tile_hi; tile += gridDim.x) {
        const int b = tile_bucket(a.bk, tile);
        const RowRef r = tile_row(a.bk, b, tile);
        switch (b) {
#define GFE_CASE(M_) case M_: score_tile_fixed<M_, kFilter>(a, s_tab, s_n, pc, p_base, r, acc, lane); break;
            GFE_CASE(1) GFE_CASE(2) GFE_CASE(3) GFE_CASE(4) GFE_CASE(5) GFE_CASE(6) GFE_CASE(7)
            GFE_CASE(8) GFE_CASE(9) GFE_CASE(10) GFE_CASE(11) GFE_CASE(12) GFE_CASE(13)
            GFE_CASE(14) GFE_CASE(15) GFE_CASE(16)
#undef GFE_CASE
            default: score_tile_long<kFilter>(a, s_tab, s_n, pc, p_base, r, acc, lane); break;
        }
    }
    // lane l of every warp holds the counts of paths l, l+32 of the chunk
#pragma unroll
    for (int h = 0; h < kMaxChunk / 32; ++h) {
        const uint32_t p = h * 32 + lane;
        if (p < pc) {
            if (acc.bad[h]) atomicAdd(&s_cnt[2 * p], acc.bad[h]);
            if (acc.good[h]) atomicAdd(&s_cnt[2 * p + 1], acc.good[h]);
        }
    }
    __syncthreads();
    // alignments without steps pass the filter and score 0 against every path: good
    const uint32_t empty_rows = (add_empty && blockIdx.x == 0) ? a.bk.count[0] : 0u;
    const size_t set_off = ((a.path_tag >> 30) & 1u) << 1;
    for (uint32_t i = threadIdx.x; i < 2 * pc; i += blockDim.x) {
        const uint32_t v = s_cnt[i] + ((i & 1u) ? empty_rows : 0u);
        if (v) atomicAdd(&a.counts[(size_t)a.count_stride * (p_base + (i >> 1)) + set_off + (i & 1u)], v);
    }
}

// K1s: the (path, alignment) pairs the bitset kernel could not settle — the alignment touches a
// node the path visits twice (pair bitsets are then only necessary, not sufficient) or the node of
// path[0] (start overhang, SURVEY.md Appendix A.3 class iv).  Same second look as the legacy
// kernel: direct comparison along the visit chain, then the suffix/prefix test; what is still
// open goes to the exact-DP worklist.
__global__ void __launch_bounds__(128)
slow_worklist_kernel(const ScoreArgs a, const uint2 *__restrict__ slow,
                     const uint32_t *__restrict__ slow_count, const uint32_t slow_cap) {
    pdl_wait();      // the scoring kernel's list and counts
    pdl_trigger();
    // volatile: a load through the read-only path may be hoisted above the wait
    const uint32_t cnt = min(*reinterpret_cast<const volatile uint32_t *>(slow_count), slow_cap);
    for (uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x; idx < cnt;
         idx += gridDim.x * blockDim.x) {
        uint2 w = slow[idx];
        if (w.x == kSlowHole) continue;     // unused tail of a warp's reserved window
        const uint32_t guess = w.x >> 31;   // the bitset kernel already counted this pair good (1) / bad (0)
        w.x &= 0x7FFFFFFFu;                 // bit 30 (accumulator set) travels on to the exact list
        const uint32_t p = entry_path(w.x);
        const size_t cb = entry_count_base(a.count_stride, w.x);
        const uint32_t wt = a.rshift ? 1u << a.rshift[w.y] : 1u;   // alignments this view row stands for
        const RowRef r = locate_row(a.bk, w.y);
        const int m = (int)r.m;
        const uint32_t po = a.poff[p];
        const int n = (int)(a.poff[p + 1] - po);
        const uint16_t *tab = a.ptab + (size_t)p * a.tab_stride;
        // always walk the visit chain: a pair the bitsets flagged for a twice-visited node needs
        // it, and for a start-overhang suspect it finds nothing at the cost of one table read
        const int v = slow_verdict(a.pstep + po, a.pnext + po, n, tab, r.ptr, m, true);
        if (v == kExact) {
            atomicSub(&a.counts[cb + guess], wt);   // the exact DP adds the right one
            const uint32_t k = atomicAdd(a.work_count, 1u);
            if (k < a.work_cap) a.work[k] = w;
            else atomicOr(a.work_count + 1, 1u);
        } else if ((v == kGood ? 1u : 0u) != guess) {          // correct the rare wrong guess
            atomicSub(&a.counts[cb + guess], wt);
            atomicAdd(&a.counts[cb + (guess ^ 1u)], wt);
        }
    }
}

// K1b: exact verdict for the worklist of ambiguous (path, alignment) pairs.  One thread runs one
// pair's DP, so a warp costs its longest pair in every register tier its lanes touch.  The list
// arrives in arbitrary order; each CTA therefore takes a tile of kExactTile pairs, counting-sorts
// it in shared memory by (register tier of m, n / 32) and hands consecutive sorted pairs to
// consecutive lanes: the lanes of a warp then run the same tier on paths of similar length.
constexpr int kExactTile = 256;
constexpr int kExactBins = 32;          // tier (0..2) * 8 + min(n / 32, 7); bin 31 = not for this kernel

__global__ void __launch_bounds__(128)
exact_worklist_kernel(const ScoreArgs a) {
    __shared__ uint2 s_item[kExactTile];
    __shared__ uint32_t s_hist[kExactBins], s_base[kExactBins + 1];
    pdl_wait();      // the second-look kernel's worklist
    const uint32_t cnt = min(*reinterpret_cast<const volatile uint32_t *>(a.work_count), a.work_cap);
    for (uint32_t tile0 = blockIdx.x * kExactTile; tile0 < cnt; tile0 += gridDim.x * kExactTile) {
        if (threadIdx.x < kExactBins) s_hist[threadIdx.x] = 0;
        __syncthreads();
        uint2 w[kExactTile / 128];
        uint32_t bin[kExactTile / 128], pos[kExactTile / 128];
#pragma unroll
        for (int k = 0; k < kExactTile / 128; ++k) {
            const uint32_t idx = tile0 + k * 128 + threadIdx.x;
            bin[k] = kExactBins;                      // nothing loaded
            if (idx < cnt) {
                w[k] = a.work[idx];
                const uint32_t m = locate_row(a.bk, w[k].y).m;
                const uint32_t n = a.poff[entry_path(w[k].x) + 1] - a.poff[entry_path(w[k].x)];
                const uint32_t tier = m <= 8 ? 0u : (m <= 16 ? 1u : 2u);
                // long alignments: exact_worklist_long_kernel
                bin[k] = m > (uint32_t)kExactRegSteps ? (uint32_t)kExactBins - 1u : tier * 8u + min(n >> 5, 7u);
                pos[k] = atomicAdd(&s_hist[bin[k]], 1u);
            }
        }
        __syncthreads();
        if (threadIdx.x < 32) {                       // exclusive prefix over the bins
            const uint32_t h = s_hist[threadIdx.x];
            uint32_t incl = h;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
                if ((int)threadIdx.x >= o) incl += t;
            }
            s_base[threadIdx.x] = incl - h;
            if (threadIdx.x == 31) s_base[32] = incl;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kExactTile / 128; ++k)
            if (bin[k] < (uint32_t)kExactBins) s_item[s_base[bin[k]] + pos[k]] = w[k];
        __syncthreads();
        const uint32_t todo = s_base[kExactBins - 1];  // everything below the "not for this kernel" bin
        for (uint32_t j = threadIdx.x; j < todo; j += blockDim.x) {
            const uint2 it = s_item[j];
            const RowRef r = locate_row(a.bk, it.y);
            const uint32_t p = entry_path(it.x);
            const uint32_t po = a.poff[p];
            const int n = (int)(a.poff[p + 1] - po);
            int tf, tr;
            exact_both(a.pstep + po, n, a.ptab + (size_t)p * a.tab_stride, a.pnext + po, r.ptr, (int)r.m, false,
                       &tf, &tr);
            const bool good = (tf == 0) || (tr == 0);
            atomicAdd(&a.counts[entry_count_base(a.count_stride, it.x) + (good ? 1 : 0)],
                      a.rshift ? 1u << a.rshift[it.y] : 1u);
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(32)
exact_worklist_long_kernel(const ScoreArgs a) {
    extern __shared__ int scratch[];
    const uint32_t cnt = min(a.work_count[0], a.work_cap);
    for (uint32_t idx = blockIdx.x; idx < cnt; idx += gridDim.x) {
        const uint2 w = a.work[idx];
        const RowRef r = locate_row(a.bk, w.y);
        const int m = (int)r.m;
        if (m <= kExactRegSteps) continue;
        const uint32_t po = a.poff[entry_path(w.x)];
        const int n = (int)(a.poff[entry_path(w.x) + 1] - po);
        int tf = warp_exact_T(a.pstep + po, n, r.ptr, m, false, scratch);
        __syncwarp();
        int tr = -1;
        if (tf != 0) tr = warp_exact_T(a.pstep + po, n, r.ptr, m, true, scratch);
        __syncwarp();
        if ((threadIdx.x & 31) == 0) {
            const bool good = (tf == 0) || (tr == 0);
            atomicAdd(&a.counts[entry_count_base(a.count_stride, w.x) + (good ? 1 : 0)], 1u);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// K2: evalPath — T_fw / T_rc of ONE path against every alignment (reference src/eval.cpp:92-93).
// Pass 1 (pair_classify) streams the table: an alignment none of whose nodes lies on the path has
// an all-mismatch DP matrix, for which the walk is m diagonal moves: T = -m (m <= n); m > n gives
// T = 0; m == 0 gives T = 0.  Everything else is queued and scored exactly by pass 2 (pair_exact),
// dense, one thread per alignment with the path in shared memory.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock)
pair_classify_kernel(const PairArgs a) {
    extern __shared__ __align__(16) unsigned char smem[];
    uint16_t *s_tab = reinterpret_cast<uint16_t *>(smem);
    __shared__ uint32_t s_cnt[2];
    for (uint32_t i = threadIdx.x; i < a.tab_entries; i += blockDim.x) s_tab[i] = a.ptab[i];
    if (threadIdx.x < 2) s_cnt[threadIdx.x] = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int n = (int)a.n;
    uint32_t my_good = 0, my_bad = 0;
    // rows without steps: T = 0 both ways (the walk only consumes the path, free)
    for (uint32_t row = blockIdx.x * blockDim.x + threadIdx.x; row < a.bk.count[0];
         row += gridDim.x * blockDim.x) {
        const uint32_t ai = a.bk.perm[a.bk.row_start[0] + row];
        if (a.t_fw) a.t_fw[ai] = 0;
        if (a.t_rc) a.t_rc[ai] = 0;
        ++my_good;
    }
    const uint32_t n_tiles = a.bk.tile_start[kNumBuckets];
    for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int b = tile_bucket(a.bk, tile);
        const RowRef r = tile_row(a.bk, b, tile);
        if (!r.valid) continue;
        const int m = (int)r.m;
        bool touch = false, pass = true;
        for (int j = 0; j < m; ++j) {
            const uint32_t e = s_tab[__ldg(r.ptr + j) >> 1];
            if (e == kTabAbsent) pass = false; else touch = true;
        }
        int t = 0;
        bool queued = false;
        if (a.filter && !pass) {
            t = INT32_MIN;
        } else if (m > n) {
            t = 0; ++my_good;
        } else if (!touch) {
            t = -m; ++my_bad;
        } else {
            queued = true;
            const uint32_t idx = atomicAdd(a.work_count, 1u);
            if (idx < a.work_cap) a.work[idx] = r.sorted;
            else atomicOr(a.work_count + 1, 1u);
        }
        if (!queued && (a.t_fw || a.t_rc)) {
            const uint32_t ai = __ldg(a.bk.perm + r.sorted);
            if (a.t_fw) a.t_fw[ai] = t;
            if (a.t_rc) a.t_rc[ai] = t;
        }
    }
    // warp-reduce the private counters, one shared atomic per warp, one global per CTA
    for (int o = 16; o; o >>= 1) {
        my_good += __shfl_down_sync(kFull, my_good, o);
        my_bad += __shfl_down_sync(kFull, my_bad, o);
    }
    if (lane == 0) {
        if (my_bad) atomicAdd(&s_cnt[0], my_bad);
        if (my_good) atomicAdd(&s_cnt[1], my_good);
    }
    __syncthreads();
    if (threadIdx.x < 2 && s_cnt[threadIdx.x]) atomicAdd(&a.counts[threadIdx.x], s_cnt[threadIdx.x]);
}

__global__ void __launch_bounds__(128)
pair_exact_kernel(const PairArgs a) {
    extern __shared__ __align__(16) unsigned char smem[];
    uint16_t *s_path = reinterpret_cast<uint16_t *>(smem);
    for (uint32_t i = threadIdx.x; i < a.n; i += blockDim.x) s_path[i] = a.pstep[i];
    __syncthreads();
    const uint32_t cnt = min(a.work_count[0], a.work_cap);
    for (uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x; idx < cnt;
         idx += gridDim.x * blockDim.x) {
        const RowRef r = locate_row(a.bk, a.work[idx]);
        const int m = (int)r.m;
        if (m > kExactRegSteps) continue;
        int tf, tr;
        exact_both(s_path, (int)a.n, a.ptab, a.pnext, r.ptr, m, true, &tf, &tr);
        const uint32_t ai = a.bk.perm[r.sorted];
        if (a.t_fw) a.t_fw[ai] = tf;
        if (a.t_rc) a.t_rc[ai] = tr;
        const int best = tf > tr ? tf : tr;
        atomicAdd(&a.counts[best < 0 ? 0 : 1], 1u);
    }
}

__global__ void __launch_bounds__(32)
pair_exact_long_kernel(const PairArgs a) {
    extern __shared__ int scratch[];
    const uint32_t cnt = min(a.work_count[0], a.work_cap);
    for (uint32_t idx = blockIdx.x; idx < cnt; idx += gridDim.x) {
        const RowRef r = locate_row(a.bk, a.work[idx]);
        const int m = (int)r.m;
        if (m <= kExactRegSteps) continue;
        int tf = warp_exact_T(a.pstep, (int)a.n, r.ptr, m, false, scratch);
        __syncwarp();
        int tr = warp_exact_T(a.pstep, (int)a.n, r.ptr, m, true, scratch);
        __syncwarp();
        if ((threadIdx.x & 31) == 0) {
            const uint32_t ai = a.bk.perm[r.sorted];
            if (a.t_fw) a.t_fw[ai] = tf;
            if (a.t_rc) a.t_rc[ai] = tr;
            const int best = tf > tr ? tf : tr;
            atomicAdd(&a.counts[best < 0 ? 0 : 1], 1u);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Inverted index of the bucketed view (see InvIndex, kernels.cuh).  One thread per bucketed row.
// ------------------------------------------------------------------------------------------------
template <bool kFill>
__global__ void __launch_bounds__(kBlock)
index_rows_kernel(const Buckets bk, const uint32_t L, const uint32_t root, uint32_t *__restrict__ cnt,
                  uint32_t *__restrict__ rows) {
    const uint32_t lo = bk.row_start[1], hi = bk.row_start[kLongBucket];
    for (uint32_t s = lo + blockIdx.x * blockDim.x + threadIdx.x; s < hi; s += gridDim.x * blockDim.x) {
        int b = 0;
#pragma unroll
        for (int k = 1; k < kNumBuckets; ++k) b += (s >= bk.row_start[k]) ? 1 : 0;
        const uint32_t r = s - bk.row_start[b];
        const uint16_t *p = bk.bstep + bk.step_base[b] + (size_t)r * row_stride(b);
        uint32_t node[kMaxBucket];
        bool outside = false, in_f = false;
#pragma unroll
        for (int j = 0; j < kMaxBucket; ++j) {
            node[j] = j < b ? (uint32_t)(__ldg(p + j) >> 1) : 0xFFFFFFFFu;
            if (j < b) {
                outside |= node[j] >= L;
                in_f |= node[j] == root;
            }
        }
        if (outside) continue;          // filtered out for every path
        if (in_f) {
            const uint32_t g = L * kIdxLens + (uint32_t)b;
            const uint32_t slot = atomicAdd(&cnt[g], 1u);
            if (kFill) rows[slot] = r;
            continue;
        }
#pragma unroll
        for (int j = 0; j < kMaxBucket; ++j) {
            if (j >= b) break;
            bool seen = false;
#pragma unroll
            for (int i = 0; i < kMaxBucket; ++i)
                if (i < j) seen |= node[i] == node[j];
            if (seen) continue;
            const uint32_t g = node[j] * kIdxLens + (uint32_t)b;
            const uint32_t slot = atomicAdd(&cnt[g], 1u);
            if (kFill) rows[slot] = r;
        }
    }
}

// exclusive scan of `groups` counters into off[0..groups] (one CTA; a one-off per universe) + cursor copy
__global__ void __launch_bounds__(1024)
index_scan_kernel(const uint32_t *cnt, uint32_t groups, uint32_t *__restrict__ off,
                  uint32_t *cursor) {   // cnt and cursor may be the same buffer (ensure_index): no __restrict__
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (uint32_t base = 0; base < groups; base += 1024) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = i < groups ? cnt[i] : 0u;
        uint32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(kFull, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) s_warp[wid] = incl;
        __syncthreads();
        if (wid == 0) {
            uint32_t w = s_warp[lane], wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(kFull, wi, o);
                if (lane >= o) wi += t;
            }
            s_warp[lane] = wi - w;
        }
        __syncthreads();
        const uint32_t excl = s_carry + s_warp[wid] + incl - v;
        if (i < groups) {
            off[i] = excl;
            cursor[i] = excl;
        }
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) off[groups] = s_carry;
}

// One warp per candidate path: base path (raw global ids) + suffix -> universe-local 16-bit steps.
__global__ void __launch_bounds__(256) assemble_paths_kernel(const AssembleArgs a) {
    const uint32_t p = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (p >= a.P) return;
    const uint4 d = a.desc[p];
    if (d.x == 0xFFFFFFFFu) return;                    // padding slot of a chunk: an empty path
    const uint32_t b0 = a.base_off[d.x], nb = a.base_off[d.x + 1] - b0;
    const uint32_t total = nb + d.z - d.w;             // d.w steps dropped at the end (the parent of a child)
    uint16_t *out = a.out + a.poff[p];
    for (uint32_t i = lane; i < total; i += 32) {
        const uint32_t s = i < nb ? a.base_steps[b0 + i] : a.suffix_steps[d.y + (i - nb)];
        const uint32_t node = s >> 1;
        uint32_t l = node < a.n_nodes ? a.map[node] : a.L;
        if (l >= a.L) { atomicOr(a.err, 1u); l = a.L; }
        out[i] = (uint16_t)((l << 1) | (s & 1u));
    }
}

__global__ void __launch_bounds__(256)
compose_items_kernel(const uint32_t *__restrict__ slot_item, const uint32_t *__restrict__ slot_aux, uint32_t P_dev,
                     const uint32_t *__restrict__ inc, const uint32_t *__restrict__ full, int32_t *__restrict__ out) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= P_dev) return;
    const uint32_t si = slot_item[k];
    if (si == 0xFFFFFFFFu) return;
    const uint32_t kind = si >> 30;
    const uint32_t *r = inc + 4 * (size_t)k;
    int32_t *o = out + 4 * (size_t)(si & kPathMask);
    if (kind == 0u) {            // the candidate itself: + its counts on the rows of its last node, its fresh set
        atomicAdd(o, (int32_t)r[0]);
        atomicAdd(o + 1, (int32_t)r[1]);
        o[2] = (int32_t)r[2];
        o[3] = (int32_t)r[3];
    } else if (kind == 1u) {     // its parent on the same rows (the last node was visited before)
        atomicAdd(o, -(int32_t)r[0]);
        atomicAdd(o + 1, -(int32_t)r[1]);
    } else {                     // full scan
        const uint32_t *f = full + 2 * (size_t)slot_aux[k];
        o[0] = (int32_t)f[0];
        o[1] = (int32_t)f[1];
        o[2] = (int32_t)r[2];
        o[3] = (int32_t)r[3];
    }
}

constexpr size_t kMaxDynSmem = 200 * 1024;

size_t score_smem_bytes(const ScoreArgs &a) {
    return (size_t)a.chunk * a.tab_stride * 2 + (size_t)a.chunk * 3 * sizeof(uint32_t);
}

}  // namespace

int configure_kernels() {
    cudaError_t e;
    e = cudaFuncSetAttribute(build_tables_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)kMaxDynSmem);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(score_paths_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)kMaxDynSmem);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(score_paths_kernel<false>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(pair_classify_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)kMaxDynSmem);
    if (e != cudaSuccess) return (int)e;
    const int rc = configure_bitset_kernels();
    if (rc != 0) return rc;
    return configure_edge_kernels();
}

void launch_length_histogram(const uint32_t *off, uint64_t A, uint32_t *hist, cudaStream_t st) {
    if (A == 0) return;
    uint64_t blocks = (A + kBlock - 1) / kBlock;
    if (blocks > 148 * 8) blocks = 148 * 8;
    length_histogram_kernel<<<(unsigned)blocks, kBlock, 0, st>>>(off, A, hist);
}

void launch_pair_collect(const uint32_t *off, const uint32_t *steps, const uint32_t *node2local,
                         uint64_t A, uint32_t L, uint32_t max_pairs, const PairDict &pd,
                         cudaStream_t st) {
    if (A == 0) return;
    uint64_t blocks = (A + kBlock - 1) / kBlock;
    if (blocks > 148 * 8) blocks = 148 * 8;
    pair_collect_kernel<<<(unsigned)blocks, kBlock, 0, st>>>(off, steps, node2local, A, L, max_pairs, pd);
}

void launch_pair_assign(const PairDict &pd, cudaStream_t st) {
    pair_assign_kernel<<<kPairHashCap / 256, 256, 0, st>>>(pd);
}

void launch_bucket_scatter(const uint32_t *off, const uint32_t *steps, const uint32_t *node2local,
                           uint64_t A, const Buckets &bk, uint16_t *bstep, uint2 *ldesc,
                           uint32_t *perm, uint32_t *cursors, const PairDict &pd, uint32_t n_pairs,
                           uint32_t pair_bias, cudaStream_t st, bool only_long) {
    if (A == 0) return;
    uint64_t blocks = (A + kBlock - 1) / kBlock;
    if (blocks > 148 * 8) blocks = 148 * 8;
    bucket_scatter_kernel<<<(unsigned)blocks, kBlock, 0, st>>>(off, steps, node2local, A, bk, bstep,
                                                               ldesc, perm, cursors, pd, n_pairs, pair_bias, only_long);
}

void launch_dedup_insert(const uint32_t *off, const uint32_t *steps, const uint32_t *node2local, uint64_t A,
                         DedupSlot *slots, uint32_t cap, cudaStream_t st) {
    if (A == 0) return;
    uint64_t blocks = (A + kBlock - 1) / kBlock;
    if (blocks > 148 * 16) blocks = 148 * 16;
    dedup_insert_kernel<<<(unsigned)blocks, kBlock, 0, st>>>(off, steps, node2local, A, slots, cap - 1);
}

void launch_dedup_count(const DedupSlot *slots, uint32_t cap, const uint32_t *off, uint32_t *gcount, cudaStream_t st) {
    dedup_count_kernel<<<std::min<uint32_t>((cap + kBlock - 1) / kBlock, 148 * 8), kBlock, 0, st>>>(slots, cap, off, gcount);
}

void launch_dedup_scatter(const DedupSlot *slots, uint32_t cap, const uint32_t *off, const uint32_t *steps,
                          const uint32_t *node2local, const Buckets &bk, uint16_t *bstep, const uint32_t *class_base,
                          uint32_t *cursors, uint8_t *rshift, const PairDict &pd, uint32_t n_pairs, uint32_t pair_bias,
                          cudaStream_t st) {
    dedup_scatter_kernel<<<std::min<uint32_t>((cap + kBlock - 1) / kBlock, 148 * 16), kBlock, 0, st>>>(
        slots, cap, off, steps, node2local, bk, bstep, class_base, cursors, rshift, pd, n_pairs, pair_bias);
}

__global__ void status_flag_kernel(uint32_t *err, const uint32_t *wc, uint32_t *out) {
    *out = ((err && (err[0] & 1u)) ? 1u : 0u) | (((wc && (wc[1] || wc[3])) || (err && err[1])) ? 2u : 0u) |
           (err ? (err[0] & 4u) : 0u);
    if (err) err[0] = err[1] = 0;   // reported: a later batch starts clean
}

// The word is SUMMED over up to GFE_PEER_MAX (16) ranks, so each condition counts ranks in a byte of its
// own and the sum still tells them apart: byte 0 = ranks with a list overflow (all ranks repeat the batch),
// byte 1 = ranks that saw a step outside the universe, byte 2 = ranks on whose shard an over-long alignment
// passed the id filter (all ranks return the same error instead of retrying).
__global__ void flag_word_kernel(const uint32_t *err, const uint32_t *wc, uint32_t *out) {
    const uint32_t e = *err;
    *out = ((wc[1] || wc[3]) ? 1u : 0u) | ((e & ~4u) ? 1u << 8 : 0u) | ((e & 4u) ? 1u << 16 : 0u);
}

__global__ void flag_or_kernel(const uint32_t *sticky, uint32_t *out) {
    if (*sticky) *out |= 1u;
}

void launch_flag_or(const uint32_t *sticky, uint32_t *out, cudaStream_t st) {
    flag_or_kernel<<<1, 1, 0, st>>>(sticky, out);
}

void launch_flag_word(const uint32_t *err, const uint32_t *wc, uint32_t *out, cudaStream_t st) {
    flag_word_kernel<<<1, 1, 0, st>>>(err, wc, out);
}

void launch_status_flag(uint32_t *err, const uint32_t *wc, uint32_t *out, cudaStream_t st) {
    status_flag_kernel<<<1, 1, 0, st>>>(err, wc, out);
}

void launch_prologue(const PrologueArgs &a, int sm_count, cudaStream_t st) {
    uint64_t most = a.psteps ? a.total : 0;
    for (int r = 0; r < 4; ++r) most = std::max<uint64_t>(most, a.zero_words[r] / 4 + 1);
    const uint64_t blocks = std::min<uint64_t>((most + 255) / 256, (uint64_t)sm_count * 8);
    prologue_kernel<<<(unsigned)std::max<uint64_t>(blocks, 1), 256, 0, st>>>(a);
}

void launch_build_tables(const uint32_t *poff, const uint16_t *pstep, uint32_t P, uint32_t L,
                         uint32_t tab_stride, uint16_t *ptab, uint16_t *pnext, uint32_t *masks,
                         uint32_t *fplane, const MaskGeom &mg, const PairDict &pd, cudaStream_t st) {
    if (P == 0) return;
    // plain launch (full stream order behind the prologue); its trigger lets the scoring kernel in early
    build_tables_kernel<<<P, kBlock, (size_t)(L + 1) * sizeof(uint32_t), st>>>(
        poff, pstep, L, tab_stride, ptab, pnext, masks, fplane, mg, pd);
}

void launch_score_paths(const ScoreArgs &a, int sm_count, uint32_t tile_lo, uint32_t tile_hi,
                        bool add_empty, cudaStream_t st) {
    if (a.P == 0) return;
    const size_t smem = score_smem_bytes(a);
    const uint32_t chunks = (a.P + a.chunk - 1) / a.chunk;
    int per_sm = 1;
    if (a.filter)
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, score_paths_kernel<true>, kBlock, smem);
    else
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, score_paths_kernel<false>, kBlock, smem);
    if (per_sm < 1) per_sm = 1;
    const uint64_t n_tiles = tile_hi - tile_lo;
    // resident CTAs = a multiple of the SM count; split between tile workers (x) and chunks (y)
    uint64_t gx = ((uint64_t)sm_count * per_sm + chunks - 1) / chunks;
    if (gx > n_tiles) gx = n_tiles;
    if (gx < 1) gx = 1;
    dim3 grid((unsigned)gx, chunks);
    if (a.filter) score_paths_kernel<true><<<grid, kBlock, smem, st>>>(a, tile_lo, tile_hi, add_empty);
    else score_paths_kernel<false><<<grid, kBlock, smem, st>>>(a, tile_lo, tile_hi, add_empty);
}

void launch_slow_worklist(const ScoreArgs &a, const uint2 *slow, const uint32_t *slow_count,
                          uint32_t slow_cap, cudaStream_t st) {
    launch_pdl(2, slow_worklist_kernel, dim3(148 * 16), dim3(128), 0, st, a, slow, slow_count, slow_cap);
}

void launch_exact_worklist(const ScoreArgs &a, cudaStream_t st) {
    launch_pdl(3, exact_worklist_kernel, dim3(148 * 4), dim3(128), 0, st, a);
}

void launch_exact_worklist_long(const ScoreArgs &a, uint32_t max_m, cudaStream_t st) {
    exact_worklist_long_kernel<<<148 * 8, 32, (size_t)(max_m + 1) * 8 * sizeof(int), st>>>(a);
}

void launch_pair_classify(const PairArgs &a, int sm_count, cudaStream_t st) {
    const uint64_t n_tiles = a.bk.tile_start[kNumBuckets];
    uint64_t gx = (uint64_t)sm_count * 8;
    if (gx > n_tiles) gx = n_tiles;
    if (gx < 1) gx = 1;
    pair_classify_kernel<<<(unsigned)gx, kBlock, (size_t)a.tab_entries * 2, st>>>(a);
}

void launch_pair_exact(const PairArgs &a, cudaStream_t st) {
    pair_exact_kernel<<<148 * 4, 128, ((size_t)a.n * 2 + 15) & ~(size_t)15, st>>>(a);
}

void launch_index_count(const Buckets &bk, uint32_t L, uint32_t root_local, uint32_t *cnt, cudaStream_t st) {
    const uint32_t n = bk.row_start[kLongBucket] - bk.row_start[1];
    if (!n) return;
    index_rows_kernel<false><<<std::min<uint32_t>((n + kBlock - 1) / kBlock, 148 * 16), kBlock, 0, st>>>(
        bk, L, root_local, cnt, nullptr);
}

void launch_index_scan(const uint32_t *cnt, uint32_t groups, uint32_t *off, uint32_t *cursor, cudaStream_t st) {
    index_scan_kernel<<<1, 1024, 0, st>>>(cnt, groups, off, cursor);
}

void launch_index_fill(const Buckets &bk, uint32_t L, uint32_t root_local, uint32_t *cursor, uint32_t *rows,
                       cudaStream_t st) {
    const uint32_t n = bk.row_start[kLongBucket] - bk.row_start[1];
    if (!n) return;
    index_rows_kernel<true><<<std::min<uint32_t>((n + kBlock - 1) / kBlock, 148 * 16), kBlock, 0, st>>>(
        bk, L, root_local, cursor, rows);
}

void launch_compose_items(const uint32_t *slot_item, const uint32_t *slot_aux, uint32_t P_dev, const uint32_t *inc,
                          const uint32_t *full, int32_t *out, cudaStream_t st) {
    if (!P_dev) return;
    compose_items_kernel<<<(P_dev + 255) / 256, 256, 0, st>>>(slot_item, slot_aux, P_dev, inc, full, out);
}

void launch_assemble_paths(const AssembleArgs &a, cudaStream_t st) {
    if (!a.P) return;
    assemble_paths_kernel<<<(a.P * 32 + 255) / 256, 256, 0, st>>>(a);
}

void launch_pair_exact_long(const PairArgs &a, uint32_t max_m, cudaStream_t st) {
    pair_exact_long_kernel<<<148 * 8, 32, (size_t)(max_m + 1) * 8 * sizeof(int), st>>>(a);
}


}  // namespace gfe
