// gfalign_b200/csrc/bitset.cu — K1: score_bitset, the path x alignment scoring kernel of
// `gfalign search` (evaluatePath with the id filter, reference src/eval.cpp:67-108) in bit-parallel
// form: one machine word carries the verdict of 32 candidate paths.
//
// Per batch the host side (build_tables_kernel) turns the candidate paths into bitsets over the
// paths of a chunk (32*W paths, W words per entry), resident in shared memory for the whole kernel:
//   N[v]  paths that contain node v                      (the id filter, src/eval.cpp:81-91)
//   T[v]  paths that visit node v more than once         (NodeTable counts > 1)
//   F[v]  paths whose first node is v
//   LM[m] paths shorter than m steps                     (m > n: always good, SURVEY.md A.3 rule 1)
//   M[q]  paths that take the oriented pair q = (x -> y) consecutively; q^1 = its reverse complement
// An alignment B = b_0..b_{m-1} is stored as its m local steps followed by the ids of its m-1
// consecutive pairs.  For the 32 paths of one word:
//   pass = AND_j N[b_j]                      every node on the path
//   cand = AND_j M[q_j]  |  AND_j M[q_j^1]   every pair of B (or of rc(B)) is taken by the path
// If no node of B is visited twice by the path, every node has one position and cand is exactly
// "B or rc(B) occurs contiguously" (rule 2).  So, with spec = (cand & OR_j T[b_j]) | OR_j F[b_j]:
//   good = pass & (LM[m] | (cand & ~OR_j T[b_j]))
//   bad  = pass & ~LM[m] & ~cand... & ~spec          (no occurrence possible, no start overhang)
//   slow = pass & ~LM[m] & ~good & spec              -> (path, row) worklist: visit chains, overhang
//                                                       test and the exact DP (kernels.cu)
// A lane owns one word (32 paths) of one row; W lanes cover a row's chunk, 32/W rows per warp step.
// All lanes of a group read the same table entry at consecutive words: conflict-free LDS.
// Counts are kept bit-sliced in registers (kPlanes planes per lane for good and for bad; eight rows
// are folded with 3:2 carry-save adders before touching the planes), so the inner loop has no
// atomics and no cross-lane traffic.  The table is brought in with one TMA bulk copy.
#include "kernels.cuh"

namespace gfe {
namespace {

__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}

// ---- TMA 1-D bulk copy global -> shared, completion on an mbarrier ------------------------------
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(dst)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

// ---- bit-sliced counters -------------------------------------------------------------------------
// 3:2 carry-save adder on 32 independent bit columns: two LOP3s.
__device__ __forceinline__ void csa(uint32_t &sum, uint32_t &carry, uint32_t a, uint32_t b, uint32_t c) {
    const uint32_t u = a ^ b;
    carry = (a & b) | (u & c);
    sum = u ^ c;
}

struct Planes {
    uint32_t p[kPlanes];
    __device__ __forceinline__ void clear() {
#pragma unroll
        for (int k = 0; k < kPlanes; ++k) p[k] = 0;
    }
    // add four / two masks (3:2 compressors first, then ripple the carry of weight 4 / 2)
    __device__ __forceinline__ void add(const uint32_t (&x)[4]) {
        uint32_t c0, c1, d0;
        csa(p[0], c0, p[0], x[0], x[1]);
        csa(p[0], c1, p[0], x[2], x[3]);
        csa(p[1], d0, p[1], c0, c1);
#pragma unroll
        for (int k = 2; k < kPlanes; ++k) {
            const uint32_t t = p[k] & d0;
            p[k] ^= d0;
            d0 = t;
        }
    }
    __device__ __forceinline__ void add(const uint32_t (&x)[2]) {
        uint32_t c0;
        csa(p[0], c0, p[0], x[0], x[1]);
#pragma unroll
        for (int k = 1; k < kPlanes; ++k) {
            const uint32_t t = p[k] & c0;
            p[k] ^= c0;
            c0 = t;
        }
    }
    // columns c, c+8, c+16, c+24 of the counter at once: planes 0-7 land in the four bytes of `lo`,
    // the planes above in the bytes of `hi` (three instructions per plane for four columns)
    __device__ __forceinline__ void value4(int c, uint32_t &lo, uint32_t &hi) const {
        static_assert(kPlanes > 8 && kPlanes <= 16, "two bytes per column");
        lo = hi = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) lo += ((p[k] >> c) & 0x01010101u) << k;
#pragma unroll
        for (int k = 8; k < kPlanes; ++k) hi += ((p[k] >> c) & 0x01010101u) << (k - 8);
    }
};

// Plain view (one group per row length): bucket of a unit by one vote — lane k compares with the
// first unit of bucket k (s_ustart: shared copy of unit_start, 0xFFFFFFFF beyond kMaxBucket).
__device__ __forceinline__ int unit_bucket(const uint32_t *s_ustart, uint32_t unit, int lane) {
    return __popc(__ballot_sync(0xFFFFFFFFu, unit >= s_ustart[lane])) - 1;   // entries 0 and 1 are 0
}

// Class-sorted view: scheduling groups live in shared memory as uint4 {unit_start, row0, nrows, m | shift << 16}.  A warp
// claims units in DEcreasing order, so the group of its next unit is found by walking down from the
// group of the current one (usually zero or one step).  Warp-uniform.
__device__ __forceinline__ uint32_t group_of(const uint4 *s_groups, uint32_t gi, uint32_t unit) {
    while (gi > 0 && s_groups[gi].x > unit) --gi;
    return gi;
}

// the rows of a unit into L1 ahead of time
__device__ __forceinline__ void prefetch_unit(const BitsetArgs &a, uint32_t unit, const uint4 g, int lane) {
    const uint32_t m = g.w & 0xFFFFu;
    const uint32_t row0 = g.y + (unit - g.x) * a.unit_rows;
    const uint32_t nrows = min(a.unit_rows, g.y + g.z - row0);
    const uint32_t rs = row_stride((int)m);
    const char *p = reinterpret_cast<const char *>(a.bk.bstep + a.bk.step_base[m] + (size_t)row0 * rs);
    const uint32_t bytes = nrows * rs * 2u;
    for (uint32_t o = (uint32_t)lane * 128u; o < bytes + 127u; o += 32u * 128u)
        asm volatile("prefetch.global.L1 [%0];" ::"l"(p + min(o, bytes - 1u)));
}

// Push the set bits of the eight masks of one warp step (paths word_base .. word_base+31, rows
// row_first + u*row_step) as (path, row) work items.  Called by the whole warp.  The warp owns a
// window [win.x, win.y) of the list, reserved kSlowChunk slots at a time with ONE atomic on the
// list counter -- the round trip to L2 is paid once per chunk, not once per call; what a warp
// leaves unused of its last window is filled with kSlowHole entries the second-look kernel skips.
constexpr uint32_t kSlowChunk = 64;

__device__ __forceinline__ void fill_slow_holes(uint2 *__restrict__ slow, uint32_t *work_count,
                                                uint32_t slow_cap, uint2 win, int lane) {
    if (lane == 0 && win.y > win.x) atomicAdd(work_count + 4, win.y - win.x);   // holes; no return value used
    for (uint32_t i = win.x + (uint32_t)lane; i < win.y; i += 32u)
        if (i < slow_cap) slow[i] = make_uint2(kSlowHole, kSlowHole);
}

__device__ __noinline__ uint2 push_slow8(uint2 *__restrict__ slow, uint32_t *work_count,
                                         uint32_t slow_cap, uint2 win, uint32_t s0, uint32_t s1, uint32_t s2,
                                         uint32_t s3, uint32_t s4, uint32_t s5, uint32_t s6,
                                         uint32_t s7, uint32_t word_base, uint32_t r0, uint32_t r1,
                                         uint32_t r2, uint32_t r3, int lane) {
    // word_base: first path of this lane's word, OR-ed with the accumulator-set tag (bit 30);
    // r0..r3: sorted row ids of this lane's four rows
    const uint32_t sm[8] = {s0, s1, s2, s3, s4, s5, s6, s7};
    const uint32_t rid[4] = {r0, r1, r2, r3};
    uint32_t n = 0;
#pragma unroll
    for (int u = 0; u < 8; ++u) n += (uint32_t)__popc(sm[u]);
    uint32_t incl = n;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= o) incl += t;
    }
    const uint32_t total = __shfl_sync(0xFFFFFFFFu, incl, 31);
    if (win.x + total > win.y) {      // warp-uniform: the window is used up
        fill_slow_holes(slow, work_count, slow_cap, win, lane);
        const uint32_t need = total > kSlowChunk ? total : kSlowChunk;
        uint32_t base = 0;
        if (lane == 0) {
            base = atomicAdd(work_count + 2, need);
        }
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        win.x = base;
        win.y = base + need;
    }
    uint32_t idx = win.x + incl - n;
#pragma unroll
    for (int u = 0; u < 8; ++u) {
        uint32_t mask = sm[u];
        while (mask) {
            const uint32_t c = (uint32_t)__ffs(mask) - 1u;
            mask &= mask - 1u;
            // masks 0-3: rows counted good for now, masks 4-7: the same rows counted bad for now
            if (idx < slow_cap)
                slow[idx] = make_uint2((word_base + c) | (u < 4 ? 0x80000000u : 0u), rid[u & 3]);
            else atomicOr(work_count + 3, 1u);
            ++idx;
        }
    }
    win.x += total;
    return win;
}

// Second, exact look at one row (runtime length m) for this lane's 32 paths; the warp calls it
// when the fast test below leaves any (path, row) open.  Two refinements over the fast test:
//  * a candidate through a twice-visited interior node is certified when none of B's triples is a
//    broken triple of the path (BT, hashed; marked by build_tables_kernel);
//  * a start overhang (SURVEY.md A.3 class iv) needs b_j == path[0] (oriented) and, unless b_j is
//    B's last step, the pair (b_j, b_j+1) taken by the path; mirrored for rc(B).
// What is still open after this goes to the (path, row) list of kernels.cu.
template <int W>
__device__ __noinline__ uint4 refine_row(const uint2 *__restrict__ nt, const uint32_t *__restrict__ tab,
                                         const uint16_t *__restrict__ rp, int m, uint32_t pair_base,
                                         uint32_t bt_base, uint32_t lm, int filter,
                                         const uint2 *__restrict__ nf) {   // nf: global {F+, F-} plane at this lane's word
    const uint32_t *mt = tab;                       // the rows hold absolute entries (pair_base + id)
    const uint32_t *bt = tab + (size_t)bt_base * W;
    uint32_t pass = 0xFFFFFFFFu, fw = 0xFFFFFFFFu, rc = 0xFFFFFFFFu, bfw = 0, brc = 0, over = 0;
    uint32_t prev_pid = 0, prev_mr = 0;
    for (int j = 0; j < m; ++j) {
        const uint32_t v = __ldg(rp + j);
        const uint2 e = nt[(v >> 1) * W];           // {N, F+|F-}
        const uint2 f = __ldg(nf + (v >> 1) * W);   // {F+, F-}
        const uint32_t ff = (v & 1u) ? f.y : f.x;   // path[0] == b_j
        const uint32_t fr = (v & 1u) ? f.x : f.y;   // path[0] == rc of b_j
        pass &= e.x;
        uint32_t mf = 0xFFFFFFFFu, mr = 0xFFFFFFFFu, pid = 0;
        if (j < m - 1) {
            pid = __ldg(rp + m + j);
            mf = mt[pid * W];
            mr = mt[(pid ^ 1u) * W];
            fw &= mf;
            rc &= mr;
        }
        if (j == m - 1) over |= ff;                 // k = 1
        else if (j >= 1) over |= ff & mf;           // k = m - j >= 2: first pair of the path
        if (j == 0) over |= fr;                     // rc, k = 1
        else if (j <= m - 2) over |= fr & prev_mr;  // rc, pair (b_j-1, b_j) reversed
        if (j >= 1 && j <= m - 2) {
            const uint32_t slot = triple_slot(prev_pid - pair_base, pid - pair_base);
            bfw |= bt[slot * W];
            brc |= bt[(slot ^ 1u) * W];
        }
        prev_pid = pid;
        prev_mr = mr;
    }
    const uint32_t in = filter ? pass : 0xFFFFFFFFu;
    const uint32_t sure = (fw & ~bfw) | (rc & ~brc);
    const uint32_t amb = (fw & bfw) | (rc & brc);
    const uint32_t open = in & ~lm & ~sure;
    // good / bad include the optimistic guesses for what goes to the second-look list:
    // an ambiguous candidate is most likely an occurrence, an overhang suspect most likely not
    return make_uint4((in & (lm | sure)) | (open & amb), open & ~amb, open & amb, open & ~amb & over);
}

// One row of length M (its u16 elements packed two per register in q) against this lane's word.
// nt / tab = shared-memory table at this lane's word (first node plane / plain entries).
template <int M, int W>
__device__ __forceinline__ void eval_row(const uint2 *__restrict__ nt, const unsigned char *__restrict__ mt,
                                         const unsigned char *__restrict__ bt, uint32_t lane_off,
                                         const uint32_t (&q)[row_stride(M) / 2], uint32_t lm,
                                         uint32_t vmask, bool filter, uint32_t &good, uint32_t &bad, uint32_t &slow,
                                         uint32_t &over_open) {
    auto elem = [&](int i) -> uint32_t { return (i & 1) ? (q[i >> 1] >> 16) : (q[i >> 1] & 0xFFFFu); };
    // node block: one LDS.64 per step -> {N, F+|F-} of this lane's 32 paths ({F+, F-} live in a
    // second plane only refine_row reads: a 16-lane row group then moves 128 B = one wavefront per
    // step instead of the two of a 16-byte quad).  A node visited
    // twice can only break the pair test when it is an INTERIOR step of B (it then joins two pairs
    // that may sit at different visits); at either end of B it joins one pair, whose other end
    // pins the visit.
    uint32_t pass = 0xFFFFFFFFu, over = 0;
#pragma unroll
    for (int j = 0; j < M; ++j) {
        // byte offset like the pair entries below: shift, one and-or LOP3, LDS.64 [R + UR + imm]
        const uint32_t o = ((elem(j) >> 1) * (8u * W)) | (2u * lane_off);
        const uint2 e = *reinterpret_cast<const uint2 *>(mt + geom_node_base() * (4u * W) + o);   // {N, F+|F-}
        pass &= e.x;
        over |= e.y;
    }
    const uint32_t in = (filter ? pass : 0xFFFFFFFFu) & vmask;   // vmask: 0 for the rows past the unit's end
    if (M == 1) {
        // a single step occurs (forward or reverse-complemented) iff its node is on the path
        good = in & (lm | pass);
        bad = in & ~(lm | pass);
        slow = 0;
        over_open = 0;
        return;
    }
    uint32_t fw = 0xFFFFFFFFu, rc = 0xFFFFFFFFu;
#pragma unroll
    for (int j = 0; j < M - 1; ++j) {
        // byte offset of entry pid at this lane's word: pid * (4W) has its low bits clear, so the
        // lane offset is OR-ed in and the warp-uniform region base stays out of the address math
        const uint32_t o = (elem(M + j) * (4u * W)) | lane_off;
        fw &= *reinterpret_cast<const uint32_t *>(mt + o);
        rc &= *reinterpret_cast<const uint32_t *>(mt + (o ^ (4u * W)));
    }
    // broken triples of the path among B's triples (one slot byte per interior step)
    uint32_t bfw = 0, brc = 0;
#pragma unroll
    for (int j = 0; j < M - 2; ++j) {
        constexpr int kByte0 = 2 * (2 * M - 1);
        const uint32_t slot = (q[(kByte0 + j) >> 2] >> (8 * ((kByte0 + j) & 3))) & 0xFFu;
        const uint32_t o = (slot * (4u * W)) | lane_off;
        bfw |= *reinterpret_cast<const uint32_t *>(bt + o);
        brc |= *reinterpret_cast<const uint32_t *>(bt + (o ^ (4u * W)));
    }
    const uint32_t sure = (fw & ~bfw) | (rc & ~brc);
    const uint32_t amb = (fw & bfw) | (rc & brc);
    const uint32_t open = in & ~lm & ~sure;
    good = in & (lm | sure);
    good |= open & amb;               // optimistic: counted good now, corrected by the second look
    slow = open & amb;                // a broken (or hash-colliding) triple: second-look kernel
    over_open = open & ~amb & over;   // start-overhang suspects: refine_row decides
    bad = open & ~amb & ~over;
}

// What a scheduling unit needs besides its rows (both scoring kernels fill one per warp).
struct UnitCtx {
    const uint2 *nt;          // node plane at this lane's word (shared memory)
    const uint2 *nf;          // {F+, F-} plane at this lane's word (global memory)
    const uint32_t *s_tab;    // table base (shared memory)
    const uint32_t *tab;      // table at this lane's word
    uint32_t pair_base, bt_base;
    int filter;
    uint2 *slow;
    uint32_t *work_count;
    uint32_t slow_cap;
    uint32_t word_base;       // first path of this lane's word | accumulator-set tag (bit 30)
};

// One scheduling unit (rows of one length M).  Four (two for M > 8) rows per lane and loop trip:
// their data is loaded first (independent LDG.64s in flight together), then scored, then ONE warp
// vote decides whether any of them needs the second look at the start overhang.  The loop is kept
// rolled and short so that the body stays in the instruction cache.
// kGather = false: the unit is rows [0, nrows) of the dense matrix at `base`, sorted ids sorted0 + r.
// kGather = true:  the unit is the rows idx[0..nrows) of the bucket matrix at `base` (inverted-index
//                  list), sorted ids sorted0 + idx[r].
template <int M, int W, bool kGather>
__device__ __forceinline__ void score_unit(const UnitCtx &c, const uint16_t *__restrict__ base,
                                           const uint32_t *__restrict__ idx, uint32_t nrows, uint32_t sorted0,
                                           uint32_t lm, Planes &G, Planes &B, uint2 &slow_win, int lane) {
    constexpr int R = 32 / W;                       // rows per warp step
    constexpr int RS = (int)row_stride(M);
    constexpr int UB = M <= 8 ? 4 : 2;   // rows scored per loop trip (their data is loaded first)
    // warp-uniform region bases (the lane's word offset is OR-ed into each entry offset)
    const uint32_t lane_off = 4u * ((uint32_t)lane % W);
    // the rows hold absolute pair entries and BT sits at entry 0: both offsets are relative to the table
    const unsigned char *mt = reinterpret_cast<const unsigned char *>(c.s_tab);
    const unsigned char *bt = mt;
    const uint32_t sub = (uint32_t)lane / W;
    const bool filter = c.filter != 0;
    // rows of trip `it0` into registers (independent LDG.64s in flight together)
    auto load_rows = [&](uint32_t it0, uint32_t (&qq)[UB][RS / 2], uint32_t (&ri)[UB]) {
#pragma unroll
        for (int k = 0; k < UB; ++k) {
            const uint32_t r = it0 + k * R + sub;
            uint32_t row = r < nrows ? r : 0u;
            if (kGather) row = __ldg(idx + row);
            ri[k] = row;
        }
#pragma unroll
        for (int k = 0; k < UB; ++k) {
            const uint2 *src = reinterpret_cast<const uint2 *>(base + (size_t)ri[k] * RS);
#pragma unroll
            for (int cc = 0; cc < RS / 4; ++cc) {
                const uint2 t = __ldg(src + cc);
                qq[k][2 * cc] = t.x;
                qq[k][2 * cc + 1] = t.y;
            }
        }
    };
    // short rows (most of them): the next trip's rows are loaded before this trip is scored, so
    // the global-load latency hides behind ~4 rows of table work
    constexpr bool kPrefetch = M <= 5;
    uint32_t qn[kPrefetch ? UB : 1][RS / 2];
    uint32_t rn[kPrefetch ? UB : 1];
    if (kPrefetch)
        load_rows(0, reinterpret_cast<uint32_t (&)[UB][RS / 2]>(qn), reinterpret_cast<uint32_t (&)[UB]>(rn));
#pragma unroll 1
    for (uint32_t it = 0; it < nrows; it += UB * R) {
        uint32_t q[UB][RS / 2], ri[UB];
        bool valid[UB];
#pragma unroll
        for (int k = 0; k < UB; ++k) valid[k] = it + k * R + sub < nrows;
        if (kPrefetch) {
#pragma unroll
            for (int k = 0; k < UB; ++k) {
                ri[k] = rn[kPrefetch ? k : 0];
#pragma unroll
                for (int cc = 0; cc < RS / 2; ++cc) q[k][cc] = qn[kPrefetch ? k : 0][cc];
            }
            if (it + UB * R < nrows)
                load_rows(it + UB * R, reinterpret_cast<uint32_t (&)[UB][RS / 2]>(qn),
                          reinterpret_cast<uint32_t (&)[UB]>(rn));
        } else {
            load_rows(it, q, ri);
        }
        uint32_t gm[UB], bm[UB], sm[UB], sb[UB], ov[UB], any_over = 0, any_slow = 0;
#pragma unroll
        for (int k = 0; k < UB; ++k) {
            eval_row<M, W>(c.nt, mt, bt, lane_off, q[k], lm, valid[k] ? 0xFFFFFFFFu : 0u, filter, gm[k], bm[k],
                           sm[k], ov[k]);
            sb[k] = 0;
            any_over |= ov[k];
        }
        if (M >= 2 && __any_sync(0xFFFFFFFFu, any_over != 0)) {
#pragma unroll
            for (int k = 0; k < UB; ++k) {
                if (!__any_sync(0xFFFFFFFFu, ov[k] != 0)) continue;
                const uint4 x = refine_row<W>(c.nt, c.tab, base + (size_t)ri[k] * RS, M, c.pair_base, c.bt_base, lm,
                                              c.filter, c.nf);
                gm[k] = valid[k] ? x.x : 0u;
                bm[k] = valid[k] ? x.y : 0u;
                sm[k] = valid[k] ? x.z : 0u;
                sb[k] = valid[k] ? x.w : 0u;
            }
        }
        G.add(gm);
        B.add(bm);
#pragma unroll
        for (int k = 0; k < UB; ++k) any_slow |= sm[k] | sb[k];
        if (__any_sync(0xFFFFFFFFu, any_slow != 0)) {
            slow_win = push_slow8(c.slow, c.work_count, c.slow_cap, slow_win, sm[0], sm[1],
                                  UB > 2 ? sm[UB > 2 ? 2 : 0] : 0u, UB > 2 ? sm[UB > 2 ? 3 : 0] : 0u, sb[0], sb[1],
                                  UB > 2 ? sb[UB > 2 ? 2 : 0] : 0u, UB > 2 ? sb[UB > 2 ? 3 : 0] : 0u, c.word_base,
                                  sorted0 + ri[0], sorted0 + ri[1], sorted0 + ri[UB > 2 ? 2 : 0],
                                  sorted0 + ri[UB > 2 ? 3 : 0], lane);
        }
    }
}

// kGroups = false: the plain view, units mapped to length buckets arithmetically (unit_start[]);
// kGroups = true:  the multiplicity-compressed view, units mapped through the group table (m, class)
template <int W, bool kGroups>
__global__ void __launch_bounds__(kBitThreads, 1)
score_bitset_kernel(const BitsetArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    uint32_t *s_tab = reinterpret_cast<uint32_t *>(smem);
    const uint32_t tab_words = a.mg.entries * W;
    uint32_t *s_cnt = s_tab + tab_words;                       // 32*W paths x {bad, good}
    uint64_t *s_bar = reinterpret_cast<uint64_t *>(s_cnt + 64 * W);

    const uint32_t chunk = blockIdx.y;
    const uint32_t p_base = chunk * 32u * W;
    const int lane = threadIdx.x & 31;
    __shared__ uint4 s_groups[kGroups ? kMaxGroups : 1];
    __shared__ uint32_t s_ustart[32];
    if (threadIdx.x == 0) mbar_init(s_bar, 1);
    if (kGroups) {
        for (uint32_t i = threadIdx.x; i < a.n_groups; i += blockDim.x)
            s_groups[i] = __ldg(reinterpret_cast<const uint4 *>(a.groups) + i);
    } else if (threadIdx.x < 32) {
        s_ustart[threadIdx.x] = threadIdx.x <= kMaxBucket ? a.unit_start[threadIdx.x] : 0xFFFFFFFFu;
    }
    for (uint32_t i = threadIdx.x; i < 64u * W; i += blockDim.x) s_cnt[i] = 0;
    __syncthreads();
    pdl_wait();      // build_tables' mask tables (everything above only touched this CTA's shared memory)
    pdl_trigger();
    if (threadIdx.x == 0) {
        // one TMA bulk copy per <= 64 KB slice of this chunk's table
        const uint32_t bytes = tab_words * 4u;
        const unsigned char *src =
            reinterpret_cast<const unsigned char *>(a.masks + (size_t)chunk * tab_words);
        mbar_expect_tx(s_bar, bytes);
        for (uint32_t o = 0; o < bytes; o += 65536u)
            bulk_g2s(smem + o, src + o, min(65536u, bytes - o), s_bar);
    }
    mbar_wait(s_bar, 0);

    const uint32_t w = (uint32_t)lane % W;
    // node block: 2 words per (node, w), {N, F+|F-} ({F+, F-} stays in global memory); the plain entries
    // (LM, pairs) live behind it at (entry * W + w)
    const uint2 *nt = reinterpret_cast<const uint2 *>(s_tab + (size_t)geom_node_base() * W) + w;
    const uint2 *nf = reinterpret_cast<const uint2 *>(a.fplane) + (size_t)chunk * (a.mg.L + 1) * W + w;
    const uint32_t *tab = s_tab + w;
    const uint32_t word_base = p_base + 32u * w;
    Planes G, B;
    G.clear();
    B.clear();
    uint32_t rows_acc = 0;
    uint2 slow_win = make_uint2(0u, 0u);   // this warp's reserved window of the second-look list
    const uint32_t n_units = a.n_units;
    uint32_t cur_shift = 0;                // the counters hold rows of weight 2^cur_shift
    UnitCtx ctx;
    ctx.nt = nt;
    ctx.nf = nf;
    ctx.s_tab = s_tab;
    ctx.tab = tab;
    ctx.pair_base = a.mg.pair_base;
    ctx.bt_base = a.mg.bt_base;
    ctx.filter = a.filter;
    ctx.slow = a.slow;
    ctx.work_count = a.work_count;
    ctx.slow_cap = a.slow_cap;
    ctx.word_base = word_base;

    auto flush = [&]() {
#pragma unroll 2
        for (int c = 0; c < 8; ++c) {
            uint32_t gl, gh, bl, bh;
            G.value4(c, gl, gh);
            B.value4(c, bl, bh);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint32_t g = ((gl >> (8 * j)) & 0xFFu) | (((gh >> (8 * j)) & 0xFFu) << 8);
                const uint32_t b = ((bl >> (8 * j)) & 0xFFu) | (((bh >> (8 * j)) & 0xFFu) << 8);
                const uint32_t col = 32 * w + c + 8 * j;
                if (b) atomicAdd(&s_cnt[2 * col], b << cur_shift);
                if (g) atomicAdd(&s_cnt[2 * col + 1], g << cur_shift);
            }
        }
        G.clear();
        B.clear();
        rows_acc = 0;
    };

    // Dynamic scheduling, two units ahead: `unit` is being scored, `next` is known to the warp and
    // its rows are being prefetched into L1, `pending` is lane 0's atomic still in flight.  Units
    // are claimed from the longest rows down to the shortest (the counter counts claims; unit =
    // n_units - 1 - claim), so the last units a CTA waits for at the end are the cheapest ones.
    uint32_t next = 0, pending = 0;
    if (lane == 0) {
        next = atomicAdd(a.unit_ctr + chunk, 1u);
        pending = atomicAdd(a.unit_ctr + chunk, 1u);
    }
    next = __shfl_sync(0xFFFFFFFFu, next, 0);
    uint32_t next_gi = 0;      // kGroups: group index of `next`; else: its bucket
    if (kGroups) next_gi = a.n_groups ? group_of(s_groups, a.n_groups - 1u, n_units - 1u - min(next, n_units - 1u)) : 0u;
    else next_gi = (uint32_t)unit_bucket(s_ustart, n_units - 1u - min(next, n_units - 1u), lane);
    for (;;) {
        if (next >= n_units) break;
        const uint32_t unit = n_units - 1u - next;
        uint32_t row0, unrows;
        int b;
        if (kGroups) {
            const uint4 gr = s_groups[next_gi];          // {unit_start, row0, nrows, m | shift << 16}
            b = (int)(gr.w & 0xFFFFu);
            row0 = gr.y + (unit - gr.x) * a.unit_rows;
            unrows = min(a.unit_rows, gr.y + gr.z - row0);
            if ((gr.w >> 16) != cur_shift) {   // another multiplicity class: settle the counters first
                flush();
                cur_shift = gr.w >> 16;
            }
        } else {
            b = (int)next_gi;
            row0 = (unit - a.unit_start[b]) * a.unit_rows;
            unrows = min(a.unit_rows, a.bk.count[b] - row0);
        }
        next = __shfl_sync(0xFFFFFFFFu, pending, 0);
        if (lane == 0) pending = atomicAdd(a.unit_ctr + chunk, 1u);
        if (next < n_units) {
            if (kGroups) {
                next_gi = group_of(s_groups, next_gi, n_units - 1u - next);
                prefetch_unit(a, n_units - 1u - next, s_groups[next_gi], lane);
            } else {
                next_gi = (uint32_t)unit_bucket(s_ustart, n_units - 1u - next, lane);
                prefetch_unit(a, n_units - 1u - next,
                              make_uint4(a.unit_start[next_gi], 0u, a.bk.count[next_gi], next_gi), lane);
            }
        }
        // rows [row0, row0 + unrows) of bucket b's dense matrix
        const uint16_t *ubase = a.bk.bstep + a.bk.step_base[b] + (size_t)row0 * row_stride(b);
        const uint32_t usorted = a.bk.row_start[b] + row0;
        const uint32_t ulm = tab[(a.mg.lm_base + b) * W];
        switch (b) {
#define GFE_CASE(M_) case M_: score_unit<M_, W, false>(ctx, ubase, nullptr, unrows, usorted, ulm, G, B, slow_win, lane); break;
            GFE_CASE(1) GFE_CASE(2) GFE_CASE(3) GFE_CASE(4) GFE_CASE(5) GFE_CASE(6) GFE_CASE(7)
            GFE_CASE(8) GFE_CASE(9) GFE_CASE(10) GFE_CASE(11) GFE_CASE(12) GFE_CASE(13)
            GFE_CASE(14) GFE_CASE(15) GFE_CASE(16)
#undef GFE_CASE
            default: break;
        }
        rows_acc += a.unit_rows;
        if (rows_acc > (1u << kPlanes) - 2u * a.unit_rows) flush();
    }
    flush();
    fill_slow_holes(a.slow, a.work_count, a.slow_cap, slow_win, lane);
    __syncthreads();
    // alignments without steps pass the filter and score 0 against every path: good
    const uint32_t empty_rows = blockIdx.x == 0 ? a.bk.count[0] : 0u;
    for (uint32_t i = threadIdx.x; i < 64u * W; i += blockDim.x) {
        const uint32_t p = p_base + (i >> 1);
        if (p >= a.P) continue;
        const uint32_t v = s_cnt[i] + ((i & 1u) ? empty_rows : 0u);
        if (v) atomicAdd(&a.counts[2 * (size_t)p_base + i], v);
    }
}

// ------------------------------------------------------------------------------------------------
// K1i: score_incr — incremental scoring for the search.  A candidate path that
// extends its parent by one step v can only change the verdict of rows that contain v (a new
// occurrence ends at the new step, a newly passing row names v) and of the rows of F (rows with the
// root node: the start-overhang class; scored afresh for every path).  A chunk is <= 32 paths that
// end in the same node v (one word per table entry, W = 1): pass 0 scores the inverted-index lists
// of v into accumulator set 0, pass 1 the lists of F into set 1.  Same table, row evaluation,
// counters and second-look list as score_bitset_kernel; rows are gathered through the index.
// ------------------------------------------------------------------------------------------------
constexpr int kIncrThreads = 128;

__global__ void __launch_bounds__(kIncrThreads)
score_incr_kernel(const IncrArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    uint32_t *s_tab = reinterpret_cast<uint32_t *>(smem);
    const uint32_t tab_words = a.mg.entries;                   // W = 1
    uint32_t *s_cnt = s_tab + tab_words;                       // 2 sets x 32 paths x {bad, good}
    uint64_t *s_bar = reinterpret_cast<uint64_t *>(s_cnt + 128);
    __shared__ uint32_t s_ustart[33], s_goff[32], s_gsize[32];

    const uint32_t chunk = blockIdx.y;
    const uint32_t p_base = chunk * 32u;
    const int lane = threadIdx.x & 31;
    if (threadIdx.x == 0) mbar_init(s_bar, 1);
    if (threadIdx.x < 128) s_cnt[threadIdx.x] = 0;
    if (threadIdx.x < 32) {
        // group k: pass (k >> 4), rows of length (k & 15) + 1 of node v (pass 0) / of F (pass 1)
        const uint32_t v = a.chunk_node[chunk];
        const uint32_t node = (threadIdx.x >> 4) ? a.mg.L : v;
        const uint32_t m = (threadIdx.x & 15u) + 1u;
        uint32_t o0 = 0, size = 0;
        if (node != 0xFFFFFFFFu) {
            o0 = a.inv.off[node * kIdxLens + m];
            size = a.inv.off[node * kIdxLens + m + 1] - o0;
        }
        const uint32_t units = (size + a.unit_rows - 1) / a.unit_rows;
        uint32_t incl = units;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += t;
        }
        s_ustart[threadIdx.x] = incl - units;
        s_goff[threadIdx.x] = o0;
        s_gsize[threadIdx.x] = size;
        if (threadIdx.x == 31) s_ustart[32] = incl;
    }
    __syncthreads();
    pdl_wait();      // build_tables' mask tables (everything above read only the index and this CTA's shared memory)
    pdl_trigger();
    if (threadIdx.x == 0) {
        const uint32_t bytes = tab_words * 4u;
        const unsigned char *src = reinterpret_cast<const unsigned char *>(a.masks + (size_t)chunk * tab_words);
        mbar_expect_tx(s_bar, bytes);
        for (uint32_t o = 0; o < bytes; o += 65536u) bulk_g2s(smem + o, src + o, min(65536u, bytes - o), s_bar);
    }
    mbar_wait(s_bar, 0);

    UnitCtx ctx;
    ctx.nt = reinterpret_cast<const uint2 *>(s_tab + geom_node_base());
    ctx.nf = reinterpret_cast<const uint2 *>(a.fplane) + (size_t)chunk * (a.mg.L + 1);
    ctx.s_tab = s_tab;
    ctx.tab = s_tab;
    ctx.pair_base = a.mg.pair_base;
    ctx.bt_base = a.mg.bt_base;
    ctx.filter = 1;
    ctx.slow = a.slow;
    ctx.work_count = a.work_count;
    ctx.slow_cap = a.slow_cap;
    ctx.word_base = p_base;
    Planes G, B;
    G.clear();
    B.clear();
    uint32_t rows_acc = 0, cur_set = 0;
    uint2 slow_win = make_uint2(0u, 0u);
    const uint32_t n_units = s_ustart[32];

    auto flush = [&](uint32_t set) {
#pragma unroll 2
        for (int c = 0; c < 8; ++c) {
            uint32_t gl, gh, bl, bh;
            G.value4(c, gl, gh);
            B.value4(c, bl, bh);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint32_t g = ((gl >> (8 * j)) & 0xFFu) | (((gh >> (8 * j)) & 0xFFu) << 8);
                const uint32_t b = ((bl >> (8 * j)) & 0xFFu) | (((bh >> (8 * j)) & 0xFFu) << 8);
                const uint32_t col = c + 8 * j;
                if (b) atomicAdd(&s_cnt[64 * set + 2 * col], b);
                if (g) atomicAdd(&s_cnt[64 * set + 2 * col + 1], g);
            }
        }
        G.clear();
        B.clear();
        rows_acc = 0;
    };

    // units are claimed in ascending order, so a warp sees pass 0 before pass 1; one claim is always in
    // flight (lane 0's atomic for the unit after next), so its round trip hides behind a unit of work
    uint32_t next = 0, pending = 0;
    if (lane == 0) {
        next = atomicAdd(a.unit_ctr + chunk, 1u);
        pending = atomicAdd(a.unit_ctr + chunk, 1u);
    }
    next = __shfl_sync(0xFFFFFFFFu, next, 0);
    for (;;) {
        const uint32_t unit = next;
        if (unit >= n_units) break;
        next = __shfl_sync(0xFFFFFFFFu, pending, 0);
        if (lane == 0) pending = atomicAdd(a.unit_ctr + chunk, 1u);
        const int g = __popc(__ballot_sync(0xFFFFFFFFu, unit >= s_ustart[lane])) - 1;
        const uint32_t set = (uint32_t)g >> 4;
        const int m = (g & 15) + 1;
        if (set != cur_set) {
            flush(cur_set);
            cur_set = set;
            ctx.word_base = p_base | (set << 30);
        }
        const uint32_t row0 = (unit - s_ustart[g]) * a.unit_rows;
        const uint32_t nrows = min(a.unit_rows, s_gsize[g] - row0);
        const uint32_t *idx = a.inv.rows + s_goff[g] + row0;
        const uint16_t *base = a.bk.bstep + a.bk.step_base[m];
        const uint32_t sorted0 = a.bk.row_start[m];
        const uint32_t lm = s_tab[a.mg.lm_base + m];
        switch (m) {
#define GFE_CASE(M_) case M_: score_unit<M_, 1, true>(ctx, base, idx, nrows, sorted0, lm, G, B, slow_win, lane); break;
            GFE_CASE(1) GFE_CASE(2) GFE_CASE(3) GFE_CASE(4) GFE_CASE(5) GFE_CASE(6) GFE_CASE(7)
            GFE_CASE(8) GFE_CASE(9) GFE_CASE(10) GFE_CASE(11) GFE_CASE(12) GFE_CASE(13)
            GFE_CASE(14) GFE_CASE(15) GFE_CASE(16)
#undef GFE_CASE
            default: break;
        }
        rows_acc += (a.unit_rows + 31u) / 32u;      // W = 1: a lane sees every 32nd row of the unit
        if (rows_acc > (1u << kPlanes) - 2u * ((a.unit_rows + 31u) / 32u)) flush(cur_set);
    }
    flush(cur_set);
    fill_slow_holes(a.slow, a.work_count, a.slow_cap, slow_win, lane);
    __syncthreads();
    // alignments without steps pass the filter and score 0 against every path: good, in the fresh set
    const uint32_t empty_rows = blockIdx.x == 0 ? a.bk.count[0] : 0u;
    for (uint32_t i = threadIdx.x; i < 128u; i += blockDim.x) {
        const uint32_t set = i >> 6, col = (i & 63u) >> 1, gb = i & 1u;
        const uint32_t v = s_cnt[i] + ((set == 1u && gb == 1u) ? empty_rows : 0u);
        if (v) atomicAdd(&a.counts[4 * (size_t)(p_base + col) + 2 * set + gb], v);
    }
}

template <int W>
void launch_w(const BitsetArgs &a, int sm_count, cudaStream_t st) {
    const size_t smem = (size_t)a.mg.entries * W * 4 + 64 * W * 4 + 16;
    const uint32_t chunks = (a.P + 32 * W - 1) / (32 * W);
    uint32_t gx = (uint32_t)sm_count / chunks;
    if (gx < 1) gx = 1;
    const uint32_t n_units = a.n_units;
    const uint32_t max_gx = (n_units + (kBitThreads / 32) - 1) / (kBitThreads / 32);
    if (gx > max_gx) gx = max_gx ? max_gx : 1;
    if (a.groups) launch_pdl(1, score_bitset_kernel<W, true>, dim3(gx, chunks), dim3(kBitThreads), smem, st, a);
    else launch_pdl(1, score_bitset_kernel<W, false>, dim3(gx, chunks), dim3(kBitThreads), smem, st, a);
}

template <int W>
cudaError_t configure_w() {
    const cudaError_t e = cudaFuncSetAttribute(score_bitset_kernel<W, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                               (int)(kBitsetSmemBudget + 1024));
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(score_bitset_kernel<W, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)(kBitsetSmemBudget + 1024));
}

}  // namespace

// cudaFuncSetAttribute applies to the CURRENT device: called from configure_kernels(), i.e. once per
// gfe_create, so every engine's device gets the large-shared-memory opt-in
int configure_bitset_kernels() {
    cudaError_t e;
    if ((e = configure_w<1>()) != cudaSuccess) return (int)e;
    if ((e = configure_w<2>()) != cudaSuccess) return (int)e;
    if ((e = configure_w<4>()) != cudaSuccess) return (int)e;
    if ((e = configure_w<8>()) != cudaSuccess) return (int)e;
    if ((e = configure_w<16>()) != cudaSuccess) return (int)e;
    if ((e = cudaFuncSetAttribute(score_incr_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)(kBitsetSmemBudget + 1024))) != cudaSuccess)
        return (int)e;
    return (int)configure_w<32>();
}

void launch_score_incr(const IncrArgs &a, int sm_count, cudaStream_t st) {
    if (a.chunks == 0) return;
    const size_t smem = (size_t)a.mg.entries * 4 + 128 * 4 + 16;
    // enough CTAs to fill the machine (4 per SM at 128 registers), shared between the chunks
    uint32_t gx = ((uint32_t)sm_count * 4 + a.chunks - 1) / a.chunks;
    if (gx < 1) gx = 1;
    if (gx > 64) gx = 64;
    launch_pdl(1, score_incr_kernel, dim3(gx, a.chunks), dim3(kIncrThreads), smem, st, a);
}

void launch_score_bitset(const BitsetArgs &a, int sm_count, cudaStream_t st) {
    if (a.P == 0) return;
    switch (a.mg.W) {
        case 1: launch_w<1>(a, sm_count, st); break;
        case 2: launch_w<2>(a, sm_count, st); break;
        case 4: launch_w<4>(a, sm_count, st); break;
        case 8: launch_w<8>(a, sm_count, st); break;
        case 16: launch_w<16>(a, sm_count, st); break;
        default: launch_w<32>(a, sm_count, st); break;
    }
}

}  // namespace gfe
