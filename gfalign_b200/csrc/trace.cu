// gfalign_b200/csrc/trace.cu — evalPath detail (traceback columns) and filter-mode membership.
//
// pair_trace_kernel: one thread per (path, alignment) pair.  Fills the reference's DP
// (src/alignments.cpp:499-509: row 0 = -j for j <= n and 0 beyond, column 0 = 0, free gap in the
// last column) keeping two decision bits per cell, then replays get_optimal_alignment
// (src/alignments.cpp:511-554: diagonal first, then up when dp[i-1][j] >= dp[i][j-1], else left)
// and writes one op code per alignment column plus the traceback score.
//   decision words: dirs[((i-1) * count + k) * W + w], 16 cells per u32 — consecutive threads write
//   consecutive words, so the n*W stores per pair are coalesced.
//
// filter mode (InAlignment::isContained + pathNodesCount, src/alignments.cpp:96-135, 459-472):
// keep[a] = every step's node is in the set && m >= min.  Two flat, fully coalesced passes:
// filter_mark_kernel streams the step array once and writes one "outside the set" bit per step
// (S/8 bytes); filter_rows_kernel reads the offsets and tests each row's bit range.
#include "kernels.cuh"

namespace gfe {

namespace {

constexpr int kTraceBlock = 128;
constexpr int kTraceReg = 32;      // alignments up to this length keep B and the DP row per thread

__global__ void max_len_kernel(const uint32_t *off, uint64_t first, uint32_t count, uint32_t *out) {
    uint32_t mx = 0;
    for (uint64_t k = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; k < count;
         k += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t m = off[first + k + 1] - off[first + k];
        mx = m > mx ? m : mx;
    }
    for (int d = 16; d; d >>= 1) {
        const uint32_t o = __shfl_xor_sync(0xFFFFFFFFu, mx, d);
        mx = o > mx ? o : mx;
    }
    if ((threadIdx.x & 31) == 0 && mx) atomicMax(out, mx);
}

__global__ void __launch_bounds__(kTraceBlock) pair_trace_kernel(TraceArgs a) {
    extern __shared__ uint32_t s_path[];
    for (uint32_t i = threadIdx.x; i < a.n; i += blockDim.x) s_path[i] = a.path[i];
    __syncthreads();
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= a.count) return;
    const uint64_t row = a.first + k;
    const uint32_t o0 = a.off[row], m = a.off[row + 1] - o0;
    const uint32_t n = a.n;
    const bool rc = a.use_rc && a.use_rc[k];
    const uint32_t *bsrc = a.steps + o0;
    // B in the chosen direction and the DP row live per thread (local memory) for short
    // alignments, in the scratch rows of the batch for long ones
    uint32_t b_loc[kTraceReg];
    int16_t r_loc[kTraceReg + 1];
    uint32_t *B = b_loc;
    int16_t *prev = r_loc;
    if (m > (uint32_t)kTraceReg) {
        B = a.long_b + (size_t)k * a.long_stride;
        prev = a.long_rows + (size_t)k * (a.long_stride + 1);
    }
    for (uint32_t j = 0; j < m; ++j) B[j] = rc ? (bsrc[m - 1 - j] ^ 1u) : bsrc[j];
    for (uint32_t j = 0; j <= m; ++j) prev[j] = (int16_t)(j <= n ? -(int)j : 0);
    const size_t rstride = (size_t)a.count * a.W;
    uint32_t *dbase = a.dirs + (size_t)k * a.W;
    for (uint32_t i = 1; i <= n; ++i) {
        const uint32_t pa = s_path[i - 1];
        int diag = prev[0];      // dp[i-1][0]
        prev[0] = 0;             // dp[i][0]: column 0 is never written by the reference
        int left = 0;
        uint32_t word = 0;
        uint32_t *drow = dbase + (size_t)(i - 1) * rstride;
        for (uint32_t j = 1; j <= m; ++j) {
            const int up = prev[j];
            const int dS = diag + (pa == B[j - 1] ? 0 : -1);
            int v = up + (j < m ? -1 : 0);
            v = v > left - 1 ? v : left - 1;
            v = dS > v ? dS : v;
            const uint32_t bits = (v == dS ? 1u : 0u) | (up >= left ? 2u : 0u);
            word |= bits << (((j - 1) & 15) * 2);
            if ((j & 15) == 0 || j == m) {
                drow[(j - 1) >> 4] = word;
                word = 0;
            }
            diag = up;
            prev[j] = (int16_t)v;
            left = v;
        }
    }
    // traceback, written from the end of the pair's region backwards
    uint8_t *ops = a.ops + a.ops_start_base[k];
    uint32_t pos = n + m;
    uint32_t ii = n, jj = m;
    int score = 0, sb_len = 0;
    while (ii != 0 || jj != 0) {
        uint8_t op;
        if (ii == 0) {
            op = 3;
            --jj;
        } else if (jj == 0) {
            op = 2;
            --ii;
        } else {
            const uint32_t w = dbase[(size_t)(ii - 1) * rstride + ((jj - 1) >> 4)];
            const uint32_t bits = (w >> (((jj - 1) & 15) * 2)) & 3u;
            if (bits & 1u) {
                const bool eq = s_path[ii - 1] == B[jj - 1];
                op = eq ? 0 : 1;
                score += eq ? 0 : -1;
                ++sb_len;
                --ii;
                --jj;
            } else if (bits & 2u) {
                op = 2;
                --ii;
                if (sb_len > 0) score -= 1;
            } else {
                op = 3;
                ++sb_len;
                --jj;
                score -= 1;
            }
        }
        ops[--pos] = op;
    }
    a.ops_first[k] = a.ops_start_base[k] + pos;
    a.ops_len[k] = n + m - pos;
    if (a.score) a.score[k] = score;
}

// exclusive start of each pair's ops region: (off[first+k] - off[first]) + k*n
__global__ void trace_bases_kernel(const uint32_t *off, uint64_t first, uint32_t count, uint32_t n,
                                   uint64_t *base) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < count) base[k] = (uint64_t)(off[first + k] - off[first]) + (uint64_t)k * n;
}

// pass 1, flat over the step array: four steps (one 128-bit load) per lane and iteration, so a warp
// covers 128 steps = four words of the "outside the set" bitmap; the lanes' 4-bit results are OR-ed
// over groups of eight lanes with three shuffles and lanes 0, 8, 16, 24 store the words.
__global__ void __launch_bounds__(256) filter_mark_kernel(FilterArgs a) {
    const uint64_t n_groups = (a.S + 127) / 128;                  // 128 steps per warp iteration
    const uint64_t warp = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> 5;
    const uint64_t n_warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    const uint32_t lane = threadIdx.x & 31;
    auto outside = [&](uint32_t step) -> uint32_t {
        const uint32_t node = step >> 1;
        return (node < a.n_nodes && ((__ldg(a.node_bits + (node >> 5)) >> (node & 31)) & 1u)) ? 0u : 1u;
    };
    const bool vec = (reinterpret_cast<uintptr_t>(a.steps) & 15u) == 0;   // caller-owned device arrays may not be
    for (uint64_t g = warp; g < n_groups; g += n_warps) {
        const uint64_t s = g * 128 + 4ull * lane;
        uint32_t nib = 0;
        if (vec && s + 4 <= a.S) {
            const uint4 v = __ldg(reinterpret_cast<const uint4 *>(a.steps + s));
            nib = outside(v.x) | (outside(v.y) << 1) | (outside(v.z) << 2) | (outside(v.w) << 3);
        } else {
            for (uint32_t k = 0; k < 4 && s + k < a.S; ++k) nib |= outside(__ldg(a.steps + s + k)) << k;
        }
        uint32_t word = nib << (4u * (lane & 7u));
        word |= __shfl_xor_sync(0xFFFFFFFFu, word, 1);
        word |= __shfl_xor_sync(0xFFFFFFFFu, word, 2);
        word |= __shfl_xor_sync(0xFFFFFFFFu, word, 4);
        const uint64_t w = g * 4 + (lane >> 3);
        if ((lane & 7u) == 0 && w * 32 < a.S) a.bad_bits[w] = word;
    }
}

// pass 2: one thread per alignment, any marked step in [off[a], off[a+1])?
__global__ void __launch_bounds__(256) filter_rows_kernel(FilterArgs a) {
    const uint64_t r = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    if (r >= a.A) return;
    const uint32_t o0 = __ldg(a.off + r), o1 = __ldg(a.off + r + 1);
    bool any = false;
    if (o1 > o0) {
        const uint32_t w0 = o0 >> 5, w1 = (o1 - 1) >> 5;
        for (uint32_t w = w0; w <= w1 && !any; ++w) {
            uint32_t bits = __ldg(a.bad_bits + w);
            if (w == w0) bits &= 0xFFFFFFFFu << (o0 & 31);
            if (w == w1) bits &= 0xFFFFFFFFu >> (31 - ((o1 - 1) & 31));
            any = bits != 0;
        }
    }
    a.keep[r] = (!any && (int64_t)(o1 - o0) >= (int64_t)a.min_nodes) ? 1 : 0;
}

}  // namespace

void launch_max_len(const uint32_t *off, uint64_t first, uint32_t count, uint32_t *out,
                    cudaStream_t st) {
    if (!count) return;
    const uint32_t blocks = (count + 255) / 256;
    max_len_kernel<<<blocks < 1184 ? blocks : 1184, 256, 0, st>>>(off, first, count, out);
}

void launch_trace_bases(const uint32_t *off, uint64_t first, uint32_t count, uint32_t n,
                        uint64_t *base, cudaStream_t st) {
    if (!count) return;
    trace_bases_kernel<<<(count + 255) / 256, 256, 0, st>>>(off, first, count, n, base);
}

void launch_pair_trace(const TraceArgs &a, cudaStream_t st) {
    if (!a.count) return;
    pair_trace_kernel<<<(a.count + kTraceBlock - 1) / kTraceBlock, kTraceBlock,
                        a.n * sizeof(uint32_t), st>>>(a);
}

void launch_filter_rows(const FilterArgs &a, int sm_count, cudaStream_t st) {
    if (!a.A) return;
    const uint64_t n_groups = (a.S + 127) / 128;
    if (n_groups) {
        uint64_t blocks = (n_groups + 7) / 8;          // 8 warps per CTA
        const uint64_t cap = (uint64_t)sm_count * 32;
        if (blocks > cap) blocks = cap;
        filter_mark_kernel<<<(unsigned)blocks, 256, 0, st>>>(a);
    }
    filter_rows_kernel<<<(unsigned)((a.A + 255) / 256), 256, 0, st>>>(a);
}

}  // namespace gfe
