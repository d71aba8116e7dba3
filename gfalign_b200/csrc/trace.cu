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
// filter_rows_kernel: filter mode (InAlignment::isContained + pathNodesCount,
// src/alignments.cpp:96-135, 459-472): keep[a] = every step's node is in the set && m >= min.
// A CTA owns 256 rows; their steps are one contiguous CSR range, streamed with coalesced loads;
// a step outside the set finds its row by binary search in the CTA's offsets (shared memory).
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

constexpr int kFilterRows = 256;

__global__ void __launch_bounds__(kFilterRows) filter_rows_kernel(FilterArgs a) {
    __shared__ uint32_t s_off[kFilterRows + 1];
    __shared__ uint32_t s_drop[kFilterRows];
    const uint64_t r0 = (uint64_t)blockIdx.x * kFilterRows;
    const uint32_t rows = (uint32_t)((a.A - r0) < (uint64_t)kFilterRows ? (a.A - r0) : kFilterRows);
    for (uint32_t t = threadIdx.x; t <= rows; t += blockDim.x) s_off[t] = a.off[r0 + t];
    s_drop[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t lo = s_off[0], hi = s_off[rows];
    for (uint32_t s = lo + threadIdx.x; s < hi; s += blockDim.x) {
        const uint32_t node = a.steps[s] >> 1;
        const bool in = node < a.n_nodes && ((a.node_bits[node >> 5] >> (node & 31)) & 1u);
        if (!in) {
            uint32_t l = 0, h = rows;   // largest r with s_off[r] <= s
            while (h - l > 1) {
                const uint32_t mid = (l + h) >> 1;
                if (s_off[mid] <= s) l = mid; else h = mid;
            }
            s_drop[l] = 1;
        }
    }
    __syncthreads();
    if (threadIdx.x < rows) {
        const uint32_t m = s_off[threadIdx.x + 1] - s_off[threadIdx.x];
        a.keep[r0 + threadIdx.x] = (!s_drop[threadIdx.x] && (int64_t)m >= (int64_t)a.min_nodes) ? 1 : 0;
    }
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

void launch_filter_rows(const FilterArgs &a, cudaStream_t st) {
    if (!a.A) return;
    filter_rows_kernel<<<(unsigned)((a.A + kFilterRows - 1) / kFilterRows), kFilterRows, 0, st>>>(a);
}

}  // namespace gfe
