// gfalign_b200/csrc/peer.cu — the one exchange step of the sharded path: the
// P x {bad, good} count records of the alignment shards are summed across the GPUs of a box by ONE
// small kernel per rank that reads every peer's record straight out of its HBM over NVLink
// (symmetric buffers mapped into every process), instead of a NCCL all-reduce (~45-60 us of launch
// and protocol latency for 8 KB on 8 GPUs; this kernel: a flag round trip plus 8 KB per peer).
//
// Buffer layout (the same on every rank; `bases[p]` is rank p's buffer as mapped here):
//   u32 data[2][n_cap]   two halves, used alternately: half (epoch & 1) carries this epoch's records
//   u32 flags[kMaxPeers] flags[q] on rank p = the last epoch rank q announced to rank p
// Protocol of epoch e (e = 1, 2, ...; all ranks call with the same e, stream-ordered behind the
// kernels that produced data[e & 1]):
//   1. block 0 publishes: fence, then flags[rank] := e on every peer (release, system scope);
//   2. every block waits until its own flags[q] >= e for all q (acquire, system scope);
//   3. out[i] = sum over q of peer q's data[e & 1][i].
// A rank can run at most one epoch ahead of a peer (its next publish needs that peer's flag), and
// then it writes the OTHER half, so no second barrier is needed; flags only grow, hence ">=".
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include "../../include/gfalign_b200.h"

namespace {

constexpr int kMaxPeers = GFE_PEER_MAX;

struct PeerArgs {
    uint32_t *base[kMaxPeers];
    int world, rank;
    uint32_t n, n_cap, epoch;
    uint32_t *out;
    unsigned long long timeout_ns;
};

__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__global__ void __launch_bounds__(512) peer_sum_kernel(PeerArgs a) {
    const size_t flag0 = 2 * (size_t)a.n_cap;
    if (blockIdx.x == 0 && threadIdx.x < a.world) {
        __threadfence_system();   // this rank's records (written by earlier kernels of the stream)
        st_release_sys(a.base[threadIdx.x] + flag0 + a.rank, a.epoch);
    }
    // Lock-step is the caller's contract (every rank calls with the same epoch); a peer that never
    // arrives (it raised, or skipped the batch) must not hang this GPU: the wait is bounded by
    // a.timeout_ns of %globaltimer, after which every output slot is poisoned with GFE_PEER_TIMEOUT
    // (no sum of real counts reaches it: counts are bounded by the number of alignments < 2^32 - 1).
    __shared__ int s_timeout;
    if (threadIdx.x == 0) s_timeout = 0;
    __syncthreads();
    // Callers test only the LAST slot (their flag word) for the poison.  Each CTA times its own wait, so a
    // peer arriving right at the deadline could be seen by some CTAs and not by others.  The CTA that writes
    // the last slot therefore gives up first; every other CTA waits 10 % + 1 ms longer.  If any CTA times
    // out, the peer was still absent after the shorter deadline too, so the last slot is poisoned whenever
    // any slot is.
    const unsigned last_owner = a.n ? ((a.n - 1) / blockDim.x) % gridDim.x : 0u;
    const unsigned long long limit_ns =
        blockIdx.x == last_owner ? a.timeout_ns : a.timeout_ns + a.timeout_ns / 10 + 1000000ull;
    if (threadIdx.x < a.world) {
        const uint32_t *f = a.base[a.rank] + flag0 + threadIdx.x;
        const unsigned long long t0 = global_ns();
        while ((int32_t)(ld_acquire_sys(f) - a.epoch) < 0) {
            __nanosleep(20);
            if (global_ns() - t0 > limit_ns) {
                s_timeout = 1;
                break;
            }
        }
    }
    __syncthreads();
    const bool dead = s_timeout != 0;
    const size_t half = (size_t)(a.epoch & 1u) * a.n_cap;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += gridDim.x * blockDim.x) {
        uint32_t s = 0;
        if (!dead)
            for (int q = 0; q < a.world; ++q) s += ld_relaxed_sys(a.base[q] + half + i);
        a.out[i] = dead ? GFE_PEER_TIMEOUT : s;
    }
}

}  // namespace

extern "C" int gfe_peer_sum_u32(const uint64_t *peer_bases, int world, int rank, uint32_t n_cap, uint32_t n,
                                uint32_t epoch, uint32_t *d_out, void *cuda_stream) {
    if (!peer_bases || !d_out || world < 1 || world > kMaxPeers || rank < 0 || rank >= world || n > n_cap ||
        epoch == 0)
        return GFE_ERR_ARG;
    PeerArgs a{};
    for (int q = 0; q < world; ++q) a.base[q] = reinterpret_cast<uint32_t *>(peer_bases[q]);
    a.world = world;
    a.rank = rank;
    a.n = n;
    a.n_cap = n_cap;
    a.epoch = epoch;
    a.out = d_out;
    {
        static const long long ms = getenv("GFE_PEER_TIMEOUT_MS") ? atoll(getenv("GFE_PEER_TIMEOUT_MS")) : 20000;
        a.timeout_ns = (unsigned long long)(ms > 0 ? ms : 20000) * 1000000ull;
    }
    unsigned blocks = (n + 2047) / 2048;
    if (blocks < 1) blocks = 1;
    if (blocks > 64) blocks = 64;
    peer_sum_kernel<<<blocks, 512, 0, (cudaStream_t)cuda_stream>>>(a);
    return cudaGetLastError() == cudaSuccess ? GFE_OK : GFE_ERR_CUDA;
}
