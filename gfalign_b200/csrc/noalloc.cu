// gfalign_b200/csrc/noalloc.cu — filter mode for callers that must not allocate device memory.
//
// Who this is for.  gfe_filter_alignments lives on an engine, and an engine grows its device
// buffers with cudaMalloc whenever a table or a request is larger than the last one.  Some hosts
// cannot afford that: a process in which another component (a framework's caching allocator, a
// database's memory manager) has already taken nearly all of HBM for itself, so that a cudaMalloc
// issued from inside this library fails or eats into a budget somebody else accounted for; and a
// service that keeps other streams busy on the same GPU, where cudaMalloc / cudaFree stall the whole
// device for a moment.  Such a caller already holds the alignment table in device memory it obtained
// its own way.  gfe_filter_alignments_static takes that table, the node set and the output array as
// device pointers and needs nothing else from the allocator: the two bitmaps the filter kernels work
// through (the node set, n_nodes bits; "this step is outside the set", one bit per step) live in
// the fixed array below, which is part of the library image.
//
// Cost and limits.  The array is kScratchWords * 4 bytes (8.125 MiB).  Every GPU on which this
// library's kernels are loaded carries it until the process exits, used or not.  It bounds the
// request: total_steps <= GFE_STATIC_FILTER_MAX_STEPS, n_nodes <= GFE_STATIC_FILTER_MAX_NODES;
// anything larger is refused with GFE_ERR_LIMIT (use the engine entry point).
//
// Sharing.  There is ONE array per GPU and every call on that GPU uses it.  It keeps whatever the
// previous call left: the node-set words are zeroed at the start of each call, the per-step words in
// range are all overwritten before they are read.  Two calls that run on the GPU at the same time
// (different streams, or different host threads on different streams) write the same words and give
// wrong answers; calls on one stream, or otherwise ordered, are fine.
//
// First call on a GPU.  `g_filter_scratch` as written in host code is only a handle the runtime
// knows the array by — it is not a pointer a kernel could use.  The first call on each device asks
// the runtime for the real device address (cudaGetSymbolAddress, with that device current) and reads
// the SM count; both are remembered per device.  That query may load the library's kernels onto the
// GPU, so make the first call on a device at start-up, before any stream capture; later calls only
// enqueue a memset and three kernels.
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>

#include <mutex>

#include "../../include/gfalign_b200.h"
#include "kernels.cuh"

static constexpr uint32_t kNodeWords = GFE_STATIC_FILTER_MAX_NODES / 32;     // node-set bitmap
static constexpr uint32_t kStepWords = GFE_STATIC_FILTER_MAX_STEPS / 32;     // one bit per step
static constexpr uint32_t kScratchWords = kNodeWords + kStepWords;

// [0, kNodeWords): node set; [kNodeWords, kScratchWords): steps outside the set.
static __device__ __align__(256) uint32_t g_filter_scratch[kScratchWords];

namespace {

constexpr int kMaxDevices = 64;

struct DeviceSetup {
    uint32_t *scratch = nullptr;    // device address of g_filter_scratch on this GPU
    int sm_count = 0;
};
DeviceSetup g_setup[kMaxDevices];
std::mutex g_setup_lock;

int fail(char *err, size_t err_len, int code, const char *fmt, ...) {
    if (err && err_len) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, err_len, fmt, ap);
        va_end(ap);
    }
    return code;
}

// bits[node >> 5] |= 1 << (node & 31) for every listed node the table knows
__global__ void __launch_bounds__(256) node_set_kernel(const uint32_t *__restrict__ nodes, uint32_t n,
                                                       uint32_t n_nodes, uint32_t *bits) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t node = __ldg(nodes + i);
        if (node < n_nodes) atomicOr(bits + (node >> 5), 1u << (node & 31));
    }
}

}  // namespace

extern "C" int gfe_filter_alignments_static(const uint32_t *d_offsets, const uint32_t *d_steps, uint64_t A,
                                            uint64_t total_steps, uint32_t n_nodes, const uint32_t *d_nodes,
                                            uint32_t n, int32_t min_nodes, uint8_t *d_keep, void *cuda_stream,
                                            char *err, size_t err_len) {
    if (err && err_len) err[0] = 0;
    if (A == 0) return GFE_OK;
    if (!d_offsets || !d_keep || (!d_steps && total_steps) || (!d_nodes && n))
        return fail(err, err_len, GFE_ERR_ARG, "gfe_filter_alignments_static: NULL device pointer");
    if (total_steps > GFE_STATIC_FILTER_MAX_STEPS)
        return fail(err, err_len, GFE_ERR_LIMIT,
                    "gfe_filter_alignments_static: %llu steps exceed the built-in scratch (%u steps); "
                    "use gfe_filter_alignments",
                    (unsigned long long)total_steps, GFE_STATIC_FILTER_MAX_STEPS);
    if (n_nodes > GFE_STATIC_FILTER_MAX_NODES)
        return fail(err, err_len, GFE_ERR_LIMIT,
                    "gfe_filter_alignments_static: %u nodes exceed the built-in scratch (%u nodes); "
                    "use gfe_filter_alignments",
                    n_nodes, GFE_STATIC_FILTER_MAX_NODES);

    int dev = -1;
    cudaError_t ce = cudaGetDevice(&dev);
    if (ce != cudaSuccess)
        return fail(err, err_len, GFE_ERR_CUDA, "gfe_filter_alignments_static: cudaGetDevice: %s (no CPU fallback)",
                    cudaGetErrorString(ce));
    if (dev < 0 || dev >= kMaxDevices)
        return fail(err, err_len, GFE_ERR_LIMIT, "gfe_filter_alignments_static: device ordinal %d not supported", dev);
    DeviceSetup ds;
    {
        std::lock_guard<std::mutex> hold(g_setup_lock);
        if (!g_setup[dev].scratch) {
            // `dev` is the current device: the address returned is this GPU's copy of the array.
            uint32_t *p = nullptr;
            ce = cudaGetSymbolAddress(reinterpret_cast<void **>(&p), g_filter_scratch);
            if (ce != cudaSuccess || !p)
                return fail(err, err_len, GFE_ERR_CUDA, "gfe_filter_alignments_static: cudaGetSymbolAddress: %s",
                            cudaGetErrorString(ce));
            int sms = 0;
            ce = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            if (ce != cudaSuccess || sms <= 0)
                return fail(err, err_len, GFE_ERR_CUDA, "gfe_filter_alignments_static: SM count query: %s",
                            cudaGetErrorString(ce));
            g_setup[dev].sm_count = sms;
            g_setup[dev].scratch = p;
        }
        ds = g_setup[dev];
    }

    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    uint32_t *node_bits = ds.scratch;
    uint32_t *bad_bits = ds.scratch + kNodeWords;
    // the node set starts empty; the last call's set is still in there
    const size_t node_words = ((size_t)n_nodes + 31) / 32;
    if (node_words) {
        ce = cudaMemsetAsync(node_bits, 0, node_words * sizeof(uint32_t), st);
        if (ce != cudaSuccess)
            return fail(err, err_len, GFE_ERR_CUDA, "gfe_filter_alignments_static: cudaMemsetAsync: %s",
                        cudaGetErrorString(ce));
    }
    if (n) {
        unsigned blocks = (n + 255) / 256;
        const unsigned cap = (unsigned)ds.sm_count * 8;
        node_set_kernel<<<blocks < cap ? blocks : cap, 256, 0, st>>>(d_nodes, n, n_nodes, node_bits);
    }
    gfe::FilterArgs a;
    a.off = d_offsets;
    a.steps = d_steps;
    a.A = A;
    a.node_bits = node_bits;
    a.n_nodes = n_nodes;
    a.min_nodes = min_nodes;
    a.keep = d_keep;
    a.S = total_steps;
    a.bad_bits = bad_bits;
    gfe::launch_filter_rows(a, ds.sm_count, st);
    ce = cudaGetLastError();
    if (ce != cudaSuccess)
        return fail(err, err_len, GFE_ERR_CUDA, "gfe_filter_alignments_static: kernel launch: %s",
                    cudaGetErrorString(ce));
    return GFE_OK;
}
