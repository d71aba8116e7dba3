// gfalign_b200/csrc/engine.cu — C ABI (include/gfalign_b200.h) over the sm_100a kernels.
// Owns the device copies of the alignment table, the universe-local view, staging buffers and
// one CUDA stream.  There is no CPU scoring path in this library: every entry point either runs
// the kernels or fails with an error status.
#include "engine.h"
#include <chrono>

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <unordered_map>

using namespace gfe;

static thread_local char g_create_error[512] = "";

static void set_error(gfe_engine *e, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    if (e) vsnprintf(e->error, sizeof(e->error), fmt, ap);
    else vsnprintf(g_create_error, sizeof(g_create_error), fmt, ap);
    va_end(ap);
}

#define CUDA_TRY(e, call)                                                                   \
    do {                                                                                    \
        cudaError_t err__ = (call);                                                         \
        if (err__ != cudaSuccess) {                                                         \
            set_error((e), "%s failed: %s (%s:%d)", #call, cudaGetErrorString(err__),       \
                      __FILE__, __LINE__);                                                  \
            return GFE_ERR_CUDA;                                                            \
        }                                                                                   \
    } while (0)

template <typename T>
static int ensure_dev(gfe_engine *e, DevBuf<T> &b, size_t n) {
    if (n <= b.cap) return GFE_OK;
    size_t want = std::max(n, b.cap + b.cap / 2);
    if (b.ptr) CUDA_TRY(e, cudaFree(b.ptr));
    b.ptr = nullptr;
    b.cap = 0;
    CUDA_TRY(e, cudaMalloc(&b.ptr, want * sizeof(T)));
    b.cap = want;
    return GFE_OK;
}

template <typename T>
static int ensure_pinned(gfe_engine *e, PinBuf<T> &b, size_t n) {
    if (n <= b.cap) return GFE_OK;
    size_t want = std::max(n, b.cap + b.cap / 2);
    if (b.ptr) CUDA_TRY(e, cudaFreeHost(b.ptr));
    b.ptr = nullptr;
    b.cap = 0;
    CUDA_TRY(e, cudaMallocHost(&b.ptr, want * sizeof(T)));
    b.cap = want;
    return GFE_OK;
}

#define TRY(x)                      \
    do {                            \
        int rc__ = (x);             \
        if (rc__ != GFE_OK) return rc__; \
    } while (0)

// next (start, stop) event pair of the timing ring; also becomes ev0/ev1 ("last launch")
static int next_event_pair(gfe_engine *e) {
    if (e->ev_used + 2 > e->ev_ring.size()) {
        for (int i = 0; i < 2; ++i) {
            cudaEvent_t ev;
            CUDA_TRY(e, cudaEventCreate(&ev));
            e->ev_ring.push_back(ev);
        }
    }
    e->ev0 = e->ev_ring[e->ev_used];
    e->ev1 = e->ev_ring[e->ev_used + 1];
    e->ev_used += 2;
    return GFE_OK;
}

extern "C" {

const char *gfe_version(void) { return "gfalign_b200 0.1 (sm_100a)"; }

const char *gfe_last_error(const gfe_engine *e) { return e ? e->error : g_create_error; }

int gfe_create(int cuda_device, gfe_engine **out) {
    if (!out) return GFE_ERR_ARG;
    *out = nullptr;
    int count = 0;
    cudaError_t err = cudaGetDeviceCount(&count);
    if (err != cudaSuccess || count == 0) {
        set_error(nullptr, "no CUDA device available (%s); gfalign_b200 has no CPU fallback",
                  err != cudaSuccess ? cudaGetErrorString(err) : "device count is 0");
        return GFE_ERR_CUDA;
    }
    if (cuda_device < 0 || cuda_device >= count) {
        set_error(nullptr, "CUDA device %d out of range (0..%d)", cuda_device, count - 1);
        return GFE_ERR_ARG;
    }
    gfe_engine *e = new (std::nothrow) gfe_engine();
    if (!e) return GFE_ERR_NOMEM;
    e->device = cuda_device;
    {
        const char *fl = getenv("GFE_FORCE_LEGACY");
        e->force_legacy = (fl && fl[0] == '1') ? 1 : 0;
        const char *dd = getenv("GFE_DEDUP");
        e->dedup_enabled = !(dd && dd[0] == '0');
    }
    const bool dbg = getenv("GFE_DEBUG_TIMING") != nullptr;
    auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t0 = now();
    err = cudaSetDevice(cuda_device);
    if (err == cudaSuccess) err = cudaFree(nullptr);   // creates the primary context here, not in the first real call
    const double t1 = now();
    if (err == cudaSuccess) err = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking);
    if (err == cudaSuccess) err = cudaDeviceGetAttribute(&e->sm_count, cudaDevAttrMultiProcessorCount, cuda_device);
    const double t2 = now();
    if (err == cudaSuccess) {
        int rc = configure_kernels();
        if (rc != 0) err = (cudaError_t)rc;
    }
    if (dbg)
        fprintf(stderr, "[gfe_create] context %.3f s, stream+attributes %.3f s, kernel configuration %.3f s\n",
                t1 - t0, t2 - t1, now() - t2);
    if (err != cudaSuccess) {
        set_error(nullptr, "CUDA initialisation failed on device %d: %s", cuda_device,
                  cudaGetErrorString(err));
        delete e;
        return GFE_ERR_CUDA;
    }
    *out = e;
    return GFE_OK;
}

void gfe_destroy(gfe_engine *e) {
    if (!e) return;
    if (getenv("GFE_DEBUG_TIMING") && e->incr_n[0])
        fprintf(stderr, "[gfe incremental] %llu batches, %llu paths, %llu chunks, %llu full-scan items; host prep %.3f s, "
                        "enqueue %.3f s, wait %.3f s\n", (unsigned long long)e->incr_n[0], (unsigned long long)e->incr_n[1],
                (unsigned long long)e->incr_n[2], (unsigned long long)e->incr_n[3], e->incr_t[0], e->incr_t[1], e->incr_t[2]);
    cudaSetDevice(e->device);
    if (e->stream) cudaStreamSynchronize(e->stream);
    e->free_all();
    for (cudaEvent_t ev : e->ev_ring) cudaEventDestroy(ev);
    if (e->ev_aux) cudaEventDestroy(e->ev_aux);
    if (e->stream) cudaStreamDestroy(e->stream);
    delete e;
}

int gfe_set_timing(gfe_engine *e, int enabled) {
    if (!e) return GFE_ERR_ARG;
    e->timing = enabled != 0;
    return GFE_OK;
}

float gfe_last_kernel_ms(const gfe_engine *e) { return e ? e->last_kernel_ms : 0.f; }
float gfe_last_aux_kernel_ms(const gfe_engine *e) { return e ? e->last_aux_ms : 0.f; }

int gfe_kernel_time_stats(gfe_engine *e, double *total_ms, uint32_t *launches) {
    if (!e || !total_ms || !launches) return GFE_ERR_ARG;
    CUDA_TRY(e, cudaSetDevice(e->device));
    CUDA_TRY(e, cudaDeviceSynchronize());
    double tot = 0;
    for (size_t i = 0; i + 1 < e->ev_used; i += 2) {
        float ms = 0.f;
        CUDA_TRY(e, cudaEventElapsedTime(&ms, e->ev_ring[i], e->ev_ring[i + 1]));
        tot += ms;
    }
    *total_ms = tot;
    *launches = (uint32_t)(e->ev_used / 2);
    e->ev_used = 0;
    return GFE_OK;
}

int gfe_device_status(gfe_engine *e, uint32_t *flags) {
    if (!e || !flags) return GFE_ERR_ARG;
    CUDA_TRY(e, cudaSetDevice(e->device));
    CUDA_TRY(e, cudaDeviceSynchronize());
    uint32_t err[2] = {0, 0}, wc[4] = {0, 0, 0, 0};
    if (e->d_err.ptr) {
        CUDA_TRY(e, cudaMemcpy(err, e->d_err.ptr, sizeof(err), cudaMemcpyDeviceToHost));
        CUDA_TRY(e, cudaMemset(e->d_err.ptr, 0, sizeof(err)));   // reported once
    }
    if (e->d_work_count.ptr)
        CUDA_TRY(e, cudaMemcpy(wc, e->d_work_count.ptr, sizeof(wc), cudaMemcpyDeviceToHost));
    // err[1]: overflow bits of every batch since the last report (the prologue of the next batch
    // folds them in before it clears the counters); wc: the batch still in the counters
    *flags = ((err[0] & 1u) ? 1u : 0u) | ((wc[1] || wc[3] || err[1]) ? 2u : 0u) | (err[0] & 4u);
    return GFE_OK;
}

int gfe_device_status_to(gfe_engine *e, uint32_t *d_flags, void *cuda_stream) {
    if (!e || !d_flags) return GFE_ERR_ARG;
    CUDA_TRY(e, cudaSetDevice(e->device));
    launch_status_flag(e->d_err.ptr, e->d_work_count.ptr, d_flags, (cudaStream_t)cuda_stream);
    e->kernel_launches += 1;
    CUDA_TRY(e, cudaGetLastError());
    return GFE_OK;
}

int gfe_set_dedup(gfe_engine *e, int enabled) {
    if (!e) return GFE_ERR_ARG;
    e->dedup_enabled = enabled != 0;
    return GFE_OK;
}

int gfe_get_dedup_stats(const gfe_engine *e, uint64_t *rows, uint64_t *distinct, uint64_t *view_rows, int *in_use) {
    if (!e) return GFE_ERR_ARG;
    if (rows) *rows = e->dedup_rows;
    if (distinct) *distinct = e->dedup_distinct;
    if (view_rows) *view_rows = e->dedup_view_rows;
    if (in_use) *in_use = (e->have_dedup && e->dedup_enabled) ? 1 : 0;
    return GFE_OK;
}

int gfe_get_counters(const gfe_engine *e, gfe_counters *out) {
    if (!e || !out) return GFE_ERR_ARG;
    out->kernel_launches = e->kernel_launches;
    out->exact_pairs = e->exact_pairs;
    out->h2d_bytes = e->h2d_bytes;
    out->d2h_bytes = e->d2h_bytes;
    out->slow_pairs = e->slow_pairs;
    return GFE_OK;
}

// ---- alignment table ----------------------------------------------------------------------------

static int validate_csr_host(gfe_engine *e, const uint32_t *offsets, const uint32_t *steps,
                             uint64_t A, uint32_t n_nodes, uint32_t *max_m) {
    if (!offsets || (A && !steps && offsets[A] != 0)) {
        set_error(e, "alignment CSR: NULL buffer");
        return GFE_ERR_ARG;
    }
    if (offsets[0] != 0) {
        set_error(e, "alignment CSR: offsets[0] must be 0");
        return GFE_ERR_ARG;
    }
    uint32_t mx = 0;
    for (uint64_t a = 0; a < A; ++a) {
        if (offsets[a + 1] < offsets[a]) {
            set_error(e, "alignment CSR: offsets not monotone at %llu", (unsigned long long)a);
            return GFE_ERR_ARG;
        }
        uint32_t m = offsets[a + 1] - offsets[a];
        // rows longer than GFE_MAX_STEPS are accepted here: with the id filter the reference drops such a row
        // before its DP unless every node of it is on the candidate path (src/eval.cpp:81-91); only then, and
        // in every unfiltered call, its dp[MAX_N][MAX_N] cannot hold the row -> GFE_ERR_LIMIT at score time
        mx = std::max(mx, m);
    }
    const uint64_t S = offsets[A];
    for (uint64_t i = 0; i < S; ++i)
        if ((steps[i] >> 1) >= n_nodes) {
            set_error(e, "alignment step %llu names node %u >= n_nodes %u", (unsigned long long)i,
                      steps[i] >> 1, n_nodes);
            return GFE_ERR_ARG;
        }
    *max_m = mx;
    return GFE_OK;
}

// rows per bucket (hist[0..kNumBuckets-1]; hist[kNumBuckets + 1] = steps of the long rows) -> geometry
static void layout_buckets(const uint32_t *hist, Buckets &bk, uint64_t *elems) {
    uint64_t row = 0, tile = 0, base = 0;
    for (int b = 0; b < kNumBuckets; ++b) {
        bk.count[b] = hist[b];
        bk.row_start[b] = (uint32_t)row;
        bk.tile_start[b] = (uint32_t)tile;
        base = (base + 7) & ~7ull;            // 16-byte aligned matrices
        bk.step_base[b] = (uint32_t)base;
        row += hist[b];
        if (b >= 1) tile += (hist[b] + kBlock - 1) / kBlock;
        base += b == kLongBucket ? (uint64_t)hist[kNumBuckets + 1]
                                 : (uint64_t)hist[b] * row_stride(b);
    }
    bk.row_start[kNumBuckets] = (uint32_t)row;
    bk.tile_start[kNumBuckets] = (uint32_t)tile;
    *elems = base + 8;
}

// Scheduling groups of the bitset kernel for one view: gcount[m * 32 + k] rows in class k of bucket m
// (plain view: everything in class 0).  One table per W.
static int build_group_tables(gfe_engine *e, int view, const uint32_t *gcount, uint64_t rows_total) {
    std::vector<UnitGroup> host((size_t)6 * kMaxGroups);
    for (int wi = 0; wi < 6; ++wi) {
        const uint32_t W = 1u << wi;
        const uint32_t step = 4u * (32u / W);
        const uint64_t warps = (uint64_t)e->sm_count * (kBitThreads / 32);
        // plain view: ~8 units per warp (dynamic balance); class-sorted view: ~2 larger units per warp, since
        // every change of class costs a warp a counter flush
        const uint64_t want = rows_total / ((view ? 2 : 8) * warps) / step * step;
        const uint32_t ur = (uint32_t)std::min<uint64_t>(std::max<uint64_t>(want, step), 8 * step);
        uint32_t unit = 0, ng = 0;
        for (int m = 1; m <= kMaxBucket; ++m) {
            uint32_t row0 = 0;
            for (int k = 0; k < 32; ++k) {
                const uint32_t n = gcount[m * 32 + k];
                if (!n) continue;
                UnitGroup g;
                g.unit_start = unit;
                g.row0 = row0;
                g.nrows = n;
                g.m = (uint16_t)m;
                g.shift = (uint16_t)k;
                host[(size_t)wi * kMaxGroups + ng++] = g;
                unit += (n + ur - 1) / ur;
                row0 += n;
            }
        }
        e->n_groups[view][wi] = ng;
        e->n_units[view][wi] = unit;
        e->unit_rows[view][wi] = ur;
    }
    TRY(ensure_dev(e, e->d_groups[view], host.size()));
    CUDA_TRY(e, cudaMemcpyAsync(e->d_groups[view].ptr, host.data(), host.size() * sizeof(UnitGroup),
                                cudaMemcpyHostToDevice, e->stream));
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));   // `host` goes out of scope
    return GFE_OK;
}

// Multiplicity-compressed copy of the scoring view (see kernels.cu); kept only when it shrinks the
// bucketed rows by at least a quarter.
static int build_dedup_view(gfe_engine *e, const PairDict &pd) {
    e->have_dedup = false;
    const uint64_t short_rows = (uint64_t)e->bk.row_start[kLongBucket] - e->bk.row_start[1];
    if (!e->dedup_enabled || !e->have_dict || short_rows < 1024) return GFE_OK;
    cudaStream_t st = e->stream;
    uint32_t cap = 1024;
    while ((uint64_t)cap < 2 * short_rows && cap < (1u << 31)) cap <<= 1;
    TRY(ensure_dev(e, e->d_dedup, cap));
    CUDA_TRY(e, cudaMemsetAsync(e->d_dedup.ptr, 0xFF, (size_t)cap * sizeof(DedupSlot), st));
    launch_dedup_insert(e->d_off, e->d_steps, e->d_node2local.ptr, e->A, e->d_dedup.ptr, cap, st);
    const size_t misc_words = (size_t)3 * (kDedupStats + 2);
    TRY(ensure_dev(e, e->d_dedup_misc, misc_words));
    CUDA_TRY(e, cudaMemsetAsync(e->d_dedup_misc.ptr, 0, misc_words * 4, st));
    uint32_t *d_gcount = e->d_dedup_misc.ptr, *d_cbase = d_gcount + (kDedupStats + 2), *d_cur = d_cbase + (kDedupStats + 2);
    launch_dedup_count(e->d_dedup.ptr, cap, e->d_off, d_gcount, st);
    std::vector<uint32_t> gcount(kDedupStats + 2);
    CUDA_TRY(e, cudaMemcpyAsync(gcount.data(), d_gcount, gcount.size() * 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(e, cudaStreamSynchronize(st));
    e->kernel_launches += 3;
    uint64_t view_rows = 0;
    for (uint32_t i = 32; i < kDedupStats; ++i) view_rows += gcount[i];
    e->dedup_rows = short_rows;
    e->dedup_view_rows = view_rows;
    e->dedup_distinct = gcount[kDedupStats];
    if (gcount[kDedupStats + 1] != short_rows) {
        set_error(e, "internal: multiplicity table holds %u of %llu rows", gcount[kDedupStats + 1],
                  (unsigned long long)short_rows);
        return GFE_ERR_CUDA;
    }
    if (view_rows * 4 > short_rows * 3) return GFE_OK;        // not worth a second copy of the view
    // geometry of the compressed view: the same empty and long rows, class-sorted short rows
    uint32_t hist[kNumBuckets + 2] = {};
    hist[0] = e->bk.count[0];
    hist[kLongBucket] = e->bk.count[kLongBucket];
    std::vector<uint32_t> cbase(kDedupStats + 2, 0);
    for (int m = 1; m <= kMaxBucket; ++m) {
        uint32_t row = 0;
        for (int k = 0; k < 32; ++k) {
            cbase[m * 32 + k] = row;
            row += gcount[m * 32 + k];
        }
        hist[m] = row;
    }
    // steps of the long rows: what the plain view holds between its long base and its end
    hist[kNumBuckets + 1] = (uint32_t)(e->bstep_elems - 8 - e->bk.step_base[kLongBucket]);
    uint64_t elems = 0;
    layout_buckets(hist, e->bk2, &elems);
    TRY(ensure_dev(e, e->d_bstep2, elems));
    TRY(ensure_dev(e, e->d_ldesc2, std::max<uint32_t>(hist[kLongBucket], 1)));
    TRY(ensure_dev(e, e->d_rshift2, (size_t)e->bk2.row_start[kNumBuckets] + 1));
    CUDA_TRY(e, cudaMemsetAsync(e->d_rshift2.ptr, 0, (size_t)e->bk2.row_start[kNumBuckets] + 1, st));
    CUDA_TRY(e, cudaMemcpyAsync(d_cbase, cbase.data(), cbase.size() * 4, cudaMemcpyHostToDevice, st));
    e->bk2.bstep = e->d_bstep2.ptr;
    e->bk2.ldesc = e->d_ldesc2.ptr;
    e->bk2.perm = nullptr;                 // the search kernels never ask which alignment a row was
    const uint32_t np = e->n_pairs, bias = geom_pair_base(e->L);
    launch_dedup_scatter(e->d_dedup.ptr, cap, e->d_off, e->d_steps, e->d_node2local.ptr, e->bk2, e->d_bstep2.ptr, d_cbase,
                         d_cur, e->d_rshift2.ptr, pd, np, bias, st);
    if (hist[kLongBucket]) {               // long rows: copied as they are
        TRY(ensure_dev(e, e->d_cursors, kNumBuckets + 1));
        CUDA_TRY(e, cudaMemsetAsync(e->d_cursors.ptr, 0, (kNumBuckets + 1) * sizeof(uint32_t), st));
        launch_bucket_scatter(e->d_off, e->d_steps, e->d_node2local.ptr, e->A, e->bk2, e->d_bstep2.ptr, e->d_ldesc2.ptr,
                              nullptr, e->d_cursors.ptr, pd, np, bias, st, true);
        e->kernel_launches += 1;
    }
    e->kernel_launches += 1;
    CUDA_TRY(e, cudaGetLastError());
    CUDA_TRY(e, cudaStreamSynchronize(st));    // cbase goes out of scope
    TRY(build_group_tables(e, 1, gcount.data(), view_rows));
    e->have_dedup = true;
    return GFE_OK;
}

// Length histogram of the alignment table on the device -> bucket geometry (host copy in e->bk).
static int build_bucket_geometry(gfe_engine *e) {
    TRY(ensure_dev(e, e->d_hist, kNumBuckets + 2));
    CUDA_TRY(e, cudaMemsetAsync(e->d_hist.ptr, 0, (kNumBuckets + 2) * sizeof(uint32_t), e->stream));
    launch_length_histogram(e->d_off, e->A, e->d_hist.ptr, e->stream);
    e->kernel_launches += e->A ? 1 : 0;
    uint32_t hist[kNumBuckets + 2];
    CUDA_TRY(e, cudaMemcpyAsync(hist, e->d_hist.ptr, sizeof(hist), cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    uint64_t elems = 0;
    layout_buckets(hist, e->bk, &elems);
    if (elems >= 0xFFFFFFFFull) {
        set_error(e, "alignment shard has too many steps for 32-bit element offsets; shard it");
        return GFE_ERR_LIMIT;
    }
    e->bstep_elems = elems;
    e->max_m = hist[kNumBuckets];
    return GFE_OK;
}

int gfe_set_alignments(gfe_engine *e, const uint32_t *offsets, const uint32_t *steps, uint64_t A,
                       uint32_t n_nodes) {
    if (!e) return GFE_ERR_ARG;
    if (A >= 0xFFFFFFFFull) {
        set_error(e, "more than 2^32-2 alignments per engine are not supported; shard them");
        return GFE_ERR_LIMIT;
    }
    uint32_t mx = 0;
    TRY(validate_csr_host(e, offsets, steps, A, n_nodes, &mx));
    CUDA_TRY(e, cudaSetDevice(e->device));
    const uint64_t S = offsets[A];
    TRY(ensure_dev(e, e->own_off, A + 1));
    TRY(ensure_dev(e, e->own_steps, std::max<uint64_t>(S, 1) + 4));
    CUDA_TRY(e, cudaMemcpyAsync(e->own_off.ptr, offsets, (A + 1) * sizeof(uint32_t),
                                cudaMemcpyHostToDevice, e->stream));
    if (S)
        CUDA_TRY(e, cudaMemcpyAsync(e->own_steps.ptr, steps, S * sizeof(uint32_t),
                                    cudaMemcpyHostToDevice, e->stream));
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    e->h2d_bytes += (A + 1 + S) * sizeof(uint32_t);
    e->d_off = e->own_off.ptr;
    e->d_steps = e->own_steps.ptr;
    e->A = A;
    e->S = S;
    e->n_nodes = n_nodes;
    e->max_m = mx;
    e->have_alignments = true;
    e->have_rowstart = false;
    e->have_universe = false;
    return build_bucket_geometry(e);
}

int gfe_set_alignments_device(gfe_engine *e, const uint32_t *d_offsets, const uint32_t *d_steps,
                              uint64_t A, uint64_t total_steps, uint32_t n_nodes) {
    if (!e || !d_offsets || (total_steps && !d_steps)) return GFE_ERR_ARG;
    if (A >= 0xFFFFFFFFull || total_steps >= 0xFFFFFFFFull) {
        set_error(e, "shard too large for 32-bit CSR offsets");
        return GFE_ERR_LIMIT;
    }
    e->d_off = d_offsets;
    e->d_steps = d_steps;
    e->A = A;
    e->S = total_steps;
    e->n_nodes = n_nodes;
    e->have_alignments = true;
    e->have_rowstart = false;
    e->have_universe = false;
    CUDA_TRY(e, cudaSetDevice(e->device));
    return build_bucket_geometry(e);   // also finds the longest alignment
}

}  // extern "C"

// with_dict = false: evalPath's view (position-table kernels only): no pair dictionary pass over the table
static int set_universe_impl(gfe_engine *e, const uint32_t *nodes, uint32_t n_universe, bool with_dict) {
    if (!e || (n_universe && !nodes)) return GFE_ERR_ARG;
    if (!e->have_alignments) {
        set_error(e, "gfe_set_universe: call gfe_set_alignments first");
        return GFE_ERR_STATE;
    }
    if (n_universe > GFE_MAX_UNIVERSE) {
        set_error(e, "universe of %u nodes exceeds the %u the 16-bit local ids can name", n_universe,
                  GFE_MAX_UNIVERSE);
        return GFE_ERR_LIMIT;
    }
    CUDA_TRY(e, cudaSetDevice(e->device));
    e->node2local.assign(e->n_nodes, n_universe);
    uint32_t L = 0;
    for (uint32_t i = 0; i < n_universe; ++i) {
        if (nodes[i] >= e->n_nodes) {
            set_error(e, "universe node %u >= n_nodes %u", nodes[i], e->n_nodes);
            return GFE_ERR_ARG;
        }
        if (e->node2local[nodes[i]] == n_universe) e->node2local[nodes[i]] = L++;
    }
    // duplicates in `nodes` leave local ids [L, n_universe) unused; the outside id stays n_universe
    e->L = n_universe;
    TRY(ensure_dev(e, e->d_node2local, std::max<uint32_t>(e->n_nodes, 1)));
    if (e->n_nodes)
        CUDA_TRY(e, cudaMemcpyAsync(e->d_node2local.ptr, e->node2local.data(),
                                    (size_t)e->n_nodes * sizeof(uint32_t), cudaMemcpyHostToDevice,
                                    e->stream));
    TRY(ensure_dev(e, e->d_bstep, e->bstep_elems));
    TRY(ensure_dev(e, e->d_ldesc, std::max<uint32_t>(e->bk.count[kLongBucket], 1)));
    TRY(ensure_dev(e, e->d_perm, std::max<uint64_t>(e->A, 1)));
    TRY(ensure_dev(e, e->d_cursors, kNumBuckets + 1));
    CUDA_TRY(e, cudaMemsetAsync(e->d_cursors.ptr, 0, (kNumBuckets + 1) * sizeof(uint32_t), e->stream));
    e->bk.bstep = e->d_bstep.ptr;
    e->bk.ldesc = e->d_ldesc.ptr;
    e->bk.perm = e->d_perm.ptr;
    // pair dictionary: distinct in-universe consecutive pairs of the alignments.  It must leave
    // room for the mask table of at least 32 paths in shared memory; otherwise (huge universes)
    // the engine keeps to the position-table kernel.
    PairDict pd{nullptr, nullptr, nullptr};
    e->have_dict = false;
    e->n_pairs = 0;
    {
        const int64_t word_budget = (int64_t)(kBitsetSmemBudget / 4) - 64 - 8;
        int64_t max_pairs = (word_budget - 2 * ((int64_t)e->L + 1) - (kMaxBucket + 2) - 2 - 3 - 2 * (int64_t)kBloom) / 2;
        max_pairs = std::min<int64_t>(max_pairs, 32766);
        if (max_pairs >= 1 && !e->force_legacy && with_dict) {
            TRY(ensure_dev(e, e->d_pair_keys, kPairHashCap));
            TRY(ensure_dev(e, e->d_pair_ids, kPairHashCap));
            TRY(ensure_dev(e, e->d_pair_ctr, 4));
            CUDA_TRY(e, cudaMemsetAsync(e->d_pair_keys.ptr, 0xFF, kPairHashCap * sizeof(uint32_t), e->stream));
            CUDA_TRY(e, cudaMemsetAsync(e->d_pair_ctr.ptr, 0, 4 * sizeof(uint32_t), e->stream));
            pd = PairDict{e->d_pair_keys.ptr, e->d_pair_ids.ptr, e->d_pair_ctr.ptr};
            launch_pair_collect(e->d_off, e->d_steps, e->d_node2local.ptr, e->A, e->L,
                                (uint32_t)max_pairs, pd, e->stream);
            launch_pair_assign(pd, e->stream);
            e->kernel_launches += (e->A ? 1 : 0) + 1;
            uint32_t ctr[4];
            CUDA_TRY(e, cudaMemcpyAsync(ctr, e->d_pair_ctr.ptr, sizeof(ctr), cudaMemcpyDeviceToHost, e->stream));
            CUDA_TRY(e, cudaStreamSynchronize(e->stream));
            e->d2h_bytes += sizeof(ctr);
            if (ctr[2] == 0 && ctr[0] <= (uint32_t)max_pairs) {
                e->have_dict = true;
                e->n_pairs = ctr[0];
            }
        }
    }
    launch_bucket_scatter(e->d_off, e->d_steps, e->d_node2local.ptr, e->A, e->bk, e->d_bstep.ptr,
                          e->d_ldesc.ptr, e->d_perm.ptr, e->d_cursors.ptr, pd,
                          e->have_dict ? e->n_pairs : 0xFFFFFFFFu,
                          e->have_dict ? geom_pair_base(e->L) : 0u, e->stream);
    e->kernel_launches += e->A ? 1 : 0;
    CUDA_TRY(e, cudaGetLastError());
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    e->h2d_bytes += (size_t)e->n_nodes * sizeof(uint32_t);
    e->tab_stride = ((e->L + 1 + 7) / 8) * 8;
    e->have_dedup = false;
    if (e->have_dict) {
        std::vector<uint32_t> gcount(kDedupStats + 2, 0);
        for (int m = 1; m <= kMaxBucket; ++m) gcount[m * 32] = e->bk.count[m];
        TRY(build_group_tables(e, 0, gcount.data(), e->A));
        TRY(build_dedup_view(e, pd));
    }
    e->have_universe = true;
    e->have_index = false;
    return GFE_OK;
}

extern "C" {

int gfe_set_universe(gfe_engine *e, const uint32_t *nodes, uint32_t n_universe) {
    return set_universe_impl(e, nodes, n_universe, true);
}

// ---- search scoring -----------------------------------------------------------------------------

static uint32_t pick_chunk(const gfe_engine *e, uint32_t P) {
    const size_t row = (size_t)e->tab_stride * 2 + 12;
    size_t budget = 48 * 1024;               // keeps several CTAs per SM for small tables
    uint32_t c = (uint32_t)std::max<size_t>(1, budget / row);
    c = std::min<uint32_t>(c, 64);
    return std::min<uint32_t>(c, std::max<uint32_t>(P, 1));
}

// Enqueue table build + scoring (+ exact fallbacks) for paths already on the device in local ids.
// d_raw_steps != nullptr: d_pstep is the engine's scratch the prologue fills from the raw steps.
static int enqueue_score(gfe_engine *e, const uint32_t *d_poff, const uint16_t *d_pstep, uint32_t P,
                         uint32_t total_steps, int filter, uint32_t *d_counts, uint32_t *d_work_count,
                         const uint32_t *d_raw_steps, uint32_t *d_bad_step, uint32_t *d_sticky,
                         cudaStream_t st) {
    TRY(ensure_dev(e, e->d_ptab, (size_t)P * e->tab_stride));
    TRY(ensure_dev(e, e->d_pnext, (size_t)total_steps + 8));
    // exact-DP worklist: sized for 1 ambiguous pair in 128 checks (cfg3 measures ~1 in 2500)
    {
        const uint64_t want = std::min<uint64_t>(std::max<uint64_t>((uint64_t)P * e->A / 128, 1u << 20),
                                                 1ull << 27);
        if (want > e->work_cap) e->work_cap = (size_t)want;
        if (e->d_work.cap < e->work_cap) {
            CUDA_TRY(e, cudaStreamSynchronize(st));
            TRY(ensure_dev(e, e->d_work, e->work_cap));
        }
    }
    // one prologue launch clears every accumulator of the batch (and localizes raw device paths)
    PrologueArgs pro{};
    pro.zero[0] = d_counts;
    pro.zero_words[0] = P * 2;
    pro.zero[1] = d_work_count;
    pro.zero_words[1] = kWorkCounters;
    pro.psteps = d_raw_steps;
    pro.map = e->d_node2local.ptr;
    pro.n_nodes = e->n_nodes;
    pro.L = e->L;
    pro.out = const_cast<uint16_t *>(d_pstep);
    pro.total = total_steps;
    pro.err = d_bad_step;
    pro.sticky = d_sticky;
    // full scans read the multiplicity-compressed view when the universe has one
    const int view = (e->have_dedup && e->dedup_enabled && e->have_dict && !e->force_legacy) ? 1 : 0;
    const Buckets &vbk = view ? e->bk2 : e->bk;
    ScoreArgs a;
    a.bk = vbk;
    a.rshift = view ? e->d_rshift2.ptr : nullptr;
    a.poff = d_poff;
    a.pstep = d_pstep;
    a.ptab = e->d_ptab.ptr;
    a.pnext = e->d_pnext.ptr;
    a.tab_stride = e->tab_stride;
    a.P = P;
    a.chunk = pick_chunk(e, P);
    a.filter = filter;
    a.counts = d_counts;
    a.work = e->d_work.ptr;
    a.work_count = d_work_count;
    a.work_cap = (uint32_t)std::min<size_t>(e->d_work.cap, 0xFFFFFFFFu);
    a.err = d_bad_step;
    const PairDict pd{e->d_pair_keys.ptr, e->d_pair_ids.ptr, e->d_pair_ctr.ptr};
    const uint32_t long_lo = vbk.tile_start[kLongBucket], long_hi = vbk.tile_start[kNumBuckets];
    if (e->have_dict && !e->force_legacy) {
        // ---- bitset kernel: 32*W paths per CTA pass, mask table in shared memory ----------------
        MaskGeom mg;
        mg.L = e->L;
        // BT | node plane {N, F+|F-} (2 words per (node, w); {F+, F-} is outside the table) | LM | pairs
        mg.bt_base = 0;
        mg.node_base = geom_node_base();
        mg.lm_base = geom_lm_base(e->L);
        mg.pair_base = geom_pair_base(e->L);   // == the bias bucket_scatter added to the rows' pair ids
        mg.n_pairs = e->n_pairs;
        mg.entries = (mg.pair_base + 2 * e->n_pairs + 2 + 3) & ~3u;
        uint32_t W = 1;
        while (W < 32 && 32 * W < P) W <<= 1;
        while (W > 1 && (size_t)mg.entries * W * 4 + 64 * W * 4 + 16 > kBitsetSmemBudget) W >>= 1;
        mg.W = W;
        const uint32_t chunks = (P + 32 * W - 1) / (32 * W);
        const size_t table_words = (size_t)chunks * mg.entries * W;
        const size_t fplane_words = (size_t)chunks * 2 * (e->L + 1) * W;   // {F+, F-}, second look only
        const size_t mask_words = table_words + fplane_words;
        {
            const uint64_t want = std::min<uint64_t>(
                std::max<uint64_t>((uint64_t)P * e->A / 64, 1u << 20), 1ull << 28);
            if (want > e->slow_cap) e->slow_cap = (size_t)want;
            if (e->d_slow.cap < e->slow_cap || e->d_masks.cap < mask_words ||
                e->d_unit_ctr.cap < chunks)
                CUDA_TRY(e, cudaStreamSynchronize(st));
            TRY(ensure_dev(e, e->d_slow, e->slow_cap));
            TRY(ensure_dev(e, e->d_masks, mask_words));
            TRY(ensure_dev(e, e->d_unit_ctr, chunks));
        }
        pro.zero[2] = e->d_masks.ptr;
        if (mask_words > 0xFFFFFFFFull) {   // beyond the prologue's 32-bit word counts
            CUDA_TRY(e, cudaMemsetAsync(e->d_masks.ptr, 0, mask_words * 4, st));
            pro.zero[2] = nullptr;
        }
        pro.zero_words[2] = pro.zero[2] ? (uint32_t)mask_words : 0;
        pro.zero[3] = e->d_unit_ctr.ptr;
        pro.zero_words[3] = chunks;
        launch_prologue(pro, e->sm_count, st);
        launch_build_tables(d_poff, d_pstep, P, e->L, e->tab_stride, e->d_ptab.ptr, e->d_pnext.ptr,
                            e->d_masks.ptr, e->d_masks.ptr + table_words, mg, pd, st);
        BitsetArgs b;
        b.bk = vbk;
        b.mg = mg;
        b.masks = e->d_masks.ptr;
        b.fplane = e->d_masks.ptr + table_words;
        b.P = P;
        b.filter = filter;
        b.counts = d_counts;
        b.unit_ctr = e->d_unit_ctr.ptr;
        b.slow = e->d_slow.ptr;
        b.work_count = d_work_count;
        b.slow_cap = (uint32_t)std::min<size_t>(e->d_slow.cap, 0xFFFFFFFFu);
        // scheduling units: a multiple of four warp steps (32/W rows each), small enough that every warp
        // of the grid gets several units; the group tables were built with the universe (build_group_tables)
        {
            int wi = 0;
            while ((1u << wi) < W) ++wi;
            b.unit_rows = e->unit_rows[view][wi];
            b.groups = view ? e->d_groups[view].ptr + (size_t)wi * kMaxGroups : nullptr;
            b.n_groups = e->n_groups[view][wi];
            b.n_units = e->n_units[view][wi];
            uint32_t unit = 0;       // plain view: the buckets' first units (the kernel maps units arithmetically)
            b.unit_start[0] = 0;
            for (int k = 1; k <= kMaxBucket; ++k) {
                b.unit_start[k] = unit;
                unit += (vbk.count[k] + b.unit_rows - 1) / b.unit_rows;
            }
            b.unit_start[kMaxBucket + 1] = unit;
            if (!view) b.n_units = unit;
        }
        if (e->timing) {
            TRY(next_event_pair(e));
            CUDA_TRY(e, cudaEventRecord(e->ev0, st));
        }
        launch_score_bitset(b, e->sm_count, st);
        if (e->timing) CUDA_TRY(e, cudaEventRecord(e->ev1, st));
        e->kernel_launches += 3;   // prologue, tables, bitset scoring
        if (long_hi > long_lo) {   // alignments longer than kMaxBucket steps: position-table kernel
            launch_score_paths(a, e->sm_count, long_lo, long_hi, false, st);
            e->kernel_launches += 1;
        }
        if (e->A) {
            launch_slow_worklist(a, e->d_slow.ptr, d_work_count + 2, b.slow_cap, st);
            e->kernel_launches += 1;
        }
    } else {
        MaskGeom mg{};
        launch_prologue(pro, e->sm_count, st);
        launch_build_tables(d_poff, d_pstep, P, e->L, e->tab_stride, e->d_ptab.ptr, e->d_pnext.ptr,
                            nullptr, nullptr, mg, pd, st);
        if (e->timing) {
            TRY(next_event_pair(e));
            CUDA_TRY(e, cudaEventRecord(e->ev0, st));
        }
        launch_score_paths(a, e->sm_count, 0, long_hi, true, st);
        if (e->timing) CUDA_TRY(e, cudaEventRecord(e->ev1, st));
        e->kernel_launches += 3;
    }
    if (e->A) {
        launch_exact_worklist(a, st);
        e->kernel_launches += 1;
        if (e->max_m > (uint32_t)kExactRegSteps) {
            launch_exact_worklist_long(a, std::min<uint32_t>(e->max_m, GFE_MAX_STEPS), st);
            e->kernel_launches += 1;
        }
    }
    CUDA_TRY(e, cudaGetLastError());
    return GFE_OK;
}

}  // extern "C"

// One exchange buffer of the sharded path as seen from this rank (gfe_exchange_create).
struct gfe_exchange {
    uint64_t bases[GFE_PEER_MAX];
    int world = 1, rank = 0;
    uint32_t n_cap = 0, epoch = 0;
    // false: every rank holds an alignment SHARD and scores every candidate (records are summed);
    // true:  every rank holds the WHOLE table and scores its share of the candidates — the speculation
    //        subtrees are dealt round-robin — and the sum of the records is an all-gather
    bool batch_shard = false;
};

// gfe_score_paths, and with `x` its sharded form: the kernels accumulate this rank's records
// straight into the exchange buffer, the status word rides along as slot 2P, the peer kernel sums
// all shards into the result block and ONE device->host copy brings back header + summed counts.
static int score_paths_impl(gfe_engine *e, gfe_exchange *x, const uint32_t *path_offsets,
                            const uint32_t *path_steps, uint32_t P, int filter, uint32_t *counts) {
    if (!e || !path_offsets || !counts) return GFE_ERR_ARG;
    if (x && x->batch_shard) x = nullptr;   // replicated table: the full scan is local (same result on every rank)
    if (!e->have_alignments || !e->have_universe) {
        set_error(e, "gfe_score_paths: alignments and universe must be set first");
        return GFE_ERR_STATE;
    }
    if (P == 0) return GFE_OK;
    if (!path_steps || path_offsets[0] != 0) {
        set_error(e, "gfe_score_paths: malformed path CSR");
        return GFE_ERR_ARG;
    }
    if (!filter && e->max_m > GFE_MAX_STEPS) {
        set_error(e, "gfe_score_paths: an alignment has %u steps; the reference's MAX_N allows %u (unfiltered scoring)",
                  e->max_m, GFE_MAX_STEPS);
        return GFE_ERR_LIMIT;
    }
    const uint32_t total = path_offsets[P];
    for (uint32_t p = 0; p < P; ++p) {
        if (path_offsets[p + 1] <= path_offsets[p]) {
            set_error(e, "gfe_score_paths: path %u is empty or offsets not monotone", p);
            return GFE_ERR_ARG;
        }
        if (path_offsets[p + 1] - path_offsets[p] > GFE_MAX_STEPS) {
            set_error(e, "gfe_score_paths: path %u has %u steps; the reference's MAX_N allows %u",
                      p, path_offsets[p + 1] - path_offsets[p], GFE_MAX_STEPS);
            return GFE_ERR_LIMIT;
        }
    }
    CUDA_TRY(e, cudaSetDevice(e->device));
    // staging: [P+1 offsets u32][total raw steps u32]; the batch prologue maps them to universe-local
    // 16-bit steps on the device (a step outside the universe raises the flag word read back below)
    const size_t off_bytes = ((size_t)(P + 1) * 4 + 15) & ~(size_t)15;
    const size_t bytes = off_bytes + (size_t)total * 4;
    // result block: [8 work counters][bad-step flag][pad to 16 words][P x {bad, good}]: ONE device->host
    // copy per batch.  The flag word is zero between calls (cleared on allocation and after an error).
    constexpr size_t kHdr = 16;
    static_assert(kWorkCounters + 1 <= kHdr, "result header");
    TRY(ensure_pinned(e, e->h_stage, bytes));
    TRY(ensure_dev(e, e->d_stage, bytes));
    TRY(ensure_dev(e, e->d_plocal, (size_t)total + 8));
    TRY(ensure_pinned(e, e->h_counts, (size_t)P * 2 + kHdr + 1));
    {
        const uint32_t *before = e->d_counts.ptr;
        TRY(ensure_dev(e, e->d_counts, (size_t)P * 2 + kHdr + 1));
        if (e->d_counts.ptr != before)
            CUDA_TRY(e, cudaMemsetAsync(e->d_counts.ptr, 0, kHdr * sizeof(uint32_t), e->stream));
    }
    memcpy(e->h_stage.ptr, path_offsets, (size_t)(P + 1) * 4);
    memcpy(e->h_stage.ptr + off_bytes, path_steps, (size_t)total * 4);
    CUDA_TRY(e, cudaMemcpyAsync(e->d_stage.ptr, e->h_stage.ptr, bytes, cudaMemcpyHostToDevice,
                                e->stream));
    const uint32_t *d_poff = reinterpret_cast<const uint32_t *>(e->d_stage.ptr);
    const uint32_t *d_raw = reinterpret_cast<const uint32_t *>(e->d_stage.ptr + off_bytes);
    uint32_t *d_hdr = e->d_counts.ptr, *d_cnt = e->d_counts.ptr + kHdr;
    if (x) {
        if (2 * (uint64_t)P + 1 > x->n_cap) {
            set_error(e, "gfe_score_paths_sharded: %u paths exceed the exchange buffer (%u slots)", P, x->n_cap);
            return GFE_ERR_ARG;
        }
        uint32_t *d_mine = reinterpret_cast<uint32_t *>(x->bases[x->rank]) + (size_t)x->n_cap * ((x->epoch + 1) & 1u);
        TRY(enqueue_score(e, d_poff, e->d_plocal.ptr, P, total, filter, d_mine, d_hdr, d_raw,
                          d_hdr + kWorkCounters, nullptr, e->stream));
        // slot 2P: this rank's flags (bad step | overflow).  Its sum tells every rank alike to repeat the
        // batch, so the ranks stay in lock step.  The header keeps the local counters for the retry below.
        launch_flag_word(d_hdr + kWorkCounters, d_hdr, d_mine + 2 * (size_t)P, e->stream);
        e->kernel_launches += 2;
        x->epoch += 1;
        const int prc = gfe_peer_sum_u32(x->bases, x->world, x->rank, x->n_cap, 2 * P + 1, x->epoch, d_cnt, e->stream);
        if (prc != GFE_OK) {
            set_error(e, "gfe_score_paths_sharded: peer sum launch failed");
            return prc;
        }
    } else {
        TRY(enqueue_score(e, d_poff, e->d_plocal.ptr, P, total, filter, d_cnt, d_hdr, d_raw, d_hdr + kWorkCounters,
                          nullptr, e->stream));
    }
    const size_t res_words = (size_t)P * 2 + kHdr + (x ? 1 : 0);
    CUDA_TRY(e, cudaMemcpyAsync(e->h_counts.ptr, e->d_counts.ptr, res_words * sizeof(uint32_t),
                                cudaMemcpyDeviceToHost, e->stream));
    uint32_t wc[kWorkCounters] = {};
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    for (int i = 0; i < kWorkCounters; ++i) wc[i] = e->h_counts.ptr[i];
    if (e->timing && e->ev0) cudaEventElapsedTime(&e->last_kernel_ms, e->ev0, e->ev1);
    e->h2d_bytes += bytes;
    e->d2h_bytes += res_words * sizeof(uint32_t);
    const uint32_t any_flags = x ? e->h_counts.ptr[kHdr + 2 * (size_t)P] : 0u;   // summed over the ranks
    if (x && any_flags == GFE_PEER_TIMEOUT) {
        set_error(e, "gfe_score_paths_sharded: a peer never reached epoch %u (GFE_PEER_TIMEOUT_MS)", x->epoch);
        return GFE_ERR_CUDA;
    }
    if (x && (any_flags >> 8) && !e->h_counts.ptr[kWorkCounters]) {
        // another rank's shard raised an error this rank's did not (flag_word_kernel: byte 2 = over-long
        // alignment, byte 1 = bad step): return the same class on every rank, nobody retries alone
        if ((any_flags >> 16) & 0xFFu) {
            set_error(e, "gfe_score_paths_sharded: on another rank's shard an alignment of more than %u steps passes "
                         "the id filter of a candidate path", GFE_MAX_STEPS);
            return GFE_ERR_LIMIT;
        }
        set_error(e, "gfe_score_paths_sharded: another rank saw a step outside the universe");
        return GFE_ERR_ARG;
    }
    if (x && any_flags && !e->h_counts.ptr[kWorkCounters]) {
        // some rank overflowed a list (or all saw a bad step, handled below): every rank repeats the
        // batch; the ranks that overflowed grow their lists first
        if (wc[3] && e->d_slow.cap < (1ull << 28)) {
            e->slow_cap = std::min<size_t>(std::max<size_t>(e->d_slow.cap * 4, (size_t)wc[2] + 1024), (size_t)1 << 28);
            TRY(ensure_dev(e, e->d_slow, e->slow_cap));
        } else if (wc[3]) {
            set_error(e, "gfe_score_paths_sharded: second-look list overflow at its maximum size");
            return GFE_ERR_LIMIT;
        }
        if (wc[1]) {
            e->work_cap = std::max<size_t>(e->d_work.cap * 2, (size_t)wc[0] + 1024);
            TRY(ensure_dev(e, e->d_work, e->work_cap));
        }
        if (++e->shard_retries > 8) {
            set_error(e, "gfe_score_paths_sharded: batch still overflows after 8 collective retries");
            return GFE_ERR_LIMIT;
        }
        const int rc = score_paths_impl(e, x, path_offsets, path_steps, P, filter, counts);
        e->shard_retries = 0;
        return rc;
    }
    if (e->h_counts.ptr[kWorkCounters] & 4u) {
        CUDA_TRY(e, cudaMemset(e->d_counts.ptr + kWorkCounters, 0, sizeof(uint32_t)));
        set_error(e, "gfe_score_paths: an alignment of more than %u steps passes the id filter of a candidate path; "
                     "the reference's dp[MAX_N][MAX_N] cannot hold it", GFE_MAX_STEPS);
        return GFE_ERR_LIMIT;
    }
    if (e->h_counts.ptr[kWorkCounters]) {
        // a step outside the universe: name it the way the host check did, and clear the flag
        CUDA_TRY(e, cudaMemset(e->d_counts.ptr + kWorkCounters, 0, sizeof(uint32_t)));
        for (uint32_t i = 0; i < total; ++i) {
            const uint32_t node = path_steps[i] >> 1;
            if (node >= e->n_nodes || e->node2local[node] >= e->L) {
                set_error(e, "gfe_score_paths: step %u names node %u outside the universe", i, node);
                return GFE_ERR_ARG;
            }
        }
        set_error(e, "gfe_score_paths: a step names a node outside the universe");
        return GFE_ERR_ARG;
    }
    if (!x && wc[3]) {
        // the list of pairs the bitsets could not settle overflowed: grow it, or halve the batch
        if (e->d_slow.cap < (1ull << 28)) {
            e->slow_cap = std::min<size_t>(std::max<size_t>(e->d_slow.cap * 4, (size_t)wc[2] + 1024),
                                           (size_t)1 << 28);
            TRY(ensure_dev(e, e->d_slow, e->slow_cap));
            return score_paths_impl(e, nullptr, path_offsets, path_steps, P, filter, counts);
        }
        if (P < 2) {
            set_error(e, "gfe_score_paths: work list overflow on a single path");
            return GFE_ERR_LIMIT;
        }
        const uint32_t h = P / 2;
        std::vector<uint32_t> off2(path_offsets + h, path_offsets + P + 1);
        for (uint32_t &o : off2) o -= path_offsets[h];
        TRY(score_paths_impl(e, nullptr, path_offsets, path_steps, h, filter, counts));
        return score_paths_impl(e, nullptr, off2.data(), path_steps + path_offsets[h], P - h, filter,
                                counts + 2 * (size_t)h);
    }
    if (!x && wc[1]) {
        // the exact-DP worklist overflowed: grow it and redo the batch (never silently wrong)
        e->work_cap = std::max<size_t>(e->d_work.cap * 2, (size_t)wc[0] + 1024);
        TRY(ensure_dev(e, e->d_work, e->work_cap));
        return score_paths_impl(e, nullptr, path_offsets, path_steps, P, filter, counts);
    }
    e->exact_pairs += wc[0];
    e->slow_pairs += wc[2] - wc[4];   // reserved slots minus the holes
    memcpy(counts, e->h_counts.ptr + 16, (size_t)P * 2 * sizeof(uint32_t));
    return GFE_OK;
}

extern "C" {

int gfe_score_paths(gfe_engine *e, const uint32_t *path_offsets, const uint32_t *path_steps,
                    uint32_t P, int filter, uint32_t *counts) {
    return score_paths_impl(e, nullptr, path_offsets, path_steps, P, filter, counts);
}

int gfe_exchange_create(const uint64_t *peer_bases, int world, int rank, uint32_t n_cap, gfe_exchange **out) {
    if (!peer_bases || !out || world < 1 || world > GFE_PEER_MAX || rank < 0 || rank >= world || n_cap == 0)
        return GFE_ERR_ARG;
    gfe_exchange *x = new (std::nothrow) gfe_exchange();
    if (!x) return GFE_ERR_NOMEM;
    for (int q = 0; q < world; ++q) x->bases[q] = peer_bases[q];
    x->world = world;
    x->rank = rank;
    x->n_cap = n_cap;
    *out = x;
    return GFE_OK;
}

void gfe_exchange_destroy(gfe_exchange *x) { delete x; }

int gfe_exchange_set_batch_sharding(gfe_exchange *x, int on) {
    if (!x) return GFE_ERR_ARG;
    x->batch_shard = on != 0;
    return GFE_OK;
}

int gfe_score_paths_sharded(gfe_engine *e, gfe_exchange *x, const uint32_t *path_offsets,
                            const uint32_t *path_steps, uint32_t P, int filter, uint32_t *counts) {
    if (!x) return GFE_ERR_ARG;
    return score_paths_impl(e, x, path_offsets, path_steps, P, filter, counts);
}

}  // extern "C"

// Internal to the library (C++ linkage, used by search.cpp): do ALL ranks of the exchange have the
// incremental scoring path for their current universe?  Whether a rank has a pair dictionary depends on
// its own shard (distinct pairs vs the dictionary capacity, GFE_FORCE_LEGACY), and the incremental and
// the full-scan entries exchange records of different shapes under the same epoch counter — so the
// choice between them must be the same on every rank.  One word per rank (1 = "I cannot") summed by the
// peer kernel; collective: every rank calls it at the same point (start of gfe_search_sharded).
int gfe_internal_incremental_consensus(gfe_engine *e, gfe_exchange *x, int *all_have) {
    if (!e || !x || !all_have) return GFE_ERR_ARG;
    if (!e->have_alignments || !e->have_universe) {
        set_error(e, "gfe_search_sharded: alignments and universe must be set first");
        return GFE_ERR_STATE;
    }
    if (x->n_cap < 1) return GFE_ERR_ARG;
    TRY(ensure_pinned(e, e->h_counts, 2));
    TRY(ensure_dev(e, e->d_counts, 2));
    cudaStream_t st = e->stream;
    uint32_t *d_mine = reinterpret_cast<uint32_t *>(x->bases[x->rank]) + (size_t)x->n_cap * ((x->epoch + 1) & 1u);
    e->h_counts.ptr[0] = (e->have_dict && !e->force_legacy) ? 0u : 1u;
    CUDA_TRY(e, cudaMemcpyAsync(d_mine, e->h_counts.ptr, sizeof(uint32_t), cudaMemcpyHostToDevice, st));
    x->epoch += 1;
    const int prc = gfe_peer_sum_u32(x->bases, x->world, x->rank, x->n_cap, 1, x->epoch, e->d_counts.ptr, st);
    if (prc != GFE_OK) {
        set_error(e, "gfe_search_sharded: peer sum launch failed");
        return prc;
    }
    e->kernel_launches += 1;
    CUDA_TRY(e, cudaMemcpyAsync(e->h_counts.ptr + 1, e->d_counts.ptr, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(e, cudaStreamSynchronize(st));
    const uint32_t without = e->h_counts.ptr[1];
    if (without == GFE_PEER_TIMEOUT) {
        set_error(e, "gfe_search_sharded: a peer never reached epoch %u (GFE_PEER_TIMEOUT_MS)", x->epoch);
        return GFE_ERR_CUDA;
    }
    *all_have = without == 0 ? 1 : 0;
    return GFE_OK;
}

// ---- incremental search scoring (the ABI comment in include/gfalign_b200.h) ------------

static int ensure_index(gfe_engine *e, uint32_t root_local, cudaStream_t st) {
    if (e->have_index && e->inv_root == root_local) return GFE_OK;
    const uint32_t groups = (e->L + 1) * (uint32_t)kIdxLens;
    TRY(ensure_dev(e, e->d_inv_off, (size_t)groups + 1));
    TRY(ensure_dev(e, e->d_inv_cursor, (size_t)groups + 1));
    CUDA_TRY(e, cudaMemsetAsync(e->d_inv_cursor.ptr, 0, ((size_t)groups + 1) * 4, st));
    launch_index_count(e->bk, e->L, root_local, e->d_inv_cursor.ptr, st);
    // counters -> exclusive offsets; the same array becomes the fill cursors (each thread reads its
    // counter before it overwrites it)
    launch_index_scan(e->d_inv_cursor.ptr, groups, e->d_inv_off.ptr, e->d_inv_cursor.ptr, st);
    uint32_t total = 0;
    CUDA_TRY(e, cudaMemcpyAsync(&total, e->d_inv_off.ptr + groups, 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(e, cudaStreamSynchronize(st));
    TRY(ensure_dev(e, e->d_inv_rows, std::max<size_t>(total, 1)));
    launch_index_fill(e->bk, e->L, root_local, e->d_inv_cursor.ptr, e->d_inv_rows.ptr, st);
    CUDA_TRY(e, cudaGetLastError());
    e->kernel_launches += 3;
    e->inv_entries = total;
    e->inv_root = root_local;
    e->have_index = true;
    return GFE_OK;
}

namespace {
constexpr uint32_t kNoNode = 0xFFFFFFFFu;
constexpr size_t kHdrWords = 16;   // [8 work counters][bad-step flag][sticky overflow][pad]
inline size_t align16(size_t x) { return (x + 15) & ~(size_t)15; }
}  // namespace

static inline double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

static int score_incremental_impl(gfe_engine *e, gfe_exchange *x, const gfe_path_batch *b, int32_t *counts) {
    if (!e || !b || !counts) return GFE_ERR_ARG;
    const double t_begin = now_s();
    if (!e->have_alignments || !e->have_universe) {
        set_error(e, "gfe_score_paths_incremental: alignments and universe must be set first");
        return GFE_ERR_STATE;
    }
    if (!e->have_dict || e->force_legacy) {
        set_error(e, "gfe_score_paths_incremental: this universe has no pair dictionary (bitset path off)");
        return GFE_ERR_STATE;
    }
    const uint32_t P = b->P;
    if (P == 0) return GFE_OK;
    if (!b->n_bases || !b->base_offsets || !b->base_steps || !b->item_base || !b->suffix_offsets || !b->mode ||
        b->base_offsets[0] != 0 || b->suffix_offsets[0] != 0 || (b->suffix_offsets[P] && !b->suffix_steps)) {
        set_error(e, "gfe_score_paths_incremental: malformed batch");
        return GFE_ERR_ARG;
    }
    if (P >= (1u << 24)) {
        set_error(e, "gfe_score_paths_incremental: more than 2^24 paths in one batch");
        return GFE_ERR_LIMIT;
    }
    CUDA_TRY(e, cudaSetDevice(e->device));
    cudaStream_t st = e->stream;
    const uint32_t n_bases = b->n_bases, base_total = b->base_offsets[n_bases], suf_total = b->suffix_offsets[P];
    const uint32_t root = b->base_steps[0] >> 1;
    // ---- host: per-base node presence, per-item last node / revisit, chunk layout -----------------
    auto local_of = [&](uint32_t step) -> uint32_t {
        const uint32_t node = step >> 1;
        return node < e->n_nodes ? e->node2local[node] : e->L;
    };
    for (uint32_t k = 0; k < n_bases; ++k) {
        if (b->base_offsets[k + 1] <= b->base_offsets[k] || (b->base_steps[b->base_offsets[k]] >> 1) != root) {
            set_error(e, "gfe_score_paths_incremental: base path %u is empty or does not start at the root node", k);
            return GFE_ERR_ARG;
        }
    }
    if (local_of(b->base_steps[0]) >= e->L) {
        set_error(e, "gfe_score_paths_incremental: the root node is outside the universe");
        return GFE_ERR_ARG;
    }
    TRY(ensure_index(e, local_of(b->base_steps[0]), st));
    const uint32_t words_per_base = (e->L + 31) / 32;
    std::vector<uint32_t> &present = e->h_present;
    present.assign((size_t)n_bases * words_per_base, 0u);
    for (uint32_t k = 0; k < n_bases; ++k)
        for (uint32_t i = b->base_offsets[k]; i < b->base_offsets[k + 1]; ++i) {
            const uint32_t l = local_of(b->base_steps[i]);
            if (l < e->L) present[(size_t)k * words_per_base + (l >> 5)] |= 1u << (l & 31);
        }
    using Slot = gfe_engine::IncrSlot;               // parent = 1: the item's parent path (negative copy)
    std::vector<Slot> &slots = e->h_slots;
    slots.clear();
    slots.reserve((size_t)P + P / 8 + 8);
    uint32_t n_full = 0;
    const bool deal = x && x->batch_shard;     // candidate-batch sharding: this rank scores its subtrees only
    for (uint32_t i = 0; i < P; ++i) {
        const uint32_t kb = b->item_base[i];
        if (kb >= n_bases || b->suffix_offsets[i + 1] < b->suffix_offsets[i]) {
            set_error(e, "gfe_score_paths_incremental: item %u is malformed", i);
            return GFE_ERR_ARG;
        }
        // batch sharding: items of other ranks' subtrees are skipped only AFTER the argument checks, so that
        // every rank returns the same GFE_ERR_* for a bad batch instead of one failing and its peers waiting
        const bool mine = !(deal && (int)(kb % (uint32_t)x->world) != x->rank);
        const uint32_t s0 = b->suffix_offsets[i], ns = b->suffix_offsets[i + 1] - s0;
        const uint32_t nb = b->base_offsets[kb + 1] - b->base_offsets[kb];
        if ((uint64_t)nb + ns > GFE_MAX_STEPS) {
            set_error(e, "gfe_score_paths_incremental: item %u has %u steps; the reference's MAX_N allows %u", i,
                      nb + ns, GFE_MAX_STEPS);
            return GFE_ERR_LIMIT;
        }
        if (b->mode[i] == 0) {
            if (mine) {
                slots.push_back(Slot{kNoNode, i, 0});
                ++n_full;
            }
            continue;
        }
        if (nb + ns < (uint32_t)kMaxBucket + 1) {
            set_error(e, "gfe_score_paths_incremental: item %u: delta mode needs a parent of >= %d steps", i, kMaxBucket);
            return GFE_ERR_ARG;
        }
        const uint32_t last = ns ? b->suffix_steps[s0 + ns - 1] : b->base_steps[b->base_offsets[kb + 1] - 1];
        const uint32_t v = local_of(last);
        if (v >= e->L) {
            set_error(e, "gfe_score_paths_incremental: item %u ends in a node outside the universe", i);
            return GFE_ERR_ARG;
        }
        if (!mine) continue;
        // does the parent (the path without its last step) already visit v?
        bool revisit = false;
        if (ns) {
            revisit = (present[(size_t)kb * words_per_base + (v >> 5)] >> (v & 31)) & 1u;
            for (uint32_t j = 0; j + 1 < ns && !revisit; ++j) revisit = local_of(b->suffix_steps[s0 + j]) == v;
        } else {
            for (uint32_t j = b->base_offsets[kb]; j + 1 < b->base_offsets[kb + 1] && !revisit; ++j)
                revisit = local_of(b->base_steps[j]) == v;
        }
        slots.push_back(Slot{v, i, 0});
        if (revisit) slots.push_back(Slot{v, i, 1});
    }
    {   // stable counting sort by key (keys are local node ids; kNoNode sorts last)
        std::vector<uint32_t> &cnt = e->h_key_count;
        cnt.assign((size_t)e->L + 2, 0u);
        for (const Slot &sl : slots) cnt[std::min<uint32_t>(sl.key, e->L) + 1]++;
        for (size_t k = 1; k < cnt.size(); ++k) cnt[k] += cnt[k - 1];
        std::vector<Slot> &sorted = e->h_slots2;
        sorted.resize(slots.size());
        for (const Slot &sl : slots) sorted[cnt[std::min<uint32_t>(sl.key, e->L)]++] = sl;
        slots.swap(sorted);
    }
    // chunks of 32 slots, never mixing keys; device path d = 32 * chunk + position
    std::vector<uint32_t> &chunk_node = e->h_chunk_node;
    std::vector<uint32_t> &slot_dev = e->h_slot_dev;
    chunk_node.clear();
    slot_dev.assign(slots.size(), 0);
    {
        uint32_t in_chunk = 32, cur_key = 0;
        for (size_t k = 0; k < slots.size(); ++k) {
            if (in_chunk == 32 || slots[k].key != cur_key) {
                chunk_node.push_back(slots[k].key);
                cur_key = slots[k].key;
                in_chunk = 0;
            }
            slot_dev[k] = (uint32_t)(chunk_node.size() - 1) * 32u + in_chunk++;
        }
    }
    const uint32_t chunks = (uint32_t)chunk_node.size(), P_dev = chunks * 32u;
    // ---- staging: [base_off][base_steps][suffix_steps][desc inc][poff inc][chunk_node][desc full][poff full]
    const size_t o_boff = 0, o_bst = align16(o_boff + ((size_t)n_bases + 1) * 4);
    const size_t o_suf = align16(o_bst + (size_t)base_total * 4);
    const size_t o_desc = align16(o_suf + (size_t)suf_total * 4);
    const size_t o_poff = align16(o_desc + (size_t)P_dev * 16);
    const size_t o_cn = align16(o_poff + ((size_t)P_dev + 1) * 4);
    const size_t o_fdesc = align16(o_cn + (size_t)chunks * 4);
    const size_t o_fpoff = align16(o_fdesc + (size_t)n_full * 16);
    const size_t o_sitem = align16(o_fpoff + ((size_t)n_full + 1) * 4);          // batch sharding: slot -> item
    const size_t o_saux = align16(o_sitem + (deal ? (size_t)P_dev * 4 : 0));
    const size_t bytes = align16(o_saux + (deal ? (size_t)P_dev * 4 : 0));
    TRY(ensure_pinned(e, e->h_stage, bytes));
    TRY(ensure_dev(e, e->d_stage, bytes));
    unsigned char *hs = e->h_stage.ptr;
    memcpy(hs + o_boff, b->base_offsets, ((size_t)n_bases + 1) * 4);
    memcpy(hs + o_bst, b->base_steps, (size_t)base_total * 4);
    if (suf_total) memcpy(hs + o_suf, b->suffix_steps, (size_t)suf_total * 4);
    uint4 *h_desc = reinterpret_cast<uint4 *>(hs + o_desc);
    uint32_t *h_poff = reinterpret_cast<uint32_t *>(hs + o_poff);
    for (uint32_t d = 0; d < P_dev; ++d) h_desc[d] = make_uint4(kNoNode, 0, 0, 0);
    std::vector<uint32_t> &len_dev = e->h_len_dev;
    len_dev.assign(P_dev, 0);
    uint4 *h_fdesc = reinterpret_cast<uint4 *>(hs + o_fdesc);
    uint32_t *h_fpoff = reinterpret_cast<uint32_t *>(hs + o_fpoff);
    uint32_t nf = 0, ftotal = 0;
    std::vector<uint32_t> &full_of_item = e->h_full_of_item;
    full_of_item.assign(P, kNoNode);
    for (size_t k = 0; k < slots.size(); ++k) {
        const Slot &sl = slots[k];
        const uint32_t i = sl.item, kb = b->item_base[i];
        const uint32_t s0 = b->suffix_offsets[i], ns = b->suffix_offsets[i + 1] - s0;
        const uint32_t nb = b->base_offsets[kb + 1] - b->base_offsets[kb];
        h_desc[slot_dev[k]] = make_uint4(kb, s0, ns, sl.parent);
        len_dev[slot_dev[k]] = nb + ns - sl.parent;
        if (sl.key == kNoNode) {
            full_of_item[i] = nf;
            h_fdesc[nf] = make_uint4(kb, s0, ns, 0);
            h_fpoff[nf] = ftotal;
            ftotal += nb + ns;
            ++nf;
        }
    }
    h_fpoff[nf] = ftotal;
    if (deal) {
        uint32_t *h_sitem = reinterpret_cast<uint32_t *>(hs + o_sitem), *h_saux = reinterpret_cast<uint32_t *>(hs + o_saux);
        for (uint32_t d = 0; d < P_dev; ++d) h_sitem[d] = h_saux[d] = kNoNode;
        for (size_t k = 0; k < slots.size(); ++k) {
            const Slot &sl = slots[k];
            const uint32_t kind = sl.key == kNoNode ? 2u : (sl.parent ? 1u : 0u);
            h_sitem[slot_dev[k]] = sl.item | (kind << 30);
            h_saux[slot_dev[k]] = kind == 2u ? full_of_item[sl.item] : 0u;
        }
    }
    uint32_t dtotal = 0;
    for (uint32_t d = 0; d < P_dev; ++d) {
        h_poff[d] = dtotal;
        dtotal += len_dev[d];
    }
    h_poff[P_dev] = dtotal;
    memcpy(hs + o_cn, chunk_node.data(), (size_t)chunks * 4);
    const double t_prep = now_s();
    CUDA_TRY(e, cudaMemcpyAsync(e->d_stage.ptr, hs, bytes, cudaMemcpyHostToDevice, st));
    // ---- device buffers ------------------------------------------------------------------------------
    // + the summed flags (sharded) / + the gathered per-item records and flags (batch sharding)
    const size_t res_words = kHdrWords + (size_t)n_full * 2 + (size_t)P_dev * 4 + 1 + (deal ? (size_t)P * 4 + 1 : 0);
    TRY(ensure_pinned(e, e->h_counts, res_words));
    {
        const uint32_t *before = e->d_counts.ptr;
        TRY(ensure_dev(e, e->d_counts, res_words));
        if (e->d_counts.ptr != before)
            CUDA_TRY(e, cudaMemsetAsync(e->d_counts.ptr, 0, kHdrWords * sizeof(uint32_t), st));
    }
    TRY(ensure_dev(e, e->d_plocal, (size_t)dtotal + ftotal + 16));
    TRY(ensure_dev(e, e->d_ptab, (size_t)std::max(P_dev, n_full) * e->tab_stride));
    TRY(ensure_dev(e, e->d_pnext, (size_t)std::max(dtotal, ftotal) + 8));
    {
        // second-look / exact lists: the incremental pass touches ~4 % of the rows per path
        const uint64_t want = std::max<uint64_t>(std::min<uint64_t>((uint64_t)P_dev * e->A / 256, 1ull << 26), 1u << 20);
        if (want > e->slow_cap) e->slow_cap = (size_t)want;
        if (want > e->work_cap) e->work_cap = (size_t)want;
        TRY(ensure_dev(e, e->d_slow, e->slow_cap));
        TRY(ensure_dev(e, e->d_work, e->work_cap));
    }
    MaskGeom mg;
    mg.W = 1;
    mg.L = e->L;
    mg.bt_base = 0;
    mg.node_base = geom_node_base();
    mg.lm_base = geom_lm_base(e->L);
    mg.pair_base = geom_pair_base(e->L);
    mg.n_pairs = e->n_pairs;
    mg.entries = (mg.pair_base + 2 * e->n_pairs + 2 + 3) & ~3u;
    const size_t table_words = (size_t)chunks * mg.entries, fplane_words = (size_t)chunks * 2 * (e->L + 1);
    if (table_words + fplane_words > 0xFFFFFFFFull) {
        set_error(e, "gfe_score_paths_incremental: batch too large for the mask tables");
        return GFE_ERR_LIMIT;
    }
    TRY(ensure_dev(e, e->d_masks, table_words + fplane_words));
    TRY(ensure_dev(e, e->d_unit_ctr, chunks));
    uint32_t *d_hdr = e->d_counts.ptr, *d_full = d_hdr + kHdrWords, *d_inc = d_full + (size_t)n_full * 2;
    uint32_t *d_sum = d_full;                 // where the host reads the (summed) records
    const size_t rec_words = (size_t)n_full * 2 + (size_t)P_dev * 4;
    uint32_t *d_items = nullptr;             // batch sharding: per-item records in the exchange buffer
    if (deal) {
        if ((size_t)P * 4 + 1 > x->n_cap) {
            set_error(e, "gfe_score_paths_incremental_sharded: batch needs %zu exchange slots, the buffer has %u",
                      (size_t)P * 4 + 1, x->n_cap);
            return GFE_ERR_LIMIT;
        }
        d_items = reinterpret_cast<uint32_t *>(x->bases[x->rank]) + (size_t)x->n_cap * ((x->epoch + 1) & 1u);
        CUDA_TRY(e, cudaMemsetAsync(d_items, 0, ((size_t)P * 4 + 1) * 4, st));
    } else if (x) {
        // sharded: this rank's records accumulate in its half of the exchange buffer, one more slot carries
        // its flags; the peer kernel sums all shards into the result block
        if (rec_words + 1 > x->n_cap) {
            set_error(e, "gfe_score_paths_incremental_sharded: batch needs %zu exchange slots, the buffer has %u",
                      rec_words + 1, x->n_cap);
            return GFE_ERR_LIMIT;
        }
        d_full = reinterpret_cast<uint32_t *>(x->bases[x->rank]) + (size_t)x->n_cap * ((x->epoch + 1) & 1u);
        d_inc = d_full + (size_t)n_full * 2;
    }
    const unsigned char *ds = e->d_stage.ptr;
    const uint32_t *d_poff = reinterpret_cast<const uint32_t *>(ds + o_poff);
    const PairDict pd{e->d_pair_keys.ptr, e->d_pair_ids.ptr, e->d_pair_ctr.ptr};
    // ---- chain A: the incremental pass over all P_dev slots ------------------------------------------
    PrologueArgs pro{};
    pro.zero[0] = d_inc;
    pro.zero_words[0] = P_dev * 4;
    pro.zero[1] = d_hdr;
    pro.zero_words[1] = kWorkCounters;
    pro.zero[2] = e->d_masks.ptr;
    pro.zero_words[2] = (uint32_t)(table_words + fplane_words);
    pro.zero[3] = e->d_unit_ctr.ptr;
    pro.zero_words[3] = chunks;
    pro.sticky = d_hdr + kWorkCounters + 1;
    launch_prologue(pro, e->sm_count, st);
    AssembleArgs as;
    as.desc = reinterpret_cast<const uint4 *>(ds + o_desc);
    as.P = P_dev;
    as.base_off = reinterpret_cast<const uint32_t *>(ds + o_boff);
    as.base_steps = reinterpret_cast<const uint32_t *>(ds + o_bst);
    as.suffix_steps = reinterpret_cast<const uint32_t *>(ds + o_suf);
    as.poff = d_poff;
    as.map = e->d_node2local.ptr;
    as.n_nodes = e->n_nodes;
    as.L = e->L;
    as.out = e->d_plocal.ptr;
    as.err = d_hdr + kWorkCounters;
    launch_assemble_paths(as, st);
    launch_build_tables(d_poff, e->d_plocal.ptr, P_dev, e->L, e->tab_stride, e->d_ptab.ptr, e->d_pnext.ptr,
                        e->d_masks.ptr, e->d_masks.ptr + table_words, mg, pd, st);
    IncrArgs ia;
    ia.bk = e->bk;
    ia.mg = mg;
    ia.inv = InvIndex{e->d_inv_off.ptr, e->d_inv_rows.ptr};
    ia.masks = e->d_masks.ptr;
    ia.fplane = e->d_masks.ptr + table_words;
    ia.chunk_node = reinterpret_cast<const uint32_t *>(ds + o_cn);
    ia.chunks = chunks;
    ia.counts = d_inc;
    ia.unit_rows = 256;
    ia.unit_ctr = e->d_unit_ctr.ptr;
    ia.slow = e->d_slow.ptr;
    ia.work_count = d_hdr;
    ia.slow_cap = (uint32_t)std::min<size_t>(e->d_slow.cap, 0xFFFFFFFFu);
    if (e->timing) {
        TRY(next_event_pair(e));
        CUDA_TRY(e, cudaEventRecord(e->ev0, st));
    }
    launch_score_incr(ia, e->sm_count, st);
    if (e->timing) CUDA_TRY(e, cudaEventRecord(e->ev1, st));
    ScoreArgs a;
    a.bk = e->bk;
    a.poff = d_poff;
    a.pstep = e->d_plocal.ptr;
    a.ptab = e->d_ptab.ptr;
    a.pnext = e->d_pnext.ptr;
    a.tab_stride = e->tab_stride;
    a.P = P_dev;
    a.chunk = pick_chunk(e, P_dev);
    a.filter = 1;
    a.counts = d_inc;
    a.work = e->d_work.ptr;
    a.work_count = d_hdr;
    a.work_cap = (uint32_t)std::min<size_t>(e->d_work.cap, 0xFFFFFFFFu);
    a.count_stride = 4;
    a.path_tag = 1u << 30;     // what the long-row kernel queues belongs to the fresh set
    a.err = d_hdr + kWorkCounters;
    e->kernel_launches += 4;
    const uint32_t long_lo = e->bk.tile_start[kLongBucket], long_hi = e->bk.tile_start[kNumBuckets];
    if (long_hi > long_lo) {   // alignments longer than kMaxBucket steps: scored afresh, set 1
        launch_score_paths(a, e->sm_count, long_lo, long_hi, false, st);
        e->kernel_launches += 1;
    }
    launch_slow_worklist(a, e->d_slow.ptr, d_hdr + 2, ia.slow_cap, st);
    launch_exact_worklist(a, st);
    e->kernel_launches += 2;
    if (e->max_m > (uint32_t)kExactRegSteps) {
        launch_exact_worklist_long(a, std::min<uint32_t>(e->max_m, GFE_MAX_STEPS), st);
        e->kernel_launches += 1;
    }
    CUDA_TRY(e, cudaGetLastError());
    // ---- chain B: the full scan of the mode-0 items -----------------------------------------------------
    if (n_full) {
        uint16_t *d_flocal = e->d_plocal.ptr + ((dtotal + 7) & ~7u);
        AssembleArgs fs = as;
        fs.desc = reinterpret_cast<const uint4 *>(ds + o_fdesc);
        fs.P = n_full;
        fs.poff = reinterpret_cast<const uint32_t *>(ds + o_fpoff);
        fs.out = d_flocal;
        launch_assemble_paths(fs, st);
        e->kernel_launches += 1;
        TRY(enqueue_score(e, fs.poff, d_flocal, n_full, ftotal, 1, d_full, d_hdr, nullptr, d_hdr + kWorkCounters,
                          d_hdr + kWorkCounters + 1, st));
    }
    if (deal) {
        // this rank's share of the candidates, composed per item on the device; the sum over the ranks
        // (everybody else wrote zeros for them) is the all-gather
        launch_compose_items(reinterpret_cast<const uint32_t *>(ds + o_sitem), reinterpret_cast<const uint32_t *>(ds + o_saux),
                             P_dev, d_inc, d_full, reinterpret_cast<int32_t *>(d_items), st);
        launch_flag_word(d_hdr + kWorkCounters, d_hdr, d_items + (size_t)P * 4, st);
        launch_flag_or(d_hdr + kWorkCounters + 1, d_items + (size_t)P * 4, st);
        e->kernel_launches += 4;
        x->epoch += 1;
        const int prc = gfe_peer_sum_u32(x->bases, x->world, x->rank, x->n_cap, P * 4 + 1, x->epoch,
                                         d_hdr + kHdrWords + rec_words + 1, st);
        if (prc != GFE_OK) {
            set_error(e, "gfe_score_paths_incremental_sharded: peer sum launch failed");
            return prc;
        }
    } else if (x) {
        launch_flag_word(d_hdr + kWorkCounters, d_hdr, d_full + rec_words, st);   // bad step | overflow of the last chain
        launch_flag_or(d_hdr + kWorkCounters + 1, d_full + rec_words, st);        // | overflow of the first chain
        e->kernel_launches += 3;
        x->epoch += 1;
        const int prc = gfe_peer_sum_u32(x->bases, x->world, x->rank, x->n_cap, (uint32_t)rec_words + 1, x->epoch, d_sum, st);
        if (prc != GFE_OK) {
            set_error(e, "gfe_score_paths_incremental_sharded: peer sum launch failed");
            return prc;
        }
    }
    CUDA_TRY(e, cudaMemcpyAsync(e->h_counts.ptr, e->d_counts.ptr, res_words * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    const double t_enq = now_s();
    CUDA_TRY(e, cudaStreamSynchronize(st));
    const double t_done = now_s();
    e->incr_t[0] += t_prep - t_begin;
    e->incr_t[1] += t_enq - t_prep;
    e->incr_t[2] += t_done - t_enq;
    e->incr_n[0] += 1;
    e->incr_n[1] += P;
    e->incr_n[2] += chunks;
    e->incr_n[3] += n_full;
    if (e->timing && e->ev0) cudaEventElapsedTime(&e->last_kernel_ms, e->ev0, e->ev1);
    e->h2d_bytes += bytes;
    e->d2h_bytes += res_words * sizeof(uint32_t);
    const uint32_t *h = e->h_counts.ptr;
    const uint32_t any_flags = deal ? h[kHdrWords + rec_words + 1 + (size_t)P * 4]
                                    : (x ? h[kHdrWords + rec_words] : 0u);     // summed over the ranks
    if (x && any_flags == GFE_PEER_TIMEOUT) {
        set_error(e, "gfe_score_paths_incremental_sharded: a peer never reached epoch %u (GFE_PEER_TIMEOUT_MS)", x->epoch);
        return GFE_ERR_CUDA;
    }
    if (h[kWorkCounters]) {
        CUDA_TRY(e, cudaMemset(e->d_counts.ptr + kWorkCounters, 0, 2 * sizeof(uint32_t)));
        if (h[kWorkCounters] & 4u) {
            set_error(e, "gfe_score_paths_incremental: an alignment of more than %u steps passes the id filter of a "
                         "candidate path; the reference's dp[MAX_N][MAX_N] cannot hold it", GFE_MAX_STEPS);
            return GFE_ERR_LIMIT;
        }
        set_error(e, "gfe_score_paths_incremental: a step names a node outside the universe");
        return GFE_ERR_ARG;
    }
    if (x && (any_flags >> 8)) {
        // raised on another rank's shard only (flag_word_kernel: byte 2 = over-long alignment, byte 1 = bad
        // step): every rank returns the same class of error, none retries alone and waits for the others
        if ((any_flags >> 16) & 0xFFu) {
            set_error(e, "gfe_score_paths_incremental_sharded: on another rank's shard an alignment of more than %u "
                         "steps passes the id filter of a candidate path", GFE_MAX_STEPS);
            return GFE_ERR_LIMIT;
        }
        set_error(e, "gfe_score_paths_incremental_sharded: another rank saw a step outside the universe");
        return GFE_ERR_ARG;
    }
    if (h[1] || h[3] || h[kWorkCounters + 1] || any_flags) {   // sharded: every rank repeats the batch together
        // a list overflowed in one of the two chains: grow both lists and redo the batch.  The whole header
        // is cleared: the failed attempt's overflow words (hdr[1], hdr[3]) would otherwise be promoted to the
        // sticky flag by the retry's first prologue and a clean retry would be repeated once more
        CUDA_TRY(e, cudaMemset(e->d_counts.ptr, 0, kHdrWords * sizeof(uint32_t)));
        if (++e->incr_retries > 6 || e->d_slow.cap >= ((size_t)1 << 28)) {
            e->incr_retries = 0;
            set_error(e, "gfe_score_paths_incremental: work lists still overflow; split the batch");
            return GFE_ERR_LIMIT;
        }
        e->slow_cap = std::min<size_t>(e->d_slow.cap * 4, (size_t)1 << 28);
        e->work_cap = std::min<size_t>(e->d_work.cap * 4, (size_t)1 << 28);
        TRY(ensure_dev(e, e->d_slow, e->slow_cap));
        TRY(ensure_dev(e, e->d_work, e->work_cap));
        const int rc = score_incremental_impl(e, x, b, counts);
        e->incr_retries = 0;
        return rc;
    }
    e->exact_pairs += h[0];
    e->slow_pairs += h[2] - h[4];
    // ---- results per item ---------------------------------------------------------------------------------
    if (deal) {
        memcpy(counts, h + kHdrWords + rec_words + 1, (size_t)P * 4 * sizeof(int32_t));
        return GFE_OK;
    }
    const uint32_t *h_full = h + kHdrWords, *h_inc = h_full + (size_t)n_full * 2;
    for (uint32_t i = 0; i < P; ++i) counts[4 * (size_t)i] = counts[4 * (size_t)i + 1] = 0;
    for (size_t k = 0; k < slots.size(); ++k) {
        const Slot &sl = slots[k];
        const uint32_t *r = h_inc + 4 * (size_t)slot_dev[k];
        int32_t *o = counts + 4 * (size_t)sl.item;
        if (sl.key == kNoNode) {
            const uint32_t *f = h_full + 2 * (size_t)full_of_item[sl.item];
            o[0] = (int32_t)f[0];
            o[1] = (int32_t)f[1];
            o[2] = (int32_t)r[2];
            o[3] = (int32_t)r[3];
        } else if (sl.parent) {
            o[0] -= (int32_t)r[0];
            o[1] -= (int32_t)r[1];
        } else {
            o[0] += (int32_t)r[0];
            o[1] += (int32_t)r[1];
            o[2] = (int32_t)r[2];
            o[3] = (int32_t)r[3];
        }
    }
    return GFE_OK;
}

extern "C" {

int gfe_score_paths_incremental(gfe_engine *e, const gfe_path_batch *batch, int32_t *counts) {
    return score_incremental_impl(e, nullptr, batch, counts);
}

int gfe_score_paths_incremental_sharded(gfe_engine *e, gfe_exchange *x, const gfe_path_batch *batch,
                                        int32_t *counts) {
    if (!x) return GFE_ERR_ARG;
    return score_incremental_impl(e, x, batch, counts);
}

int gfe_score_paths_device(gfe_engine *e, const uint32_t *d_path_offsets,
                           const uint32_t *d_path_steps, uint32_t P, uint32_t total_path_steps,
                           int filter, uint32_t *d_counts, void *cuda_stream) {
    if (!e || !d_path_offsets || !d_path_steps || !d_counts) return GFE_ERR_ARG;
    if (!e->have_alignments || !e->have_universe) {
        set_error(e, "gfe_score_paths_device: alignments and universe must be set first");
        return GFE_ERR_STATE;
    }
    if (P == 0) return GFE_OK;
    CUDA_TRY(e, cudaSetDevice(e->device));
    cudaStream_t st = (cudaStream_t)cuda_stream;   // NULL = the CUDA default stream
    TRY(ensure_dev(e, e->d_plocal, (size_t)total_path_steps + 8));
    if (!e->d_err.ptr) {
        TRY(ensure_dev(e, e->d_err, 4));
        CUDA_TRY(e, cudaMemsetAsync(e->d_err.ptr, 0, 4 * sizeof(uint32_t), st));
    }
    if (!e->d_work_count.ptr) {
        // zero-filled on allocation: the prologue folds the previous batch's overflow flags into the
        // sticky word before it clears the counters
        TRY(ensure_dev(e, e->d_work_count, kWorkCounters));
        CUDA_TRY(e, cudaMemsetAsync(e->d_work_count.ptr, 0, kWorkCounters * sizeof(uint32_t), st));
    }
    return enqueue_score(e, d_path_offsets, e->d_plocal.ptr, P, total_path_steps, filter, d_counts,
                         e->d_work_count.ptr, d_path_steps, e->d_err.ptr, e->d_err.ptr + 1, st);
}

// ---- evalPath scoring ---------------------------------------------------------------------------

int gfe_score_pairs(gfe_engine *e, const uint32_t *path_steps, uint32_t n, int filter,
                    int32_t *t_fw, int32_t *t_rc, uint32_t *bad, uint32_t *good) {
    if (!e || !path_steps || !bad || !good) return GFE_ERR_ARG;
    if (!e->have_alignments) {
        set_error(e, "gfe_score_pairs: call gfe_set_alignments first");
        return GFE_ERR_STATE;
    }
    if (n == 0 || n > GFE_MAX_STEPS) {
        set_error(e, "gfe_score_pairs: path has %u steps; allowed 1..%u", n, GFE_MAX_STEPS);
        return n == 0 ? GFE_ERR_ARG : GFE_ERR_LIMIT;
    }
    if (e->max_m > GFE_MAX_STEPS) {
        set_error(e, "gfe_score_pairs: an alignment has %u steps; the reference's MAX_N allows %u", e->max_m, GFE_MAX_STEPS);
        return GFE_ERR_LIMIT;
    }
    // universe = the distinct nodes of the path, in first-visit order
    std::vector<uint32_t> uni;
    uni.reserve(n);
    {
        std::unordered_map<uint32_t, uint32_t> seen;
        for (uint32_t i = 0; i < n; ++i) {
            const uint32_t node = path_steps[i] >> 1;
            if (node >= e->n_nodes) {
                set_error(e, "gfe_score_pairs: step %u names node %u >= n_nodes %u", i, node,
                          e->n_nodes);
                return GFE_ERR_ARG;
            }
            if (seen.emplace(node, (uint32_t)uni.size()).second) uni.push_back(node);
        }
    }
    TRY(set_universe_impl(e, uni.data(), (uint32_t)uni.size(), false));
    const size_t off_bytes = 16;
    const size_t bytes = off_bytes + (size_t)n * 2;
    TRY(ensure_pinned(e, e->h_stage, bytes));
    TRY(ensure_dev(e, e->d_stage, bytes));
    uint32_t *po = reinterpret_cast<uint32_t *>(e->h_stage.ptr);
    po[0] = 0;
    po[1] = n;
    uint16_t *ls = reinterpret_cast<uint16_t *>(e->h_stage.ptr + off_bytes);
    for (uint32_t i = 0; i < n; ++i)
        ls[i] = (uint16_t)((e->node2local[path_steps[i] >> 1] << 1) | (path_steps[i] & 1u));
    cudaStream_t st = e->stream;
    CUDA_TRY(e, cudaMemcpyAsync(e->d_stage.ptr, e->h_stage.ptr, bytes, cudaMemcpyHostToDevice, st));
    const uint32_t *d_poff = reinterpret_cast<const uint32_t *>(e->d_stage.ptr);
    const uint16_t *d_pstep = reinterpret_cast<const uint16_t *>(e->d_stage.ptr + off_bytes);
    TRY(ensure_dev(e, e->d_ptab, (size_t)e->tab_stride));
    TRY(ensure_dev(e, e->d_counts, 2));
    TRY(ensure_pinned(e, e->h_counts, 2));
    TRY(ensure_dev(e, e->d_pair_work, std::max<uint64_t>(e->A, 1)));
    TRY(ensure_dev(e, e->d_work_count, kWorkCounters));
    const bool want_scores = t_fw || t_rc;
    if (want_scores) {
        TRY(ensure_dev(e, e->d_tfw, std::max<uint64_t>(e->A, 1)));
        TRY(ensure_dev(e, e->d_trc, std::max<uint64_t>(e->A, 1)));
    }
    CUDA_TRY(e, cudaMemsetAsync(e->d_counts.ptr, 0, 2 * sizeof(uint32_t), st));
    CUDA_TRY(e, cudaMemsetAsync(e->d_work_count.ptr, 0, kWorkCounters * sizeof(uint32_t), st));
    TRY(ensure_dev(e, e->d_pnext, (size_t)n + 8));
    {
        MaskGeom mg{};
        PairDict pd{nullptr, nullptr, nullptr};
        launch_build_tables(d_poff, d_pstep, 1, e->L, e->tab_stride, e->d_ptab.ptr, e->d_pnext.ptr,
                            nullptr, nullptr, mg, pd, st);
    }
    PairArgs a;
    a.bk = e->bk;
    a.pstep = d_pstep;
    a.n = n;
    a.ptab = e->d_ptab.ptr;
    a.pnext = e->d_pnext.ptr;
    a.tab_entries = e->tab_stride;
    a.filter = filter;
    a.t_fw = want_scores ? e->d_tfw.ptr : nullptr;
    a.t_rc = want_scores ? e->d_trc.ptr : nullptr;
    a.counts = e->d_counts.ptr;
    a.work = e->d_pair_work.ptr;
    a.work_count = e->d_work_count.ptr;
    a.work_cap = (uint32_t)std::min<uint64_t>(e->d_pair_work.cap, 0xFFFFFFFFu);
    if (e->timing) {
        TRY(next_event_pair(e));
        CUDA_TRY(e, cudaEventRecord(e->ev0, st));
    }
    launch_pair_classify(a, e->sm_count, st);
    if (e->timing) CUDA_TRY(e, cudaEventRecord(e->ev1, st));
    e->kernel_launches += 2;
    if (e->A) {
        launch_pair_exact(a, st);
        e->kernel_launches += 1;
        if (e->max_m > (uint32_t)kExactRegSteps) {
            launch_pair_exact_long(a, std::min<uint32_t>(e->max_m, GFE_MAX_STEPS), st);
            e->kernel_launches += 1;
        }
    }
    if (e->timing) {       // the exact DP behind the classify pass: ev1 .. aux
        if (!e->ev_aux) CUDA_TRY(e, cudaEventCreate(&e->ev_aux));
        CUDA_TRY(e, cudaEventRecord(e->ev_aux, st));
    }
    CUDA_TRY(e, cudaGetLastError());
    CUDA_TRY(e, cudaMemcpyAsync(e->h_counts.ptr, e->d_counts.ptr, 2 * sizeof(uint32_t),
                                cudaMemcpyDeviceToHost, st));
    CUDA_TRY(e, cudaMemcpyAsync(e->h_wc, e->d_work_count.ptr, 2 * sizeof(uint32_t),
                                cudaMemcpyDeviceToHost, st));
    if (t_fw && e->A)
        CUDA_TRY(e, cudaMemcpyAsync(t_fw, e->d_tfw.ptr, e->A * sizeof(int32_t),
                                    cudaMemcpyDeviceToHost, st));
    if (t_rc && e->A)
        CUDA_TRY(e, cudaMemcpyAsync(t_rc, e->d_trc.ptr, e->A * sizeof(int32_t),
                                    cudaMemcpyDeviceToHost, st));
    CUDA_TRY(e, cudaStreamSynchronize(st));
    if (e->timing && e->ev0) cudaEventElapsedTime(&e->last_kernel_ms, e->ev0, e->ev1);
    if (e->timing && e->ev_aux) cudaEventElapsedTime(&e->last_aux_ms, e->ev1, e->ev_aux);
    e->exact_pairs += e->h_wc[0];
    e->h2d_bytes += bytes;
    e->d2h_bytes += 16 + ((t_fw ? 1 : 0) + (t_rc ? 1 : 0)) * e->A * sizeof(int32_t);
    *bad = e->h_counts.ptr[0];
    *good = e->h_counts.ptr[1];
    return GFE_OK;
}

// ---- edge support -------------------------------------------------------------------------------

namespace {
struct PairKeyHash {
    size_t operator()(uint64_t k) const { return (size_t)(k * 0x9E3779B97F4A7C15ull >> 16); }
};
inline uint64_t pair_key(uint32_t from, uint32_t to) { return ((uint64_t)from << 32) | to; }

// representative slot of every directed pair: the first slot (2i = edge i, 2i+1 = its reverse
// complement) that carries the pair.  Host side of the sharded flow only
// (gfe_edge_weights_from_counts); the single-GPU path keeps everything on the device.
void build_rep(const uint32_t *edge_from, const uint32_t *edge_to, uint32_t E,
               std::unordered_map<uint64_t, uint32_t, PairKeyHash> &rep) {
    rep.reserve((size_t)E * 2 + 1);
    for (uint32_t i = 0; i < E; ++i) {
        rep.emplace(pair_key(edge_from[i], edge_to[i]), 2 * i);
        rep.emplace(pair_key(edge_to[i] ^ 1u, edge_from[i] ^ 1u), 2 * i + 1);
    }
}

// bitmap over the step array, bit s = step s starts an alignment; built lazily once per table
int ensure_rowstart(gfe_engine *e, cudaStream_t st) {
    if (e->have_rowstart) return GFE_OK;
    const size_t words = (size_t)(e->S >> 5) + 2;
    TRY(ensure_dev(e, e->d_rowstart, words));
    CUDA_TRY(e, cudaMemsetAsync(e->d_rowstart.ptr, 0, words * 4, st));
    launch_rowstart(e->d_off, e->A, e->d_rowstart.ptr, st);
    e->kernel_launches += e->A ? 1 : 0;
    CUDA_TRY(e, cudaGetLastError());
    e->have_rowstart = true;
    return GFE_OK;
}

// uploads the link records, builds the device pair table and streams the alignments through it
int edge_count_on_device(gfe_engine *e, const uint32_t *edge_from, const uint32_t *edge_to, uint32_t E,
                         cudaStream_t st) {
    if (E >= 0x7FFFFFFFu / 3) {
        set_error(e, "gfe_edge_support: %u link records exceed the table limit", E);
        return GFE_ERR_LIMIT;
    }
    const uint32_t space = 2 * e->n_nodes;
    for (uint32_t i = 0; i < E; ++i)
        if (edge_from[i] >= space || edge_to[i] >= space) {
            set_error(e, "edge %u names a node >= n_nodes %u", i, e->n_nodes);
            return GFE_ERR_ARG;
        }
    e->edge_cap = edge_table_capacity(E);
    TRY(ensure_dev(e, e->d_edge_ft, std::max<size_t>((size_t)E * 2, 1)));
    TRY(ensure_dev(e, e->d_edge_table, e->edge_cap));
    if (E) {
        CUDA_TRY(e, cudaMemcpyAsync(e->d_edge_ft.ptr, edge_from, (size_t)E * 4, cudaMemcpyHostToDevice, st));
        CUDA_TRY(e, cudaMemcpyAsync(e->d_edge_ft.ptr + E, edge_to, (size_t)E * 4, cudaMemcpyHostToDevice, st));
    }
    const uint32_t steps2 = 2 * e->n_nodes;
    TRY(ensure_dev(e, e->d_edge_adj, std::max<size_t>(steps2, 1)));
    TRY(ensure_dev(e, e->d_edge_deg, (size_t)steps2 + 1));
    TRY(ensure_dev(e, e->d_edge_cnt, (size_t)E * 2 + 1));
    EdgeBuild eb;
    eb.table = e->d_edge_table.ptr;
    eb.cap = e->edge_cap;
    eb.deg = e->d_edge_deg.ptr;
    eb.adj = e->d_edge_adj.ptr;
    eb.steps2 = steps2;
    eb.cursor = e->d_edge_deg.ptr + steps2;
    launch_edge_build(e->d_edge_ft.ptr, e->d_edge_ft.ptr + E, E, eb, st);
    CUDA_TRY(e, cudaMemsetAsync(e->d_edge_cnt.ptr, 0, ((size_t)E * 2 + 1) * 4, st));
    EdgeStreamArgs a;
    TRY(ensure_rowstart(e, st));
    a.steps = e->d_steps;
    a.rowstart = e->d_rowstart.ptr;
    a.S = e->S;
    a.adj = e->d_edge_adj.ptr;
    a.cnt = e->d_edge_cnt.ptr;
    a.steps2 = steps2;
    a.n_arcs = 2 * E;
    a.table = e->d_edge_table.ptr;
    a.cap = e->edge_cap;
    if (e->timing) {
        TRY(next_event_pair(e));
        CUDA_TRY(e, cudaEventRecord(e->ev0, st));
    }
    if (E) launch_edge_stream(a, e->sm_count, st);
    if (e->timing) CUDA_TRY(e, cudaEventRecord(e->ev1, st));
    e->kernel_launches += (E ? 3 : 0) + ((E && e->S >= 2) ? 1 : 0) + 1;
    e->h2d_bytes += (uint64_t)E * 8;
    CUDA_TRY(e, cudaGetLastError());
    return GFE_OK;
}
}  // namespace

int gfe_edge_traversals_device(gfe_engine *e, const uint32_t *edge_from, const uint32_t *edge_to,
                               uint32_t E, uint32_t *d_counts, void *cuda_stream) {
    if (!e || (E && (!edge_from || !edge_to)) || !d_counts) return GFE_ERR_ARG;
    if (!e->have_alignments) {
        set_error(e, "gfe_edge_support: call gfe_set_alignments first");
        return GFE_ERR_STATE;
    }
    CUDA_TRY(e, cudaSetDevice(e->device));
    cudaStream_t st = (cudaStream_t)cuda_stream;   // NULL = the CUDA default stream
    TRY(edge_count_on_device(e, edge_from, edge_to, E, st));
    CUDA_TRY(e, cudaMemsetAsync(d_counts, 0, (size_t)E * 2 * sizeof(uint32_t), st));
    launch_edge_scatter(e->d_edge_table.ptr, e->edge_cap, e->d_edge_cnt.ptr, d_counts, st);
    e->kernel_launches += 1;
    CUDA_TRY(e, cudaGetLastError());
    // the caller's host arrays may go away after the call
    CUDA_TRY(e, cudaStreamSynchronize(st));
    return GFE_OK;
}

int gfe_edge_weights_from_counts(const uint32_t *edge_from, const uint32_t *edge_to, uint32_t E,
                                 const uint32_t *counts, uint32_t *weights) {
    if (E && (!edge_from || !edge_to || !counts || !weights)) return GFE_ERR_ARG;
    std::unordered_map<uint64_t, uint32_t, PairKeyHash> rep;
    build_rep(edge_from, edge_to, E, rep);
    for (uint32_t i = 0; i < E; ++i) {
        const uint32_t x = edge_from[i], y = edge_to[i];
        const bool hairpin = (y == (x ^ 1u));   // its reverse complement is itself
        uint64_t k = counts[rep[pair_key(x, y)]];
        if (!hairpin) k += counts[rep[pair_key(y ^ 1u, x ^ 1u)]];
        // reference src/alignments.cpp:371-396: the twin entries are incremented together; a
        // self-twin edge is inserted with weight 1 and then incremented twice per traversal.
        weights[i] = (uint32_t)(hairpin ? (k ? 2 * k - 1 : 0) : k);
    }
    return GFE_OK;
}

int gfe_edge_support(gfe_engine *e, const uint32_t *edge_from, const uint32_t *edge_to, uint32_t E,
                     uint32_t *weights) {
    if (!e || (E && (!weights || !edge_from || !edge_to))) return GFE_ERR_ARG;
    if (!e->have_alignments) {
        set_error(e, "gfe_edge_support: call gfe_set_alignments first");
        return GFE_ERR_STATE;
    }
    CUDA_TRY(e, cudaSetDevice(e->device));
    cudaStream_t st = e->stream;
    TRY(edge_count_on_device(e, edge_from, edge_to, E, st));
    if (E) {
        TRY(ensure_dev(e, e->d_edge_counts, E));
        launch_edge_weights(e->d_edge_ft.ptr, e->d_edge_ft.ptr + E, E, e->d_edge_table.ptr, e->edge_cap,
                            e->d_edge_cnt.ptr, e->d_edge_counts.ptr, st);
        e->kernel_launches += 1;
        CUDA_TRY(e, cudaGetLastError());
        CUDA_TRY(e, cudaMemcpyAsync(weights, e->d_edge_counts.ptr, (size_t)E * 4, cudaMemcpyDeviceToHost, st));
    }
    CUDA_TRY(e, cudaStreamSynchronize(st));
    if (e->timing && e->ev0) cudaEventElapsedTime(&e->last_kernel_ms, e->ev0, e->ev1);
    e->d2h_bytes += (uint64_t)E * 4;
    return GFE_OK;
}

// ---- evalPath detail: traceback columns ----------------------------------------------------------

int gfe_trace_pairs(gfe_engine *e, const uint32_t *path_steps, uint32_t n, uint64_t first,
                    uint32_t count, const uint8_t *use_rc, uint8_t *ops, uint64_t ops_capacity,
                    uint64_t *ops_first, uint32_t *ops_len, int32_t *score) {
    if (!e || !path_steps || !ops_first || !ops_len || (!ops && ops_capacity)) return GFE_ERR_ARG;
    if (!e->have_alignments) {
        set_error(e, "gfe_trace_pairs: call gfe_set_alignments first");
        return GFE_ERR_STATE;
    }
    if (n == 0 || n > GFE_MAX_STEPS) {
        set_error(e, "gfe_trace_pairs: path has %u steps; allowed 1..%u", n, GFE_MAX_STEPS);
        return n == 0 ? GFE_ERR_ARG : GFE_ERR_LIMIT;
    }
    if (e->max_m > GFE_MAX_STEPS) {
        set_error(e, "gfe_trace_pairs: an alignment has %u steps; the reference's MAX_N allows %u", e->max_m, GFE_MAX_STEPS);
        return GFE_ERR_LIMIT;
    }
    if (first > e->A || count > e->A - first) {
        set_error(e, "gfe_trace_pairs: alignments [%llu, +%u) out of range (A = %llu)",
                  (unsigned long long)first, count, (unsigned long long)e->A);
        return GFE_ERR_ARG;
    }
    if (count == 0) return GFE_OK;
    for (uint32_t i = 0; i < n; ++i)
        if ((path_steps[i] >> 1) >= e->n_nodes) {
            set_error(e, "gfe_trace_pairs: step %u names node %u >= n_nodes %u", i,
                      path_steps[i] >> 1, e->n_nodes);
            return GFE_ERR_ARG;
        }
    cudaStream_t st = e->stream;
    // total ops bytes of the request = steps in range + count * n
    uint32_t range[2];
    CUDA_TRY(e, cudaMemcpyAsync(&range[0], e->d_off + first, 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(e, cudaMemcpyAsync(&range[1], e->d_off + first + count, 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(e, cudaStreamSynchronize(st));
    const uint64_t need = (uint64_t)(range[1] - range[0]) + (uint64_t)count * n;
    if (ops_capacity < need) {
        set_error(e, "gfe_trace_pairs: ops buffer holds %llu bytes, %llu needed",
                  (unsigned long long)ops_capacity, (unsigned long long)need);
        return GFE_ERR_ARG;
    }
    TRY(ensure_dev(e, e->d_tr_path, n));
    CUDA_TRY(e, cudaMemcpyAsync(e->d_tr_path.ptr, path_steps, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    e->h2d_bytes += (uint64_t)n * 4;
    const size_t budget_words = (size_t)1 << 28;   // 1 GiB of decision words per sub-batch
    uint32_t done = 0;
    uint64_t ops_done = 0;   // bytes of ops regions before the current sub-batch
    while (done < count) {
        // sub-batch size: start from what fits with W = 1, shrink once the real max m is known
        uint32_t sub = (uint32_t)std::min<uint64_t>(count - done, std::max<size_t>(budget_words / n, 1));
        uint32_t max_m = 0, W = 1;
        for (;;) {
            TRY(ensure_dev(e, e->d_tr_misc, 4));
            CUDA_TRY(e, cudaMemsetAsync(e->d_tr_misc.ptr, 0, 4, st));
            launch_max_len(e->d_off, first + done, sub, e->d_tr_misc.ptr, st);
            e->kernel_launches += 1;
            CUDA_TRY(e, cudaMemcpyAsync(&max_m, e->d_tr_misc.ptr, 4, cudaMemcpyDeviceToHost, st));
            CUDA_TRY(e, cudaStreamSynchronize(st));
            W = std::max<uint32_t>((max_m + 15) / 16, 1);
            const size_t per_pair = (size_t)n * W + (max_m > 32 ? (size_t)max_m * 2 : 0);
            const uint32_t fit = (uint32_t)std::max<size_t>(budget_words / per_pair, 1);
            if (fit >= sub) break;
            sub = fit;
        }
        const uint64_t f0 = first + done;
        uint32_t sr[2];
        CUDA_TRY(e, cudaMemcpyAsync(&sr[0], e->d_off + f0, 4, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(e, cudaMemcpyAsync(&sr[1], e->d_off + f0 + sub, 4, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(e, cudaStreamSynchronize(st));
        const uint64_t sub_ops = (uint64_t)(sr[1] - sr[0]) + (uint64_t)sub * n;
        TRY(ensure_dev(e, e->d_tr_dirs, (size_t)n * sub * W));
        TRY(ensure_dev(e, e->d_tr_base, sub));
        TRY(ensure_dev(e, e->d_tr_first, sub));
        TRY(ensure_dev(e, e->d_tr_len, sub));
        TRY(ensure_dev(e, e->d_tr_score, sub));
        TRY(ensure_dev(e, e->d_tr_ops, sub_ops));
        TRY(ensure_dev(e, e->d_tr_rc, sub));
        if (max_m > 32) {
            TRY(ensure_dev(e, e->d_tr_longb, (size_t)sub * max_m));
            TRY(ensure_dev(e, e->d_tr_longrows, (size_t)sub * (max_m + 1)));
        }
        if (use_rc) {
            CUDA_TRY(e, cudaMemcpyAsync(e->d_tr_rc.ptr, use_rc + done, sub, cudaMemcpyHostToDevice, st));
            e->h2d_bytes += sub;
        }
        launch_trace_bases(e->d_off, f0, sub, n, e->d_tr_base.ptr, st);
        TraceArgs a;
        a.off = e->d_off;
        a.steps = e->d_steps;
        a.path = e->d_tr_path.ptr;
        a.n = n;
        a.first = f0;
        a.count = sub;
        a.use_rc = use_rc ? e->d_tr_rc.ptr : nullptr;
        a.W = W;
        a.dirs = e->d_tr_dirs.ptr;
        a.long_b = e->d_tr_longb.ptr;
        a.long_rows = e->d_tr_longrows.ptr;
        a.long_stride = max_m;
        a.ops_start_base = e->d_tr_base.ptr;
        a.ops = e->d_tr_ops.ptr;
        a.ops_first = e->d_tr_first.ptr;
        a.ops_len = e->d_tr_len.ptr;
        a.score = e->d_tr_score.ptr;
        if (e->timing) {
            TRY(next_event_pair(e));
            CUDA_TRY(e, cudaEventRecord(e->ev0, st));
        }
        launch_pair_trace(a, st);
        if (e->timing) CUDA_TRY(e, cudaEventRecord(e->ev1, st));
        e->kernel_launches += 2;
        CUDA_TRY(e, cudaGetLastError());
        CUDA_TRY(e, cudaMemcpyAsync(ops + ops_done, e->d_tr_ops.ptr, sub_ops, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(e, cudaMemcpyAsync(ops_first + done, e->d_tr_first.ptr, (size_t)sub * 8, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(e, cudaMemcpyAsync(ops_len + done, e->d_tr_len.ptr, (size_t)sub * 4, cudaMemcpyDeviceToHost, st));
        if (score)
            CUDA_TRY(e, cudaMemcpyAsync(score + done, e->d_tr_score.ptr, (size_t)sub * 4, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(e, cudaStreamSynchronize(st));
        if (e->timing && e->ev0) cudaEventElapsedTime(&e->last_kernel_ms, e->ev0, e->ev1);
        e->d2h_bytes += sub_ops + (uint64_t)sub * (score ? 16 : 12);
        for (uint32_t k = 0; k < sub; ++k) ops_first[done + k] += ops_done;   // sub-batch -> request
        done += sub;
        ops_done += sub_ops;
    }
    return GFE_OK;
}

// ---- filter mode: membership of every alignment in a node set ------------------------------------

int gfe_filter_alignments(gfe_engine *e, const uint32_t *nodes, uint32_t n, int32_t min_nodes,
                          uint8_t *keep) {
    if (!e || (!nodes && n) || !keep) return GFE_ERR_ARG;
    if (!e->have_alignments) {
        set_error(e, "gfe_filter_alignments: call gfe_set_alignments first");
        return GFE_ERR_STATE;
    }
    if (e->A == 0) return GFE_OK;
    const size_t words = ((size_t)e->n_nodes + 31) / 32 + 1;
    std::vector<uint32_t> bits(words, 0);
    for (uint32_t i = 0; i < n; ++i) {
        if (nodes[i] >= e->n_nodes) {
            set_error(e, "gfe_filter_alignments: node %u >= n_nodes %u", nodes[i], e->n_nodes);
            return GFE_ERR_ARG;
        }
        bits[nodes[i] >> 5] |= 1u << (nodes[i] & 31);
    }
    cudaStream_t st = e->stream;
    TRY(ensure_dev(e, e->d_tr_misc, words));
    TRY(ensure_dev(e, e->d_tr_rc, e->A));
    CUDA_TRY(e, cudaMemcpyAsync(e->d_tr_misc.ptr, bits.data(), words * 4, cudaMemcpyHostToDevice, st));
    FilterArgs a;
    a.off = e->d_off;
    a.steps = e->d_steps;
    a.A = e->A;
    a.node_bits = e->d_tr_misc.ptr;
    a.n_nodes = e->n_nodes;
    a.min_nodes = min_nodes;
    a.keep = e->d_tr_rc.ptr;
    a.S = e->S;
    TRY(ensure_dev(e, e->d_tr_dirs, (size_t)((e->S + 31) / 32) + 1));
    a.bad_bits = e->d_tr_dirs.ptr;
    if (e->timing) {
        TRY(next_event_pair(e));
        CUDA_TRY(e, cudaEventRecord(e->ev0, st));
    }
    launch_filter_rows(a, e->sm_count, st);
    if (e->timing) CUDA_TRY(e, cudaEventRecord(e->ev1, st));
    e->kernel_launches += 2;
    CUDA_TRY(e, cudaGetLastError());
    CUDA_TRY(e, cudaMemcpyAsync(keep, e->d_tr_rc.ptr, e->A, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(e, cudaStreamSynchronize(st));
    if (e->timing && e->ev0) cudaEventElapsedTime(&e->last_kernel_ms, e->ev0, e->ev1);
    e->h2d_bytes += words * 4;
    e->d2h_bytes += e->A;
    return GFE_OK;
}


}  // extern "C"
