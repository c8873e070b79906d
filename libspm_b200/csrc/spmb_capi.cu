// spmb_capi.cu -- implementation of the C ABI declared in include/spmb.h.
// Host-side plumbing only: owns device buffers, builds pattern tables, enqueues the kernels of
// pack.cuh / exact.cuh / myers.cuh on the context's stream, brings the hit list back.
// There is no CPU search path in this file: if no CUDA device is present every entry point
// fails with SPMB_ERR_NO_DEVICE.
#include <cuda_runtime.h>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "../../include/spmb.h"
#include "exact.cuh"
#include "hits.cuh"
#include "hit_order.hpp"
#include "jst.cuh"
#include "layout.cuh"
#include "myers.cuh"
#include "myers_cut.cuh"
#include "myers_chain.cuh"
#include "pack.cuh"
#include "pigeon.cuh"
#include "upload.cuh"
#include "stream.cuh"

// ------------------------------------------------------------------------------------------
// objects
// ------------------------------------------------------------------------------------------
struct spmb_context {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    cudaStream_t aux_stream = nullptr;     // small JST buckets run here beside the large ones (jst_enqueue)
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    cudaStream_t copy_stream = nullptr;    // counters + short hit lists come back here, behind the launch's event:
                                           // a later launch already queued on `stream` does not delay them
    std::string err;
    uint32_t launches = 0;
    int sm_count = 0;
    unsigned long long *h_count = nullptr; // pinned (8 slots)
    spmb_hit *h_spec = nullptr;            // pinned: the first SPEC_HITS hits travel with the counters (one round trip)
    bool myers_cutoff = true;              // spmb_context_set_myers_cutoff
    uint32_t shard = 0, n_shards = 1;      // searches cover super-tiles [n_st*shard/n, n_st*(shard+1)/n)
    // stores / patterns / hit lists / streams / JSTs keep a raw pointer to their context: the
    // context outlives them (a destroy() with live children is deferred until the last one goes)
    uint32_t children = 0;
    bool closed = false;
    UploadPool *upload = nullptr;          // host-haystack upload pipeline (upload.cuh), created on first use
    int upload_threads = -1;               // spmb_context_set_upload_threads (-1: SPMB_HOST_THREADS / available cpus - 1)
    // grow-only device scratch for the journal's temporaries (event table, prefix arrays): the stream-ordered
    // pool took 5-15 ms to hand out ~1.6 GB in seven pieces on every call; released by spmb_context_trim
    uint8_t *scratch = nullptr;
    size_t scratch_bytes = 0;
    // free list of large device blocks (store planes / lane chunks / upload intermediates), by size class:
    // a store that is re-created every call (the host-haystack call shape) gets its blocks back without a
    // driver call; the stream-ordered pool spent 3-6 ms per call on the same five ~0.8 GB requests
    struct Block { void *p; size_t bytes; };
    std::vector<Block> blocks;
    size_t blocks_bytes = 0, blocks_cap = 0;
};

enum { SPEC_HITS = 4096 };
// raw hits from which the next launch on the same hit list orders them itself (enqueue_order): the ~13 small kernels
// cost ~0.12 ms of stream time, which only pays once the list is long (measured on C5 shards: 10 k raw hits per rank
// at N=8 0.32 ms per step ordered from finish, 0.38 ms ordered by the launch; 80 k at N=1 1.73 against 1.27 ms)
enum { SPMB_ORDER_AT_LAUNCH_MIN = 16384 };

static void context_free(spmb_context *ctx)
{
    cudaSetDevice(ctx->device);
    delete ctx->upload;
    if (ctx->scratch) cudaFree(ctx->scratch);
    for (auto &b : ctx->blocks) cudaFreeAsync(b.p, ctx->stream);
    ctx->blocks.clear();
    if (ctx->copy_stream) { cudaStreamSynchronize(ctx->copy_stream); cudaStreamDestroy(ctx->copy_stream); }
    if (ctx->aux_stream) { cudaStreamSynchronize(ctx->aux_stream); cudaStreamDestroy(ctx->aux_stream); }
    if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
    if (ctx->ev_join) cudaEventDestroy(ctx->ev_join);
    if (ctx->own_stream) { cudaStreamSynchronize(ctx->stream); cudaStreamDestroy(ctx->stream); }
    if (ctx->h_count) cudaFreeHost(ctx->h_count);
    if (ctx->h_spec) cudaFreeHost(ctx->h_spec);
    cudaGetLastError();
    delete ctx;
}

static void ctx_adopt(spmb_context *ctx) { if (ctx) ctx->children++; }
static void ctx_release(spmb_context *ctx)
{
    if (!ctx) return;
    if (ctx->children) ctx->children--;
    if (ctx->closed && ctx->children == 0) context_free(ctx);
}


// a child's counted pointer to its context (released after the child's device memory is freed)
struct CtxRef {
    spmb_context *p = nullptr;
    CtxRef() = default;
    CtxRef(const CtxRef &) = delete;
    CtxRef &operator=(const CtxRef &) = delete;
    CtxRef &operator=(spmb_context *c) { ctx_adopt(c); ctx_release(p); p = c; return *this; }
    ~CtxRef() { ctx_release(p); }
    operator spmb_context *() const { return p; }
    spmb_context *operator->() const { return p; }
};

struct spmb_store {
    CtxRef ctx;
    SpmbLayout lay{};
    uint2 *planes = nullptr;
    size_t bytes = 0;
    // lane-chunk copy of the same symbols for the Myers kernels, derived from the planes on the
    // first Myers search of this store (same size as the planes)
    mutable uint2 *lanes = nullptr;
    // windows mode
    uint32_t *row_len = nullptr, *row_seq = nullptr, *row_own = nullptr;
    uint32_t seq_base = 0;          // row_seq[r] = seq_base + r (window buckets of a JST)
    // 4-bit alphabets (spmb_store_create_wide): this store holds bits 1..0 of every symbol, `hi` bits 3..2
    spmb_store *hi = nullptr;
    uint32_t sym_bits = 2;
};

struct spmb_pattern {
    CtxRef ctx;
    int alg = 0;
    uint32_t n_pat = 0, k = 0;
    std::vector<uint32_t> len, off;   // per pattern
    std::vector<uint8_t> ranks;       // host copy, concatenated
    std::vector<uint32_t> filt;       // exact: the min(m,8) symbols of the filter window, packed
    std::vector<uint32_t> filt_off;   // exact: offset of the filter window inside the needle (exact_window_offset)
    uint32_t max_m = 0, min_m = 0;
    int NW = 0;                       // myers: 32-bit words
    uint32_t sym_bits = 2;            // 2 (dna4) or 4 (spmb_pattern_create_wide)
    std::vector<uint32_t> filt_hi;    // exact, 4-bit alphabets: bits 3..2 of the first min(m,8) symbols
    uint32_t warm = 0;                // myers: warm-up columns (multiple of 32)
    uint8_t *d_ranks = nullptr;
    uint32_t *d_eq = nullptr;
    MyersPatMeta *d_meta = nullptr;
    // pigeonhole seed filter (SPMB_ALG_PIGEONHOLE, SPMB_ALG_MYERS_FILTERED)
    PigeonHost pg;
    bool filtered = false;            // the filter is in use (q long enough)
    uint32_t filter_step = 1;         // text positions = 0 (mod step) are scanned (filtered Myers: 1, 2 or 4)
    uint32_t *d_off = nullptr;        // needle offsets into d_ranks
    // Myers sets the one-lane kernels cannot take in one launch (needles that differ in ceil(m/32); needles longer than
    // 256 symbols): one child set per word count, hits carry the caller's pattern ids (MyersPatMeta.pattern_id)
    std::vector<spmb_pattern *> subs;  // children of 1..8 words, searched directly
    spmb_pattern *long_prefix = nullptr; // child over the first 256 symbols of the long needles: the candidate filter
    std::vector<uint32_t> ids;        // child: the caller's pattern id of every needle (empty = identity)
    uint32_t n_long = 0;              // needles longer than 256 symbols
    int chain_wpl = 0;                // words per lane of the lane-chained verification (myers_chain.cuh)
    uint32_t *d_chain_eq = nullptr;
    MyersPatMeta *d_chain_meta = nullptr;
    uint32_t *d_pg_bitmap = nullptr, *d_pg_bitmap2 = nullptr, *d_pg_ufirst = nullptr, *d_pg_pat = nullptr, *d_pg_soff = nullptr;
    unsigned long long *d_pg_tab = nullptr;
};

struct spmb_hits {
    CtxRef ctx;
    spmb_hit *d_hits = nullptr;
    unsigned long long *d_count = nullptr; // [0] hits, [1] filter candidates of the current launch, [2] their max
    unsigned long long cap = 0;
    PigeonCand *d_cand = nullptr;          // candidate buffer of the seed filter (grown on demand)
    unsigned long long cand_cap = 0;
    unsigned long long *d_dd_keys = nullptr; // duplicate-suppression table of the filtered Myers path (pigeon.cuh pg_emit_min)
    std::vector<spmb_hit> host;
    uint64_t count = 0;
    const spmb_store *pend_store = nullptr;
    const spmb_pattern *pend_pat = nullptr;
    struct spmb_jst *pend_jst = nullptr;   // spmb_jst_search_launch
    cudaStream_t collect_stream = nullptr; // the stream hits_collect last read this list on (copy stream or the context's)
    bool spec_done = false;                // hits_collect already brought the (short) raw list to `host`
    cudaEvent_t done = nullptr;            // recorded behind the kernels of spmb_search_launch
    bool done_armed = false;
    // ordering enqueued with the search itself (a list that was long last time): buffers over `ord_cap` slots
    bool expect_many = false;              // the last spmb_search_finish on this list saw >= SPMB_ORDER_AT_LAUNCH_MIN raw hits
    bool ordered = false;                  // the pending launch sorted (and deduplicated) its hits into ord_out
    int ord_drop = 0;
    unsigned long long ord_cap = 0;
    unsigned long long *ord_k0 = nullptr, *ord_k1 = nullptr;
    uint32_t *ord_i0 = nullptr, *ord_i1 = nullptr, *ord_isel = nullptr;
    uint8_t *ord_flag = nullptr;
    spmb_hit *ord_out = nullptr;
    void *ord_tmp = nullptr;
    size_t ord_tmp_bytes = 0;
    spmb_hit *pinned = nullptr;            // pinned landing area of long lists; `view` points here when in use
    unsigned long long pinned_cap = 0;
    const spmb_hit *view = nullptr;
};

enum { SPMB_STREAM_SLOTS = 3 }; // chunks in flight per stream (one being collected, two in the upload pool's queue)

struct spmb_stream {
    CtxRef ctx;
    const spmb_pattern *pat = nullptr;
    uint64_t offset = 0;            // symbols consumed so far
    uint32_t keep = 0;              // symbols carried from feed to feed (window_size-1; Myers: at least 2m for `errors`)
    uint32_t tail_len = 0;          // symbols in d_tail right now
    uint2 *d_tail = nullptr;        // the carried symbols as lane-chunk words, on the device
    uint32_t *d_err = nullptr;      // stream_score_kernel's output
    std::vector<uint8_t> host_tail; // the same symbols for capture() (copied from the caller's chunks, never searched)
    // two chunks may be in flight (spmb_stream_submit): the upload of one runs beside the search of the other
    struct Slot {
        uint2 *d_words = nullptr;   // [tail | chunk] as lane-chunk words
        size_t words_bytes = 0;
        uint8_t *d_raw = nullptr;   // landing area of short chunks (1 B per symbol)
        size_t raw_bytes = 0;
        spmb_store *store = nullptr;
        spmb_hits *hits = nullptr;
        uint64_t t = 0, n = 0, total = 0;
        uint32_t launches = 0;
        bool busy = false, empty = false;
        // handed in by spmb_stream_submit, turned into device work by spmb_stream_collect
        const uint8_t *host = nullptr;  // the chunk behind the symbols that complete the tail's partial word
        uint64_t bulk = 0;
        uint32_t a = 0;
        StreamHead hd;
        bool job = false;               // the bulk is a job of the upload pool
        uint64_t job_seq = 0;
    } slot[SPMB_STREAM_SLOTS];
    int next_submit = 0, next_collect = 0;
    bool any_busy() const { for (const Slot &x : slot) if (x.busy) return true; return false; }
    // anchored (prefix) streams: the state is the short prefix itself, on the host
    std::vector<uint8_t> tail;
    std::vector<uint8_t> scratch;
};

// ------------------------------------------------------------------------------------------
// helpers
// ------------------------------------------------------------------------------------------
static int bits_for(unsigned long long max_value);
static int set_err(spmb_context *ctx, int code, const char *what, cudaError_t e = cudaSuccess)
{
    if (ctx) {
        ctx->err = what ? what : "";
        if (e != cudaSuccess) { ctx->err += ": "; ctx->err += cudaGetErrorString(e); }
    }
    return code;
}

#define CU(call)                                                                      \
    do {                                                                              \
        cudaError_t e__ = (call);                                                     \
        if (e__ != cudaSuccess) {                                                     \
            cudaGetLastError();                                                       \
            return set_err(ctx, e__ == cudaErrorMemoryAllocation ? SPMB_ERR_OOM : SPMB_ERR_CUDA, #call, e__); \
        }                                                                             \
    } while (0)

// every entry point: select the context's device and drop any stale (non-sticky) error state so that
// a later cudaGetLastError() reports only this call's launches
#define ENTER(ctx)                                   \
    do {                                             \
        if ((ctx)->closed) return SPMB_ERR_ARG;      \
        CU(cudaSetDevice((ctx)->device));            \
        cudaGetLastError();                          \
    } while (0)

extern "C" int spmb_version(void) { return SPMB_VERSION; }

extern "C" const char *spmb_status_string(int s)
{
    switch (s) {
    case SPMB_OK: return "ok";
    case SPMB_ERR_ARG: return "bad argument";
    case SPMB_ERR_CUDA: return "CUDA error";
    case SPMB_ERR_OOM: return "out of device memory";
    case SPMB_ERR_UNSUPPORTED: return "unsupported pattern / geometry";
    case SPMB_ERR_NO_DEVICE: return "no CUDA device (this library has no CPU path)";
    default: return "unknown";
    }
}

// ------------------------------------------------------------------------------------------
// context
// ------------------------------------------------------------------------------------------
extern "C" int spmb_context_create(int device, void *stream, spmb_context **out)
{
    if (!out) return SPMB_ERR_ARG;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return SPMB_ERR_NO_DEVICE; }
    if (device < 0 || device >= ndev) return SPMB_ERR_ARG;
    spmb_context *ctx = new (std::nothrow) spmb_context();
    if (!ctx) return SPMB_ERR_OOM;
    ctx->device = device;
    if (cudaSetDevice(device) != cudaSuccess) { delete ctx; return SPMB_ERR_CUDA; }
    if (stream) ctx->stream = (cudaStream_t)stream;
    else {
        if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) { delete ctx; return SPMB_ERR_CUDA; }
        ctx->own_stream = true;
    }
    if (cudaStreamCreateWithFlags(&ctx->aux_stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming) != cudaSuccess) {
        cudaGetLastError();
        if (ctx->aux_stream) cudaStreamDestroy(ctx->aux_stream);
        ctx->aux_stream = nullptr;
    }
    {   // highest priority: the small ordering kernels of a finished pass go ahead of the next pass's search kernels
        int prio_lo = 0, prio_hi = 0;
        if (cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi) != cudaSuccess) { cudaGetLastError(); prio_hi = 0; }
        if (cudaStreamCreateWithPriority(&ctx->copy_stream, cudaStreamNonBlocking, prio_hi) != cudaSuccess) { cudaGetLastError(); ctx->copy_stream = nullptr; }
    }
    cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device);
    {
        cudaMemPool_t pool;
        if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
            unsigned long long thr = ~0ull;
            cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
        }
        cudaGetLastError();
    }
    if (cudaMallocHost(&ctx->h_count, 8 * sizeof(unsigned long long)) != cudaSuccess ||
        cudaMallocHost(&ctx->h_spec, SPEC_HITS * sizeof(spmb_hit)) != cudaSuccess) {
        if (ctx->h_count) cudaFreeHost(ctx->h_count);
        cudaGetLastError();
        delete ctx;
        return SPMB_ERR_CUDA;
    }
    *out = ctx;
    return SPMB_OK;
}

extern "C" int spmb_context_destroy(spmb_context *ctx)
{
    if (!ctx || ctx->closed) return SPMB_OK;
    ctx->closed = true;
    if (ctx->children == 0) context_free(ctx);
    return SPMB_OK;
}

extern "C" void *spmb_context_stream(const spmb_context *ctx) { return ctx ? (void *)ctx->stream : nullptr; }
extern "C" const char *spmb_last_error(const spmb_context *ctx) { return ctx ? ctx->err.c_str() : "null context"; }
extern "C" uint32_t spmb_last_launch_count(const spmb_context *ctx) { return ctx ? ctx->launches : 0; }

extern "C" int spmb_context_set_myers_cutoff(spmb_context *ctx, int enabled)
{
    if (!ctx) return SPMB_ERR_ARG;
    ctx->myers_cutoff = enabled != 0;
    return SPMB_OK;
}

extern "C" int spmb_context_set_upload_threads(spmb_context *ctx, int n_threads)
{
    if (!ctx || n_threads < -1) return SPMB_ERR_ARG;
    if (ctx->upload && ctx->upload->n_workers != n_threads) { // the pool is re-created with the new thread count on the next upload
        if (ctx->upload->in_flight()) return set_err(ctx, SPMB_ERR_ARG, "chunks of a stream are still in flight");
        ENTER(ctx);
        CU(cudaStreamSynchronize(ctx->stream));
        delete ctx->upload;
        ctx->upload = nullptr;
    }
    ctx->upload_threads = n_threads;
    return SPMB_OK;
}

extern "C" int spmb_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

extern "C" int spmb_context_set_shard(spmb_context *ctx, uint32_t shard, uint32_t n_shards)
{
    if (!ctx || n_shards == 0 || shard >= n_shards) return SPMB_ERR_ARG;
    ctx->shard = shard;
    ctx->n_shards = n_shards;
    return SPMB_OK;
}

// ---- large blocks: size classes of 1/16 of a power of two (<= 6 % slack), exact-class reuse, all use is
// ordered on the context's stream (so is the reuse)
static size_t big_class(size_t bytes)
{
    size_t p2 = 1;
    while (p2 * 2 <= bytes) p2 *= 2;
    size_t g = std::max<size_t>(p2 / 16, (size_t)2 << 20);
    return (bytes + g - 1) / g * g;
}
enum : size_t { BIG_MIN = (size_t)16 << 20 };

static cudaError_t big_alloc(spmb_context *ctx, size_t bytes, void **out)
{
    if (bytes < BIG_MIN) return cudaMallocAsync(out, bytes ? bytes : 1, ctx->stream);
    const size_t cls = big_class(bytes);
    for (size_t i = 0; i < ctx->blocks.size(); ++i)
        if (ctx->blocks[i].bytes == cls) {
            *out = ctx->blocks[i].p;
            ctx->blocks_bytes -= cls;
            ctx->blocks.erase(ctx->blocks.begin() + i);
            return cudaSuccess;
        }
    cudaError_t e = cudaMallocAsync(out, cls, ctx->stream);
    if (e != cudaSuccess && !ctx->blocks.empty()) { // make room: give the cached blocks back and retry
        cudaGetLastError();
        for (auto &b : ctx->blocks) cudaFreeAsync(b.p, ctx->stream);
        ctx->blocks.clear();
        ctx->blocks_bytes = 0;
        e = cudaMallocAsync(out, cls, ctx->stream);
    }
    return e;
}

static void big_free(spmb_context *ctx, void *p, size_t bytes)
{
    if (!p) return;
    if (bytes >= BIG_MIN && !ctx->closed) {
        const size_t cls = big_class(bytes);
        if (ctx->blocks_cap == 0) {
            size_t fr = 0, tot = 0;
            if (cudaMemGetInfo(&fr, &tot) != cudaSuccess) { cudaGetLastError(); tot = 0; }
            ctx->blocks_cap = tot / 4 + 1;
        }
        if (ctx->blocks.size() < 16 && ctx->blocks_bytes + cls <= ctx->blocks_cap) {
            ctx->blocks.push_back({p, cls});
            ctx->blocks_bytes += cls;
            return;
        }
    }
    cudaFreeAsync(p, ctx->stream);
}

static void big_release_all(spmb_context *ctx)
{
    for (auto &b : ctx->blocks) cudaFreeAsync(b.p, ctx->stream);
    ctx->blocks.clear();
    ctx->blocks_bytes = 0;
}

// the context's scratch block, at least `bytes` long; contents are only valid within one call
static cudaError_t ctx_scratch(spmb_context *ctx, size_t bytes, uint8_t **out)
{
    if (bytes > ctx->scratch_bytes) {
        cudaError_t e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) return e;
        if (ctx->scratch) cudaFree(ctx->scratch);
        ctx->scratch = nullptr;
        ctx->scratch_bytes = 0;
        size_t want = bytes + bytes / 8;
        e = cudaMalloc(&ctx->scratch, want);
        if (e != cudaSuccess) { cudaGetLastError(); want = bytes; e = cudaMalloc(&ctx->scratch, want); }
        if (e != cudaSuccess) { ctx->scratch = nullptr; return e; }
        ctx->scratch_bytes = want;
    }
    *out = ctx->scratch;
    return cudaSuccess;
}

extern "C" int spmb_context_trim(spmb_context *ctx)
{
    if (!ctx) return SPMB_ERR_ARG;
    ENTER(ctx);
    CU(cudaStreamSynchronize(ctx->stream));
    if (ctx->scratch) cudaFree(ctx->scratch);
    ctx->scratch = nullptr;
    ctx->scratch_bytes = 0;
    delete ctx->upload;
    ctx->upload = nullptr;
    big_release_all(ctx);
    CU(cudaStreamSynchronize(ctx->stream));
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, ctx->device) == cudaSuccess) cudaMemPoolTrimTo(pool, 0);
    cudaGetLastError();
    return SPMB_OK;
}

extern "C" int spmb_context_sync(spmb_context *ctx)
{
    if (!ctx) return SPMB_ERR_ARG;
    ENTER(ctx);
    CU(cudaStreamSynchronize(ctx->stream));
    return SPMB_OK;
}

extern "C" int spmb_device_info(spmb_context *ctx, int *sm_count, int *sm_clock_khz, uint64_t *mem_bytes)
{
    if (!ctx) return SPMB_ERR_ARG;
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, ctx->device));
    if (sm_count) *sm_count = prop.multiProcessorCount;
    if (sm_clock_khz) { int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, ctx->device); *sm_clock_khz = khz; }
    if (mem_bytes) *mem_bytes = (uint64_t)prop.totalGlobalMem;
    return SPMB_OK;
}

// ------------------------------------------------------------------------------------------
// store
// ------------------------------------------------------------------------------------------
static int store_alloc(spmb_context *ctx, uint64_t n, uint32_t row_len, uint32_t mode, spmb_store **out)
{
    if (row_len == 0) row_len = spmb_auto_row_len(n);
    if (row_len % 32 != 0 || row_len > 65536) return set_err(ctx, SPMB_ERR_ARG, "row_len must be a multiple of 32, <= 65536");
    if (n >= (1ull << 41)) return set_err(ctx, SPMB_ERR_UNSUPPORTED, "sequence too long");
    spmb_store *s = new (std::nothrow) spmb_store();
    if (!s) return SPMB_ERR_OOM;
    s->ctx = ctx;
    s->lay = spmb_make_layout(n, row_len, mode);
    s->bytes = (size_t)s->lay.n_cols_alloc * sizeof(uint2);
    cudaError_t e = big_alloc(ctx, s->bytes, (void **)&s->planes);
    if (e != cudaSuccess) { cudaGetLastError(); delete s; return set_err(ctx, SPMB_ERR_OOM, "cudaMallocAsync(store)", e); }
    *out = s;
    return SPMB_OK;
}

// pack_chunks_kernel walks its blocks with a stride of the grid: two resident CTAs per SM, a few rounds of slack
static unsigned pack_grid(const spmb_context *ctx, const SpmbLayout &lay)
{
    const uint64_t blocks = lay.n_cols_alloc / 1024;
    return (unsigned)std::min<uint64_t>(blocks, (uint64_t)ctx->sm_count * 2);
}

template <class Source>
static int store_pack(spmb_context *ctx, spmb_store *s, const Source &src)
{
    // every 1024-column block of the allocation is written (padding blocks become zeros)
    unsigned blocks = (unsigned)(s->lay.n_cols_alloc / 1024);
    pack_kernel<Source><<<blocks, 1024, 0, ctx->stream>>>(src, s->lay, s->planes);
    ctx->launches++;
    CU(cudaGetLastError());
    return SPMB_OK;
}

static int store_from_device_bits(spmb_context *ctx, const uint8_t *dev_ranks, uint64_t n, uint32_t row_len,
                                  uint32_t shift, spmb_store **out)
{
    spmb_store *s = nullptr;
    int rc = store_alloc(ctx, n, row_len, SPMB_MODE_PLAIN, &s);
    if (rc) return rc;
    ByteSource src{dev_ranks, n, shift};
    rc = store_pack(ctx, s, src);
    if (rc) { spmb_store_destroy(s); return rc; }
    *out = s;
    return SPMB_OK;
}

extern "C" int spmb_store_create_from_device(spmb_context *ctx, const uint8_t *dev_ranks, uint64_t n,
                                             uint32_t row_len, spmb_store **out)
{
    if (!ctx || !out || (!dev_ranks && n)) return SPMB_ERR_ARG;
    ENTER(ctx);
    return store_from_device_bits(ctx, dev_ranks, n, row_len, 0, out);
}

// 4-bit alphabets: two plane pairs in the same geometry (bits 1..0 and bits 3..2 of every symbol)
extern "C" int spmb_store_create_wide_from_device(spmb_context *ctx, const uint8_t *dev_ranks, uint64_t n, uint32_t bits,
                                                  uint32_t row_len, spmb_store **out)
{
    if (!ctx || !out || (!dev_ranks && n)) return SPMB_ERR_ARG;
    if (bits == 2) return spmb_store_create_from_device(ctx, dev_ranks, n, row_len, out);
    if (bits != 3 && bits != 4) return set_err(ctx, SPMB_ERR_ARG, "symbol width must be 2, 3 or 4 bits");
    ENTER(ctx);
    spmb_store *lo = nullptr, *hi = nullptr;
    int rc = store_from_device_bits(ctx, dev_ranks, n, row_len, 0, &lo);
    if (rc) return rc;
    rc = store_from_device_bits(ctx, dev_ranks, n, lo->lay.row_len, 2, &hi);
    if (rc) { spmb_store_destroy(lo); return rc; }
    lo->hi = hi;
    lo->sym_bits = 4;
    hi->sym_bits = 4;
    *out = lo;
    return SPMB_OK;
}

extern "C" int spmb_store_create_wide(spmb_context *ctx, const uint8_t *host_ranks, uint64_t n, uint32_t bits,
                                      uint32_t row_len, spmb_store **out)
{
    if (!ctx || !out || (!host_ranks && n)) return SPMB_ERR_ARG;
    if (bits == 2) return spmb_store_create(ctx, host_ranks, n, row_len, out);
    if (bits != 3 && bits != 4) return set_err(ctx, SPMB_ERR_ARG, "symbol width must be 2, 3 or 4 bits");
    ENTER(ctx);
    uint8_t *d = nullptr;
    CU(big_alloc(ctx, n ? n : 1, (void **)&d));
    cudaError_t e = n ? cudaMemcpyAsync(d, host_ranks, n, cudaMemcpyHostToDevice, ctx->stream) : cudaSuccess;
    int rc = SPMB_OK;
    if (e != cudaSuccess) rc = set_err(ctx, SPMB_ERR_CUDA, "H2D haystack", e);
    if (rc == SPMB_OK) rc = spmb_store_create_wide_from_device(ctx, d, n, bits, row_len, out);
    big_free(ctx, d, n ? n : 1);
    e = cudaStreamSynchronize(ctx->stream); // the caller's buffer may be pinned: consumed before we return
    if (rc == SPMB_OK && e != cudaSuccess) {
        spmb_store_destroy(*out);
        *out = nullptr;
        rc = set_err(ctx, SPMB_ERR_CUDA, "store_create_wide", e);
    }
    return rc;
}

extern "C" uint32_t spmb_store_symbol_bits(const spmb_store *s) { return s ? s->sym_bits : 0; }

static int ensure_upload(spmb_context *ctx)
{
    if (ctx->upload) return SPMB_OK;
    UploadPool *up = new (std::nothrow) UploadPool();
    if (!up) return SPMB_ERR_OOM;
    cudaError_t e = up->init(ctx->device, ctx->upload_threads);
    if (e != cudaSuccess) { delete up; cudaGetLastError(); return set_err(ctx, SPMB_ERR_CUDA, "upload pool", e); }
    ctx->upload = up;
    return SPMB_OK;
}

// long host haystacks: chunked upload, raw over the copy engine and 2-bit packed by host threads at once
static int store_create_pipelined(spmb_context *ctx, const uint8_t *host_ranks, uint64_t n, uint32_t row_len,
                                  spmb_store **out)
{
    const bool trace = UploadPool::env_u64("SPMB_UPLOAD_TRACE", 0) != 0;
    auto tnow = [] { return std::chrono::steady_clock::now(); };
    auto tms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
        return std::chrono::duration<double, std::milli>(b - a).count();
    };
    const auto t0 = tnow();
    int rcu = ensure_upload(ctx);
    if (rcu) return rcu;
    spmb_store *s = nullptr;
    int rc = store_alloc(ctx, n, row_len, SPMB_MODE_PLAIN, &s);
    if (rc) return rc;
    uint2 *d_chunks = nullptr;
    const size_t chunk_bytes = (n + 31) / 32 * sizeof(uint2);
    cudaError_t e = big_alloc(ctx, chunk_bytes, (void **)&d_chunks);
    if (e != cudaSuccess) { cudaGetLastError(); spmb_store_destroy(s); return set_err(ctx, SPMB_ERR_OOM, "cudaMallocAsync(chunks)", e); }
    const auto t1 = tnow();
    e = ctx->upload->run(host_ranks, n, d_chunks, ctx->stream, &ctx->launches);
    const auto t2 = tnow();
    if (e == cudaSuccess) {
        // the lane-chunk copy the Myers / journal kernels read is the same words in another order: write it
        // now instead of deriving it from the planes later (skipped when memory is short)
        uint2 *l = nullptr;
        if (big_alloc(ctx, s->bytes, (void **)&l) != cudaSuccess) { cudaGetLastError(); l = nullptr; }
        ChunkSource src{d_chunks, n};
        pack_chunks_kernel<ChunkSource><<<pack_grid(ctx, s->lay), 1024, 0, ctx->stream>>>(src, s->lay, s->planes, l);
        s->lanes = l;
        ctx->launches++;
        e = cudaGetLastError();
    }
    big_free(ctx, d_chunks, chunk_bytes);
    const auto t3 = tnow();
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (trace)
        fprintf(stderr, "[spmb store_create] allocs %.2f ms  upload %.2f  enqueue pack %.2f  sync %.2f\n", tms(t0, t1), tms(t1, t2), tms(t2, t3),
                tms(t3, tnow()));
    if (e != cudaSuccess) { cudaGetLastError(); spmb_store_destroy(s); return set_err(ctx, SPMB_ERR_CUDA, "store upload", e); }
    *out = s;
    return SPMB_OK;
}

extern "C" int spmb_store_create(spmb_context *ctx, const uint8_t *host_ranks, uint64_t n, uint32_t row_len,
                                 spmb_store **out)
{
    if (!ctx || !out || (!host_ranks && n)) return SPMB_ERR_ARG;
    ENTER(ctx);
    if (n >= UploadPool::env_u64("SPMB_UPLOAD_MIN", 32ull << 20) && n > 0)
        return store_create_pipelined(ctx, host_ranks, n, row_len, out);
    uint8_t *d = nullptr;
    CU(big_alloc(ctx, n, (void **)&d));
    cudaError_t e = cudaMemcpyAsync(d, host_ranks, n, cudaMemcpyHostToDevice, ctx->stream);
    int rc = SPMB_OK;
    if (e != cudaSuccess) rc = set_err(ctx, SPMB_ERR_CUDA, "H2D haystack", e);
    if (rc == SPMB_OK) rc = spmb_store_create_from_device(ctx, d, n, row_len, out);
    big_free(ctx, d, n);
    // the caller's buffer may be pinned (truly asynchronous copy): do not return before it is consumed
    e = cudaStreamSynchronize(ctx->stream);
    if (rc == SPMB_OK && e != cudaSuccess) {
        spmb_store_destroy(*out);
        *out = nullptr;
        rc = set_err(ctx, SPMB_ERR_CUDA, "store_create", e);
    }
    return rc;
}

extern "C" int spmb_upload_last_stats(const spmb_context *ctx, uint64_t stats[8])
{
    if (!ctx || !stats) return SPMB_ERR_ARG;
    for (int i = 0; i < 8; ++i) stats[i] = 0;
    if (ctx->upload) {
        stats[0] = ctx->upload->last_raw;
        stats[1] = ctx->upload->last_packed;
        stats[2] = ctx->upload->job_chunk;
        stats[3] = (uint64_t)ctx->upload->n_workers;
        stats[4] = ctx->upload->last_ev_raw;
        stats[5] = ctx->upload->last_ev_packed;
        stats[6] = ctx->upload->last_ev_chunk;
    }
    return SPMB_OK;
}

static int store_ensure_lanes(spmb_context *ctx, const spmb_store *s);

// ref_bias: the events' ref_pos are coordinates of a longer sequence of which `base` holds [ref_bias, ...) (the
// shards of the multi-GPU driver pass slices of ONE caller-owned event table); probe: *probe_der = derived position
// of event `probe_index` (where its inserted segment begins), or the derived length when probe_index == n_ev
static int journaled_pair(spmb_context *ctx, const spmb_store *base, const spmb_journal_event *ev, uint64_t n_ev,
                          const uint8_t *pool, uint64_t pool_len, uint32_t row_len, uint64_t ref_bias,
                          uint64_t probe_index, uint64_t *probe_der, uint32_t pool_shift, spmb_store **out);

int spmb_store_create_journaled_ex(spmb_context *ctx, const spmb_store *base, const spmb_journal_event *ev, uint64_t n_ev,
                                   const uint8_t *pool, uint64_t pool_len, uint32_t row_len, uint64_t ref_bias,
                                   uint64_t probe_index, uint64_t *probe_der, spmb_store **out)
{
    if (!ctx || !base || !out || (n_ev && !ev) || (pool_len && !pool)) return SPMB_ERR_ARG;
    if (base->lay.mode != SPMB_MODE_PLAIN) return set_err(ctx, SPMB_ERR_ARG, "journal base must be a plain store");
    if (base->sym_bits != 4) return journaled_pair(ctx, base, ev, n_ev, pool, pool_len, row_len, ref_bias, probe_index, probe_der, 0, out);
    // a 4-bit alphabet (spm::dna5 / spm::dna15) is two plane pairs of the same geometry: the journal is applied to
    // each, with bits 1..0 / 3..2 of the pool's symbols
    if (!base->hi) return set_err(ctx, SPMB_ERR_ARG, "journal base: incomplete 4-bit store");
    spmb_store *lo = nullptr, *hi = nullptr;
    int rc = journaled_pair(ctx, base, ev, n_ev, pool, pool_len, row_len, ref_bias, probe_index, probe_der, 0, &lo);
    if (rc) return rc;
    rc = journaled_pair(ctx, base->hi, ev, n_ev, pool, pool_len, lo->lay.row_len, ref_bias, probe_index, nullptr, 2, &hi);
    if (rc) { spmb_store_destroy(lo); return rc; }
    lo->hi = hi;
    lo->sym_bits = 4;
    hi->sym_bits = 4;
    *out = lo;
    return SPMB_OK;
}

// one plane pair (2 bits per symbol) of a journaled store
static int journaled_pair(spmb_context *ctx, const spmb_store *base, const spmb_journal_event *ev, uint64_t n_ev,
                          const uint8_t *pool, uint64_t pool_len, uint32_t row_len, uint64_t ref_bias,
                          uint64_t probe_index, uint64_t *probe_der, uint32_t pool_shift, spmb_store **out)
{
    ENTER(ctx);
    cudaStream_t sm = ctx->stream;
    const size_t ne = n_ev ? n_ev : 1;
    spmb_journal_event *d_ev = nullptr;
    long long *d_delta = nullptr, *d_shift = nullptr, *d_total = nullptr;
    uint64_t *d_der = nullptr;
    int64_t *d_blk = nullptr;
    int *d_bad = nullptr;
    uint8_t *d_pool = nullptr;
    uint2 *d_poolw = nullptr, *d_chunks = nullptr; // the pool as 2-bit words; the derived sequence as linear lane-chunk words
    JournalEvRec *d_jrec = nullptr;                // per event: what the fast walk needs (journal_rec_kernel)
    size_t chunk_bytes = 0;
    void *d_tmp = nullptr;
    size_t d_tmp_bytes = 0;
    spmb_store *s = nullptr;
    int rc = SPMB_OK;
    cudaError_t e = cudaSuccess;
    size_t blk_bytes = 0;
    auto cleanup = [&]() { // everything else lives in the context's scratch block
        big_free(ctx, d_blk, blk_bytes); d_blk = nullptr;
        big_free(ctx, d_chunks, chunk_bytes); d_chunks = nullptr;
    };
#define JCU(call) do { if (e == cudaSuccess) e = (call); } while (0)
    const bool trace = UploadPool::env_u64("SPMB_UPLOAD_TRACE", 0) != 0;
    auto tnow = [] { return std::chrono::steady_clock::now(); };
    auto tms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
        return std::chrono::duration<double, std::milli>(b - a).count();
    };
    const auto tj0 = tnow();
    {
        size_t tmp_bytes = 0;
        JCU(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, (long long *)nullptr, (long long *)nullptr, (int64_t)ne, sm));
        auto up = [](size_t x) { return (x + 255) / 256 * 256; };
        const size_t o_ev = 0, o_delta = o_ev + up(ne * sizeof(spmb_journal_event)), o_shift = o_delta + up(ne * 8),
                     o_total = o_shift + up(ne * 8), o_bad = o_total + 256, o_pool = o_bad + 256,
                     o_poolw = o_pool + up(pool_len + 1), o_sym = o_poolw + up(((pool_len + 31) / 32 + 2) * sizeof(uint2)),
                     o_tmp = o_sym + up(ne * sizeof(JournalEvRec)), o_end = o_tmp + up(tmp_bytes + 1);
        uint8_t *sc = nullptr;
        JCU(ctx_scratch(ctx, o_end, &sc));
        if (e != cudaSuccess) { cudaGetLastError(); return set_err(ctx, SPMB_ERR_OOM, "journal scratch", e); }
        d_ev = (spmb_journal_event *)(sc + o_ev);
        d_delta = (long long *)(sc + o_delta);
        d_shift = (long long *)(sc + o_shift);
        d_der = (uint64_t *)d_shift;            // journal_der_kernel turns the shifts into derived positions in place
        d_total = (long long *)(sc + o_total);
        d_bad = (int *)(sc + o_bad);
        d_pool = sc + o_pool;
        d_poolw = (uint2 *)(sc + o_poolw);
        d_jrec = (JournalEvRec *)(sc + o_sym);
        d_tmp = sc + o_tmp;
        d_tmp_bytes = tmp_bytes;
    }
    JCU(cudaMemsetAsync(d_total, 0, 16, sm));
    JCU(cudaMemsetAsync(d_bad, 0, 16, sm));
    const auto tj1 = tnow();
    JCU(cudaMemsetAsync(d_poolw, 0, ((pool_len + 31) / 32 + 2) * sizeof(uint2), sm));
    if (pool_len) {
        JCU(cudaMemcpyAsync(d_pool, pool, pool_len, cudaMemcpyHostToDevice, sm));
        if (e == cudaSuccess) { // inserted runs are fetched like base runs: 32 symbols per word pair
            bytes_to_chunks_kernel<<<(unsigned)(((pool_len + 31) / 32 + 255) / 256), 256, 0, sm>>>(d_pool, pool_len, d_poolw, pool_shift);
            ctx->launches++;
        }
    }
    // long event tables take the upload queue (copy engine + host threads that shrink 24-byte events to 8)
    if (e == cudaSuccess && n_ev * sizeof(spmb_journal_event) >= UploadPool::env_u64("SPMB_UPLOAD_MIN", 32ull << 20) &&
        ensure_upload(ctx) == SPMB_OK)
        e = ctx->upload->run_events(ev, n_ev, pool_len, d_ev, sm, &ctx->launches);
    else if (n_ev) JCU(cudaMemcpyAsync(d_ev, ev, n_ev * sizeof(spmb_journal_event), cudaMemcpyHostToDevice, sm));
    const auto tj2 = tnow();
    if (e == cudaSuccess && n_ev) {
        unsigned blocks = (unsigned)((n_ev + 255) / 256);
        if (ref_bias) { journal_bias_kernel<<<blocks, 256, 0, sm>>>(d_ev, n_ev, ref_bias, d_bad); ctx->launches++; }
        journal_delta_kernel<<<blocks, 256, 0, sm>>>(d_ev, n_ev, base->lay.n, pool_len, d_delta, d_bad);
        size_t tmp_bytes = d_tmp_bytes;
        JCU(cub::DeviceScan::ExclusiveSum(d_tmp, tmp_bytes, d_delta, d_shift, (int64_t)n_ev, sm));
        journal_der_kernel<<<blocks, 256, 0, sm>>>(d_ev, n_ev, d_shift, d_delta, d_total);
        ctx->launches += 3;
        JCU(cudaGetLastError());
    }
    // derived length (and the validation verdict) must reach the host before the store is sized
    JCU(cudaMemcpyAsync(&ctx->h_count[1], d_total, 8, cudaMemcpyDeviceToHost, sm));
    JCU(cudaMemcpyAsync(&ctx->h_count[2], d_bad, 4, cudaMemcpyDeviceToHost, sm));
    if (probe_der && probe_index < n_ev) JCU(cudaMemcpyAsync(&ctx->h_count[5], d_der + probe_index, 8, cudaMemcpyDeviceToHost, sm));
    JCU(cudaStreamSynchronize(sm));
    if (e != cudaSuccess) { cleanup(); cudaGetLastError(); return set_err(ctx, SPMB_ERR_CUDA, "journal upload/prefix", e); }
    if ((int)(ctx->h_count[2] & 0xffffffffu) != 0) {
        cleanup();
        return set_err(ctx, SPMB_ERR_ARG, "journal events must be sorted, non-overlapping, inside the reference and the pool");
    }
    const auto tj3 = tnow();
    const long long total_shift = (long long)ctx->h_count[1];
    const uint64_t n_der = (uint64_t)((long long)base->lay.n + total_shift);
    if (probe_der) *probe_der = probe_index < n_ev ? (uint64_t)ctx->h_count[5] : n_der;
    rc = store_alloc(ctx, n_der, row_len, SPMB_MODE_PLAIN, &s);
    if (rc) { cleanup(); return rc; }
    const uint64_t n_blk = (n_der >> 10) + 2;
    blk_bytes = n_blk * 8;
    JCU(big_alloc(ctx, blk_bytes, (void **)&d_blk));
    if (e == cudaSuccess) {
        journal_blkidx_kernel<<<(unsigned)((n_blk + 255) / 256), 256, 0, sm>>>(d_der, n_ev, n_blk, d_blk);
        ctx->launches++;
        // the gather reads the base through its lane chunks; the derived store gets its own for free
        rc = store_ensure_lanes(ctx, base);
        uint2 *l = nullptr;
        if (rc == SPMB_OK && big_alloc(ctx, s->bytes, (void **)&l) != cudaSuccess) { cudaGetLastError(); l = nullptr; }
        chunk_bytes = ((n_der + 31) / 32 + 2) * sizeof(uint2);
        if (rc == SPMB_OK) JCU(big_alloc(ctx, chunk_bytes, (void **)&d_chunks));
        if (rc == SPMB_OK && e == cudaSuccess) {
            if (n_ev) {
                journal_rec_kernel<<<(unsigned)((n_ev + 255) / 256), 256, 0, sm>>>(d_ev, d_der, n_ev, d_poolw, d_jrec);
                ctx->launches++;
            }
            JournalDeriveArgs ja{base->lanes, base->lay, d_ev, d_der, d_blk, d_poolw, d_jrec, n_ev, n_der, d_chunks};
            const uint64_t warps = (n_der + 1023) >> 10;
            const unsigned grid = (unsigned)std::min<uint64_t>((warps + 7) / 8, (uint64_t)ctx->sm_count * 32);
            if (n_der) journal_derive_kernel<<<grid, 256, 0, sm>>>(ja);
            ChunkSource src{d_chunks, n_der};
            pack_chunks_kernel<ChunkSource><<<pack_grid(ctx, s->lay), 1024, 0, sm>>>(src, s->lay, s->planes, l);
            ctx->launches += 2;
            JCU(cudaGetLastError());
            s->lanes = l;
        } else if (l) {
            big_free(ctx, l, s->bytes);
        }
    }
    const auto tj4 = tnow();
    cleanup();
    JCU(cudaStreamSynchronize(sm));
    if (trace)
        fprintf(stderr, "[spmb journal] allocs %.2f ms  events+pool up %.2f  prefix (to host) %.2f  enqueue materialise %.2f  free+sync %.2f\n",
                tms(tj0, tj1), tms(tj1, tj2), tms(tj2, tj3), tms(tj3, tj4), tms(tj4, tnow()));
#undef JCU
    if (e != cudaSuccess || rc) {
        spmb_store_destroy(s);
        cudaGetLastError();
        return rc ? rc : set_err(ctx, SPMB_ERR_CUDA, "journal materialise", e);
    }
    *out = s;
    return SPMB_OK;
}

extern "C" int spmb_store_create_journaled(spmb_context *ctx, const spmb_store *base,
                                           const spmb_journal_event *ev, uint64_t n_ev,
                                           const uint8_t *pool, uint64_t pool_len, uint32_t row_len,
                                           spmb_store **out)
{
    return spmb_store_create_journaled_ex(ctx, base, ev, n_ev, pool, pool_len, row_len, 0, 0, nullptr, out);
}

extern "C" int spmb_store_destroy(spmb_store *s)
{
    if (!s) return SPMB_OK;
    if (s->hi) spmb_store_destroy(s->hi);
    if (s->ctx) cudaSetDevice(s->ctx->device);
    if (s->ctx) big_free(s->ctx, s->planes, s->bytes); else cudaFree(s->planes);
    if (s->lanes) { if (s->ctx) big_free(s->ctx, s->lanes, s->bytes); else cudaFree(s->lanes); }
    cudaFree(s->row_len);
    cudaFree(s->row_seq);
    cudaFree(s->row_own);
    delete s;
    return SPMB_OK;
}

extern "C" uint64_t spmb_store_length(const spmb_store *s) { return s ? s->lay.n : 0; }
extern "C" uint32_t spmb_store_row_len(const spmb_store *s) { return s ? s->lay.row_len : 0; }
extern "C" uint64_t spmb_store_device_bytes(const spmb_store *s)
{
    return s ? s->bytes * (s->lanes ? 2 : 1) + spmb_store_device_bytes(s->hi) : 0;
}

// the Myers kernels read the lane-chunk layout: build it once per store, stream-ordered
static int store_ensure_lanes(spmb_context *ctx, const spmb_store *s)
{
    if (s->lanes) return SPMB_OK;
    uint2 *l = nullptr;
    cudaError_t e = big_alloc(ctx, s->bytes, (void **)&l);
    if (e != cudaSuccess) { cudaGetLastError(); return set_err(ctx, SPMB_ERR_OOM, "cudaMallocAsync(lanes)", e); }
    const uint64_t n_groups = s->lay.n_cols_alloc / 32;
    lanes_kernel<<<(unsigned)((n_groups + 7) / 8), 256, 0, ctx->stream>>>(s->planes, n_groups, l);
    ctx->launches++;
    e = cudaGetLastError();
    if (e != cudaSuccess) { big_free(ctx, l, s->bytes); return set_err(ctx, SPMB_ERR_CUDA, "lanes_kernel launch", e); }
    s->lanes = l;
    return SPMB_OK;
}

extern "C" int spmb_store_read(spmb_context *ctx, const spmb_store *s, uint64_t begin, uint64_t n, uint8_t *host_out)
{
    if (!ctx || !s || (!host_out && n)) return SPMB_ERR_ARG;
    if (s->lay.mode != SPMB_MODE_PLAIN) return set_err(ctx, SPMB_ERR_ARG, "read: plain stores only");
    if (begin + n > s->lay.n) return set_err(ctx, SPMB_ERR_ARG, "read range outside the sequence");
    if (n == 0) return SPMB_OK;
    ENTER(ctx);
    uint8_t *d = nullptr;
    CU(cudaMalloc(&d, n));
    unsigned blocks = (unsigned)((n + 255) / 256);
    unpack_kernel<<<blocks, 256, 0, ctx->stream>>>(s->planes, s->lay, begin, n, d);
    ctx->launches++;
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(host_out, d, n, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d);
    if (e != cudaSuccess) return set_err(ctx, SPMB_ERR_CUDA, "store_read", e);
    if (s->hi) { // 4-bit alphabet: bits 3..2 come from the second plane pair
        std::vector<uint8_t> top(n);
        int rc = spmb_store_read(ctx, s->hi, begin, n, top.data());
        if (rc) return rc;
        for (uint64_t i = 0; i < n; ++i) host_out[i] |= (uint8_t)(top[i] << 2);
    }
    return SPMB_OK;
}

// ------------------------------------------------------------------------------------------
// patterns
// ------------------------------------------------------------------------------------------
// Eq tables + meta of a Myers needle set (direct kernel, windowed verification, prefix matcher)
static cudaError_t pattern_upload_myers(spmb_context *ctx, spmb_pattern *p, uint32_t k)
{
    const uint32_t n_pat = p->n_pat;
    const int NWP = p->NW <= 4 ? 4 : 8;
    const int NSYM = p->sym_bits == 4 ? 16 : 4;
    std::vector<uint32_t> eq((size_t)n_pat * NSYM * NWP, 0);
    std::vector<MyersPatMeta> meta(n_pat);
    for (uint32_t i = 0; i < n_pat; ++i) {
        const uint32_t m = p->len[i];
        const uint8_t *pr = p->ranks.data() + p->off[i];
        uint32_t *e4 = eq.data() + (size_t)i * NSYM * NWP;
        for (uint32_t j = 0; j < (uint32_t)p->NW * 32; ++j) {
            for (int c = 0; c < NSYM; ++c) {
                bool match = (j < m) ? (pr[j] == c) : true; // wildcard padding rows
                if (match) e4[c * NWP + j / 32] |= 1u << (j % 32);
            }
        }
        meta[i].m = m; meta[i].k = k; meta[i].pad = (uint32_t)p->NW * 32 - m;
        meta[i].pattern_id = p->ids.empty() ? i : p->ids[i];
    }
    p->warm = (p->max_m + k - 1 + 31) / 32 * 32;
    cudaError_t e = cudaMalloc(&p->d_eq, eq.size() * 4);
    if (e == cudaSuccess) e = cudaMalloc(&p->d_meta, meta.size() * sizeof(MyersPatMeta));
    if (e == cudaSuccess) e = cudaMemcpyAsync(p->d_eq, eq.data(), eq.size() * 4, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(p->d_meta, meta.data(), meta.size() * sizeof(MyersPatMeta), cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream); // host vectors go out of scope
    return e;
}

template <class T>
static cudaError_t upload_vec(spmb_context *ctx, T **d, const std::vector<T> &v)
{
    cudaError_t e = cudaMalloc(d, (v.size() ? v.size() : 1) * sizeof(T));
    if (e == cudaSuccess && !v.empty()) e = cudaMemcpyAsync(*d, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, ctx->stream);
    return e;
}

// seed table of the pigeonhole filter -> device
static cudaError_t pattern_upload_pigeon(spmb_context *ctx, spmb_pattern *p,
                                         const std::vector<std::pair<uint32_t, uint32_t>> &seeds, uint32_t q)
{
    p->pg.build(p->ranks, p->off, seeds, q);
    cudaError_t e = upload_vec(ctx, &p->d_pg_bitmap, p->pg.bitmap);
    if (e == cudaSuccess) e = upload_vec(ctx, &p->d_pg_bitmap2, p->pg.bitmap2);
    if (e == cudaSuccess) e = upload_vec(ctx, &p->d_pg_tab, p->pg.tab);
    if (e == cudaSuccess) e = upload_vec(ctx, &p->d_pg_ufirst, p->pg.ufirst);
    if (e == cudaSuccess) e = upload_vec(ctx, &p->d_pg_pat, p->pg.seed_pat);
    if (e == cudaSuccess) e = upload_vec(ctx, &p->d_pg_soff, p->pg.seed_off);
    if (e == cudaSuccess) e = upload_vec(ctx, &p->d_off, p->off);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    return e;
}

// The exact kernel rejects alignments on ONE window of 8 needle symbols (exact.cuh).  Which window: the first one
// costs nothing on random text, but real genomes are full of low-complexity runs (poly-A, dinucleotide repeats,
// satellites) -- a needle that starts inside one would send every alignment of every such run to the slow
// warp-cooperative verification.  So the host takes the LEAST REPETITIVE window: most distinct 3-mers, then most
// distinct 2-mers, then the leftmost (SPMB_EXACT_WINDOW=first restores offset 0 for A/B runs).
static uint32_t exact_window_offset(const uint8_t *pat, uint32_t m)
{
    if (m <= 8) return 0;
    static const bool first = [] { const char *e = getenv("SPMB_EXACT_WINDOW"); return e && e[0] == 'f'; }();
    if (first) return 0;
    uint32_t best = 0;
    int best_score = -1;
    const uint32_t last = std::min<uint32_t>(m - 8, 4096);
    for (uint32_t o = 0; o <= last; ++o) {
        uint32_t k3[6], k2[7];
        int d3 = 0, d2 = 0;
        for (uint32_t j = 0; j + 1 < 8; ++j) {
            const uint32_t c2 = ((uint32_t)pat[o + j] << 8) | pat[o + j + 1];
            bool dup = false;
            for (int t = 0; t < d2; ++t) dup |= k2[t] == c2;
            if (!dup) k2[d2++] = c2;
            if (j + 2 < 8) {
                const uint32_t c3 = (c2 << 8) | pat[o + j + 2];
                dup = false;
                for (int t = 0; t < d3; ++t) dup |= k3[t] == c3;
                if (!dup) k3[d3++] = c3;
            }
        }
        const int score = d3 * 16 + d2;
        if (score > best_score) { best_score = score; best = o; }
        if (score == 6 * 16 + 7) break; // all distinct: nothing better to find
    }
    return best;
}

// shortest q for which the seed filter replaces the direct Myers kernel: 4^q must dwarf the number of
// seeds (a 12-mer has 1.7e7 values)
static const uint32_t SPMB_FILTER_MIN_Q = 12;

static const uint32_t SPMB_CHAIN_PREFIX = 256; // symbols of a long needle the one-lane kernels filter on

static int pattern_build(spmb_context *ctx, int alg, const uint8_t *ranks, const uint32_t *lengths, uint32_t n_pat,
                         uint32_t k, double error_rate, spmb_pattern **out, uint32_t sym_bits = 2, const uint32_t *ids = nullptr);

// A Myers set with needles of different word counts, or longer than 256 symbols: one child set per word count
// (searched by the one-lane kernels as before), and for the long needles a child over their first 256 symbols
// whose hits are the candidates of the lane-chained verification (myers_chain.cuh).  The reference takes any
// needle (myers_matcher.hpp:37-43).
static int pattern_build_buckets(spmb_context *ctx, spmb_pattern *p, uint32_t k)
{
    std::vector<uint32_t> group[9]; // [1..8]: needles of that many words; [0]: longer ones
    for (uint32_t i = 0; i < p->n_pat; ++i) {
        if (p->len[i] == 0) continue;
        const uint32_t nw = (p->len[i] + 31) / 32;
        group[nw > 8 ? 0 : nw].push_back(i);
    }
    for (int g = 1; g <= 8; ++g) {
        if (group[g].empty()) continue;
        std::vector<uint8_t> r;
        std::vector<uint32_t> l, id;
        for (uint32_t i : group[g]) {
            r.insert(r.end(), p->ranks.begin() + p->off[i], p->ranks.begin() + p->off[i] + p->len[i]);
            l.push_back(p->len[i]);
            id.push_back(p->ids.empty() ? i : p->ids[i]);
        }
        spmb_pattern *c = nullptr;
        int rc = pattern_build(ctx, SPMB_ALG_MYERS, r.data(), l.data(), (uint32_t)l.size(), k, 0.0, &c, p->sym_bits, id.data());
        if (rc) return rc;
        p->subs.push_back(c);
    }
    if (group[0].empty()) return SPMB_OK;
    if (p->sym_bits != 2) return set_err(ctx, SPMB_ERR_UNSUPPORTED, "4-bit alphabets: Myers needles up to 256 symbols");
    const uint32_t words = (p->max_m + 31) / 32;
    const int wpl = words <= 32 ? 1 : words <= 64 ? 2 : words <= 128 ? 4 : words <= 256 ? 8 : 0;
    if (!wpl) return set_err(ctx, SPMB_ERR_UNSUPPORTED, "Myers needle longer than 8192 symbols");
    p->chain_wpl = wpl;
    p->n_long = (uint32_t)group[0].size();
    const uint32_t NWT = 32u * (uint32_t)wpl;
    std::vector<uint32_t> eq((size_t)p->n_long * 4 * NWT, 0);
    std::vector<MyersPatMeta> meta(p->n_long);
    std::vector<uint8_t> r;
    std::vector<uint32_t> l, id;
    for (uint32_t q = 0; q < p->n_long; ++q) {
        const uint32_t i = group[0][q];
        const uint8_t *pr = p->ranks.data() + p->off[i];
        uint32_t *e = eq.data() + (size_t)q * 4 * NWT;
        for (uint32_t j = 0; j < p->len[i]; ++j) e[(pr[j] & 3u) * NWT + j / 32] |= 1u << (j % 32);
        meta[q].m = p->len[i]; meta[q].k = k; meta[q].pad = 0;
        meta[q].pattern_id = p->ids.empty() ? i : p->ids[i];
        r.insert(r.end(), pr, pr + SPMB_CHAIN_PREFIX);
        l.push_back(SPMB_CHAIN_PREFIX);
        id.push_back(q); // a candidate names its needle by the index into the chain tables
    }
    int rc = pattern_build(ctx, SPMB_ALG_MYERS, r.data(), l.data(), p->n_long, k, 0.0, &p->long_prefix, 2, id.data());
    if (rc) return rc;
    cudaError_t e = upload_vec(ctx, &p->d_chain_eq, eq);
    if (e == cudaSuccess) e = upload_vec(ctx, &p->d_chain_meta, meta);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) return set_err(ctx, SPMB_ERR_CUDA, "chain table upload", e);
    return SPMB_OK;
}

static int pattern_build(spmb_context *ctx, int alg, const uint8_t *ranks, const uint32_t *lengths, uint32_t n_pat,
                         uint32_t k, double error_rate, spmb_pattern **out, uint32_t sym_bits, const uint32_t *ids)
{
    if (!ctx || !out || !lengths || n_pat == 0) return SPMB_ERR_ARG;
    if (sym_bits == 3) sym_bits = 4; // dna5 travels in the 4-bit layout
    if (sym_bits != 2 && sym_bits != 4) return set_err(ctx, SPMB_ERR_ARG, "symbol width must be 2, 3 or 4 bits");
    if (sym_bits == 4 && alg != SPMB_ALG_HORSPOOL && alg != SPMB_ALG_SHIFTOR && alg != SPMB_ALG_MYERS)
        return set_err(ctx, SPMB_ERR_UNSUPPORTED, "4-bit alphabets: Horspool, Shift-Or and Myers only");
    if (alg < SPMB_ALG_HORSPOOL || alg > SPMB_ALG_MYERS_PREFIX) return SPMB_ERR_ARG;
    if (alg == SPMB_ALG_PIGEONHOLE && !(error_rate >= 0.0 && error_rate < 1.0)) return SPMB_ERR_ARG;
    ENTER(ctx);
    const bool myers_like = (alg == SPMB_ALG_MYERS || alg == SPMB_ALG_MYERS_FILTERED || alg == SPMB_ALG_MYERS_PREFIX);
    spmb_pattern *p = new (std::nothrow) spmb_pattern();
    if (!p) return SPMB_ERR_OOM;
    p->ctx = ctx; p->alg = alg; p->n_pat = n_pat; p->k = myers_like ? k : 0;
    p->sym_bits = sym_bits;
    uint64_t total = 0;
    p->max_m = 0; p->min_m = 0xffffffffu;
    for (uint32_t i = 0; i < n_pat; ++i) {
        p->off.push_back((uint32_t)total);
        p->len.push_back(lengths[i]);
        total += lengths[i];
        p->max_m = std::max(p->max_m, lengths[i]);
        p->min_m = std::min(p->min_m, lengths[i]);
    }
    if (total && !ranks) { delete p; return SPMB_ERR_ARG; }
    if (total >= (1ull << 31)) { delete p; return set_err(ctx, SPMB_ERR_UNSUPPORTED, "pattern set too large"); }
    p->ranks.assign(ranks, ranks + total);
    for (auto &c : p->ranks) c &= (sym_bits == 4 ? 15 : 3);
    if (ids) p->ids.assign(ids, ids + n_pat);

    cudaError_t e = cudaMalloc(&p->d_ranks, total ? total : 1);
    if (e == cudaSuccess && total) e = cudaMemcpyAsync(p->d_ranks, p->ranks.data(), total, cudaMemcpyHostToDevice, ctx->stream);

    if (myers_like) {
        // an empty needle matches nothing (window_size 0): keep NW = 0 and never launch
        const bool too_long = p->max_m > SPMB_CHAIN_PREFIX;
        if (too_long && alg == SPMB_ALG_MYERS_PREFIX) {
            spmb_pattern_destroy(p);
            return set_err(ctx, SPMB_ERR_UNSUPPORTED, "anchored Myers: needles up to 256 symbols");
        }
        p->NW = (int)((std::min<uint32_t>(p->max_m, SPMB_CHAIN_PREFIX) + 31) / 32);
        // the filtered path verifies needle by needle and reads the score at row m-1, so a set may mix
        // word counts there; the direct kernel wants one word count per launch
        bool mixed = false;
        for (uint32_t i = 0; i < n_pat; ++i)
            if (p->len[i] && (int)((p->len[i] + 31) / 32) != p->NW) mixed = true;
        std::vector<std::pair<uint32_t, uint32_t>> seeds;
        uint32_t q = 0;
        if (alg == SPMB_ALG_MYERS_FILTERED && p->min_m > 0 && !too_long) {
            const uint32_t L = p->min_m / (k + 1); // pigeonhole piece length
            if (L >= SPMB_FILTER_MIN_Q) {
                p->filtered = true;
                // sample every step-th text position and index the step sub-seeds of length L-step+1 of each
                // piece instead (pigeon.cuh): worth it while the sub-seeds stay selective (> 16 symbols) and
                // the shared-memory bitmap (2^20 bits, two per seed) does not fill up
                const uint64_t pieces = (uint64_t)n_pat * (k + 1);
                uint32_t step = 1;
                if (L >= 20 + 3 && pieces * 4 <= 200000) step = 4;
                else if (L >= 18 + 1 && pieces * 2 <= 200000) step = 2;
                if (const char *fs = getenv("SPMB_FILTER_STEP")) { // A/B switch for profiling
                    const uint32_t v = (uint32_t)atoi(fs);
                    if ((v == 1 || v == 2 || v == 4) && L >= 16 + v) step = v;
                }
                p->filter_step = step;
                q = L - step + 1;
                for (uint32_t i = 0; i < n_pat; ++i)
                    for (uint32_t piece = 0; piece <= k; ++piece)
                        for (uint32_t d = 0; d < step; ++d) seeds.push_back({i, piece * L + d});
            }
        }
        if ((too_long || mixed) && !(p->filtered || alg == SPMB_ALG_MYERS_PREFIX)) {
            if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
            int rc = (e == cudaSuccess) ? pattern_build_buckets(ctx, p, k) : set_err(ctx, SPMB_ERR_CUDA, "pattern upload", e);
            if (rc) { spmb_pattern_destroy(p); return rc; }
        } else if (p->NW > 0) {
            if (e == cudaSuccess) e = pattern_upload_myers(ctx, p, k);
            if (e == cudaSuccess && p->filtered) e = pattern_upload_pigeon(ctx, p, seeds, q);
            // A/B switch: the lane-chained scan (myers_chain_scan_kernel) over whole rows instead of the one-lane kernels
            static const bool chain_ab = getenv("SPMB_MYERS_CHAIN") != nullptr;
            if (e == cudaSuccess && chain_ab && alg == SPMB_ALG_MYERS && sym_bits == 2 && !ids) {
                const uint32_t NWT = 32;
                std::vector<uint32_t> ceq((size_t)n_pat * 4 * NWT, 0);
                std::vector<MyersPatMeta> cm(n_pat);
                for (uint32_t i = 0; i < n_pat; ++i) {
                    const uint8_t *pr = p->ranks.data() + p->off[i];
                    for (uint32_t j = 0; j < p->len[i]; ++j) ceq[(size_t)i * 4 * NWT + (pr[j] & 3u) * NWT + j / 32] |= 1u << (j % 32);
                    cm[i].m = p->len[i]; cm[i].k = k; cm[i].pad = 0; cm[i].pattern_id = i;
                }
                e = upload_vec(ctx, &p->d_chain_eq, ceq);
                if (e == cudaSuccess) e = upload_vec(ctx, &p->d_chain_meta, cm);
                if (e == cudaSuccess) p->chain_wpl = 1;
            }
        }
    } else if (alg == SPMB_ALG_PIGEONHOLE) {
        // SeqAn2 _patternInit(Pattern<TIndex, Pigeonhole<TSpec>>, errorRate) (UNVERIFIED-UPSTREAM; pinned by
        // the golden vectors of pigeonhole_matcher_test.cpp): delta = len / (floor(rate*len) + 1) per needle,
        // deltas < 3 ignored, q = the smallest one, q-grams every q symbols
        uint32_t q = 0xffffffffu;
        for (uint32_t i = 0; i < n_pat; ++i) {
            const uint32_t len = p->len[i];
            if (len == 0) continue;
            const uint32_t errors = (uint32_t)(error_rate * (double)len);
            const uint32_t delta = len / (errors + 1);
            if (delta < 3) continue;
            q = std::min(q, delta);
        }
        if (q == 0xffffffffu) {
            // no needle yields a usable q-gram: nothing can be found (window_size 0)
            p->max_m = 0;
        } else {
            std::vector<std::pair<uint32_t, uint32_t>> seeds;
            for (uint32_t i = 0; i < n_pat; ++i)
                for (uint32_t o = 0; o + q <= p->len[i]; o += q) seeds.push_back({i, o});
            p->filtered = true;
            if (e == cudaSuccess) e = pattern_upload_pigeon(ctx, p, seeds, q);
        }
    } else {
        for (uint32_t i = 0; i < n_pat; ++i) {
            uint32_t fs = 0, fh = 0;
            const uint32_t J = std::min<uint32_t>(p->len[i], 8);
            const uint32_t fo = exact_window_offset(p->ranks.data() + p->off[i], p->len[i]);
            for (uint32_t j = 0; j < J; ++j) {
                fs |= (uint32_t)(p->ranks[p->off[i] + fo + j] & 3) << (2 * j);
                fh |= (uint32_t)(p->ranks[p->off[i] + fo + j] >> 2) << (2 * j);
            }
            p->filt.push_back(fs);
            p->filt_hi.push_back(fh);
            p->filt_off.push_back(fo);
        }
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) { spmb_pattern_destroy(p); return set_err(ctx, SPMB_ERR_CUDA, "pattern upload", e); }
    *out = p;
    return SPMB_OK;
}

extern "C" int spmb_pattern_set_create(spmb_context *ctx, int alg, const uint8_t *ranks, const uint32_t *lengths,
                                       uint32_t n_pat, uint32_t k, spmb_pattern **out)
{
    if (alg == SPMB_ALG_PIGEONHOLE) return ctx ? set_err(ctx, SPMB_ERR_ARG, "use spmb_pigeonhole_create") : SPMB_ERR_ARG;
    return pattern_build(ctx, alg, ranks, lengths, n_pat, k, 0.0, out);
}

extern "C" int spmb_pigeonhole_create(spmb_context *ctx, const uint8_t *ranks, const uint32_t *lengths,
                                      uint32_t n_needles, double error_rate, spmb_pattern **out)
{
    return pattern_build(ctx, SPMB_ALG_PIGEONHOLE, ranks, lengths, n_needles, 0, error_rate, out);
}

extern "C" int spmb_pattern_filter_info(const spmb_pattern *p, uint32_t info[5])
{
    if (!p || !info) return SPMB_ERR_ARG;
    info[0] = p->pg.q; info[1] = (uint32_t)p->pg.seed_pat.size(); info[2] = (uint32_t)p->pg.n_unique;
    info[3] = p->pg.B; info[4] = p->filtered ? p->filter_step : 0u;
    return SPMB_OK;
}

extern "C" int spmb_pattern_create(spmb_context *ctx, int alg, const uint8_t *ranks, uint32_t m, uint32_t k,
                                   spmb_pattern **out)
{
    return spmb_pattern_set_create(ctx, alg, ranks, &m, 1, k, out);
}

extern "C" int spmb_pattern_create_wide(spmb_context *ctx, int alg, const uint8_t *ranks, uint32_t m, uint32_t k,
                                        uint32_t bits, spmb_pattern **out)
{
    if (alg == SPMB_ALG_PIGEONHOLE) return ctx ? set_err(ctx, SPMB_ERR_UNSUPPORTED, "4-bit alphabets: no pigeonhole matcher") : SPMB_ERR_ARG;
    return pattern_build(ctx, alg, ranks, &m, 1, k, 0.0, out, bits);
}

extern "C" uint32_t spmb_pattern_symbol_bits(const spmb_pattern *p) { return p ? p->sym_bits : 0; }

extern "C" int spmb_pattern_destroy(spmb_pattern *p)
{
    if (!p) return SPMB_OK;
    if (p->ctx) cudaSetDevice(p->ctx->device);
    cudaFree(p->d_ranks);
    cudaFree(p->d_eq);
    cudaFree(p->d_meta);
    cudaFree(p->d_off);
    cudaFree(p->d_pg_bitmap); cudaFree(p->d_pg_bitmap2); cudaFree(p->d_pg_tab);
    cudaFree(p->d_pg_ufirst); cudaFree(p->d_pg_pat); cudaFree(p->d_pg_soff);
    cudaFree(p->d_chain_eq); cudaFree(p->d_chain_meta);
    for (spmb_pattern *c : p->subs) spmb_pattern_destroy(c);
    spmb_pattern_destroy(p->long_prefix);
    delete p;
    return SPMB_OK;
}

extern "C" uint32_t spmb_pattern_window_size(const spmb_pattern *p)
{
    if (!p || p->max_m == 0) return 0;
    if (p->alg == SPMB_ALG_PIGEONHOLE) return p->pg.q;
    return p->max_m + p->k;
}
extern "C" uint32_t spmb_pattern_count(const spmb_pattern *p) { return p ? p->n_pat : 0; }

// ------------------------------------------------------------------------------------------
// hits
// ------------------------------------------------------------------------------------------
static int hits_reserve(spmb_context *ctx, spmb_hits *h, unsigned long long cap)
{
    if (cap <= h->cap) return SPMB_OK;
    cudaFree(h->d_hits);
    h->d_hits = nullptr;
    h->cap = 0;
    CU(cudaMalloc(&h->d_hits, cap * sizeof(spmb_hit)));
    h->cap = cap;
    return SPMB_OK;
}

static int cand_reserve(spmb_context *ctx, spmb_hits *h, unsigned long long cap)
{
    if (cap <= h->cand_cap) return SPMB_OK;
    cudaFree(h->d_cand);
    h->d_cand = nullptr;
    h->cand_cap = 0;
    static_assert(sizeof(spmb_hit) >= sizeof(PigeonCand), "the candidate buffer holds either record");
    CU(cudaMalloc((void **)&h->d_cand, cap * sizeof(spmb_hit)));
    h->cand_cap = cap;
    return SPMB_OK;
}

extern "C" int spmb_hits_create(spmb_context *ctx, uint64_t capacity_hint, spmb_hits **out)
{
    if (!ctx || !out) return SPMB_ERR_ARG;
    ENTER(ctx);
    spmb_hits *h = new (std::nothrow) spmb_hits();
    if (!h) return SPMB_ERR_OOM;
    h->ctx = ctx;
    cudaError_t e = cudaMalloc(&h->d_count, 4 * sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMemset(h->d_count, 0, 4 * sizeof(unsigned long long));
    if (e != cudaSuccess) { delete h; return set_err(ctx, SPMB_ERR_OOM, "cudaMalloc(hit counter)", e); }
    int rc = hits_reserve(ctx, h, capacity_hint ? capacity_hint : (1ull << 16));
    if (rc) { cudaFree(h->d_count); delete h; return rc; }
    if (cudaEventCreateWithFlags(&h->done, cudaEventDisableTiming) != cudaSuccess) { cudaGetLastError(); h->done = nullptr; }
    *out = h;
    return SPMB_OK;
}

extern "C" int spmb_hits_destroy(spmb_hits *h)
{
    if (!h) return SPMB_OK;
    if (h->ctx) cudaSetDevice(h->ctx->device);
    cudaFree(h->d_hits);
    cudaFree(h->d_count);
    cudaFree(h->d_dd_keys);
    cudaFree(h->d_cand);
    cudaFree(h->ord_k0); cudaFree(h->ord_k1); cudaFree(h->ord_i0); cudaFree(h->ord_i1); cudaFree(h->ord_isel);
    cudaFree(h->ord_flag); cudaFree(h->ord_out); cudaFree(h->ord_tmp);
    if (h->pinned) cudaFreeHost(h->pinned);
    if (h->done) cudaEventDestroy(h->done);
    delete h;
    return SPMB_OK;
}

// the multi-GPU driver hands its merged list over through a hit object of device 0
int spmb_hits_set_host(spmb_hits *h, std::vector<spmb_hit> &&list)
{
    if (!h) return SPMB_ERR_ARG;
    h->host = std::move(list);
    h->count = h->host.size();
    h->view = nullptr;
    h->pend_store = nullptr;
    h->pend_pat = nullptr;
    return SPMB_OK;
}

extern "C" uint64_t spmb_hits_count(const spmb_hits *h) { return h ? h->count : 0; }
extern "C" const spmb_hit *spmb_hits_data(const spmb_hits *h)
{
    if (!h) return nullptr;
    if (h->view) return h->count ? h->view : nullptr;
    return h->host.empty() ? nullptr : h->host.data();
}

// ------------------------------------------------------------------------------------------
// search
// ------------------------------------------------------------------------------------------
template <int MODE>
static cudaError_t launch_myers(spmb_context *ctx, const spmb_store *s, const spmb_pattern *p, const MyersArgs &a_in)
{
    (void)s;
    if (a_in.st1 <= a_in.st0) return cudaSuccess;
    // needles of two words and more run the cut-off form (myers_cut.cuh); SPMB_MYERS_FULL=1 keeps the
    // full recurrence for A/B measurements
    static const bool full = [] { const char *e = getenv("SPMB_MYERS_FULL"); return e && *e && *e != '0'; }();
    const bool cut = p->NW >= 2 && !full && ctx->myers_cutoff;
    const int split_env = [] { const char *e = getenv("SPMB_MYERS_SPLIT"); return e ? atoi(e) : 0; }(); // read per launch
    MyersArgs a = a_in;
    a.split = 1;
    if (cut && MODE == SPMB_MODE_PLAIN) {
        // pieces per row (SPMB_MYERS_SPLIT): measured on B200 at 1 / 2 / 3 pieces (6 / 12 / 18 warps per
        // scheduler), the gain per column (6-8 %) only pays for the extra warm-up, so the default is 1
        uint32_t sp = split_env > 0 ? (uint32_t)split_env : 1u;
        const uint32_t max_sp = a.lay.row_len / (4 * (a.warm + 32)); // keep the warm-up below a quarter of a piece
        if (sp > max_sp) sp = max_sp;
        if (sp < 1) sp = 1;
        a.split = sp;
    }
    dim3 grid((unsigned)(((a.st1 - a.st0) * a.split + 3) / 4), p->n_pat);
    if (a.lanes_hi) {
        // 4-bit alphabets (plain stores only): the cut-off kernel for two words and more, the full one for one
        if (MODE != SPMB_MODE_PLAIN) return cudaErrorInvalidValue;
#define SPMB_MYERS_WIDE_CASE(nw)                                                                                  \
    case nw:                                                                                                      \
        if (cut) myers_cut_kernel<(nw < 2 ? 2 : nw), SPMB_MODE_PLAIN, true><<<grid, 128, 0, ctx->stream>>>(a);      \
        else myers_kernel<nw, SPMB_MODE_PLAIN, true><<<grid, 128, 0, ctx->stream>>>(a);                           \
        break;
        switch (p->NW) {
        SPMB_MYERS_WIDE_CASE(1) SPMB_MYERS_WIDE_CASE(2) SPMB_MYERS_WIDE_CASE(3) SPMB_MYERS_WIDE_CASE(4)
        SPMB_MYERS_WIDE_CASE(5) SPMB_MYERS_WIDE_CASE(6) SPMB_MYERS_WIDE_CASE(7) SPMB_MYERS_WIDE_CASE(8)
#undef SPMB_MYERS_WIDE_CASE
        default: return cudaErrorInvalidValue;
        }
        ctx->launches++;
        return cudaGetLastError();
    }
#define SPMB_MYERS_CASE(nw)                                                                       \
    case nw:                                                                                      \
        if (cut) myers_cut_kernel<(nw < 2 ? 2 : nw), MODE><<<grid, 128, 0, ctx->stream>>>(a);      \
        else myers_kernel<nw, MODE><<<grid, 128, 0, ctx->stream>>>(a);                            \
        break;
    switch (p->NW) {
    SPMB_MYERS_CASE(1) SPMB_MYERS_CASE(2) SPMB_MYERS_CASE(3) SPMB_MYERS_CASE(4)
    SPMB_MYERS_CASE(5) SPMB_MYERS_CASE(6) SPMB_MYERS_CASE(7) SPMB_MYERS_CASE(8)
#undef SPMB_MYERS_CASE
    default: return cudaErrorInvalidValue;
    }
    ctx->launches++;
    return cudaGetLastError();
}

static PigeonDev pigeon_dev(const spmb_pattern *p)
{
    PigeonDev d;
    d.bitmap = p->d_pg_bitmap; d.bitmap2 = p->d_pg_bitmap2; d.tab = reinterpret_cast<const ulonglong2 *>(p->d_pg_tab);
    d.ufirst = p->d_pg_ufirst; d.seed_pat = p->d_pg_pat; d.seed_off = p->d_pg_soff;
    d.q = p->pg.q; d.kq = p->pg.kq; d.B = p->pg.B; d.T = p->pg.T; d.B2 = p->pg.B2;
    d.himask = p->pg.himask(); d.lomask = p->pg.lomask();
    return d;
}

template <int MODE, bool WIDE, int STEP>
static cudaError_t launch_filter_kernel(spmb_context *ctx, const PigeonFilterArgs &fa, unsigned grid, size_t smem)
{
    static bool attr_set[16] = {false};
    if (ctx->device < 16 && !attr_set[ctx->device]) {
        cudaError_t e = cudaFuncSetAttribute(pigeon_filter_kernel<MODE, WIDE, STEP>, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024 + 32 * PG_QCAP * 4);
        if (e != cudaSuccess) return e;
        attr_set[ctx->device] = true;
    } else if (ctx->device >= 16) {
        cudaFuncSetAttribute(pigeon_filter_kernel<MODE, WIDE, STEP>, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024 + 32 * PG_QCAP * 4);
    }
    pigeon_filter_kernel<MODE, WIDE, STEP><<<grid, 1024, smem, ctx->stream>>>(fa);
    return cudaGetLastError();
}

template <int MODE>
static cudaError_t launch_verify(spmb_context *ctx, const spmb_pattern *p, const PigeonVerifyArgs &va)
{
    const unsigned grid = (unsigned)ctx->sm_count * 8;
    if (p->alg == SPMB_ALG_PIGEONHOLE) { pigeon_seed_kernel<MODE><<<grid, 128, 0, ctx->stream>>>(va); return cudaGetLastError(); }
    switch (p->NW) {
    case 1: pigeon_verify_kernel<1, MODE><<<grid, 128, 0, ctx->stream>>>(va); break;
    case 2: pigeon_verify_kernel<2, MODE><<<grid, 128, 0, ctx->stream>>>(va); break;
    case 3: pigeon_verify_kernel<3, MODE><<<grid, 128, 0, ctx->stream>>>(va); break;
    case 4: pigeon_verify_kernel<4, MODE><<<grid, 128, 0, ctx->stream>>>(va); break;
    case 5: pigeon_verify_kernel<5, MODE><<<grid, 128, 0, ctx->stream>>>(va); break;
    case 6: pigeon_verify_kernel<6, MODE><<<grid, 128, 0, ctx->stream>>>(va); break;
    case 7: pigeon_verify_kernel<7, MODE><<<grid, 128, 0, ctx->stream>>>(va); break;
    case 8: pigeon_verify_kernel<8, MODE><<<grid, 128, 0, ctx->stream>>>(va); break;
    default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

static int bits_for(unsigned long long max_value);
enum { SPMB_DD_BITS = 17 }; // slots of the duplicate-suppression table (1.5 MB, cleared before every launch)

// seed filter + (seed emission | windowed Myers verification) over the shard [st0, st1) of a store
template <int MODE>
static int enqueue_filtered(spmb_context *ctx, const spmb_store *s, const spmb_pattern *p, spmb_hits *h,
                            uint64_t st0, uint64_t st1, uint32_t seq_id, const HitSink &sink, bool counters_fresh)
{
    if (st1 <= st0 || p->pg.seed_pat.empty()) return SPMB_OK;
    int rc = store_ensure_lanes(ctx, s);
    if (rc) return rc;
    rc = cand_reserve(ctx, h, 1ull << 16);
    if (rc) return rc;
    if (!counters_fresh) CU(cudaMemsetAsync(h->d_count + 1, 0, sizeof(unsigned long long), ctx->stream));
    const uint64_t per_st = 32ull * s->lay.row_len;
    PigeonFilterArgs fa;
    fa.lanes = s->lanes; fa.lay = s->lay; fa.pg = pigeon_dev(p);
    fa.row_len = s->row_len;
    fa.st0 = st0; fa.st1 = st1;
    if (MODE == SPMB_MODE_PLAIN && p->alg == SPMB_ALG_MYERS_FILTERED && st0 > 0) {
        // a hit owned by this shard may hang on a seed that begins up to m+k symbols before it
        const uint64_t back = ((uint64_t)p->max_m + p->k + per_st - 1) / per_st;
        fa.st0 = st0 > back ? st0 - back : 0;
    }
    fa.cand = h->d_cand; fa.cand_count = h->d_count + 1; fa.cand_cap = h->cand_cap;
    // one persistent CTA per SM; every warp takes an equal contiguous share of the chunk-columns (at least ~8 each)
    const uint64_t chunk_cols = (fa.st1 - fa.st0) * (s->lay.row_len / 32);
    unsigned grid = (unsigned)std::min<uint64_t>((uint64_t)ctx->sm_count, (chunk_cols + 255) / 256);
    if (grid == 0) grid = 1;
    const size_t smem = ((size_t)4 << (p->pg.B - 5)) + 32 * PG_QCAP * 4; // bitmap + per-warp survivor queues
    cudaError_t e;
    if (p->pg.kq <= 16) e = launch_filter_kernel<MODE, false, 1>(ctx, fa, grid, smem);
    else if (p->filter_step == 4) e = launch_filter_kernel<MODE, true, 4>(ctx, fa, grid, smem);
    else if (p->filter_step == 2) e = launch_filter_kernel<MODE, true, 2>(ctx, fa, grid, smem);
    else e = launch_filter_kernel<MODE, true, 1>(ctx, fa, grid, smem);
    if (e != cudaSuccess) return set_err(ctx, SPMB_ERR_CUDA, "pigeon_filter_kernel launch", e);
    PigeonVerifyArgs va;
    va.lanes = s->lanes; va.lay = s->lay; va.pg = fa.pg;
    va.cand = h->d_cand; va.counters = h->d_count; va.cand_cap = h->cand_cap;
    va.pat_ranks = p->d_ranks; va.pat_off = p->d_off; va.eq = p->d_eq; va.meta = p->d_meta;
    va.row_len = s->row_len; va.row_seq = s->row_seq; va.row_own = s->row_own;
    va.seq_id = seq_id;
    va.own_lo = st0 * per_st;
    va.own_hi = std::min<uint64_t>(s->lay.n, st1 * per_st);
    va.sink = sink;
    va.dd_keys = nullptr; va.dd_err = nullptr; va.dd_bits = SPMB_DD_BITS; va.dd_end_bits = 0; va.dd_nseq = 1;
    static const bool dd_off = getenv("SPMB_NO_DEDUP") != nullptr; // A/B switch
    if (p->alg == SPMB_ALG_MYERS_FILTERED && !dd_off && h->cap < (1ull << 31) && p->k < 255) {
        const uint64_t n_seq = (MODE == SPMB_MODE_PLAIN) ? 1 : (uint64_t)s->seq_base + s->lay.n_st * 32;
        const uint64_t max_end = (MODE == SPMB_MODE_PLAIN) ? s->lay.n : s->lay.row_len;
        const int eb = bits_for(max_end + 1);
        if (bits_for((uint64_t)p->n_pat * n_seq) + eb <= 62) {
            if (!h->d_dd_keys && cudaMalloc(&h->d_dd_keys, (size_t)12 << SPMB_DD_BITS) != cudaSuccess) { // 8-byte keys, 4-byte minima
                cudaGetLastError();
                h->d_dd_keys = nullptr;
            }
            if (h->d_dd_keys) {
                CU(cudaMemsetAsync(h->d_dd_keys, 0, (size_t)12 << SPMB_DD_BITS, ctx->stream));
                va.dd_keys = h->d_dd_keys; va.dd_err = reinterpret_cast<uint32_t *>(h->d_dd_keys + ((size_t)1 << SPMB_DD_BITS));
                va.dd_end_bits = (uint32_t)eb; va.dd_nseq = n_seq;
            }
        }
    }
    e = launch_verify<MODE>(ctx, p, va);
    if (e != cudaSuccess) return set_err(ctx, SPMB_ERR_CUDA, "pigeon verify launch", e);
    ctx->launches += 2;
    if (va.dd_keys) { // every record that claimed a table slot takes the slot's minimum
        const unsigned fgrid = (unsigned)std::min<unsigned long long>((sink.cap + 255) / 256, (unsigned long long)ctx->sm_count * 4);
        pigeon_fixup_kernel<<<fgrid ? fgrid : 1, 256, 0, ctx->stream>>>(sink.hits, sink.count, sink.cap, va.dd_err);
        ctx->launches++;
        CU(cudaGetLastError());
    }
    return SPMB_OK;
}

template <int NWX>
static void launch_prefix_nw(spmb_context *ctx, const PrefixArgs &pa, unsigned grid)
{
    prefix_kernel<NWX><<<grid, 128, 0, ctx->stream>>>(pa);
}

static int enqueue_prefix(spmb_context *ctx, const spmb_store *s, const spmb_pattern *p, const uint64_t *d_anchors,
                          uint64_t n_anchors, uint32_t seq_id, const HitSink &sink)
{
    if (s->lay.mode != SPMB_MODE_PLAIN) return set_err(ctx, SPMB_ERR_UNSUPPORTED, "prefix matcher: plain stores only");
    if (p->NW == 0 || n_anchors == 0) return SPMB_OK;
    int rc = store_ensure_lanes(ctx, s);
    if (rc) return rc;
    PrefixArgs pa;
    pa.lanes = s->lanes; pa.lay = s->lay; pa.anchors = d_anchors; pa.n_anchors = n_anchors;
    pa.eq = p->d_eq; pa.meta = p->d_meta; pa.n_pat = p->n_pat; pa.seq_id = seq_id; pa.sink = sink;
    const uint64_t tasks = n_anchors * p->n_pat;
    const unsigned grid = (unsigned)std::min<uint64_t>((tasks + 127) / 128, (uint64_t)ctx->sm_count * 16);
    switch (p->NW) {
    case 1: launch_prefix_nw<1>(ctx, pa, grid); break;
    case 2: launch_prefix_nw<2>(ctx, pa, grid); break;
    case 3: launch_prefix_nw<3>(ctx, pa, grid); break;
    case 4: launch_prefix_nw<4>(ctx, pa, grid); break;
    case 5: launch_prefix_nw<5>(ctx, pa, grid); break;
    case 6: launch_prefix_nw<6>(ctx, pa, grid); break;
    case 7: launch_prefix_nw<7>(ctx, pa, grid); break;
    case 8: launch_prefix_nw<8>(ctx, pa, grid); break;
    default: return set_err(ctx, SPMB_ERR_UNSUPPORTED, "prefix matcher: needle too long");
    }
    ctx->launches++;
    CU(cudaGetLastError());
    return SPMB_OK;
}

// the one-lane Myers kernels over the super-tiles [st0, st1) of a store
static int enqueue_myers_direct(spmb_context *ctx, const spmb_store *s, const spmb_pattern *p, const HitSink &sink,
                                uint64_t st0, uint64_t st1, uint32_t seq_id, bool row_own)
{
    if (p->n_pat > 65535) return set_err(ctx, SPMB_ERR_UNSUPPORTED, "more than 65535 patterns in one Myers set");
    int rc = store_ensure_lanes(ctx, s);
    if (rc) return rc;
    if (s->hi) { rc = store_ensure_lanes(ctx, s->hi); if (rc) return rc; }
    MyersArgs a;
    a.lanes = s->lanes;
    a.lanes_hi = s->hi ? s->hi->lanes : nullptr;
    a.mul4 = 4;
    a.eq_stride = (p->NW <= 4 ? 4 : 8) * 4;
    a.lay = s->lay;
    a.eq = p->d_eq;
    a.meta = p->d_meta;
    a.row_len = s->row_len; a.row_seq = s->row_seq; a.row_own = row_own ? s->row_own : nullptr;
    a.warm = p->warm;
    a.seq_id = seq_id;
    a.st0 = st0; a.st1 = st1;
    a.sink = sink;
    cudaError_t e = (s->lay.mode == SPMB_MODE_PLAIN) ? launch_myers<SPMB_MODE_PLAIN>(ctx, s, p, a)
                                                      : launch_myers<SPMB_MODE_WINDOWS>(ctx, s, p, a);
    if (e != cudaSuccess) return set_err(ctx, SPMB_ERR_CUDA, "myers_kernel launch", e);
    return SPMB_OK;
}

template <int MODE>
static cudaError_t launch_chain_verify(spmb_context *ctx, int wpl, const ChainArgs &ca)
{
    const unsigned grid = (unsigned)ctx->sm_count * 8;
    switch (wpl) {
    case 1: myers_chain_verify_kernel<1, MODE><<<grid, 128, 0, ctx->stream>>>(ca); break;
    case 2: myers_chain_verify_kernel<2, MODE><<<grid, 128, 0, ctx->stream>>>(ca); break;
    case 4: myers_chain_verify_kernel<4, MODE><<<grid, 128, 0, ctx->stream>>>(ca); break;
    case 8: myers_chain_verify_kernel<8, MODE><<<grid, 128, 0, ctx->stream>>>(ca); break;
    default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

// needles longer than 256 symbols (myers_chain.cuh): the one-lane kernels search their 256-symbol prefixes into the
// candidate buffer, one warp per candidate then runs the whole needle with its words chained across the lanes
static int enqueue_chain(spmb_context *ctx, const spmb_store *s, const spmb_pattern *p, spmb_hits *h,
                         uint64_t st0, uint64_t st1, uint32_t seq_id, const HitSink &sink)
{
    if (st1 <= st0) return SPMB_OK;
    int rc = cand_reserve(ctx, h, 1ull << 16);
    if (rc) return rc;
    CU(cudaMemsetAsync(h->d_count + 1, 0, sizeof(unsigned long long), ctx->stream));
    const uint64_t per_st = 32ull * s->lay.row_len;
    // a hit owned by this shard hangs on the prefix occurrence that ends m - 256 symbols before it
    uint64_t f0 = st0;
    if (s->lay.mode == SPMB_MODE_PLAIN && st0 > 0) {
        const uint64_t back = ((uint64_t)p->max_m - SPMB_CHAIN_PREFIX + per_st - 1) / per_st;
        f0 = st0 > back ? st0 - back : 0;
    }
    HitSink cand_sink{reinterpret_cast<spmb_hit *>(h->d_cand), h->d_count + 1, h->cand_cap};
    rc = enqueue_myers_direct(ctx, s, p->long_prefix, cand_sink, f0, st1, seq_id, false);
    if (rc) return rc;
    ChainArgs ca;
    memset(&ca, 0, sizeof(ca));
    ca.lanes = s->lanes; ca.lay = s->lay;
    ca.eq = p->d_chain_eq; ca.meta = p->d_chain_meta;
    ca.cand = reinterpret_cast<const spmb_hit *>(h->d_cand);
    ca.counters = h->d_count; ca.cand_cap = h->cand_cap;
    ca.prefix_m = SPMB_CHAIN_PREFIX;
    ca.row_len = s->row_len; ca.row_own = s->row_own;
    ca.seq_base = 0;
    ca.n_needles = p->n_long;
    ca.seq_id = seq_id;
    ca.own_lo = st0 * per_st;
    ca.own_hi = std::min<uint64_t>(s->lay.n, st1 * per_st);
    ca.sink = sink;
    ca.seq_base = s->seq_base; // rows of a window bucket carry consecutive window indices: row = seq_id - seq_base
    cudaError_t e = (s->lay.mode == SPMB_MODE_PLAIN) ? launch_chain_verify<SPMB_MODE_PLAIN>(ctx, p->chain_wpl, ca)
                                                      : launch_chain_verify<SPMB_MODE_WINDOWS>(ctx, p->chain_wpl, ca);
    if (e != cudaSuccess) return set_err(ctx, SPMB_ERR_CUDA, "myers_chain_verify_kernel launch", e);
    ctx->launches++;
    return SPMB_OK;
}

static int enqueue_search(spmb_context *ctx, const spmb_store *s, const spmb_pattern *p, spmb_hits *h,
                          bool reset = true, uint32_t seq_id = 0)
{
    if (reset) CU(cudaMemsetAsync(h->d_count, 0, 4 * sizeof(unsigned long long), ctx->stream));
    HitSink sink{h->d_hits, h->d_count, h->cap};
    if ((s->sym_bits == 4) != (p->sym_bits == 4))
        return set_err(ctx, SPMB_ERR_ARG, "store and pattern differ in alphabet width (2-bit vs 4-bit)");
    if (s->sym_bits == 4 && (s->lay.mode != SPMB_MODE_PLAIN || !s->hi))
        return set_err(ctx, SPMB_ERR_UNSUPPORTED, "4-bit alphabets: plain stores only");
    if (p->max_m == 0 || s->lay.n == 0) return SPMB_OK; // empty needle / haystack: no hits
    const uint64_t st0 = s->lay.n_st * ctx->shard / ctx->n_shards;
    const uint64_t st1 = s->lay.n_st * (ctx->shard + 1) / ctx->n_shards;
    if (p->alg == SPMB_ALG_MYERS_PREFIX) {
        if (ctx->shard != 0) return SPMB_OK; // the single anchor 0 belongs to shard 0
        return enqueue_prefix(ctx, s, p, nullptr, 1, seq_id, sink);
    }
    if (p->filtered) { // SPMB_ALG_PIGEONHOLE, SPMB_ALG_MYERS_FILTERED with a usable q
        return (s->lay.mode == SPMB_MODE_PLAIN) ? enqueue_filtered<SPMB_MODE_PLAIN>(ctx, s, p, h, st0, st1, seq_id, sink, reset)
                                                : enqueue_filtered<SPMB_MODE_WINDOWS>(ctx, s, p, h, st0, st1, seq_id, sink, reset);
    }
    if (p->alg == SPMB_ALG_PIGEONHOLE) return SPMB_OK; // no usable q-gram
    if (p->alg == SPMB_ALG_MYERS || p->alg == SPMB_ALG_MYERS_FILTERED) {
        if (p->subs.empty() && !p->long_prefix && p->chain_wpl == 1 && p->d_chain_eq && s->lay.mode == SPMB_MODE_PLAIN && !s->hi) {
            // SPMB_MYERS_CHAIN=1: one warp per (row, needle), words chained across the lanes (A/B measurement)
            int rc = store_ensure_lanes(ctx, s);
            if (rc) return rc;
            ChainArgs ca;
            memset(&ca, 0, sizeof(ca));
            ca.lanes = s->lanes; ca.lay = s->lay; ca.eq = p->d_chain_eq; ca.meta = p->d_chain_meta;
            ca.st0 = st0; ca.st1 = st1; ca.n_needles = p->n_pat; ca.warm = p->warm; ca.seq_id = seq_id; ca.sink = sink;
            const unsigned long long tasks = (unsigned long long)(st1 - st0) * 32 * p->n_pat;
            const unsigned grid = (unsigned)std::min<unsigned long long>((tasks + 3) / 4, (unsigned long long)ctx->sm_count * 16);
            if (grid) myers_chain_scan_kernel<1><<<grid, 128, 0, ctx->stream>>>(ca);
            ctx->launches++;
            CU(cudaGetLastError());
            return SPMB_OK;
        }
        if (p->subs.empty() && !p->long_prefix) return enqueue_myers_direct(ctx, s, p, sink, st0, st1, seq_id, true);
        for (const spmb_pattern *c : p->subs) {
            int rc = enqueue_myers_direct(ctx, s, c, sink, st0, st1, seq_id, true);
            if (rc) return rc;
        }
        if (p->long_prefix) return enqueue_chain(ctx, s, p, h, st0, st1, seq_id, sink);
        return SPMB_OK;
    }
    // exact searchers (Horspool and Shift-Or share the bit-sliced kernel)
    if ((uint64_t)p->max_m > 31ull * s->lay.row_len)
        return set_err(ctx, SPMB_ERR_UNSUPPORTED, "needle longer than 31 rows of this store");
    for (uint32_t i = 0; i < p->n_pat; ++i) {
        if (p->len[i] == 0) continue;
        ExactArgs a;
        a.planes = reinterpret_cast<const uint4 *>(s->planes);
        a.lay = s->lay;
        a.tile0 = st0 * (s->lay.row_len / 32);
        a.n_tiles = st1 * (s->lay.row_len / 32);
        if (a.n_tiles <= a.tile0) continue;
        a.pat = p->d_ranks + p->off[i];
        a.row_len = s->row_len; a.row_seq = s->row_seq; a.row_own = s->row_own;
        a.m = p->len[i];
        a.J = std::min<uint32_t>(a.m, 8);
        a.fo = p->filt_off[i];
        a.filt_syms = p->filt[i];
        a.planes_hi = s->hi ? reinterpret_cast<const uint4 *>(s->hi->planes) : nullptr;
        a.filt_syms_hi = s->hi ? p->filt_hi[i] : 0;
        a.seq_id = seq_id;
        a.pattern_id = i;
        a.sink = sink;
        unsigned blocks = (unsigned)((a.n_tiles - a.tile0 + 255) / 256);
        const unsigned grid = std::min(blocks, 8u * (unsigned)ctx->sm_count); // grid-stride kernel
        if (s->hi) exact_kernel<SPMB_MODE_PLAIN, true><<<grid, 256, 0, ctx->stream>>>(a);
        else if (s->lay.mode == SPMB_MODE_PLAIN) exact_kernel<SPMB_MODE_PLAIN><<<grid, 256, 0, ctx->stream>>>(a);
        else exact_kernel<SPMB_MODE_WINDOWS><<<grid, 256, 0, ctx->stream>>>(a);
        ctx->launches++;
        CU(cudaGetLastError());
    }
    return SPMB_OK;
}

static int bits_for(unsigned long long max_value);

// key layout of the device-side hit ordering; false = the key does not fit 64 bits
static bool hit_key_spec(const spmb_store *s, const spmb_pattern *p, HitKeySpec &spec, int &total_bits, int &drop)
{
    const uint64_t n_seq = (s->lay.mode == SPMB_MODE_PLAIN) ? 1 : s->lay.n_st * 32;
    const uint64_t max_end = (s->lay.mode == SPMB_MODE_PLAIN) ? s->lay.n : s->lay.row_len;
    const int b_pat = bits_for(p->n_pat), b_seq = n_seq > 1 ? bits_for(n_seq) : 0, b_end = bits_for(max_end);
    drop = 0;
    if (p->alg == SPMB_ALG_PIGEONHOLE) {
        spec = {{1, 2, 0, 4}, {b_seq, b_end, b_pat, bits_for(p->max_m)}};
    } else {
        const int b_err = bits_for(p->k);
        spec = {{0, 1, 2, 3}, {b_pat, b_seq, b_end, b_err}};
        if (p->alg == SPMB_ALG_MYERS_FILTERED) drop = b_err;
    }
    if (b_seq == 0) { for (int f = 0; f < 4; ++f) if (spec.field[f] == 1) spec.field[f] = -1; }
    total_bits = 0;
    for (int f = 0; f < 4; ++f) if (spec.field[f] >= 0) total_bits += spec.bits[f];
    return total_bits <= 64;
}

// The ordering of a long hit list, enqueued right behind the search kernels on the context's stream: a later
// spmb_search_finish then only copies (count, ordered hits) back.  Ordering from finish instead means launching
// the sort while the NEXT pass's kernels hold the SMs (the seed filter is one persistent CTA per SM for ~1 ms):
// measured on C5, 1.7 ms per finish against 0.8 ms on an idle device.  Buffers span the whole hit buffer.
static int enqueue_order(spmb_context *ctx, spmb_hits *h, const spmb_store *s, const spmb_pattern *p)
{
    HitKeySpec spec;
    int total_bits = 0, drop = 0;
    if (!hit_key_spec(s, p, spec, total_bits, drop) || total_bits >= 64 || h->cap == 0 || h->cap >= (1ull << 31)) return SPMB_OK;
    const int cap = (int)h->cap;
    size_t tb = 0, tb2 = 0;
    if (h->ord_cap != h->cap) { // (re)size: the hit buffer grew
        cudaFree(h->ord_k0); cudaFree(h->ord_k1); cudaFree(h->ord_i0); cudaFree(h->ord_i1); cudaFree(h->ord_isel);
        cudaFree(h->ord_flag); cudaFree(h->ord_out); cudaFree(h->ord_tmp);
        h->ord_k0 = h->ord_k1 = nullptr; h->ord_i0 = h->ord_i1 = h->ord_isel = nullptr; h->ord_flag = nullptr;
        h->ord_out = nullptr; h->ord_tmp = nullptr; h->ord_cap = 0;
        cub::DeviceRadixSort::SortPairs(nullptr, tb, h->ord_k0, h->ord_k1, h->ord_i0, h->ord_i1, cap, 0, 64, ctx->stream);
        cub::DeviceSelect::Flagged(nullptr, tb2, h->ord_i1, h->ord_flag, h->ord_isel, (uint32_t *)nullptr, cap, ctx->stream);
        h->ord_tmp_bytes = std::max(tb, tb2) + 16;
        cudaError_t e = cudaMalloc(&h->ord_k0, cap * 8ull);
        if (e == cudaSuccess) e = cudaMalloc(&h->ord_k1, cap * 8ull);
        if (e == cudaSuccess) e = cudaMalloc(&h->ord_i0, cap * 4ull);
        if (e == cudaSuccess) e = cudaMalloc(&h->ord_i1, cap * 4ull);
        if (e == cudaSuccess) e = cudaMalloc(&h->ord_isel, cap * 4ull);
        if (e == cudaSuccess) e = cudaMalloc(&h->ord_flag, cap);
        if (e == cudaSuccess) e = cudaMalloc(&h->ord_out, cap * sizeof(spmb_hit));
        if (e == cudaSuccess) e = cudaMalloc(&h->ord_tmp, h->ord_tmp_bytes);
        if (e != cudaSuccess) { cudaGetLastError(); return SPMB_OK; } // no memory for it: finish orders the list as before
        h->ord_cap = h->cap;
    }
    cudaStream_t sm = ctx->stream;
    uint32_t *n_sel = reinterpret_cast<uint32_t *>(h->d_count + 3); // low half of counter slot 3 (zeroed with the others)
    const unsigned blocks = ((unsigned)cap + 255) / 256;
    hit_keys_dev_kernel<<<blocks, 256, 0, sm>>>(h->d_hits, h->d_count, (uint32_t)cap, spec, total_bits, h->ord_k0, h->ord_i0);
    size_t tmp_bytes = h->ord_tmp_bytes;
    cudaError_t e = cub::DeviceRadixSort::SortPairs(h->ord_tmp, tmp_bytes, h->ord_k0, h->ord_k1, h->ord_i0, h->ord_i1, cap, 0,
                                                    total_bits + 1, sm);
    if (e == cudaSuccess && drop) {
        hit_flags_dev_kernel<<<blocks, 256, 0, sm>>>(h->ord_k1, h->d_count, (uint32_t)cap, drop, h->ord_flag);
        tmp_bytes = h->ord_tmp_bytes;
        e = cub::DeviceSelect::Flagged(h->ord_tmp, tmp_bytes, h->ord_i1, h->ord_flag, h->ord_isel, n_sel, cap, sm);
        hit_gather_dev_kernel<<<blocks, 256, 0, sm>>>(h->d_hits, h->ord_isel, h->d_count, n_sel, 1, (uint32_t)cap, h->ord_out);
        ctx->launches += 5;
    } else if (e == cudaSuccess) {
        hit_gather_dev_kernel<<<blocks, 256, 0, sm>>>(h->d_hits, h->ord_i1, h->d_count, n_sel, 0, (uint32_t)cap, h->ord_out);
        ctx->launches += 3;
    }
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) { cudaGetLastError(); return set_err(ctx, SPMB_ERR_CUDA, "device hit ordering (launch)", e); }
    h->ordered = true;
    h->ord_drop = drop;
    return SPMB_OK;
}

extern "C" int spmb_search_launch(spmb_context *ctx, const spmb_store *s, const spmb_pattern *p, spmb_hits *h)
{
    if (!ctx || !s || !p || !h) return SPMB_ERR_ARG;
    ENTER(ctx);
    ctx->launches = 0;
    h->pend_store = s;
    h->pend_pat = p;
    h->ordered = false;
    h->view = nullptr;
    int rc = enqueue_search(ctx, s, p, h);
    if (rc == SPMB_OK && h->expect_many && p->max_m && s->lay.n) rc = enqueue_order(ctx, h, s, p);
    h->done_armed = rc == SPMB_OK && h->done && ctx->copy_stream &&
                    cudaEventRecord(h->done, ctx->stream) == cudaSuccess;
    return rc;
}

static bool hit_less(const spmb_hit &a, const spmb_hit &b)
{
    if (a.pattern_id != b.pattern_id) return a.pattern_id < b.pattern_id;
    if (a.seq_id != b.seq_id) return a.seq_id < b.seq_id;
    return a.end < b.end;
}

// order of the reference's callback sequence for the seed matcher: haystack position first
static bool seed_hit_less(const spmb_hit &a, const spmb_hit &b)
{
    if (a.seq_id != b.seq_id) return a.seq_id < b.seq_id;
    if (a.end != b.end) return a.end < b.end;
    if (a.pattern_id != b.pattern_id) return a.pattern_id < b.pattern_id;
    return a.reserved < b.reserved;
}

// sort a raw device hit list into the order handed back; the filtered Myers path reports an end once
// per seed window it lies in -- keep the smallest error count (the exact one, pigeon.cuh).  Lists of 256 records
// and more go through the packed-key ordering of hit_order.hpp (counting sort over the needle id for the large sets,
// LSD radix sort otherwise), shorter ones and keys that do not fit 64 bits through std::sort.
static void hits_order(std::vector<spmb_hit> &v, int alg)
{
    // buffers kept between calls: a fresh output vector per call cost 0.12 ms of page faults per 13 k hits
    static thread_local HitOrderScratch sc;
    if (alg == SPMB_ALG_PIGEONHOLE) { std::sort(v.begin(), v.end(), seed_hit_less); return; }
    if (alg == SPMB_ALG_MYERS_FILTERED) {
        if (v.size() >= 256 && hits_order_packed(v, true, true, sc)) return;
        std::sort(v.begin(), v.end(), [](const spmb_hit &a, const spmb_hit &b) {
            if (hit_less(a, b)) return true;
            if (hit_less(b, a)) return false;
            return a.errors < b.errors;
        });
        v.erase(std::unique(v.begin(), v.end(), [](const spmb_hit &a, const spmb_hit &b) {
                    return a.pattern_id == b.pattern_id && a.seq_id == b.seq_id && a.end == b.end;
                }), v.end());
        return;
    }
    if (v.size() < 256 || !hits_order_packed(v, false, false, sc)) std::sort(v.begin(), v.end(), hit_less);
}

static int bits_for(unsigned long long max_value)
{
    int b = 1;
    while (b < 64 && (max_value >> b)) ++b;
    return b;
}

// Long hit lists are ordered (and, for the filtered Myers path, deduplicated) on the device before
// they travel: returns SPMB_OK with *done = true and `out` filled, or *done = false when the list is
// short or its key does not fit 64 bits (the caller then sorts on the host).
static int hits_order_device(spmb_context *ctx, spmb_hits *h, const spmb_store *s, const spmb_pattern *p,
                             unsigned long long cnt, uint64_t n_seq, std::vector<spmb_hit> &out, bool *done)
{
    *done = false;
    if (cnt < SPMB_ORDER_AT_LAUNCH_MIN || cnt >= (1ull << 31)) return SPMB_OK; // shorter lists: host radix sort (~10 ns per hit)
    const int alg = p->alg;
    const uint64_t max_end = (s->lay.mode == SPMB_MODE_PLAIN) ? s->lay.n : s->lay.row_len;
    HitKeySpec spec;
    const int b_pat = bits_for(p->n_pat), b_seq = n_seq > 1 ? bits_for(n_seq) : 0, b_end = bits_for(max_end);
    int drop = 0;
    if (alg == SPMB_ALG_PIGEONHOLE) {
        spec = {{1, 2, 0, 4}, {b_seq, b_end, b_pat, bits_for(p->max_m)}};
    } else {
        const int b_err = bits_for(p->k);
        spec = {{0, 1, 2, 3}, {b_pat, b_seq, b_end, b_err}};
        if (alg == SPMB_ALG_MYERS_FILTERED) drop = b_err;
    }
    if (b_seq == 0) { for (int f = 0; f < 4; ++f) if (spec.field[f] == 1) spec.field[f] = -1; }
    int total_bits = 0;
    for (int f = 0; f < 4; ++f) if (spec.field[f] >= 0) total_bits += spec.bits[f];
    if (total_bits > 64) return SPMB_OK;

    // on the stream that looked at the counters (the copy stream behind this launch's event): the sort of pass i
    // then runs beside the kernels of pass i+1 instead of behind them
    cudaStream_t sm = h->collect_stream ? h->collect_stream : ctx->stream;
    const uint32_t n = (uint32_t)cnt;
    unsigned long long *k0 = nullptr, *k1 = nullptr;
    uint32_t *i0 = nullptr, *i1 = nullptr, *isel = nullptr, *d_nsel = nullptr;
    uint8_t *flag = nullptr;
    spmb_hit *d_out = nullptr;
    void *tmp = nullptr;
    cudaError_t e = cudaSuccess;
#define HCU(call) do { if (e == cudaSuccess) e = (call); } while (0)
    HCU(cudaMallocAsync(&k0, n * 8ull, sm)); HCU(cudaMallocAsync(&k1, n * 8ull, sm));
    HCU(cudaMallocAsync(&i0, n * 4ull, sm)); HCU(cudaMallocAsync(&i1, n * 4ull, sm));
    HCU(cudaMallocAsync(&d_out, n * sizeof(spmb_hit), sm));
    const unsigned blocks = (n + 255) / 256;
    if (e == cudaSuccess) hit_keys_kernel<<<blocks, 256, 0, sm>>>(h->d_hits, n, spec, k0, i0);
    size_t tb = 0, tb2 = 0;
    HCU(cub::DeviceRadixSort::SortPairs(nullptr, tb, k0, k1, i0, i1, (int)n, 0, total_bits, sm));
    if (drop) {
        HCU(cudaMallocAsync(&flag, n, sm)); HCU(cudaMallocAsync(&isel, n * 4ull, sm)); HCU(cudaMallocAsync(&d_nsel, 4, sm));
        HCU(cub::DeviceSelect::Flagged(nullptr, tb2, i1, flag, isel, d_nsel, (int)n, sm));
    }
    HCU(cudaMallocAsync(&tmp, std::max(tb, tb2) + 16, sm));
    HCU(cub::DeviceRadixSort::SortPairs(tmp, tb, k0, k1, i0, i1, (int)n, 0, total_bits, sm));
    uint32_t n_out = n;
    if (drop) {
        if (e == cudaSuccess) hit_flags_kernel<<<blocks, 256, 0, sm>>>(k1, n, drop, flag);
        HCU(cub::DeviceSelect::Flagged(tmp, tb2, i1, flag, isel, d_nsel, (int)n, sm));
        if (e == cudaSuccess) hit_gather_kernel<<<blocks, 256, 0, sm>>>(h->d_hits, isel, d_nsel, n, d_out);
        HCU(cudaMemcpyAsync(&ctx->h_count[4], d_nsel, 4, cudaMemcpyDeviceToHost, sm));
        HCU(cudaStreamSynchronize(sm));
        if (e == cudaSuccess) n_out = (uint32_t)(ctx->h_count[4] & 0xffffffffu);
        ctx->launches += 5;
    } else {
        if (e == cudaSuccess) hit_gather_kernel<<<blocks, 256, 0, sm>>>(h->d_hits, i1, nullptr, n, d_out);
        ctx->launches += 3;
    }
    HCU(cudaGetLastError());
    if (e == cudaSuccess) {
        out.resize(n_out);
        if (n_out) HCU(cudaMemcpyAsync(out.data(), d_out, n_out * sizeof(spmb_hit), cudaMemcpyDeviceToHost, sm));
        HCU(cudaStreamSynchronize(sm));
    }
#undef HCU
    cudaFreeAsync(k0, sm); cudaFreeAsync(k1, sm); cudaFreeAsync(i0, sm); cudaFreeAsync(i1, sm); cudaFreeAsync(isel, sm);
    cudaFreeAsync(d_nsel, sm); cudaFreeAsync(flag, sm); cudaFreeAsync(d_out, sm); cudaFreeAsync(tmp, sm);
    if (e != cudaSuccess) { cudaGetLastError(); return set_err(ctx, SPMB_ERR_CUDA, "device hit sort", e); }
    *done = true;
    return SPMB_OK;
}

// wait for the enqueued search, regrow the hit / candidate buffers and re-run (`again`) while one of
// them overflowed (the counters keep counting, so they hold the exact sizes needed), then copy the raw
// hits to `raw`
template <class Again>
static int hits_collect(spmb_context *ctx, spmb_hits *h, std::vector<spmb_hit> &raw, Again again,
                        unsigned long long *count_only = nullptr)
{
    h->spec_done = false;
    for (int attempt = 0; attempt < 5; ++attempt) {
        // the first look at a launched search goes through the copy stream, ordered behind the launch's event
        // only: launches of other hit lists that were queued meanwhile keep running
        cudaStream_t cs = ctx->stream;
        if (h->done_armed) {
            h->done_armed = false;
            if (cudaStreamWaitEvent(ctx->copy_stream, h->done, 0) == cudaSuccess) cs = ctx->copy_stream;
            else cudaGetLastError();
        }
        h->collect_stream = cs;
        CU(cudaMemcpyAsync(ctx->h_count, h->d_count, 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, cs));
        // short lists (the common case) ride along with the counters: one synchronisation instead of two
        const unsigned long long n_spec = std::min<unsigned long long>(h->cap, SPEC_HITS);
        if (n_spec) CU(cudaMemcpyAsync(ctx->h_spec, h->d_hits, n_spec * sizeof(spmb_hit), cudaMemcpyDeviceToHost, cs));
        CU(cudaStreamSynchronize(cs));
        const unsigned long long cnt = ctx->h_count[0];
        const unsigned long long cand = std::max(ctx->h_count[1], ctx->h_count[2]);
        if (cnt <= h->cap && cand <= h->cand_cap) {
            if (cnt <= n_spec) {
                raw.assign(ctx->h_spec, ctx->h_spec + cnt);
                h->spec_done = true;
                if (count_only) *count_only = cnt;
                return SPMB_OK;
            }
            if (count_only) { *count_only = cnt; return SPMB_OK; } // the hits stay on the device for now
            raw.resize(cnt);
            if (cnt) { // on the stream of the first look: a later launch queued on the context's stream does not delay it
                CU(cudaMemcpyAsync(raw.data(), h->d_hits, cnt * sizeof(spmb_hit), cudaMemcpyDeviceToHost, cs));
                CU(cudaStreamSynchronize(cs));
            }
            return SPMB_OK;
        }
        int rc = SPMB_OK;
        if (cand > h->cand_cap) rc = cand_reserve(ctx, h, cand + cand / 8 + 1024);
        // a truncated candidate list also truncates the hit count: size the hit buffer on the next pass
        if (rc == SPMB_OK && cnt > h->cap) rc = hits_reserve(ctx, h, cnt + cnt / 8 + 1024);
        if (rc) return rc;
        rc = again();
        if (rc) return rc;
    }
    return set_err(ctx, SPMB_ERR_CUDA, "hit buffer kept overflowing");
}

extern "C" int spmb_search_finish(spmb_context *ctx, spmb_hits *h)
{
    if (!ctx || !h || !h->pend_store) return SPMB_ERR_ARG;
    ENTER(ctx);
    const spmb_store *s = h->pend_store;
    const spmb_pattern *p = h->pend_pat;
    unsigned long long cnt = 0;
    h->view = nullptr;
    int rc = hits_collect(ctx, h, h->host, [&]() { h->ordered = false; return enqueue_search(ctx, s, p, h); }, &cnt);
    if (rc) return rc;
    bool done = false;
    if (h->ordered && !h->spec_done && cnt <= h->ord_cap && cnt <= h->cap) {
        // the launch ordered the list itself: ctx->h_count[3] holds the number of records to fetch
        const unsigned long long n_out = h->ord_drop ? (ctx->h_count[3] & 0xffffffffull) : cnt;
        if (n_out > h->pinned_cap) {
            if (h->pinned) cudaFreeHost(h->pinned);
            h->pinned = nullptr; h->pinned_cap = 0;
            const unsigned long long want = std::max<unsigned long long>(h->cap, n_out);
            if (cudaMallocHost(&h->pinned, want * sizeof(spmb_hit)) == cudaSuccess) h->pinned_cap = want;
            else cudaGetLastError();
        }
        cudaStream_t cs = h->collect_stream ? h->collect_stream : ctx->stream;
        if (h->pinned_cap >= n_out) {
            if (n_out) CU(cudaMemcpyAsync(h->pinned, h->ord_out, n_out * sizeof(spmb_hit), cudaMemcpyDeviceToHost, cs));
            CU(cudaStreamSynchronize(cs));
            h->host.clear();
            h->view = h->pinned;
            h->count = n_out;
        } else {
            h->host.resize(n_out);
            if (n_out) CU(cudaMemcpyAsync(h->host.data(), h->ord_out, n_out * sizeof(spmb_hit), cudaMemcpyDeviceToHost, cs));
            CU(cudaStreamSynchronize(cs));
            h->count = n_out;
        }
        h->ordered = false;
        h->expect_many = cnt >= SPMB_ORDER_AT_LAUNCH_MIN;
        h->pend_store = nullptr;
        h->pend_pat = nullptr;
        return SPMB_OK;
    }
    h->ordered = false;
    h->expect_many = cnt >= SPMB_ORDER_AT_LAUNCH_MIN;
    const uint64_t n_seq = (s->lay.mode == SPMB_MODE_PLAIN) ? 1 : s->lay.n_st * 32;
    rc = hits_order_device(ctx, h, s, p, cnt, n_seq, h->host, &done);
    if (rc) return rc;
    if (!done) {
        h->host.resize(cnt);
        if (cnt && !h->spec_done) {
            cudaStream_t cs = h->collect_stream ? h->collect_stream : ctx->stream;
            CU(cudaMemcpyAsync(h->host.data(), h->d_hits, cnt * sizeof(spmb_hit), cudaMemcpyDeviceToHost, cs));
            CU(cudaStreamSynchronize(cs));
        }
        hits_order(h->host, p->alg);
    }
    h->count = h->host.size();
    h->pend_store = nullptr;
    h->pend_pat = nullptr;
    return SPMB_OK;
}

extern "C" int spmb_search_anchored(spmb_context *ctx, const spmb_store *s, const spmb_pattern *p,
                                    const uint64_t *anchors, uint64_t n_anchors, spmb_hits *h)
{
    if (!ctx || !s || !p || !h || (n_anchors && !anchors)) return SPMB_ERR_ARG;
    if (p->alg != SPMB_ALG_MYERS_PREFIX) return set_err(ctx, SPMB_ERR_ARG, "anchored search takes a SPMB_ALG_MYERS_PREFIX pattern");
    if (s->sym_bits != 2) return set_err(ctx, SPMB_ERR_UNSUPPORTED, "anchored search is dna4 (2-bit) only");
    ENTER(ctx);
    ctx->launches = 0;
    h->host.clear();
    h->count = 0;
    uint64_t *d_anchors = nullptr;
    CU(cudaMallocAsync(&d_anchors, (n_anchors ? n_anchors : 1) * 8, ctx->stream));
    cudaError_t e = n_anchors ? cudaMemcpyAsync(d_anchors, anchors, n_anchors * 8, cudaMemcpyHostToDevice, ctx->stream) : cudaSuccess;
    int rc = (e == cudaSuccess) ? SPMB_OK : set_err(ctx, SPMB_ERR_CUDA, "H2D anchors", e);
    auto run = [&]() -> int {
        CU(cudaMemsetAsync(h->d_count, 0, 4 * sizeof(unsigned long long), ctx->stream));
        HitSink sink{h->d_hits, h->d_count, h->cap};
        return enqueue_prefix(ctx, s, p, d_anchors, n_anchors, 0, sink);
    };
    if (rc == SPMB_OK) rc = run();
    if (rc == SPMB_OK) rc = hits_collect(ctx, h, h->host, run);
    cudaFreeAsync(d_anchors, ctx->stream);
    cudaStreamSynchronize(ctx->stream);
    if (rc) return rc;
    hits_order(h->host, p->alg);
    h->count = h->host.size();
    return SPMB_OK;
}

extern "C" int spmb_search(spmb_context *ctx, const spmb_store *s, const spmb_pattern *p, spmb_hits *h)
{
    int rc = spmb_search_launch(ctx, s, p, h);
    if (rc) return rc;
    return spmb_search_finish(ctx, h);
}

// ------------------------------------------------------------------------------------------
// allocation-free filtered search (spmb.h: no counterpart in the reference).
// The seed filter's candidate list, its counters and the duplicate-suppression table
// live in ONE file-scope device array instead of buffers of a spmb_hits object: nothing is allocated,
// grown or freed here.  Every GPU that loads this library holds the array until the process exits.
// ------------------------------------------------------------------------------------------
enum : unsigned long long {
    SPMB_NA_COUNTER_BYTES = 256,                                              // 4 counters, padded
    SPMB_NA_DD_BYTES = 12ull << SPMB_DD_BITS,                                 // 8-byte keys + 4-byte minima
    SPMB_NA_CAND_BYTES = (unsigned long long)SPMB_NOALLOC_MAX_CANDIDATES * sizeof(PigeonCand),
    SPMB_NA_BYTES = SPMB_NA_COUNTER_BYTES + SPMB_NA_DD_BYTES + SPMB_NA_CAND_BYTES // 3.5 MiB + 256 B
};
static_assert(SPMB_NA_DD_BYTES % 256 == 0, "the candidate records start on a 256-byte boundary");
static_assert(SPMB_NA_BYTES == 256 + (3ull << 20) + (1ull << 19), "spmb.h states 3.5 MiB + 256 B");

// In host code this name is only a handle for cudaGetSymbolAddress -- never a pointer a kernel may be given.
static __device__ __align__(256) unsigned char spmb_noalloc_scratch[SPMB_NA_BYTES];

// the array's address on the context's device (looked up once per device, with that device current)
static int noalloc_scratch(spmb_context *ctx, unsigned char **out)
{
    enum { MAX_DEV = 64 };
    static std::mutex mu;
    static unsigned char *addr[MAX_DEV] = {nullptr};
    if (ctx->device < 0 || ctx->device >= MAX_DEV) return set_err(ctx, SPMB_ERR_UNSUPPORTED, "no-alloc search: device ordinal beyond 63");
    std::lock_guard<std::mutex> g(mu);
    if (!addr[ctx->device]) {
        void *p = nullptr;
        CU(cudaGetSymbolAddress(&p, spmb_noalloc_scratch));
        if (!p) return set_err(ctx, SPMB_ERR_CUDA, "cudaGetSymbolAddress(spmb_noalloc_scratch) gave no address");
        addr[ctx->device] = static_cast<unsigned char *>(p);
    }
    *out = addr[ctx->device];
    return SPMB_OK;
}

extern "C" int spmb_search_filtered_noalloc(spmb_context *ctx, const spmb_store *s, const spmb_pattern *p,
                                            spmb_hit *dev_hits, uint64_t capacity, unsigned long long *dev_count,
                                            uint64_t *count_out)
{
    if (!ctx || !s || !p || !dev_count || !count_out || (!dev_hits && capacity)) return SPMB_ERR_ARG;
    ENTER(ctx);
    ctx->launches = 0;
    *count_out = 0;
    if (p->alg != SPMB_ALG_MYERS_FILTERED || !p->filtered || p->pg.seed_pat.empty() || p->NW < 1 || p->NW > 8)
        return set_err(ctx, SPMB_ERR_UNSUPPORTED, "no-alloc search: takes a SPMB_ALG_MYERS_FILTERED set whose seed filter is in use "
                                                  "(spmb_pattern_filter_info: step != 0)");
    if (s->lay.mode != SPMB_MODE_PLAIN || s->sym_bits != 2 || p->sym_bits != 2 || s->hi)
        return set_err(ctx, SPMB_ERR_UNSUPPORTED, "no-alloc search: plain 2-bit stores only");
    if (s->lay.n > SPMB_NOALLOC_MAX_SYMBOLS)
        return set_err(ctx, SPMB_ERR_UNSUPPORTED, "no-alloc search: haystack longer than SPMB_NOALLOC_MAX_SYMBOLS");
    if (capacity > SPMB_NOALLOC_MAX_HITS)
        return set_err(ctx, SPMB_ERR_UNSUPPORTED, "no-alloc search: hit capacity beyond SPMB_NOALLOC_MAX_HITS");
    if (!s->lanes)
        return set_err(ctx, SPMB_ERR_ARG, "no-alloc search: the store has no lane-chunk copy yet (it is derived, with an allocation, "
                                          "by the first Myers / filtered spmb_search on the store: run one at set-up)");
    unsigned char *scratch = nullptr;
    int rc = noalloc_scratch(ctx, &scratch);
    if (rc) return rc;
    unsigned long long *counters = reinterpret_cast<unsigned long long *>(scratch);
    unsigned long long *dd_keys = reinterpret_cast<unsigned long long *>(scratch + SPMB_NA_COUNTER_BYTES);
    PigeonCand *cand = reinterpret_cast<PigeonCand *>(scratch + SPMB_NA_COUNTER_BYTES + SPMB_NA_DD_BYTES);
    // the array keeps the last call's contents: counters and table start at zero (the candidate records are
    // only read below the counter, so they need no clearing)
    CU(cudaMemsetAsync(scratch, 0, SPMB_NA_COUNTER_BYTES + SPMB_NA_DD_BYTES, ctx->stream));
    CU(cudaMemsetAsync(dev_count, 0, sizeof(unsigned long long), ctx->stream));
    if (s->lay.n == 0 || p->max_m == 0) { CU(cudaStreamSynchronize(ctx->stream)); return SPMB_OK; }
    const uint64_t st0 = s->lay.n_st * ctx->shard / ctx->n_shards;
    const uint64_t st1 = s->lay.n_st * (ctx->shard + 1) / ctx->n_shards;
    if (st1 <= st0) { CU(cudaStreamSynchronize(ctx->stream)); return SPMB_OK; }
    const uint64_t per_st = 32ull * s->lay.row_len;
    HitSink sink{dev_hits, dev_count, capacity};
    PigeonFilterArgs fa;
    fa.lanes = s->lanes; fa.lay = s->lay; fa.pg = pigeon_dev(p);
    fa.row_len = s->row_len;
    fa.st0 = st0; fa.st1 = st1;
    if (st0 > 0) { // a hit owned by this shard may hang on a seed that begins up to m+k symbols before it
        const uint64_t back = ((uint64_t)p->max_m + p->k + per_st - 1) / per_st;
        fa.st0 = st0 > back ? st0 - back : 0;
    }
    fa.cand = cand; fa.cand_count = counters + 1; fa.cand_cap = SPMB_NOALLOC_MAX_CANDIDATES;
    const uint64_t chunk_cols = (fa.st1 - fa.st0) * (s->lay.row_len / 32);
    unsigned grid = (unsigned)std::min<uint64_t>((uint64_t)ctx->sm_count, (chunk_cols + 255) / 256);
    if (grid == 0) grid = 1;
    const size_t smem = ((size_t)4 << (p->pg.B - 5)) + 32 * PG_QCAP * 4;
    cudaError_t e;
    if (p->pg.kq <= 16) e = launch_filter_kernel<SPMB_MODE_PLAIN, false, 1>(ctx, fa, grid, smem);
    else if (p->filter_step == 4) e = launch_filter_kernel<SPMB_MODE_PLAIN, true, 4>(ctx, fa, grid, smem);
    else if (p->filter_step == 2) e = launch_filter_kernel<SPMB_MODE_PLAIN, true, 2>(ctx, fa, grid, smem);
    else e = launch_filter_kernel<SPMB_MODE_PLAIN, true, 1>(ctx, fa, grid, smem);
    if (e != cudaSuccess) return set_err(ctx, SPMB_ERR_CUDA, "pigeon_filter_kernel launch (no-alloc)", e);
    PigeonVerifyArgs va;
    va.lanes = s->lanes; va.lay = s->lay; va.pg = fa.pg;
    va.cand = cand; va.counters = counters; va.cand_cap = SPMB_NOALLOC_MAX_CANDIDATES;
    va.pat_ranks = p->d_ranks; va.pat_off = p->d_off; va.eq = p->d_eq; va.meta = p->d_meta;
    va.row_len = s->row_len; va.row_seq = s->row_seq; va.row_own = s->row_own;
    va.seq_id = 0;
    va.own_lo = st0 * per_st;
    va.own_hi = std::min<uint64_t>(s->lay.n, st1 * per_st);
    va.sink = sink;
    va.dd_keys = nullptr; va.dd_err = nullptr; va.dd_bits = SPMB_DD_BITS; va.dd_end_bits = 0; va.dd_nseq = 1;
    if (p->k < 255) { // the table's key: (needle, end); else every finder reports and the caller removes the duplicates
        const int eb = bits_for(s->lay.n + 1);
        if (bits_for((uint64_t)p->n_pat) + eb <= 62) {
            va.dd_keys = dd_keys; va.dd_err = reinterpret_cast<uint32_t *>(dd_keys + ((size_t)1 << SPMB_DD_BITS));
            va.dd_end_bits = (uint32_t)eb;
        }
    }
    e = launch_verify<SPMB_MODE_PLAIN>(ctx, p, va);
    if (e != cudaSuccess) return set_err(ctx, SPMB_ERR_CUDA, "pigeon verify launch (no-alloc)", e);
    ctx->launches += 2;
    if (va.dd_keys && capacity) {
        const unsigned fgrid = (unsigned)std::min<unsigned long long>((capacity + 255) / 256, (unsigned long long)ctx->sm_count * 4);
        pigeon_fixup_kernel<<<fgrid ? fgrid : 1, 256, 0, ctx->stream>>>(dev_hits, dev_count, capacity, va.dd_err);
        ctx->launches++;
        CU(cudaGetLastError());
    }
    // counters back through the context's pinned slots (allocated with the context)
    CU(cudaMemcpyAsync(ctx->h_count, counters, 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaMemcpyAsync(ctx->h_count + 4, dev_count, sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    const unsigned long long n_cand = std::max(ctx->h_count[1], ctx->h_count[2]);
    *count_out = ctx->h_count[4];
    if (n_cand > SPMB_NOALLOC_MAX_CANDIDATES) {
        char msg[160];
        snprintf(msg, sizeof msg, "no-alloc search: %llu filter candidates, the scratch array holds %u (use spmb_search)", n_cand,
                 (unsigned)SPMB_NOALLOC_MAX_CANDIDATES);
        return set_err(ctx, SPMB_ERR_OOM, msg);
    }
    if (ctx->h_count[4] > capacity) {
        char msg[160];
        snprintf(msg, sizeof msg, "no-alloc search: %llu hit records, the caller's buffer holds %llu", ctx->h_count[4],
                 (unsigned long long)capacity);
        return set_err(ctx, SPMB_ERR_OOM, msg);
    }
    return SPMB_OK;
}

extern "C" int spmb_search_host_wide(spmb_context *ctx, const uint8_t *host_ranks, uint64_t n, uint32_t bits,
                                     const spmb_pattern *p, spmb_hits *h)
{
    if (!ctx || !p || !h || (!host_ranks && n)) return SPMB_ERR_ARG;
    spmb_store *s = nullptr;
    int rc = spmb_store_create_wide(ctx, host_ranks, n, bits, 0, &s);
    if (rc) return rc;
    uint32_t pack_launches = ctx->launches;
    rc = spmb_search(ctx, s, p, h);
    ctx->launches += pack_launches;
    spmb_store_destroy(s);
    return rc;
}

extern "C" int spmb_search_host(spmb_context *ctx, const uint8_t *host_ranks, uint64_t n, const spmb_pattern *p,
                                spmb_hits *h)
{
    if (!ctx || !p || !h || (!host_ranks && n)) return SPMB_ERR_ARG;
    spmb_store *s = nullptr;
    int rc = spmb_store_create(ctx, host_ranks, n, 0, &s);
    if (rc) return rc;
    uint32_t pack_launches = ctx->launches;
    rc = spmb_search(ctx, s, p, h);
    ctx->launches += pack_launches;
    spmb_store_destroy(s);
    return rc;
}


// ------------------------------------------------------------------------------------------
// journaled string tree
// ------------------------------------------------------------------------------------------
struct JstBucket {
    spmb_store *store = nullptr;
    uint64_t first = 0, count = 0;
};

struct spmb_jst {
    CtxRef ctx;
    const spmb_store *ref = nullptr;
    JstHost H;
    spmb_journal_event *d_var = nullptr;
    uint8_t *d_pool = nullptr;
    JstFlatHost flat;
    bool have_flat = false;
    JstWindow *d_win = nullptr;
    uint32_t *d_win_vars = nullptr;
    std::vector<JstBucket> buckets;
    uint64_t stats[4] = {0, 0, 0, 0};
    double ms[4] = {0, 0, 0, 0};
    uint32_t in_flight = 0;            // spmb_jst_search_launch calls whose finish is still due
};

static const uint32_t JST_REF_SEQ = 0xffffffffu;

static void jst_drop_flat(spmb_jst *j)
{
    for (auto &b : j->buckets) spmb_store_destroy(b.store);
    j->buckets.clear();
    cudaFree(j->d_win); j->d_win = nullptr;
    cudaFree(j->d_win_vars); j->d_win_vars = nullptr;
    j->have_flat = false;
}

extern "C" int spmb_jst_create(spmb_context *ctx, const spmb_store *reference, const spmb_journal_event *variants,
                               uint64_t n_var, const uint8_t *pool, uint64_t pool_len, const uint64_t *coverage,
                               uint32_t n_hap, spmb_jst **out)
{
    if (!ctx || !reference || !out || n_hap == 0 || (n_var && (!variants || !coverage)) || (pool_len && !pool))
        return SPMB_ERR_ARG;
    if (reference->lay.mode != SPMB_MODE_PLAIN) return set_err(ctx, SPMB_ERR_ARG, "JST reference must be a plain store");
    if (reference->sym_bits != 2) return set_err(ctx, SPMB_ERR_UNSUPPORTED, "the JST is dna4 (2-bit) only");
    if (n_var >= (1ull << 32)) return set_err(ctx, SPMB_ERR_UNSUPPORTED, "more than 2^32 variants");
    ENTER(ctx);
    spmb_jst *j = new (std::nothrow) spmb_jst();
    if (!j) return SPMB_ERR_OOM;
    j->ctx = ctx; j->ref = reference;
    JstHost &H = j->H;
    H.n_hap = n_hap; H.wpr = (n_hap + 63) / 64; H.n_ref = reference->lay.n;
    H.var.assign(variants, variants + n_var);
    H.pool.assign(pool, pool + pool_len);
    H.cov.assign(coverage, coverage + n_var * H.wpr);
    if (n_hap % 64) {
        const uint64_t mask = (1ull << (n_hap % 64)) - 1;
        for (uint64_t v = 0; v < n_var; ++v) H.cov[v * H.wpr + H.wpr - 1] &= mask;
    }
    H.max_del = 0;
    for (uint64_t v = 0; v < n_var; ++v) {
        const spmb_journal_event &x = H.var[v];
        if ((v && x.ref_pos < H.var[v - 1].ref_pos) || x.ref_pos + x.del_len > H.n_ref ||
            (uint64_t)x.ins_off + x.ins_len > pool_len) {
            delete j;
            return set_err(ctx, SPMB_ERR_ARG, "JST variants must be sorted by ref_pos and lie inside the reference / pool");
        }
        H.max_del = std::max(H.max_del, x.del_len);
    }
    {   // every haplotype's carried variants must be a valid journal: the flattening, the coordinate shifts and the
        // clean-coverage test all count a carried variant, so an overlapping one would shift later hits silently.
        // Threads split the coverage words (64 haplotypes each); one pass over the matrix.
        unsigned nt = std::min<unsigned>(std::max(1u, std::thread::hardware_concurrency()), 16u);
        const uint32_t n_grp = (H.wpr + 7) / 8; // 8 words = one cache line of a coverage row
        nt = std::min<unsigned>(nt, n_grp);
        if (n_var * (uint64_t)H.wpr < (1u << 16)) nt = 1;
        std::atomic<int> bad{0};
        auto check = [&](unsigned t) {
            std::vector<uint64_t> last_end(512);
            for (uint32_t g = t; g < n_grp && !bad.load(std::memory_order_relaxed); g += nt) {
                std::fill(last_end.begin(), last_end.end(), 0);
                const uint32_t w0 = g * 8, w1 = std::min(H.wpr, w0 + 8);
                for (uint64_t v = 0; v < n_var; ++v) {
                    const uint64_t p = H.var[v].ref_pos, e = p + H.var[v].del_len;
                    for (uint32_t w = w0; w < w1; ++w) {
                        uint64_t bits = H.cov[v * H.wpr + w];
                        while (bits) {
                            const int b = (int)(w - w0) * 64 + __builtin_ctzll(bits);
                            bits &= bits - 1;
                            if (p < last_end[b]) { bad.store(1); return; }
                            last_end[b] = e;
                        }
                    }
                }
            }
        };
        std::vector<std::thread> th;
        for (unsigned t = 1; t < nt; ++t) th.emplace_back(check, t);
        check(0);
        for (auto &x : th) x.join();
        if (bad.load()) {
            delete j;
            return set_err(ctx, SPMB_ERR_ARG, "JST: the variants carried by one haplotype overlap (or a replacement precedes an insertion at the same position)");
        }
    }
    H.build_ckpt();
    cudaError_t e = cudaMalloc(&j->d_var, (n_var ? n_var : 1) * sizeof(spmb_journal_event));
    if (e == cudaSuccess) e = cudaMalloc(&j->d_pool, pool_len ? pool_len : 1);
    if (e == cudaSuccess && n_var) e = cudaMemcpyAsync(j->d_var, H.var.data(), n_var * sizeof(spmb_journal_event), cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess && pool_len) e = cudaMemcpyAsync(j->d_pool, H.pool.data(), pool_len, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) { spmb_jst_destroy(j); return set_err(ctx, SPMB_ERR_CUDA, "JST upload", e); }
    *out = j;
    return SPMB_OK;
}

extern "C" int spmb_jst_destroy(spmb_jst *j)
{
    if (!j) return SPMB_OK;
    if (j->ctx) cudaSetDevice(j->ctx->device);
    jst_drop_flat(j);
    cudaFree(j->d_var);
    cudaFree(j->d_pool);
    delete j;
    return SPMB_OK;
}

extern "C" int spmb_jst_last_stats(const spmb_jst *j, uint64_t stats[4], double ms[4])
{
    if (!j) return SPMB_ERR_ARG;
    for (int i = 0; i < 4; ++i) { if (stats) stats[i] = j->stats[i]; if (ms) ms[i] = j->ms[i]; }
    return SPMB_OK;
}

// flatten for window size W (cached) and materialise the windows on the device, bucketed by length
static int jst_prepare(spmb_context *ctx, spmb_jst *j, uint32_t W)
{
    if (j->have_flat && j->flat.W == W) { j->ms[0] = 0; j->ms[1] = 0; return SPMB_OK; }
    jst_drop_flat(j);
    jst_flatten(j->H, W, j->flat);
    if (j->flat.overflow) return set_err(ctx, SPMB_ERR_UNSUPPORTED, "JST: a branch context exceeds 2^32-1 symbols / the window-variant list 2^32-1 entries");
    j->ms[0] = j->flat.flatten_ms;
    std::vector<JstWindow> &wins = j->flat.wins;
    const uint64_t nw = wins.size();
    j->stats[0] = nw; j->stats[1] = j->flat.total_symbols;
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0)); CU(cudaEventCreate(&e1));
    CU(cudaEventRecord(e0, ctx->stream));
    static const bool dbg = getenv("SPMB_DEBUG_TIMING") != nullptr;
    const auto th0 = std::chrono::steady_clock::now();
    auto lap = [&](const char *what) {
        if (!dbg) return;
        cudaStreamSynchronize(ctx->stream);
        fprintf(stderr, "[spmb jst_prepare] %-28s %8.2f ms\n", what,
                std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - th0).count());
    };
    if (nw) {
        int rcl = store_ensure_lanes(ctx, j->ref); // the window gather reads the reference's lane chunks
        if (rcl) { cudaEventDestroy(e0); cudaEventDestroy(e1); return rcl; }
        // windows sorted by length; a row holds len symbols + up to 31 pad columns for the Myers score delay
        {   // stable counting sort on the length (a comparison sort of millions of 32-byte records costs 0.2 s)
            uint32_t max_len = 0;
            for (const JstWindow &w : wins) max_len = std::max(max_len, w.len);
            std::vector<uint64_t> start((size_t)max_len + 2, 0);
            for (const JstWindow &w : wins) start[w.len + 1]++;
            for (size_t i = 1; i < start.size(); ++i) start[i] += start[i - 1];
            std::vector<JstWindow> sorted(nw);
            for (const JstWindow &w : wins) sorted[start[w.len]++] = w;
            wins.swap(sorted);
        }
        lap("sort windows by length");
        // geometric length classes: a bucket takes every window up to 1.25x its shortest row
        struct Range { uint64_t first, last; uint32_t RL; };
        std::vector<Range> ranges;
        for (uint64_t first = 0; first < nw;) {
            const uint32_t rl_min = (wins[first].len + 32 + 31) / 32 * 32;
            const uint32_t rl_cap = std::max<uint32_t>(rl_min, (uint32_t)((uint64_t)rl_min * 5 / 4 / 32 * 32));
            uint64_t last = first;
            while (last < nw && (wins[last].len + 32 + 31) / 32 * 32 <= rl_cap) ++last;
            ranges.push_back({first, last, (wins[last - 1].len + 32 + 31) / 32 * 32});
            first = last;
        }
        // Inside a bucket the 32-window groups (one super-tile each: the rows a warp walks together, so a group stays
        // sorted and homogeneous) are dealt into 64 stripes: every contiguous 1/N of a bucket -- what shard r of N
        // searches, spmb_context_set_shard -- then holds short and long windows alike.  In sorted order the last shard
        // of every bucket got its longest windows: C4 on 8 GPUs, kernels 0.150 ms on rank 0 and 0.186 ms on rank 7.
        {
            uint64_t G = 64;
            if (const char *e = getenv("SPMB_JST_STRIPES")) { const long v = atol(e); if (v >= 1) G = (uint64_t)v; } // tests: small trees
            std::vector<JstWindow> tmp;
            for (const Range &rg : ranges) {
                const uint64_t groups = (rg.last - rg.first) / 32; // whole groups; a partial last one stays where it is
                if (groups <= G) continue;
                tmp.assign(wins.begin() + rg.first, wins.begin() + rg.first + groups * 32);
                uint64_t o = rg.first;
                for (uint64_t st = 0; st < G; ++st)
                    for (uint64_t t = st; t < groups; t += G) {
                        std::copy(tmp.begin() + t * 32, tmp.begin() + (t + 1) * 32, wins.begin() + o);
                        o += 32;
                    }
            }
        }
        lap("deal groups into stripes");
        CU(cudaMalloc(&j->d_win, nw * sizeof(JstWindow)));
        CU(cudaMalloc(&j->d_win_vars, (j->flat.win_vars.size() + 1) * sizeof(uint32_t)));
        CU(cudaMemcpyAsync(j->d_win, wins.data(), nw * sizeof(JstWindow), cudaMemcpyHostToDevice, ctx->stream));
        CU(cudaMemcpyAsync(j->d_win_vars, j->flat.win_vars.data(), j->flat.win_vars.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
        lap("upload window table");
        for (const Range &rg : ranges) {
            const uint64_t first = rg.first, last = rg.last;
            const uint32_t RL = rg.RL;
            const uint64_t count = last - first;
            const uint64_t rows = (count + 31) / 32 * 32;
            JstBucket b;
            b.first = first; b.count = count;
            int rc = store_alloc(ctx, rows * RL, RL, SPMB_MODE_WINDOWS, &b.store);
            if (rc) { cudaEventDestroy(e0); cudaEventDestroy(e1); return rc; }
            b.store->seq_base = (uint32_t)first;
            j->buckets.push_back(b);
            std::vector<uint32_t> rl(rows, 0), ro(rows, 0), rs(rows, 0);
            for (uint64_t r = 0; r < count; ++r) {
                rl[r] = wins[first + r].len; ro[r] = wins[first + r].own_lo; rs[r] = (uint32_t)(first + r);
            }
            CU(cudaMalloc(&b.store->row_len, rows * 4));
            CU(cudaMalloc(&b.store->row_own, rows * 4));
            CU(cudaMalloc(&b.store->row_seq, rows * 4));
            CU(cudaMemcpy(b.store->row_len, rl.data(), rows * 4, cudaMemcpyHostToDevice));
            CU(cudaMemcpy(b.store->row_own, ro.data(), rows * 4, cudaMemcpyHostToDevice));
            CU(cudaMemcpy(b.store->row_seq, rs.data(), rows * 4, cudaMemcpyHostToDevice));
            lap("bucket rows up");
            {
                // windows are gathered from the reference's lane chunks; the bucket's own lane chunks (what the
                // windows-mode Myers kernels read) fall out of the same pass
                uint2 *l = nullptr;
                if (big_alloc(ctx, b.store->bytes, (void **)&l) != cudaSuccess) { cudaGetLastError(); l = nullptr; }
                WindowChunkSource src{j->ref->lanes, j->ref->lay, j->d_var, j->d_pool, j->d_win + first, j->d_win_vars, count, b.store->lay};
                pack_chunks_kernel<WindowChunkSource><<<pack_grid(ctx, b.store->lay), 1024, 0, ctx->stream>>>(
                    src, b.store->lay, b.store->planes, l);
                ctx->launches++;
                b.store->lanes = l;
                cudaError_t ek = cudaGetLastError();
                if (ek != cudaSuccess) { cudaEventDestroy(e0); cudaEventDestroy(e1); return set_err(ctx, SPMB_ERR_CUDA, "window materialisation", ek); }
            }
            lap("bucket pack kernel");
        }
    }
    CU(cudaEventRecord(e1, ctx->stream));
    CU(cudaEventSynchronize(e1));
    float t = 0;
    cudaEventElapsedTime(&t, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    j->ms[1] = t;
    j->stats[2] = j->buckets.size();
    j->have_flat = true;
    return SPMB_OK;
}

static int jst_enqueue(spmb_context *ctx, spmb_jst *j, const spmb_pattern *p, spmb_hits *h)
{
    // One scan of the reference + the branch-context windows, one launch per length bucket.  The reference scan
    // and the buckets of a few long windows fill a fraction of the device for tens of microseconds each (C4: the
    // scan + 4 of 7 buckets, 190 of 1010 us): the direct kernels only share the hit sink (atomics), so those
    // launches go to a second stream beside the large buckets.
    const bool direct = !p->filtered && !p->long_prefix && p->alg != SPMB_ALG_PIGEONHOLE && p->alg != SPMB_ALG_MYERS_PREFIX &&
                        ctx->aux_stream && !j->buckets.empty() && !getenv("SPMB_JST_SERIAL");
    cudaStream_t main_stream = ctx->stream;
    if (!direct) {
        int rc = enqueue_search(ctx, j->ref, p, h, true, JST_REF_SEQ);
        if (rc) return rc;
        for (auto &b : j->buckets) { rc = enqueue_search(ctx, b.store, p, h, false, 0); if (rc) return rc; }
        return SPMB_OK;
    }
    if (cudaMemsetAsync(h->d_count, 0, 4 * sizeof(unsigned long long), main_stream) != cudaSuccess ||
        cudaEventRecord(ctx->ev_fork, main_stream) != cudaSuccess ||
        cudaStreamWaitEvent(ctx->aux_stream, ctx->ev_fork, 0) != cudaSuccess)
        return set_err(ctx, SPMB_ERR_CUDA, "jst fork", cudaGetLastError());
    ctx->stream = ctx->aux_stream;
    int rc = enqueue_search(ctx, j->ref, p, h, false, JST_REF_SEQ);
    ctx->stream = main_stream;
    for (auto &b : j->buckets) {
        if (rc) break;
        const bool small = b.store->lanes && b.store->lay.n_st < 16ull * (uint64_t)ctx->sm_count; // under 4 CTAs per SM
        if (small) ctx->stream = ctx->aux_stream;
        rc = enqueue_search(ctx, b.store, p, h, false, 0);
        ctx->stream = main_stream;
    }
    if (cudaEventRecord(ctx->ev_join, ctx->aux_stream) != cudaSuccess ||
        cudaStreamWaitEvent(main_stream, ctx->ev_join, 0) != cudaSuccess)
        return set_err(ctx, SPMB_ERR_CUDA, "jst join", cudaGetLastError());
    return rc;
}

// host half of a JST search: every device hit expanded to the haplotypes it stands for, ordered
static void jst_expand(spmb_jst *j, const spmb_pattern *p, spmb_hits *h, std::vector<spmb_hit> &raw, uint32_t W)
{
    const auto t0 = std::chrono::steady_clock::now();
    if (p->alg == SPMB_ALG_MYERS_FILTERED) hits_order(raw, p->alg); // one record per (window, end)
    j->stats[3] = raw.size();
    const JstHost &H = j->H;
    std::sort(raw.begin(), raw.end(), [](const spmb_hit &a, const spmb_hit &b) {
        if (a.seq_id != b.seq_id) return a.seq_id < b.seq_id;
        if (a.end != b.end) return a.end < b.end;
        return a.pattern_id < b.pattern_id;
    });
    std::vector<uint64_t> cov;
    std::vector<int64_t> shift;
    std::vector<spmb_hit> &out = h->host;
    out.clear();
    size_t i = 0;
    while (i < raw.size()) {
        size_t e = i;
        uint64_t base_pos = 0;
        if (raw[i].seq_id == JST_REF_SEQ) {
            // all reference hits with the same end share the haplotype set and the shifts
            while (e < raw.size() && raw[e].seq_id == JST_REF_SEQ && raw[e].end == raw[i].end) ++e;
            const uint64_t E = raw[i].end;
            H.clean_cov(E, W, cov);
            const uint64_t hi = std::lower_bound(H.var.begin(), H.var.end(), E,
                                                 [](const spmb_journal_event &x, uint64_t v) { return x.ref_pos < v; }) - H.var.begin();
            H.shifts_before(hi, shift);
        } else {
            while (e < raw.size() && raw[e].seq_id == raw[i].seq_id) ++e;
            const JstWindow &wd = j->flat.wins[raw[i].seq_id];
            const uint32_t *ids = j->flat.win_vars.data() + wd.var_off;
            H.window_cov(ids, wd.n_var, W, cov);
            H.shifts_before(ids[0], shift);
            base_pos = wd.ref_start;
        }
        for (uint32_t w = 0; w < H.wpr; ++w) {
            uint64_t bits = cov[w];
            while (bits) {
                const uint32_t hap = w * 64 + (uint32_t)__builtin_ctzll(bits);
                bits &= bits - 1;
                if (hap >= H.n_hap) continue;
                for (size_t q = i; q < e; ++q) {
                    spmb_hit x = raw[q];
                    x.end = (uint64_t)((int64_t)(raw[q].end + base_pos) + shift[hap]);
                    x.seq_id = hap;
                    out.push_back(x);
                }
            }
        }
        i = e;
    }
    hits_order(out, p->alg);
    h->count = out.size();
    h->view = nullptr;
    h->pend_store = nullptr; h->pend_pat = nullptr; h->pend_jst = nullptr;
    j->ms[3] = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

extern "C" int spmb_jst_search(spmb_context *ctx, spmb_jst *j, const spmb_pattern *p, spmb_hits *h)
{
    if (!ctx || !j || !p || !h) return SPMB_ERR_ARG;
    ENTER(ctx);
    ctx->launches = 0;
    h->count = 0;
    h->host.clear();
    h->view = nullptr;
    const uint32_t W = spmb_pattern_window_size(p);
    j->stats[3] = 0; j->ms[2] = j->ms[3] = 0;
    if (W == 0) return SPMB_OK;
    int rc = jst_prepare(ctx, j, W);
    if (rc) return rc;

    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0)); CU(cudaEventCreate(&e1));
    std::vector<spmb_hit> raw;
    auto run = [&]() -> int {
        CU(cudaEventRecord(e0, ctx->stream));
        int r = jst_enqueue(ctx, j, p, h);
        if (r) return r;
        CU(cudaEventRecord(e1, ctx->stream));
        return SPMB_OK;
    };
    rc = run();
    if (rc == SPMB_OK) rc = hits_collect(ctx, h, raw, run);
    if (rc) { cudaEventDestroy(e0); cudaEventDestroy(e1); return rc; }
    float t = 0;
    cudaEventElapsedTime(&t, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    j->ms[2] = t;
    jst_expand(j, p, h, raw, W);
    return SPMB_OK;
}

// The same search in two halves, like spmb_search_launch / spmb_search_finish: launch enqueues the kernels of the
// reference scan and of every window bucket and returns; finish waits for them, brings the device hits back and
// expands them on the host.  With two or three hit lists in rotation the host half of a search (D2H, expansion,
// ordering: 0.25 of the 1.09 ms of a C4 search) runs under the kernels of the next one.
extern "C" int spmb_jst_search_launch(spmb_context *ctx, spmb_jst *j, const spmb_pattern *p, spmb_hits *h)
{
    if (!ctx || !j || !p || !h) return SPMB_ERR_ARG;
    ENTER(ctx);
    ctx->launches = 0;
    h->view = nullptr;
    h->ordered = false;
    h->pend_store = nullptr;
    h->pend_pat = p;
    h->pend_jst = j;
    h->done_armed = false;
    const uint32_t W = spmb_pattern_window_size(p);
    if (W == 0) { j->in_flight++; return SPMB_OK; }
    // another window size rebuilds the window stores the pending searches still read
    if (j->in_flight && !(j->have_flat && j->flat.W == W)) {
        h->pend_jst = nullptr; h->pend_pat = nullptr;
        return set_err(ctx, SPMB_ERR_ARG, "spmb_jst_search_launch: searches in flight on this tree use another window size "
                                          "(finish them first)");
    }
    int rc = jst_prepare(ctx, j, W);
    if (rc == SPMB_OK) rc = jst_enqueue(ctx, j, p, h);
    if (rc) { h->pend_jst = nullptr; h->pend_pat = nullptr; return rc; }
    j->in_flight++;
    h->done_armed = h->done && ctx->copy_stream && cudaEventRecord(h->done, ctx->stream) == cudaSuccess;
    return SPMB_OK;
}

extern "C" int spmb_jst_search_finish(spmb_context *ctx, spmb_hits *h)
{
    if (!ctx || !h || !h->pend_jst || !h->pend_pat) return SPMB_ERR_ARG;
    ENTER(ctx);
    spmb_jst *j = h->pend_jst;
    const spmb_pattern *p = h->pend_pat;
    h->count = 0;
    h->host.clear();
    const uint32_t W = spmb_pattern_window_size(p);
    j->stats[3] = 0; j->ms[2] = j->ms[3] = 0; // the kernels of a split search are not timed here (the caller brackets the launch)
    if (j->in_flight) j->in_flight--;
    h->pend_jst = nullptr; h->pend_pat = nullptr;
    if (W == 0) return SPMB_OK;
    std::vector<spmb_hit> raw;
    int rc = hits_collect(ctx, h, raw, [&]() { return jst_enqueue(ctx, j, p, h); });
    if (rc) return rc;
    jst_expand(j, p, h, raw, W);
    return SPMB_OK;
}

// ------------------------------------------------------------------------------------------
// resumable search: the carried state is the last `keep` symbols of the stream, resident on the device
// ------------------------------------------------------------------------------------------
static uint32_t stream_keep(const spmb_pattern *p)
{
    const uint32_t ws = spmb_pattern_window_size(p);
    uint32_t keep = ws ? ws - 1 : 0;
    // Myers: 2m symbols determine the score at the current column (capture().errors, stream.cuh)
    if (p->alg == SPMB_ALG_MYERS || p->alg == SPMB_ALG_MYERS_FILTERED) keep = std::max(keep, 2 * p->max_m);
    return keep;
}

extern "C" int spmb_stream_create(spmb_context *ctx, const spmb_pattern *p, spmb_stream **out)
{
    if (!ctx || !p || !out) return SPMB_ERR_ARG;
    if (p->sym_bits != 2) return set_err(ctx, SPMB_ERR_UNSUPPORTED, "streams are dna4 (2-bit) only");
    if (p->alg == SPMB_ALG_MYERS_PREFIX && p->n_pat != 1) return set_err(ctx, SPMB_ERR_UNSUPPORTED, "an anchored stream takes a single pattern");
    ENTER(ctx);
    spmb_stream *st = new (std::nothrow) spmb_stream();
    if (!st) return SPMB_ERR_OOM;
    st->ctx = ctx; st->pat = p;
    st->keep = stream_keep(p);
    cudaError_t e = cudaMalloc(&st->d_tail, ((size_t)st->keep / 32 + 2) * sizeof(uint2));
    if (e == cudaSuccess) e = cudaMemset(st->d_tail, 0, ((size_t)st->keep / 32 + 2) * sizeof(uint2));
    if (e == cudaSuccess) e = cudaMalloc(&st->d_err, (p->n_pat ? p->n_pat : 1) * sizeof(uint32_t));
    for (int i = 0; i < SPMB_STREAM_SLOTS && e == cudaSuccess; ++i) {
        int rc = spmb_hits_create(ctx, 1 << 12, &st->slot[i].hits);
        if (rc) { spmb_stream_destroy(st); return rc; }
    }
    if (e != cudaSuccess) { spmb_stream_destroy(st); return set_err(ctx, SPMB_ERR_OOM, "stream state", e); }
    *out = st;
    return SPMB_OK;
}

static void stream_slot_release(spmb_stream *st, int i)
{
    spmb_stream::Slot &S = st->slot[i];
    if (S.store) { spmb_store_destroy(S.store); S.store = nullptr; }
    S.busy = false;
}

extern "C" int spmb_stream_destroy(spmb_stream *st)
{
    if (!st) return SPMB_OK;
    if (st->ctx) { cudaSetDevice(st->ctx->device); cudaStreamSynchronize(st->ctx->stream); }
    // chunks handed in but never collected: their upload jobs still write into this stream's buffers
    for (int i = 0, si = st->next_collect; i < SPMB_STREAM_SLOTS; ++i, si = (si + 1) % SPMB_STREAM_SLOTS)
        if (st->slot[si].busy && st->slot[si].job && st->ctx && st->ctx->upload) {
            st->ctx->upload->wait(st->slot[si].job_seq, st->ctx->stream, true, nullptr);
            st->slot[si].job = false;
        }
    for (int i = 0; i < SPMB_STREAM_SLOTS; ++i) {
        stream_slot_release(st, i);
        if (st->slot[i].d_words && st->ctx) big_free(st->ctx, st->slot[i].d_words, st->slot[i].words_bytes);
        cudaFree(st->slot[i].d_raw);
        spmb_hits_destroy(st->slot[i].hits);
    }
    cudaFree(st->d_tail);
    cudaFree(st->d_err);
    delete st;
    return SPMB_OK;
}

// the anchored matcher scans at most m + k + 1 symbols from the stream start (myers_prefix_matcher_restorable.hpp:47-50):
// the whole state is that short prefix, kept on the host
static int stream_feed_prefix(spmb_context *ctx, spmb_stream *st, const uint8_t *host_ranks, uint64_t n, spmb_hits *h)
{
    const uint32_t ws = spmb_pattern_window_size(st->pat);
    const size_t clamp = ws ? (size_t)ws + 1 : 0;
    const size_t tl = st->tail.size();
    h->host.clear();
    h->count = 0;
    h->view = nullptr;
    if (st->offset >= clamp || n == 0) { st->offset += n; return SPMB_OK; }
    const size_t take = std::min<size_t>(n, clamp - tl);
    st->scratch.resize(tl + take);
    if (tl) memcpy(st->scratch.data(), st->tail.data(), tl);
    memcpy(st->scratch.data() + tl, host_ranks, take);
    int rc = spmb_search_host(ctx, st->scratch.data(), st->scratch.size(), st->pat, h);
    if (rc) return rc;
    size_t w = 0;
    for (size_t i = 0; i < h->host.size(); ++i)
        if (h->host[i].end > tl) { spmb_hit x = h->host[i]; x.end -= tl; h->host[w++] = x; }
    h->host.resize(w);
    h->count = w;
    st->tail = st->scratch;
    st->offset += n;
    return SPMB_OK;
}

// hand one chunk in.  The bulk of a long chunk goes to the upload pool as an asynchronous job (the packing threads
// and the copy engine start on it at once and run on into the next chunk's job); everything that depends on the
// previous chunk's state -- [device tail | chunk] -> store -> search kernels -> new tail -- is enqueued by
// spmb_stream_collect, in order.  The host buffer must stay valid until that chunk has been collected.
extern "C" int spmb_stream_submit(spmb_context *ctx, spmb_stream *st, const uint8_t *host_ranks, uint64_t n)
{
    if (!ctx || !st || (!host_ranks && n)) return SPMB_ERR_ARG;
    if (st->pat->alg == SPMB_ALG_MYERS_PREFIX) return set_err(ctx, SPMB_ERR_UNSUPPORTED, "anchored streams are fed with spmb_stream_feed");
    ENTER(ctx);
    const int si = st->next_submit;
    spmb_stream::Slot &S = st->slot[si];
    if (S.busy) return set_err(ctx, SPMB_ERR_ARG, "three chunks are in flight: collect one first");
    const uint64_t t = st->tail_len, total = t + n;
    S.t = t; S.n = n; S.total = total; S.launches = 0;
    S.job = false; S.host = nullptr; S.bulk = 0; S.a = 0;
    S.empty = (total == 0 || st->pat->max_m == 0);
    // host mirror of the tail (serialised by capture(); the searches never read it)
    {
        std::vector<uint8_t> &ht = st->host_tail;
        const size_t nk = (size_t)std::min<uint64_t>(st->keep, total);
        if (n >= nk) ht.assign(host_ranks + (n - nk), host_ranks + n);
        else {
            const size_t from_old = nk - (size_t)n;
            ht.erase(ht.begin(), ht.end() - from_old);
            ht.insert(ht.end(), host_ranks, host_ranks + n);
        }
        for (auto &c : ht) c &= 3;
    }
    if (total) {
        // ---- room for [tail | chunk] as lane-chunk words ----
        const size_t words = (size_t)((total + 31) / 32) + 2;
        if (words * sizeof(uint2) > S.words_bytes) {
            if (S.d_words) big_free(ctx, S.d_words, S.words_bytes);
            S.d_words = nullptr; S.words_bytes = 0;
            const size_t want = words * sizeof(uint2) + (words * sizeof(uint2)) / 8;
            CU(big_alloc(ctx, want, (void **)&S.d_words));
            S.words_bytes = want;
            CU(cudaStreamSynchronize(ctx->stream)); // the movers' streams may touch a fresh stream-ordered block only now
        }
        uint64_t a = (32 - (t & 31u)) & 31u; // symbols that complete the tail's partial word
        if (a > n) a = n;
        S.a = (uint32_t)a;
        memset(&S.hd, 0, sizeof(S.hd));
        if (a) memcpy(S.hd.sym, host_ranks, (size_t)a);
        S.bulk = n - a;
        S.host = host_ranks + a;
        // the pool stays warm from feed to feed, so a stream takes it from 4 Mi symbols on (a one-shot upload: 32 Mi)
        const uint64_t pool_min = UploadPool::env_u64("SPMB_STREAM_UPLOAD_MIN", std::min<uint64_t>(4ull << 20, UploadPool::env_u64("SPMB_UPLOAD_MIN", 32ull << 20)));
        if (S.bulk >= pool_min && ensure_upload(ctx) == SPMB_OK) {
            // about two chunks per mover and feed: the movers run on into the next feed's job, so a feed needs no fine
            // cut to keep them level, and every chunk costs three driver calls that all movers' threads contend for
            // (512 Ki-symbol chunks: 64 Mi-symbol feeds at 0.52 of the one-shot rate).  The copy engine's share of RAW
            // chunks is capped: a one-shot upload finds its share (~16 % here) by racing the packing threads, but in a
            // stream the raw mover reaches the next job while they are still in the previous one and would take half of
            // it at four times the PCIe bytes per symbol (traced: 8 of 17 chunks, packing threads idle).
            const uint64_t movers = (uint64_t)ctx->upload->n_workers + 1;
            uint64_t chunk = S.bulk / (2 * movers);
            if (chunk < (1u << 20)) chunk = 1u << 20;
            static const uint64_t raw_pct = UploadPool::env_u64("SPMB_STREAM_RAW_PCT", 15);
            if (!ctx->upload->submit_text(S.host, S.bulk, S.d_words + (t + a) / 32, chunk, &S.job_seq, raw_pct))
                return set_err(ctx, SPMB_ERR_CUDA, "stream upload: job ring full");
            S.job = true;
        }
    }
    st->next_submit = (st->next_submit + 1) % SPMB_STREAM_SLOTS;
    S.busy = true;
    st->tail_len = (uint32_t)std::min<uint64_t>(st->keep, total);
    st->offset += n;
    return SPMB_OK;
}

// the device side of the oldest chunk in flight, in stream order behind the previous chunk's
static int stream_enqueue(spmb_context *ctx, spmb_stream *st, spmb_stream::Slot &S)
{
    cudaStream_t sm = ctx->stream;
    const uint64_t t = S.t, total = S.total;
    ctx->launches = 0;
    if (S.job) {
        cudaError_t e = ctx->upload->wait(S.job_seq, sm, false, &ctx->launches);
        S.job = false;
        if (e != cudaSuccess) { cudaGetLastError(); return set_err(ctx, SPMB_ERR_CUDA, "stream upload", e); }
    } else if (S.bulk) {
        if (S.bulk > S.raw_bytes) {
            cudaFree(S.d_raw);
            S.d_raw = nullptr; S.raw_bytes = 0;
            CU(cudaMalloc(&S.d_raw, S.bulk + S.bulk / 8 + 64));
            S.raw_bytes = S.bulk + S.bulk / 8 + 64;
        }
        CU(cudaMemcpyAsync(S.d_raw, S.host, S.bulk, cudaMemcpyHostToDevice, sm));
        const uint64_t groups = (S.bulk + 31) / 32;
        bytes_to_chunks_kernel<<<(unsigned)((groups + 255) / 256), 256, 0, sm>>>(S.d_raw, S.bulk, S.d_words + (t + S.a) / 32);
        ctx->launches++;
        CU(cudaGetLastError());
    }
    if (t) CU(cudaMemcpyAsync(S.d_words, st->d_tail, ((size_t)(t + 31) / 32) * sizeof(uint2), cudaMemcpyDeviceToDevice, sm));
    if (S.a) {
        stream_head_kernel<<<1, 32, 0, sm>>>(S.d_words, t, S.hd, S.a);
        ctx->launches++;
    }
    // ---- the words as a store (planes for the exact kernel, lane chunks for the others) ----
    if (S.store) { spmb_store_destroy(S.store); S.store = nullptr; }
    int rc = store_alloc(ctx, total, 0, SPMB_MODE_PLAIN, &S.store);
    if (rc) return rc;
    {
        uint2 *l = nullptr;
        const bool want_lanes = !(st->pat->alg == SPMB_ALG_HORSPOOL || st->pat->alg == SPMB_ALG_SHIFTOR);
        if (want_lanes && big_alloc(ctx, S.store->bytes, (void **)&l) != cudaSuccess) { cudaGetLastError(); l = nullptr; }
        ChunkSource src{S.d_words, total};
        pack_chunks_kernel<ChunkSource><<<pack_grid(ctx, S.store->lay), 1024, 0, sm>>>(src, S.store->lay, S.store->planes, l);
        S.store->lanes = l;
        ctx->launches++;
        CU(cudaGetLastError());
    }
    // ---- the state for the next chunk ----
    const uint32_t nk = (uint32_t)std::min<uint64_t>(st->keep, total);
    if (nk) {
        stream_carry_kernel<<<(nk / 32 + 1 + 63) / 64, 64, 0, sm>>>(S.d_words, total, nk, st->d_tail);
        ctx->launches++;
        CU(cudaGetLastError());
    }
    const uint32_t pre = ctx->launches;
    if (!S.empty) {
        rc = spmb_search_launch(ctx, S.store, st->pat, S.hits); // resets ctx->launches
        ctx->launches += pre;
        if (rc) return rc;
    }
    S.launches = ctx->launches;
    return SPMB_OK;
}

// hits of the oldest chunk in flight; ends are relative to that chunk's start, like the reference's fresh Finder on
// every chunk (myers_matcher_restorable_test.cpp:66-69)
extern "C" int spmb_stream_collect(spmb_context *ctx, spmb_stream *st, spmb_hits *h)
{
    if (!ctx || !st || !h) return SPMB_ERR_ARG;
    ENTER(ctx);
    const int si = st->next_collect;
    spmb_stream::Slot &S = st->slot[si];
    if (!S.busy) return set_err(ctx, SPMB_ERR_ARG, "no chunk in flight");
    st->next_collect = (st->next_collect + 1) % SPMB_STREAM_SLOTS;
    h->host.clear();
    h->count = 0;
    h->view = nullptr;
    h->pend_store = nullptr; h->pend_pat = nullptr;
    int rc = SPMB_OK;
    if (S.total) rc = stream_enqueue(ctx, st, S);
    if (rc == SPMB_OK && !S.empty) {
        rc = spmb_search_finish(ctx, S.hits);
        if (rc == SPMB_OK) {
            const spmb_hit *d = spmb_hits_data(S.hits);
            const uint64_t cnt = spmb_hits_count(S.hits);
            h->host.reserve(cnt);
            // hits ending inside the carried context were reported by an earlier feed
            for (uint64_t i = 0; i < cnt; ++i)
                if (d[i].end > S.t) { spmb_hit x = d[i]; x.end -= S.t; h->host.push_back(x); }
            h->count = h->host.size();
        }
    } else if (rc == SPMB_OK && S.total) {
        // nothing to search for, but the caller's buffer must have left before we return
        cudaError_t e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) rc = set_err(ctx, SPMB_ERR_CUDA, "stream sync", e);
    }
    if (rc != SPMB_OK && S.job) { // leave no job of a failed chunk behind
        ctx->upload->wait(S.job_seq, ctx->stream, true, nullptr);
        S.job = false;
    }
    ctx->launches = S.launches;
    stream_slot_release(st, si);
    return rc;
}

extern "C" int spmb_stream_feed(spmb_context *ctx, spmb_stream *st, const uint8_t *host_ranks, uint64_t n, spmb_hits *h)
{
    if (!ctx || !st || !h || (!host_ranks && n)) return SPMB_ERR_ARG;
    if (st->pat->alg == SPMB_ALG_MYERS_PREFIX) return stream_feed_prefix(ctx, st, host_ranks, n, h);
    if (st->any_busy()) return set_err(ctx, SPMB_ERR_ARG, "chunks submitted with spmb_stream_submit are still in flight");
    int rc = spmb_stream_submit(ctx, st, host_ranks, n);
    if (rc) return rc;
    return spmb_stream_collect(ctx, st, h);
}

// the score of every needle at the current column (Myers streams; 0 for the exact searchers)
static int stream_errors(spmb_stream *st, std::vector<uint32_t> &err)
{
    const spmb_pattern *p = st->pat;
    err.assign(p->n_pat, 0);
    if (!(p->alg == SPMB_ALG_MYERS || p->alg == SPMB_ALG_MYERS_FILTERED) || p->NW == 0) return SPMB_OK;
    spmb_context *ctx = st->ctx;
    ENTER(ctx);
    CU(cudaMemsetAsync(st->d_err, 0, p->n_pat * sizeof(uint32_t), ctx->stream)); // empty needles of a bucketed set
    std::vector<const spmb_pattern *> sets(p->subs.begin(), p->subs.end());
    if (sets.empty() && !p->long_prefix) sets.push_back(p);
    for (const spmb_pattern *q : sets) {
        StreamScoreArgs a{st->d_tail, st->tail_len, q->d_eq, q->d_meta, q->n_pat, st->d_err};
        const unsigned grid = (q->n_pat + 63) / 64;
        switch (q->NW) {
        case 1: stream_score_kernel<1><<<grid, 64, 0, ctx->stream>>>(a); break;
        case 2: stream_score_kernel<2><<<grid, 64, 0, ctx->stream>>>(a); break;
        case 3: stream_score_kernel<3><<<grid, 64, 0, ctx->stream>>>(a); break;
        case 4: stream_score_kernel<4><<<grid, 64, 0, ctx->stream>>>(a); break;
        case 5: stream_score_kernel<5><<<grid, 64, 0, ctx->stream>>>(a); break;
        case 6: stream_score_kernel<6><<<grid, 64, 0, ctx->stream>>>(a); break;
        case 7: stream_score_kernel<7><<<grid, 64, 0, ctx->stream>>>(a); break;
        case 8: stream_score_kernel<8><<<grid, 64, 0, ctx->stream>>>(a); break;
        default: return set_err(ctx, SPMB_ERR_UNSUPPORTED, "stream score: needle too long");
        }
    }
    if (p->long_prefix) { // needles longer than 256 symbols: one warp each
        StreamScoreArgs a{st->d_tail, st->tail_len, p->d_chain_eq, p->d_chain_meta, p->n_long, st->d_err};
        const unsigned grid = (p->n_long + 3) / 4;
        switch (p->chain_wpl) {
        case 1: stream_score_chain_kernel<1><<<grid, 128, 0, ctx->stream>>>(a); break;
        case 2: stream_score_chain_kernel<2><<<grid, 128, 0, ctx->stream>>>(a); break;
        case 4: stream_score_chain_kernel<4><<<grid, 128, 0, ctx->stream>>>(a); break;
        default: stream_score_chain_kernel<8><<<grid, 128, 0, ctx->stream>>>(a); break;
        }
    }
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(err.data(), st->d_err, p->n_pat * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return SPMB_OK;
}

// state blob: {u64 offset, u32 tail_len, u32 n_patterns, u32 errors[n_patterns], u8 tail[tail_len]}
extern "C" int spmb_stream_capture(const spmb_stream *st_c, void *buf, size_t *size)
{
    if (!st_c || !size) return SPMB_ERR_ARG;
    spmb_stream *st = const_cast<spmb_stream *>(st_c);
    const bool prefix = st->pat->alg == SPMB_ALG_MYERS_PREFIX;
    const std::vector<uint8_t> &tl = prefix ? st->tail : st->host_tail;
    const uint32_t np = st->pat->n_pat;
    const size_t need = 16 + 4 * (size_t)np + tl.size();
    if (!buf) { *size = need; return SPMB_OK; }
    if (*size < need) { *size = need; return SPMB_ERR_ARG; }
    std::vector<uint32_t> err(np, 0);
    if (!prefix) {
        if (st->any_busy()) return SPMB_ERR_ARG; // collect first: the state is that of the last chunk
        int rc = stream_errors(st, err);
        if (rc) return rc;
    }
    char *o = (char *)buf;
    const uint64_t off = st->offset;
    const uint32_t tlen = (uint32_t)tl.size();
    memcpy(o, &off, 8); memcpy(o + 8, &tlen, 4); memcpy(o + 12, &np, 4);
    if (np) memcpy(o + 16, err.data(), 4 * (size_t)np);
    if (tlen) memcpy(o + 16 + 4 * (size_t)np, tl.data(), tlen);
    *size = need;
    return SPMB_OK;
}

extern "C" int spmb_stream_restore(spmb_stream *st, const void *buf, size_t size)
{
    if (!st || !buf || size < 16) return SPMB_ERR_ARG;
    const char *in = (const char *)buf;
    uint64_t off; uint32_t tlen, np;
    memcpy(&off, in, 8); memcpy(&tlen, in + 8, 4); memcpy(&np, in + 12, 4);
    if (np != st->pat->n_pat || size < 16 + 4 * (size_t)np + tlen) return SPMB_ERR_ARG;
    const uint8_t *tb = (const uint8_t *)in + 16 + 4 * (size_t)np;
    if (st->pat->alg == SPMB_ALG_MYERS_PREFIX) {
        st->offset = off;
        st->tail.assign(tb, tb + tlen);
        return SPMB_OK;
    }
    if (tlen > st->keep || st->any_busy()) return SPMB_ERR_ARG;
    spmb_context *ctx = st->ctx;
    ENTER(ctx);
    std::vector<uint64_t> words((size_t)st->keep / 32 + 2, 0);
    if (tlen) spmb_host_pack_chunks(tb, tlen, words.data()); // format conversion only (hostpack.cpp)
    CU(cudaMemcpyAsync(st->d_tail, words.data(), words.size() * 8, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    st->host_tail.assign(tb, tb + tlen);
    for (auto &c : st->host_tail) c &= 3;
    st->tail_len = tlen;
    st->offset = off;
    return SPMB_OK;
}

// the score of needle `pattern_index` at the current column of a Myers stream (capture().errors of the reference)
extern "C" int spmb_stream_errors(spmb_stream *st, uint32_t pattern_index, uint32_t *errors)
{
    if (!st || !errors || pattern_index >= st->pat->n_pat) return SPMB_ERR_ARG;
    if (st->pat->alg == SPMB_ALG_MYERS_PREFIX || st->any_busy()) return SPMB_ERR_ARG;
    std::vector<uint32_t> err;
    int rc = stream_errors(st, err);
    if (rc) return rc;
    *errors = err[pattern_index];
    return SPMB_OK;
}

extern "C" uint64_t spmb_stream_offset(const spmb_stream *st) { return st ? st->offset : 0; }

// ------------------------------------------------------------------------------------------
// measurement helpers
// ------------------------------------------------------------------------------------------
extern "C" int spmb_measure_int_peak(spmb_context *ctx, double *lop3_gops, double *iadd3_gops)
{
    if (!ctx) return SPMB_ERR_ARG;
    ENTER(ctx);
    uint32_t *d = nullptr;
    CU(cudaMalloc(&d, 4));
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0));
    CU(cudaEventCreate(&e1));
    const int iters = 4000, blocks = ctx->sm_count * 8, threads = 256;
    double res[2] = {0, 0};
    for (int op = 0; op < 2; ++op) {
        float best = 1e30f;
        for (int rep = 0; rep < 4; ++rep) {
            cudaEventRecord(e0, ctx->stream);
            if (op == 0) int_peak_kernel<0><<<blocks, threads, 0, ctx->stream>>>(d, iters, 0x5a5a5a5au + rep);
            else int_peak_kernel<1><<<blocks, threads, 0, ctx->stream>>>(d, iters, 0x5a5a5a5au + rep);
            cudaEventRecord(e1, ctx->stream);
            cudaError_t e = cudaEventSynchronize(e1);
            if (e != cudaSuccess) { cudaFree(d); return set_err(ctx, SPMB_ERR_CUDA, "int peak probe", e); }
            float ms = 0;
            cudaEventElapsedTime(&ms, e0, e1);
            if (rep > 0 && ms < best) best = ms;
        }
        double ops = (double)blocks * threads * (double)iters * 64.0;
        res[op] = ops / (best * 1e-3) / 1e9;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d);
    if (lop3_gops) *lop3_gops = res[0];
    if (iadd3_gops) *iadd3_gops = res[1];
    return SPMB_OK;
}
