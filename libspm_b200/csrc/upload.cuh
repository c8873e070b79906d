// upload.cuh -- host haystack -> device, two movers over one chunk queue.
//
// The haystack of seqan_pattern_base::operator() (libspm/libspm/matcher/seqan_pattern_base.hpp:40-47)
// is host memory, 1 byte per symbol.  PCIe is the narrowest link on the way to the search kernels
// (~50 GB/s against 6.5 TB/s of HBM), and the device keeps 2 bits per symbol, so every chunk that a
// host core packs before it crosses the link costs a quarter of the PCIe time.  The sequence is cut
// into chunks; the calling thread feeds RAW chunks from the head of the queue to the copy engine
// (staging slot -> bytes_to_chunks_kernel), worker threads take chunks from the TAIL, pack them with
// hostpack.cpp into pinned buffers and copy the packed words straight to their final place.  Both
// ends claim from one atomic {head, tail} word, so the split follows the machine: no host cores ->
// everything goes raw, a fast host -> most of it goes packed.
// Both paths fill the same intermediate: d_chunks[p/32] = the 32 symbols at p as one lane-chunk word
// (pack.cuh); one pack_chunks_kernel<ChunkSource> pass then builds the strided bit planes.
#pragma once
#include <cuda_runtime.h>
#include <sched.h>

#include "../../include/spmb.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>

extern "C" void spmb_host_pack_chunks(const uint8_t *ranks, uint64_t n, uint64_t *out);

struct ChunkSource {
    const uint2 *__restrict__ in;
    uint64_t n;
    __device__ __forceinline__ uint2 load_chunk(uint64_t p0) const { return __ldg(in + (p0 >> 5)); }
};

// 4 symbols (first in the lowest byte) -> 8 bits, first symbol on top
__device__ __forceinline__ uint32_t spmb_fold4(uint32_t w)
{
    const uint32_t r = __byte_perm(w, 0, 0x0123) & 0x03030303u;
    return (r | (r >> 6) | (r >> 12) | (r >> 18)) & 0xffu;
}

// raw staging slot (1 B/symbol, n_sym symbols, 16-byte aligned) -> lane-chunk words
__global__ void __launch_bounds__(256) bytes_to_chunks_kernel(const uint8_t *__restrict__ in, uint64_t n_sym,
                                                              uint2 *__restrict__ out)
{
    const uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t p0 = g * 32;
    if (p0 >= n_sym) return;
    uint32_t w[8];
    if (p0 + 32 <= n_sym) {
        const uint4 *q = reinterpret_cast<const uint4 *>(in + p0);
        const uint4 a = __ldg(q), b = __ldg(q + 1);
        w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w;
        w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
    } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            uint32_t v = 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint64_t p = p0 + 4 * i + j;
                if (p < n_sym) v |= (uint32_t)in[p] << (8 * j);
            }
            w[i] = v;
        }
    }
    uint2 o;
    o.x = (spmb_fold4(w[0]) << 24) | (spmb_fold4(w[1]) << 16) | (spmb_fold4(w[2]) << 8) | spmb_fold4(w[3]);
    o.y = (spmb_fold4(w[4]) << 24) | (spmb_fold4(w[5]) << 16) | (spmb_fold4(w[6]) << 8) | spmb_fold4(w[7]);
    out[g] = o;
}

// ---- journal events (24 B each in the C ABI) cross the link the same way: a host thread rewrites a chunk
// of events as 8-byte records, the device expands it back; chunks that do not fit the short form, and the
// chunks the copy engine claims, travel as they are.  Short form: the chunk is cut into blocks of
// EV_BLOCK events, each with a header {base_pos, base_off} (position of its first event, pool offset of
// its first inserting event); a record holds the position and pool-offset deltas and the two lengths
// (< 256).  An event that inserts nothing never reads the pool: its offset is only checked against the
// pool length (like the device validation would) and comes back as base_off.
enum { EV_BLOCK = 1024 };
struct SpmbPackedEvent {
    uint32_t dpos;
    uint16_t doff;
    uint8_t del_len, ins_len;
};
struct SpmbPackedEventHeader {
    uint64_t base_pos, base_off;
};

// slot = [ceil(n / EV_BLOCK) headers][n records]
__global__ void __launch_bounds__(256) expand_events_kernel(const uint8_t *__restrict__ slot, uint32_t n,
                                                            spmb_journal_event *__restrict__ out)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t n_blk = (n + EV_BLOCK - 1) / EV_BLOCK;
    const SpmbPackedEventHeader h = reinterpret_cast<const SpmbPackedEventHeader *>(slot)[i / EV_BLOCK];
    const SpmbPackedEvent r = reinterpret_cast<const SpmbPackedEvent *>(slot + n_blk * sizeof(SpmbPackedEventHeader))[i];
    spmb_journal_event e;
    e.ref_pos = h.base_pos + r.dpos;
    e.del_len = r.del_len;
    e.ins_len = r.ins_len;
    e.ins_off = h.base_off + r.doff;
    out[i] = e;
}

static inline size_t spmb_packed_events_bytes(uint32_t n)
{
    return (size_t)((n + EV_BLOCK - 1) / EV_BLOCK) * sizeof(SpmbPackedEventHeader) + (size_t)n * sizeof(SpmbPackedEvent);
}

// host side of the above (hostpack.cpp; AVX2 where the CPU has it): 0 = some record does not fit, the chunk
// then travels raw
static_assert(sizeof(SpmbPackedEvent) == 8 && sizeof(SpmbPackedEventHeader) == 16, "short event form");
extern "C" int spmb_host_pack_events(const spmb_journal_event *ev, uint32_t n, uint64_t pool_len, uint8_t *out);

static inline int spmb_host_cpus()
{
    cpu_set_t set;
    if (sched_getaffinity(0, sizeof(set), &set) == 0) {
        int c = CPU_COUNT(&set);
        if (c > 0) return c;
    }
    unsigned h = std::thread::hardware_concurrency();
    return h ? (int)h : 1;
}

// knobs (environment, read once per context): SPMB_HOST_THREADS = packing threads (0 = raw only;
// default: available cpus - 1, at most 31), SPMB_UPLOAD_CHUNK = symbols per chunk (default 4 Mi),
// SPMB_UPLOAD_MIN = shortest haystack that takes this path (default 32 Mi symbols)
struct UploadPool {
    enum { SLOTS = 4 };
    int device = 0;
    uint64_t chunk_sym = 4ull << 20;            // capacity of a staging buffer, and the chunk size of one-shot uploads
    uint64_t job_chunk = 4ull << 20;            // chunk size of the current text job (<= chunk_sym; stream feeds use smaller ones)
    uint64_t min_sym = 32ull << 20;
    int n_workers = 0;
    bool trace = false;                         // SPMB_UPLOAD_TRACE=1: per-job timings on stderr
    uint64_t raw_pct_text = 100, raw_pct_events = 10; // SPMB_UPLOAD_RAW_PCT / SPMB_EVENT_RAW_PCT: most chunks (in %) the copy engine may take raw

    struct Worker {
        std::thread th;
        cudaStream_t stream = nullptr;
        uint64_t *pinned[2] = {nullptr, nullptr};
        uint8_t *d_slot[2] = {nullptr, nullptr};   // device landing area of a packed event chunk
        cudaEvent_t ev[2] = {nullptr, nullptr};
        bool used[2] = {false, false};
    };
    std::vector<Worker> workers;
    uint8_t *d_slot[SLOTS] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t slot_ev[SLOTS] = {nullptr, nullptr, nullptr, nullptr};
    bool ready = false;

    // job state
    std::mutex mu;
    std::condition_variable cv_job, cv_done;
    uint64_t gen = 0;
    bool quit = false;
    int running = 0;
    const uint8_t *src = nullptr;
    uint64_t n = 0, n_chunks = 0;
    uint2 *d_chunks = nullptr;
    const spmb_journal_event *ev_src = nullptr; // events job (src == nullptr)
    spmb_journal_event *d_ev = nullptr;
    uint64_t ev_chunk = 0;                      // events per chunk
    uint64_t ev_pool_len = 0;
    std::atomic<uint64_t> ends{0};  // head << 32 | tail: chunks [head, tail) are unclaimed
    std::atomic<int> cuda_err{0};
    std::atomic<uint64_t> packed_chunks{0};
    std::atomic<uint64_t> unpackable_chunks{0};
    std::atomic<uint32_t> worker_launches{0};
    uint64_t last_ev_raw = 0, last_ev_packed = 0, last_ev_chunk = 0;
    // figures of the last upload (spmb_upload_last_stats)
    uint64_t last_raw = 0, last_packed = 0;

    static uint64_t env_u64(const char *name, uint64_t dflt)
    {
        const char *v = getenv(name);
        if (!v || !*v) return dflt;
        char *end = nullptr;
        unsigned long long x = strtoull(v, &end, 10);
        return end && end != v ? (uint64_t)x : dflt;
    }

    bool claim_head(uint64_t &idx)
    {
        uint64_t v = ends.load(std::memory_order_relaxed);
        for (;;) {
            const uint64_t h = v >> 32, t = v & 0xffffffffull;
            if (h >= t) return false;
            if (ends.compare_exchange_weak(v, ((h + 1) << 32) | t, std::memory_order_acq_rel)) { idx = h; return true; }
        }
    }
    bool claim_tail(uint64_t &idx)
    {
        uint64_t v = ends.load(std::memory_order_relaxed);
        for (;;) {
            const uint64_t h = v >> 32, t = v & 0xffffffffull;
            if (h >= t) return false;
            if (ends.compare_exchange_weak(v, (h << 32) | (t - 1), std::memory_order_acq_rel)) { idx = t - 1; return true; }
        }
    }

    cudaError_t init(int dev, int threads_override = -1)
    {
        device = dev;
        raw_pct_text = env_u64("SPMB_UPLOAD_RAW_PCT", raw_pct_text);
        raw_pct_events = env_u64("SPMB_EVENT_RAW_PCT", raw_pct_events);
        chunk_sym = env_u64("SPMB_UPLOAD_CHUNK", chunk_sym);
        chunk_sym = (chunk_sym + 1023) / 1024 * 1024;
        if (chunk_sym < 1024) chunk_sym = 1024;
        min_sym = env_u64("SPMB_UPLOAD_MIN", min_sym);
        trace = env_u64("SPMB_UPLOAD_TRACE", 0) != 0;
        // SPMB_UPLOAD_SPIN=0: threads sleep while they wait for a slot instead of polling
        const unsigned ev_flags = cudaEventDisableTiming | (env_u64("SPMB_UPLOAD_SPIN", 1) ? 0u : (unsigned)cudaEventBlockingSync);
        int dflt = spmb_host_cpus() - 1;
        if (dflt > 31) dflt = 31;
        if (dflt < 0) dflt = 0;
        n_workers = (int)env_u64("SPMB_HOST_THREADS", (uint64_t)dflt);
        if (threads_override >= 0) n_workers = threads_override; // spmb_context_set_upload_threads
        if (n_workers > 256) n_workers = 256;
        cudaError_t e = cudaSuccess;
        for (int s = 0; s < SLOTS && e == cudaSuccess; ++s) {
            e = cudaMalloc(&d_slot[s], chunk_sym);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&slot_ev[s], ev_flags);
        }
        workers.resize(n_workers);
        for (int w = 0; w < n_workers && e == cudaSuccess; ++w) {
            Worker &W = workers[w];
            e = cudaStreamCreateWithFlags(&W.stream, cudaStreamNonBlocking);
            for (int b = 0; b < 2 && e == cudaSuccess; ++b) {
                e = cudaHostAlloc((void **)&W.pinned[b], chunk_sym / 4, cudaHostAllocDefault);
                if (e == cudaSuccess) e = cudaMalloc(&W.d_slot[b], chunk_sym / 4);
                if (e == cudaSuccess) e = cudaEventCreateWithFlags(&W.ev[b], ev_flags);
            }
        }
        if (e != cudaSuccess) return e;
        for (int w = 0; w < n_workers; ++w) workers[w].th = std::thread([this, w] { worker_main(w); });
        ready = true;
        return cudaSuccess;
    }

    void worker_main(int w)
    {
        Worker &W = workers[w];
        cudaSetDevice(device);
        uint64_t seen = 0;
        for (;;) {
            {
                std::unique_lock<std::mutex> lk(mu);
                cv_job.wait(lk, [&] { return quit || gen != seen; });
                if (quit) return;
                seen = gen;
            }
            int b = 0;
            uint64_t idx;
            cudaError_t e = cudaSuccess;
            auto now = [] { return std::chrono::steady_clock::now(); };
            auto us = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point c) {
                return (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(c - a).count();
            };
            const auto tw0 = now();
            uint64_t t_wait = 0, t_pack = 0, t_api = 0;
            while (e == cudaSuccess && cuda_err.load(std::memory_order_relaxed) == 0 && claim_tail(idx)) {
                const auto ta = now();
                if (W.used[b]) e = cudaEventSynchronize(W.ev[b]);
                const auto tb = now();
                t_wait += us(ta, tb);
                if (src) {
                    const uint64_t p0 = idx * job_chunk;
                    const uint64_t len = (n - p0 < job_chunk) ? n - p0 : job_chunk;
                    spmb_host_pack_chunks(src + p0, len, W.pinned[b]);
                    t_pack += us(tb, now());
                    if (e == cudaSuccess)
                        e = cudaMemcpyAsync(d_chunks + p0 / 32, W.pinned[b], (len + 31) / 32 * 8, cudaMemcpyHostToDevice, W.stream);
                } else {
                    const uint64_t i0 = idx * ev_chunk;
                    const uint32_t len = (uint32_t)((n - i0 < ev_chunk) ? n - i0 : ev_chunk);
                    const bool fits = spmb_host_pack_events(ev_src + i0, len, ev_pool_len, (uint8_t *)W.pinned[b]);
                    t_pack += us(tb, now());
                    if (fits) {
                        if (e == cudaSuccess)
                            e = cudaMemcpyAsync(W.d_slot[b], W.pinned[b], spmb_packed_events_bytes(len), cudaMemcpyHostToDevice, W.stream);
                        if (e == cudaSuccess) {
                            expand_events_kernel<<<(len + 255) / 256, 256, 0, W.stream>>>(W.d_slot[b], len, d_ev + i0);
                            e = cudaGetLastError();
                            worker_launches.fetch_add(1, std::memory_order_relaxed);
                        }
                    } else if (e == cudaSuccess) {
                        e = cudaMemcpyAsync(d_ev + i0, ev_src + i0, (size_t)len * sizeof(spmb_journal_event), cudaMemcpyHostToDevice, W.stream);
                        unpackable_chunks.fetch_add(1, std::memory_order_relaxed);
                    }
                }
                if (e == cudaSuccess) e = cudaEventRecord(W.ev[b], W.stream);
                W.used[b] = true;
                b ^= 1;
                packed_chunks.fetch_add(1, std::memory_order_relaxed);
            }
            const auto tc = now();
            if (e == cudaSuccess) e = cudaStreamSynchronize(W.stream);
            if (trace) {
                t_api = us(tw0, tc) - t_wait - t_pack;
                fprintf(stderr, "[spmb upload] worker %d: total %.2f ms  pack %.2f  wait-slot %.2f  api %.2f  drain %.2f\n", w,
                        us(tw0, now()) / 1e6, t_pack / 1e6, t_wait / 1e6, t_api / 1e6, us(tc, now()) / 1e6);
            }
            W.used[0] = W.used[1] = false;
            if (e != cudaSuccess) cuda_err.store((int)e);
            {
                std::lock_guard<std::mutex> lk(mu);
                if (--running == 0) cv_done.notify_all();
            }
        }
    }

    // the calling thread's half of a job: start the workers, feed head chunks to the copy engine (at most
    // SLOTS-1 ahead of the link, so the split stays adaptive), wait for the workers, drain `stream`
    template <class RawChunk>
    cudaError_t drive(cudaStream_t stream, uint64_t *raw_out, RawChunk raw_chunk, uint64_t raw_pct = 100)
    {
        cuda_err.store(0);
        packed_chunks.store(0);
        unpackable_chunks.store(0);
        worker_launches.store(0);
        ends.store(n_chunks & 0xffffffffull); // head 0, tail n_chunks
        // the destination comes from the stream-ordered allocator: the workers' streams may touch it only
        // after the allocation point of `stream`
        const auto t0 = std::chrono::steady_clock::now();
        cudaError_t e = cudaStreamSynchronize(stream);
        if (e != cudaSuccess) return e;
        const auto t1 = std::chrono::steady_clock::now();
        if (n_workers) {
            std::lock_guard<std::mutex> lk(mu);
            running = n_workers;
            ++gen;
            cv_job.notify_all();
        }
        uint64_t idx, raw = 0;
        int s = 0;
        bool used[SLOTS] = {false, false, false, false};
        while (e == cudaSuccess && cuda_err.load(std::memory_order_relaxed) == 0) {
            if (used[s]) e = cudaEventSynchronize(slot_ev[s]);
            // a raw chunk costs the link 3-4x the bytes of a packed one: with packing threads at work the copy
            // engine's share is capped (raw_pct of the chunks), or a fast link starves the packed copies
            if (n_workers && raw * 100 >= raw_pct * n_chunks) break;
            if (e != cudaSuccess || !claim_head(idx)) break;
            e = raw_chunk(idx, s);
            if (e == cudaSuccess) e = cudaEventRecord(slot_ev[s], stream);
            used[s] = true;
            s = (s + 1) % SLOTS;
            ++raw;
        }
        if (e != cudaSuccess) { // stop the workers early: nothing left to claim
            cuda_err.store((int)e);
            ends.store(0);
        }
        const auto t2 = std::chrono::steady_clock::now();
        if (n_workers) {
            std::unique_lock<std::mutex> lk(mu);
            cv_done.wait(lk, [&] { return running == 0; });
        }
        const auto t3 = std::chrono::steady_clock::now();
        // the caller's buffer may be reused as soon as we return: the raw copies must have left it
        cudaError_t e2 = cudaStreamSynchronize(stream);
        if (e == cudaSuccess) e = e2;
        if (e == cudaSuccess && cuda_err.load() != 0) e = (cudaError_t)cuda_err.load();
        *raw_out = raw;
        if (trace) {
            auto ms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point c) {
                return std::chrono::duration<double, std::milli>(c - a).count();
            };
            fprintf(stderr, "[spmb upload] %s job: %llu chunks (%llu raw)  pre-sync %.2f ms  raw loop %.2f  wait workers %.2f  drain %.2f\n",
                    src ? "text" : "events", (unsigned long long)n_chunks, (unsigned long long)raw, ms(t0, t1), ms(t1, t2), ms(t2, t3),
                    ms(t3, std::chrono::steady_clock::now()));
        }
        return e;
    }

    // fills dst[0, ceil(n/32)) (device, allocated in `stream`'s order) from host_ranks[0, n).  Returns when
    // every byte of host_ranks has been consumed and dst is complete.
    // chunk = 0: chunk_sym; else the chunk size of this job (multiple of 1024, <= chunk_sym): a short job (one feed
    // of a stream) is cut finer so that every mover gets several chunks
    cudaError_t run(const uint8_t *host_ranks, uint64_t n_sym, uint2 *dst, cudaStream_t stream, uint32_t *launches, uint64_t chunk = 0)
    {
        src = host_ranks;
        ev_src = nullptr;
        n = n_sym;
        d_chunks = dst;
        job_chunk = (chunk == 0 || chunk > chunk_sym) ? chunk_sym : std::max<uint64_t>(1024, chunk / 1024 * 1024);
        n_chunks = (n_sym + job_chunk - 1) / job_chunk;
        cudaError_t e = drive(stream, &last_raw, [&](uint64_t idx, int s) -> cudaError_t {
            const uint64_t p0 = idx * job_chunk;
            const uint64_t len = (n - p0 < job_chunk) ? n - p0 : job_chunk;
            cudaError_t r = cudaMemcpyAsync(d_slot[s], src + p0, len, cudaMemcpyHostToDevice, stream);
            if (r != cudaSuccess) return r;
            const uint64_t groups = (len + 31) / 32;
            bytes_to_chunks_kernel<<<(unsigned)((groups + 255) / 256), 256, 0, stream>>>(d_slot[s], len, d_chunks + p0 / 32);
            if (launches) ++*launches;
            return cudaGetLastError();
        }, raw_pct_text);
        last_packed = packed_chunks.load();
        return e;
    }

    // same queue for the journal's event table: dst[0, n_ev) <- ev[0, n_ev)
    cudaError_t run_events(const spmb_journal_event *ev, uint64_t n_ev, uint64_t pool_len, spmb_journal_event *dst,
                           cudaStream_t stream, uint32_t *launches)
    {
        ev_pool_len = pool_len;
        src = nullptr;
        ev_src = ev;
        d_ev = dst;
        n = n_ev;
        // whole blocks that fit a worker's pinned buffer (chunk_sym / 4 bytes), at most 64 Ki events
        ev_chunk = (chunk_sym / 4) / (EV_BLOCK * sizeof(SpmbPackedEvent) + sizeof(SpmbPackedEventHeader)) * EV_BLOCK;
        if (ev_chunk > 65536) ev_chunk = 65536;
        if (ev_chunk == 0) ev_chunk = 16;       // tiny test geometries: one partial block
        n_chunks = (n_ev + ev_chunk - 1) / ev_chunk;
        cudaError_t e = drive(stream, &last_ev_raw, [&](uint64_t idx, int) -> cudaError_t {
            const uint64_t i0 = idx * ev_chunk;
            const uint64_t len = (n - i0 < ev_chunk) ? n - i0 : ev_chunk;
            return cudaMemcpyAsync(d_ev + i0, ev_src + i0, len * sizeof(spmb_journal_event), cudaMemcpyHostToDevice, stream);
        }, raw_pct_events);
        last_ev_packed = packed_chunks.load() - unpackable_chunks.load();
        last_ev_raw += unpackable_chunks.load();
        last_ev_chunk = ev_chunk;
        if (launches) *launches += worker_launches.load();
        return e;
    }

    ~UploadPool()
    {
        {
            std::lock_guard<std::mutex> lk(mu);
            quit = true;
            cv_job.notify_all();
        }
        for (auto &W : workers) if (W.th.joinable()) W.th.join();
        cudaSetDevice(device);
        for (auto &W : workers) {
            for (int b = 0; b < 2; ++b) {
                if (W.pinned[b]) cudaFreeHost(W.pinned[b]);
                if (W.d_slot[b]) cudaFree(W.d_slot[b]);
                if (W.ev[b]) cudaEventDestroy(W.ev[b]);
            }
            if (W.stream) cudaStreamDestroy(W.stream);
        }
        for (int s = 0; s < SLOTS; ++s) {
            if (d_slot[s]) cudaFree(d_slot[s]);
            if (slot_ev[s]) cudaEventDestroy(slot_ev[s]);
        }
        cudaGetLastError();
    }
};
