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

// raw staging slot (1 B/symbol, n_sym symbols, 16-byte aligned) -> lane-chunk words; shift = which two bits of a
// symbol byte are packed (0: dna4 and the low pair of a 4-bit alphabet, 2: its high pair)
__global__ void __launch_bounds__(256) bytes_to_chunks_kernel(const uint8_t *__restrict__ in, uint64_t n_sym,
                                                              uint2 *__restrict__ out, uint32_t shift = 0)
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
    if (shift) {
#pragma unroll
        for (int i = 0; i < 8; ++i) w[i] = (w[i] >> shift) & 0x03030303u;
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
//
// Jobs are ASYNCHRONOUS (round 2): submit_*() puts a job into a ring and returns; the packing threads and one raw-mover
// thread (which feeds the copy engine from the head of the chunk queue) take the jobs in order and move on to the
// next job the moment the current one has no chunk left to claim -- a stream of feeds (spmb_stream_submit) keeps every
// mover busy across feed boundaries instead of ramping the pool up and down per feed (measured before: 64 Mi-symbol
// feeds at 0.53 of the one-shot rate).  wait() blocks until every copy of the job has been ENQUEUED, makes the caller's
// stream wait for the movers' events, and (host_sync) for the copies themselves.  The source buffer of a job must
// stay valid until its wait() returned.
struct UploadPool {
    enum { SLOTS = 4, RING = 4 };
    int device = 0;
    uint64_t chunk_sym = 4ull << 20;            // capacity of a staging buffer, and the chunk size of one-shot uploads
    uint64_t job_chunk = 4ull << 20;            // chunk size of the last text job (<= chunk_sym; stream feeds use smaller ones)
    uint64_t min_sym = 32ull << 20;
    int n_workers = 0;
    bool trace = false;                         // SPMB_UPLOAD_TRACE=1: per-job timings on stderr
    uint64_t raw_pct_text = 100, raw_pct_events = 10; // SPMB_UPLOAD_RAW_PCT / SPMB_EVENT_RAW_PCT: most chunks (in %) the copy engine may take raw
    uint64_t spin_us = 200;                     // SPMB_UPLOAD_IDLE_SPIN_US: an idle mover polls this long for the next job before it sleeps

    struct Job {
        const uint8_t *src = nullptr;           // text job
        uint2 *d_chunks = nullptr;
        const spmb_journal_event *ev_src = nullptr; // events job (src == nullptr)
        spmb_journal_event *d_ev = nullptr;
        uint64_t n = 0, n_chunks = 0, chunk = 0, ev_pool_len = 0, raw_pct = 100;
        bool caller_waits = true;               // a synchronous run(): the caller's thread sleeps until the job is done
        std::atomic<uint64_t> ends{0};          // head << 32 | tail: chunks [head, tail) are unclaimed
        std::atomic<int> cuda_err{0};
        std::atomic<uint64_t> packed_chunks{0}, unpackable_chunks{0}, raw_chunks{0};
        std::atomic<uint32_t> launches{0};
        int movers_left = 0;                    // guarded by mu
        bool issued = false;                    // guarded by mu: every mover has left the job, all its copies are enqueued
        std::chrono::steady_clock::time_point t_submit;
    };
    Job jobs[RING];
    std::atomic<uint64_t> submitted{0};         // jobs handed in so far; job s lives in jobs[s % RING]
    uint64_t retired = 0;                       // jobs waited for (in order)

    struct Worker {
        std::thread th;
        cudaStream_t stream = nullptr;
        uint64_t *pinned[2] = {nullptr, nullptr};
        uint8_t *d_slot[2] = {nullptr, nullptr};   // device landing area of a packed event chunk
        cudaEvent_t ev[2] = {nullptr, nullptr};
        bool used[2] = {false, false};
        cudaEvent_t job_ev[RING] = {nullptr, nullptr, nullptr, nullptr}; // behind this mover's last copy of a job
    };
    std::vector<Worker> workers;
    // the raw mover: staging slots on the device, its own stream
    std::thread raw_th;
    cudaStream_t raw_stream = nullptr;
    uint8_t *d_slot[SLOTS] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t slot_ev[SLOTS] = {nullptr, nullptr, nullptr, nullptr};
    bool slot_used[SLOTS] = {false, false, false, false};
    cudaEvent_t raw_job_ev[RING] = {nullptr, nullptr, nullptr, nullptr};
    bool ready = false;

    std::mutex mu;
    std::condition_variable cv_job, cv_done;
    bool quit = false;
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

    static bool claim_head(Job &J, uint64_t &idx)
    {
        uint64_t v = J.ends.load(std::memory_order_relaxed);
        for (;;) {
            const uint64_t h = v >> 32, t = v & 0xffffffffull;
            if (h >= t) return false;
            if (J.ends.compare_exchange_weak(v, ((h + 1) << 32) | t, std::memory_order_acq_rel)) { idx = h; return true; }
        }
    }
    static bool claim_tail(Job &J, uint64_t &idx)
    {
        uint64_t v = J.ends.load(std::memory_order_relaxed);
        for (;;) {
            const uint64_t h = v >> 32, t = v & 0xffffffffull;
            if (h >= t) return false;
            if (J.ends.compare_exchange_weak(v, (h << 32) | (t - 1), std::memory_order_acq_rel)) { idx = t - 1; return true; }
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
        spin_us = env_u64("SPMB_UPLOAD_IDLE_SPIN_US", spin_us);
        // SPMB_UPLOAD_SPIN=0: threads sleep while they wait for a slot instead of polling
        const unsigned ev_flags = cudaEventDisableTiming | (env_u64("SPMB_UPLOAD_SPIN", 1) ? 0u : (unsigned)cudaEventBlockingSync);
        int dflt = spmb_host_cpus() - 1;
        if (dflt > 31) dflt = 31;
        if (dflt < 0) dflt = 0;
        n_workers = (int)env_u64("SPMB_HOST_THREADS", (uint64_t)dflt);
        if (threads_override >= 0) n_workers = threads_override; // spmb_context_set_upload_threads
        if (n_workers > 256) n_workers = 256;
        // the copy engine's share of RAW event chunks (24 B per event against 8 packed): 10 % with 15 packing threads
        // (measured, round 1), more the fewer threads a device has -- the multi-GPU driver gives each of 8 devices 3
        // threads, and a 10 % cap left the links idle while those packed (journal phase 15.8 ms at 8 devices)
        if (!getenv("SPMB_EVENT_RAW_PCT")) {
            const int pct = 100 - 6 * n_workers;
            raw_pct_events = (uint64_t)(pct < 10 ? 10 : pct);
        }
        cudaError_t e = cudaStreamCreateWithFlags(&raw_stream, cudaStreamNonBlocking);
        for (int s = 0; s < SLOTS && e == cudaSuccess; ++s) {
            e = cudaMalloc(&d_slot[s], chunk_sym);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&slot_ev[s], cudaEventDisableTiming);
        }
        for (int r = 0; r < RING && e == cudaSuccess; ++r) e = cudaEventCreateWithFlags(&raw_job_ev[r], cudaEventDisableTiming);
        workers.resize(n_workers);
        for (int w = 0; w < n_workers && e == cudaSuccess; ++w) {
            Worker &W = workers[w];
            e = cudaStreamCreateWithFlags(&W.stream, cudaStreamNonBlocking);
            for (int b = 0; b < 2 && e == cudaSuccess; ++b) {
                e = cudaHostAlloc((void **)&W.pinned[b], chunk_sym / 4, cudaHostAllocDefault);
                if (e == cudaSuccess) e = cudaMalloc(&W.d_slot[b], chunk_sym / 4);
                if (e == cudaSuccess) e = cudaEventCreateWithFlags(&W.ev[b], ev_flags);
            }
            for (int r = 0; r < RING && e == cudaSuccess; ++r) e = cudaEventCreateWithFlags(&W.job_ev[r], cudaEventDisableTiming);
        }
        if (e != cudaSuccess) return e;
        for (int w = 0; w < n_workers; ++w) workers[w].th = std::thread([this, w] { mover_main(w); });
        raw_th = std::thread([this] { mover_main(-1); });
        ready = true;
        return cudaSuccess;
    }

    // wait for job `seq` to be submitted (a short poll first: a stream of feeds hands in the next job within
    // microseconds, and a condition-variable wake-up costs 30-60 us per mover and feed); false = quit
    bool next_job(uint64_t seq, bool spin)
    {
        if (submitted.load(std::memory_order_acquire) > seq) return true;
        const auto t0 = std::chrono::steady_clock::now();
        while (spin && std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now() - t0).count() < (long long)spin_us) {
            if (submitted.load(std::memory_order_acquire) > seq) return true;
#if defined(__x86_64__)
            __builtin_ia32_pause();
#endif
        }
        std::unique_lock<std::mutex> lk(mu);
        cv_job.wait(lk, [&] { return quit || submitted.load(std::memory_order_acquire) > seq; });
        return !quit || submitted.load(std::memory_order_acquire) > seq;
    }

    void leave_job(Job &J)
    {
        std::lock_guard<std::mutex> lk(mu);
        if (--J.movers_left == 0) { J.issued = true; cv_done.notify_all(); }
    }

    // w >= 0: packing thread w (claims from the tail); w < 0: the raw mover (claims from the head)
    void mover_main(int w)
    {
        cudaSetDevice(device);
        auto now = [] { return std::chrono::steady_clock::now(); };
        auto ns = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point c) {
            return (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(c - a).count();
        };
        int b = 0, s = 0; // next pinned buffer (packing thread) / staging slot (raw mover): carried from job to job
        for (uint64_t seq = 0;; ++seq) {
            if (!next_job(seq, w >= 0)) return;
            Job &J = jobs[seq % RING];
            const int r = (int)(seq % RING);
            cudaError_t e = cudaSuccess;
            uint64_t idx;
            const auto tw0 = now();
            uint64_t t_wait = 0, t_pack = 0, n_mine = 0;
            if (w >= 0) {
                Worker &W = workers[w];
                while (e == cudaSuccess && J.cuda_err.load(std::memory_order_relaxed) == 0 && claim_tail(J, idx)) {
                    const auto ta = now();
                    if (W.used[b]) e = cudaEventSynchronize(W.ev[b]);
                    const auto tb = now();
                    t_wait += ns(ta, tb);
                    if (J.src) {
                        const uint64_t p0 = idx * J.chunk;
                        const uint64_t len = (J.n - p0 < J.chunk) ? J.n - p0 : J.chunk;
                        spmb_host_pack_chunks(J.src + p0, len, W.pinned[b]);
                        t_pack += ns(tb, now());
                        if (e == cudaSuccess)
                            e = cudaMemcpyAsync(J.d_chunks + p0 / 32, W.pinned[b], (len + 31) / 32 * 8, cudaMemcpyHostToDevice, W.stream);
                    } else {
                        const uint64_t i0 = idx * J.chunk;
                        const uint32_t len = (uint32_t)((J.n - i0 < J.chunk) ? J.n - i0 : J.chunk);
                        const bool fits = spmb_host_pack_events(J.ev_src + i0, len, J.ev_pool_len, (uint8_t *)W.pinned[b]);
                        t_pack += ns(tb, now());
                        if (fits) {
                            if (e == cudaSuccess)
                                e = cudaMemcpyAsync(W.d_slot[b], W.pinned[b], spmb_packed_events_bytes(len), cudaMemcpyHostToDevice, W.stream);
                            if (e == cudaSuccess) {
                                expand_events_kernel<<<(len + 255) / 256, 256, 0, W.stream>>>(W.d_slot[b], len, J.d_ev + i0);
                                e = cudaGetLastError();
                                J.launches.fetch_add(1, std::memory_order_relaxed);
                            }
                        } else if (e == cudaSuccess) {
                            e = cudaMemcpyAsync(J.d_ev + i0, J.ev_src + i0, (size_t)len * sizeof(spmb_journal_event), cudaMemcpyHostToDevice, W.stream);
                            J.unpackable_chunks.fetch_add(1, std::memory_order_relaxed);
                        }
                    }
                    if (e == cudaSuccess) e = cudaEventRecord(W.ev[b], W.stream);
                    W.used[b] = true;
                    b ^= 1;
                    ++n_mine;
                    J.packed_chunks.fetch_add(1, std::memory_order_relaxed);
                }
                if (e == cudaSuccess) e = cudaEventRecord(W.job_ev[r], W.stream);
            } else {
                while (e == cudaSuccess && J.cuda_err.load(std::memory_order_relaxed) == 0) {
                    // text chunks land in a staging slot: wait for the slot (this also keeps the raw mover at the
                    // link's pace, so the split with the packing threads follows the machine).  Event chunks are
                    // copied in place and their raw share is capped, so nothing is waited for: they are short
                    // (1.5 MB), and a sleeping wait per chunk cost more than the copy (raw-only: 78 -> 17 ms).
                    if (J.src) {
                        // The raw mover spends its life in this wait.  While the caller sleeps in run() it polls (the
                        // caller's core is free); behind an asynchronous submit the caller's thread is at work too --
                        // packing threads + raw mover + caller are one thread more than the default thread count
                        // leaves cores, and a packing thread that lost its core for a time slice held a whole stream
                        // feed back by 0.3-0.6 ms (traced) -- so there it sleeps between looks.
                        const auto ta = now();
                        if (slot_used[s]) {
                            if (J.caller_waits) e = cudaEventSynchronize(slot_ev[s]);
                            else {
                                cudaError_t q;
                                while ((q = cudaEventQuery(slot_ev[s])) == cudaErrorNotReady) std::this_thread::sleep_for(std::chrono::microseconds(20));
                                e = q;
                            }
                        }
                        t_wait += ns(ta, now());
                    }
                    // a raw chunk costs the link 3-4x the bytes of a packed one: with packing threads at work the copy
                    // engine's share is capped (raw_pct of the chunks), or a fast link starves the packed copies
                    if (n_workers && n_mine * 100 >= J.raw_pct * J.n_chunks) break;
                    if (e != cudaSuccess || !claim_head(J, idx)) break;
                    if (J.src) {
                        const uint64_t p0 = idx * J.chunk;
                        const uint64_t len = (J.n - p0 < J.chunk) ? J.n - p0 : J.chunk;
                        e = cudaMemcpyAsync(d_slot[s], J.src + p0, len, cudaMemcpyHostToDevice, raw_stream);
                        if (e == cudaSuccess) {
                            const uint64_t groups = (len + 31) / 32;
                            bytes_to_chunks_kernel<<<(unsigned)((groups + 255) / 256), 256, 0, raw_stream>>>(d_slot[s], len, J.d_chunks + p0 / 32);
                            e = cudaGetLastError();
                            J.launches.fetch_add(1, std::memory_order_relaxed);
                        }
                        if (e == cudaSuccess) e = cudaEventRecord(slot_ev[s], raw_stream);
                        slot_used[s] = true;
                        s = (s + 1) % SLOTS;
                    } else {
                        const uint64_t i0 = idx * J.chunk;
                        const uint64_t len = (J.n - i0 < J.chunk) ? J.n - i0 : J.chunk;
                        e = cudaMemcpyAsync(J.d_ev + i0, J.ev_src + i0, len * sizeof(spmb_journal_event), cudaMemcpyHostToDevice, raw_stream);
                    }
                    ++n_mine;
                }
                J.raw_chunks.store(n_mine);
                if (e == cudaSuccess) e = cudaEventRecord(raw_job_ev[r], raw_stream);
            }
            if (e != cudaSuccess) { // stop the other movers early: nothing left to claim
                J.cuda_err.store((int)e);
                J.ends.store(0);
            }
            if (trace)
                fprintf(stderr, "[spmb upload] job %llu mover %d: %llu chunks  total %.3f ms  pack %.3f  wait-slot %.3f  (since submit %.3f)\n",
                        (unsigned long long)seq, w, (unsigned long long)n_mine, ns(tw0, now()) / 1e6, t_pack / 1e6, t_wait / 1e6,
                        ns(J.t_submit, now()) / 1e6);
            leave_job(J);
        }
    }

    // ---- job hand-in ----
    Job *begin_job(uint64_t &seq)
    {
        // the ring holds RING jobs that have not been waited for
        seq = submitted.load(std::memory_order_relaxed);
        if (seq - retired >= RING) return nullptr;
        Job &J = jobs[seq % RING];
        J.src = nullptr; J.d_chunks = nullptr; J.ev_src = nullptr; J.d_ev = nullptr;
        J.cuda_err.store(0); J.packed_chunks.store(0); J.unpackable_chunks.store(0); J.raw_chunks.store(0); J.launches.store(0);
        J.movers_left = n_workers + 1;
        J.issued = false;
        J.t_submit = std::chrono::steady_clock::now();
        return &J;
    }
    void commit_job(Job &J)
    {
        J.ends.store(J.n_chunks & 0xffffffffull); // head 0, tail n_chunks
        {
            std::lock_guard<std::mutex> lk(mu);
            submitted.fetch_add(1, std::memory_order_release);
        }
        cv_job.notify_all();
    }

    // host_ranks[0, n_sym) -> dst[0, ceil(n/32)) (device memory every stream may touch already).  chunk = 0: chunk_sym;
    // else the chunk size of this job (multiple of 1024, <= chunk_sym).  false: the ring is full (wait() first).
    bool submit_text(const uint8_t *host_ranks, uint64_t n_sym, uint2 *dst, uint64_t chunk, uint64_t *seq_out, uint64_t raw_pct = 0,
                     bool caller_waits = false)
    {
        uint64_t seq;
        Job *J = begin_job(seq);
        if (!J) return false;
        J->src = host_ranks; J->n = n_sym; J->d_chunks = dst;
        J->chunk = (chunk == 0 || chunk > chunk_sym) ? chunk_sym : std::max<uint64_t>(1024, chunk / 1024 * 1024);
        J->n_chunks = (n_sym + J->chunk - 1) / J->chunk;
        J->raw_pct = raw_pct ? std::min(raw_pct, raw_pct_text) : raw_pct_text;
        J->caller_waits = caller_waits;
        job_chunk = J->chunk;
        commit_job(*J);
        *seq_out = seq;
        return true;
    }

    // every copy of job `seq` is enqueued; `stream` waits for them; host_sync: they have completed (the source
    // buffer is free again).  Jobs are waited for in the order they were submitted.
    cudaError_t wait(uint64_t seq, cudaStream_t stream, bool host_sync, uint32_t *launches)
    {
        Job &J = jobs[seq % RING];
        const int r = (int)(seq % RING);
        {
            std::unique_lock<std::mutex> lk(mu);
            cv_done.wait(lk, [&] { return J.issued; });
        }
        cudaError_t e = J.cuda_err.load() ? (cudaError_t)J.cuda_err.load() : cudaSuccess;
        for (int w = 0; w < n_workers && e == cudaSuccess; ++w) e = cudaStreamWaitEvent(stream, workers[w].job_ev[r], 0);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(stream, raw_job_ev[r], 0);
        if (e == cudaSuccess && host_sync) e = cudaStreamSynchronize(stream);
        if (e != cudaSuccess) { // leave no copy of this job in flight
            for (auto &W : workers) cudaStreamSynchronize(W.stream);
            cudaStreamSynchronize(raw_stream);
            cudaGetLastError();
        }
        if (launches) *launches += J.launches.load();
        if (J.src) { last_raw = J.raw_chunks.load(); last_packed = J.packed_chunks.load(); }
        else {
            last_ev_packed = J.packed_chunks.load() - J.unpackable_chunks.load();
            last_ev_raw = J.raw_chunks.load() + J.unpackable_chunks.load();
            last_ev_chunk = J.chunk;
        }
        if (trace)
            fprintf(stderr, "[spmb upload] job %llu (%s): %llu chunks (%llu raw)  submit -> waited %.3f ms\n", (unsigned long long)seq,
                    J.src ? "text" : "events", (unsigned long long)J.n_chunks, (unsigned long long)J.raw_chunks.load(),
                    std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - J.t_submit).count());
        ++retired;
        return e;
    }

    // fills dst[0, ceil(n/32)) (device, allocated in `stream`'s order) from host_ranks[0, n).  Returns when
    // every byte of host_ranks has been consumed and dst is complete.
    cudaError_t run(const uint8_t *host_ranks, uint64_t n_sym, uint2 *dst, cudaStream_t stream, uint32_t *launches, uint64_t chunk = 0)
    {
        // the destination comes from the stream-ordered allocator: the movers' streams may touch it only
        // after the allocation point of `stream`
        cudaError_t e = cudaStreamSynchronize(stream);
        if (e != cudaSuccess) return e;
        uint64_t seq;
        if (!submit_text(host_ranks, n_sym, dst, chunk, &seq, 0, true)) return cudaErrorNotReady; // jobs of a stream still in flight
        return wait(seq, stream, true, launches);
    }

    // same queue for the journal's event table: dst[0, n_ev) <- ev[0, n_ev)
    cudaError_t run_events(const spmb_journal_event *ev, uint64_t n_ev, uint64_t pool_len, spmb_journal_event *dst,
                           cudaStream_t stream, uint32_t *launches)
    {
        cudaError_t e = cudaStreamSynchronize(stream);
        if (e != cudaSuccess) return e;
        uint64_t seq;
        Job *J = begin_job(seq);
        if (!J) return cudaErrorNotReady;
        J->ev_pool_len = pool_len;
        J->ev_src = ev; J->d_ev = dst; J->n = n_ev;
        // whole blocks that fit a worker's pinned buffer (chunk_sym / 4 bytes), at most 64 Ki events
        J->chunk = (chunk_sym / 4) / (EV_BLOCK * sizeof(SpmbPackedEvent) + sizeof(SpmbPackedEventHeader)) * EV_BLOCK;
        if (J->chunk > 65536) J->chunk = 65536;
        if (J->chunk == 0) J->chunk = 16;       // tiny test geometries: one partial block
        J->n_chunks = (n_ev + J->chunk - 1) / J->chunk;
        J->raw_pct = raw_pct_events;
        J->caller_waits = true;
        commit_job(*J);
        return wait(seq, stream, true, launches);
    }

    uint64_t in_flight() const { return submitted.load(std::memory_order_relaxed) - retired; }

    ~UploadPool()
    {
        {
            std::lock_guard<std::mutex> lk(mu);
            quit = true;
        }
        cv_job.notify_all();
        for (auto &W : workers) if (W.th.joinable()) W.th.join();
        if (raw_th.joinable()) raw_th.join();
        cudaSetDevice(device);
        for (auto &W : workers) {
            if (W.stream) cudaStreamSynchronize(W.stream);
            for (int b = 0; b < 2; ++b) {
                if (W.pinned[b]) cudaFreeHost(W.pinned[b]);
                if (W.d_slot[b]) cudaFree(W.d_slot[b]);
                if (W.ev[b]) cudaEventDestroy(W.ev[b]);
            }
            for (int r = 0; r < RING; ++r) if (W.job_ev[r]) cudaEventDestroy(W.job_ev[r]);
            if (W.stream) cudaStreamDestroy(W.stream);
        }
        if (raw_stream) cudaStreamSynchronize(raw_stream);
        for (int s = 0; s < SLOTS; ++s) {
            if (d_slot[s]) cudaFree(d_slot[s]);
            if (slot_ev[s]) cudaEventDestroy(slot_ev[s]);
        }
        for (int r = 0; r < RING; ++r) if (raw_job_ev[r]) cudaEventDestroy(raw_job_ev[r]);
        if (raw_stream) cudaStreamDestroy(raw_stream);
        cudaGetLastError();
    }
};
