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

#include <atomic>
#include <condition_variable>
#include <cstdlib>
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
    enum { SLOTS = 3 };
    int device = 0;
    uint64_t chunk_sym = 4ull << 20;
    uint64_t min_sym = 32ull << 20;
    int n_workers = 0;

    struct Worker {
        std::thread th;
        cudaStream_t stream = nullptr;
        uint64_t *pinned[2] = {nullptr, nullptr};
        cudaEvent_t ev[2] = {nullptr, nullptr};
        bool used[2] = {false, false};
    };
    std::vector<Worker> workers;
    uint8_t *d_slot[SLOTS] = {nullptr, nullptr, nullptr};
    cudaEvent_t slot_ev[SLOTS] = {nullptr, nullptr, nullptr};
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
    std::atomic<uint64_t> ends{0};  // head << 32 | tail: chunks [head, tail) are unclaimed
    std::atomic<int> cuda_err{0};
    std::atomic<uint64_t> packed_chunks{0};
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

    cudaError_t init(int dev)
    {
        device = dev;
        chunk_sym = env_u64("SPMB_UPLOAD_CHUNK", chunk_sym);
        chunk_sym = (chunk_sym + 1023) / 1024 * 1024;
        if (chunk_sym < 1024) chunk_sym = 1024;
        min_sym = env_u64("SPMB_UPLOAD_MIN", min_sym);
        int dflt = spmb_host_cpus() - 1;
        if (dflt > 31) dflt = 31;
        if (dflt < 0) dflt = 0;
        n_workers = (int)env_u64("SPMB_HOST_THREADS", (uint64_t)dflt);
        if (n_workers > 256) n_workers = 256;
        cudaError_t e = cudaSuccess;
        for (int s = 0; s < SLOTS && e == cudaSuccess; ++s) {
            e = cudaMalloc(&d_slot[s], chunk_sym);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&slot_ev[s], cudaEventDisableTiming | cudaEventBlockingSync);
        }
        workers.resize(n_workers);
        for (int w = 0; w < n_workers && e == cudaSuccess; ++w) {
            Worker &W = workers[w];
            e = cudaStreamCreateWithFlags(&W.stream, cudaStreamNonBlocking);
            for (int b = 0; b < 2 && e == cudaSuccess; ++b) {
                e = cudaHostAlloc((void **)&W.pinned[b], chunk_sym / 4, cudaHostAllocDefault);
                if (e == cudaSuccess) e = cudaEventCreateWithFlags(&W.ev[b], cudaEventDisableTiming | cudaEventBlockingSync);
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
            while (e == cudaSuccess && cuda_err.load(std::memory_order_relaxed) == 0 && claim_tail(idx)) {
                if (W.used[b]) e = cudaEventSynchronize(W.ev[b]);
                const uint64_t p0 = idx * chunk_sym;
                const uint64_t len = (n - p0 < chunk_sym) ? n - p0 : chunk_sym;
                spmb_host_pack_chunks(src + p0, len, W.pinned[b]);
                if (e == cudaSuccess)
                    e = cudaMemcpyAsync(d_chunks + p0 / 32, W.pinned[b], (len + 31) / 32 * 8, cudaMemcpyHostToDevice, W.stream);
                if (e == cudaSuccess) e = cudaEventRecord(W.ev[b], W.stream);
                W.used[b] = true;
                b ^= 1;
                packed_chunks.fetch_add(1, std::memory_order_relaxed);
            }
            if (e == cudaSuccess) e = cudaStreamSynchronize(W.stream);
            W.used[0] = W.used[1] = false;
            if (e != cudaSuccess) cuda_err.store((int)e);
            {
                std::lock_guard<std::mutex> lk(mu);
                if (--running == 0) cv_done.notify_all();
            }
        }
    }

    // fills dst[0, ceil(n/32)) (device, visible to `stream` and already allocated in its order) from
    // host_ranks[0, n).  Returns when every byte of host_ranks has been consumed and all work of the
    // packed path is complete; the raw path's last kernels are ordered on `stream`.
    cudaError_t run(const uint8_t *host_ranks, uint64_t n_sym, uint2 *dst, cudaStream_t stream, uint32_t *launches)
    {
        src = host_ranks;
        n = n_sym;
        d_chunks = dst;
        n_chunks = (n_sym + chunk_sym - 1) / chunk_sym;
        cuda_err.store(0);
        packed_chunks.store(0);
        ends.store(n_chunks & 0xffffffffull); // head 0, tail n_chunks
        // dst comes from the stream-ordered allocator: the workers' streams may touch it only after
        // the allocation point of `stream`
        cudaError_t e = cudaStreamSynchronize(stream);
        if (e != cudaSuccess) return e;
        if (n_workers) {
            std::lock_guard<std::mutex> lk(mu);
            running = n_workers;
            ++gen;
            cv_job.notify_all();
        }
        uint64_t idx, raw = 0;
        int s = 0;
        bool used[SLOTS] = {false, false, false};
        while (e == cudaSuccess && cuda_err.load(std::memory_order_relaxed) == 0) {
            if (used[s]) e = cudaEventSynchronize(slot_ev[s]); // at most SLOTS-1 raw chunks ahead of the link
            if (e != cudaSuccess || !claim_head(idx)) break;
            const uint64_t p0 = idx * chunk_sym;
            const uint64_t len = (n - p0 < chunk_sym) ? n - p0 : chunk_sym;
            e = cudaMemcpyAsync(d_slot[s], src + p0, len, cudaMemcpyHostToDevice, stream);
            if (e == cudaSuccess) {
                const uint64_t groups = (len + 31) / 32;
                bytes_to_chunks_kernel<<<(unsigned)((groups + 255) / 256), 256, 0, stream>>>(d_slot[s], len, d_chunks + p0 / 32);
                if (launches) ++*launches;
                e = cudaGetLastError();
            }
            if (e == cudaSuccess) e = cudaEventRecord(slot_ev[s], stream);
            used[s] = true;
            s = (s + 1) % SLOTS;
            ++raw;
        }
        if (e != cudaSuccess) { // stop the workers early: nothing left to claim
            cuda_err.store((int)e);
            ends.store(0);
        }
        if (n_workers) {
            std::unique_lock<std::mutex> lk(mu);
            cv_done.wait(lk, [&] { return running == 0; });
        }
        // the caller's buffer may be reused as soon as we return: the raw copies must have left it
        cudaError_t e2 = cudaStreamSynchronize(stream);
        if (e == cudaSuccess) e = e2;
        if (e == cudaSuccess && cuda_err.load() != 0) e = (cudaError_t)cuda_err.load();
        last_raw = raw;
        last_packed = packed_chunks.load();
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
