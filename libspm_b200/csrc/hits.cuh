// hits.cuh -- the compacted hit buffer kernels append to, and the device-side ordering of long hit lists.
#pragma once
#include <stdint.h>
#include "../../include/spmb.h"

struct HitSink {
    spmb_hit *hits;             // device buffer
    unsigned long long *count;  // total hits produced (may exceed cap: finish() then grows and re-runs)
    unsigned long long cap;
};

#ifdef __CUDACC__
#include <cooperative_groups.h>
// next free slot of a compacted buffer, warp-aggregated: the lanes that reach the append together (a divergent
// branch of a scan loop) share ONE atomicAdd -- on a hit-dense text all 32 lanes of a warp report in the same
// column, and 32 atomics on one address per column would serialise in L2
__device__ __forceinline__ unsigned long long spmb_append_slot(unsigned long long *count)
{
    const cooperative_groups::coalesced_group g = cooperative_groups::coalesced_threads();
    unsigned long long base = 0;
    if (g.thread_rank() == 0) base = atomicAdd(count, (unsigned long long)g.size());
    return g.shfl(base, 0) + g.thread_rank();
}
#endif

// Hit order from atomics is nondeterministic; short lists are sorted on the host, long ones here:
// every hit becomes one 64-bit key (its fields concatenated in comparison order), CUB radix-sorts
// (key, index) pairs over the bits in use, and the hits are gathered in that order.  For the filtered
// Myers path the last key field is the error count and runs that agree on everything else collapse
// to their first element (the smallest error count = the exact one, pigeon.cuh).
struct HitKeySpec {
    int field[4]; // 0 pattern_id, 1 seq_id, 2 end, 3 errors, 4 reserved; -1 = unused
    int bits[4];
};

__device__ __forceinline__ unsigned long long hit_field(const spmb_hit &h, int f)
{
    switch (f) {
    case 0: return h.pattern_id;
    case 1: return h.seq_id;
    case 2: return h.end;
    case 3: return h.errors;
    default: return h.reserved;
    }
}

__global__ void hit_keys_kernel(const spmb_hit *__restrict__ hits, uint32_t n, HitKeySpec spec,
                                unsigned long long *__restrict__ keys, uint32_t *__restrict__ idx)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const spmb_hit h = hits[i];
    unsigned long long k = 0;
#pragma unroll
    for (int f = 0; f < 4; ++f)
        if (spec.field[f] >= 0) k = (k << spec.bits[f]) | hit_field(h, spec.field[f]);
    keys[i] = k;
    idx[i] = i;
}

// flag[i] = 1 if sorted element i starts a new run of keys >> drop_bits
__global__ void hit_flags_kernel(const unsigned long long *__restrict__ keys, uint32_t n, int drop_bits,
                                 uint8_t *__restrict__ flag)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    flag[i] = (i == 0) || ((keys[i] >> drop_bits) != (keys[i - 1] >> drop_bits));
}

__global__ void hit_gather_kernel(const spmb_hit *__restrict__ hits, const uint32_t *__restrict__ idx,
                                  const uint32_t *__restrict__ n_ptr, uint32_t n_max, spmb_hit *__restrict__ out)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t n = n_ptr ? min(*n_ptr, n_max) : n_max;
    if (i >= n) return;
    out[i] = hits[idx[i]];
}

// ---- the same ordering enqueued WITH the search (spmb_search_launch on a hit list that saw a long list last
// time): the list length is only known on the device, so the kernels run over the whole buffer (cap slots) and
// read the count themselves; empty slots get a key above every real one (bit `pad_bit`).
__global__ void hit_keys_dev_kernel(const spmb_hit *__restrict__ hits, const unsigned long long *__restrict__ count,
                                    uint32_t cap, HitKeySpec spec, int pad_bit, unsigned long long *__restrict__ keys,
                                    uint32_t *__restrict__ idx)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= cap) return;
    const unsigned long long n = min(*count, (unsigned long long)cap);
    unsigned long long k = 1ull << pad_bit;
    if (i < n) {
        const spmb_hit h = hits[i];
        k = 0;
#pragma unroll
        for (int f = 0; f < 4; ++f)
            if (spec.field[f] >= 0) k = (k << spec.bits[f]) | hit_field(h, spec.field[f]);
    }
    keys[i] = k;
    idx[i] = i;
}

__global__ void hit_flags_dev_kernel(const unsigned long long *__restrict__ keys, const unsigned long long *__restrict__ count,
                                     uint32_t cap, int drop_bits, uint8_t *__restrict__ flag)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= cap) return;
    const unsigned long long n = min(*count, (unsigned long long)cap);
    flag[i] = i < n && ((i == 0) || ((keys[i] >> drop_bits) != (keys[i - 1] >> drop_bits)));
}

// out[i] = hits[idx[i]] for i < n, n = *n_sel (deduplicated lists: use_sel) or min(*count, cap).  *n_sel is only
// written by the deduplication (DeviceSelect) -- without it the host takes the length from *count
__global__ void hit_gather_dev_kernel(const spmb_hit *__restrict__ hits, const uint32_t *__restrict__ idx,
                                      const unsigned long long *__restrict__ count, uint32_t *__restrict__ n_sel,
                                      int use_sel, uint32_t cap, spmb_hit *__restrict__ out)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t n = use_sel ? min(*n_sel, cap) : (uint32_t)min(*count, (unsigned long long)cap);
    if (i < n) out[i] = hits[idx[i]];
}
