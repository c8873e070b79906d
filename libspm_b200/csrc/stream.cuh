// stream.cuh -- device side of the resumable searchers (spmb_stream_*).
//
// spm::restorable_myers_matcher / restorable_shiftor_matcher carry {VP, VN, errors} / prefSufMatch from one call to
// the next (libspm/libspm/matcher/myers_matcher_restorable.hpp:57-63,72-82, shiftor_matcher_restorable.hpp:44-63), so
// a caller can stream a haystack of any size.  Here the carried state is the last `keep` symbols of the stream,
// held ON THE DEVICE as lane-chunk words (32 symbols per 8-byte word, first symbol on top -- pack.cuh): a feed
// uploads only its new chunk behind them, and the kernels below do the splicing.
#pragma once
#include "layout.cuh"
#include "pack.cuh"
#include "pigeon.cuh"
#include "myers_chain.cuh"

// up to 31 symbols passed by value: they complete the partial word at the end of the carried tail
struct StreamHead {
    uint8_t sym[32];
};

// words[t / 32] holds the symbols [32*(t/32), t) of the sequence in its top bits; append `a` symbols (a <= 32 - t % 32)
__global__ void stream_head_kernel(uint2 *__restrict__ words, uint64_t t, StreamHead head, uint32_t a)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const uint64_t w = t >> 5;
    const uint32_t o = (uint32_t)(t & 31u);
    unsigned long long v = o ? (((unsigned long long)words[w].x << 32) | words[w].y) : 0ull;
    if (o) v &= ~(~0ull >> (2 * o)); // keep the first o symbols
    for (uint32_t j = 0; j < a; ++j) v |= (unsigned long long)(head.sym[j] & 3u) << (62 - 2 * (o + j));
    words[w] = make_uint2((uint32_t)(v >> 32), (uint32_t)v);
}

// the 32 symbols at position p of a packed linear sequence (words beyond `n_words` read as zero)
__device__ __forceinline__ unsigned long long stream_fetch32(const uint2 *__restrict__ words, uint64_t n_words, uint64_t p)
{
    const uint64_t w = p >> 5;
    const uint32_t o = (uint32_t)(p & 31u);
    const uint2 a = w < n_words ? words[w] : make_uint2(0, 0);
    const unsigned long long A = ((unsigned long long)a.x << 32) | a.y;
    if (o == 0) return A;
    const uint2 b = w + 1 < n_words ? words[w + 1] : make_uint2(0, 0);
    const unsigned long long B = ((unsigned long long)b.x << 32) | b.y;
    return (A << (2 * o)) | (B >> (64 - 2 * o));
}

// tail[0, keep) = seq[total - keep, total): the state carried to the next feed (symbols past `keep` are zero)
__global__ void stream_carry_kernel(const uint2 *__restrict__ words, uint64_t total, uint32_t keep, uint2 *__restrict__ tail)
{
    const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w * 32 >= keep) return;
    const uint64_t n_words = (total + 31) >> 5;
    unsigned long long v = stream_fetch32(words, n_words, total - keep + 32ull * w);
    const uint32_t left = keep - w * 32;
    if (left < 32) v &= ~(~0ull >> (2 * left));
    tail[w] = make_uint2((uint32_t)(v >> 32), (uint32_t)v);
}

// raw bytes (1 per symbol, device) appended at symbol position t of a packed sequence whose word t/32 is already
// complete up to t (t % 32 == 0, or stream_head_kernel ran): thread g writes word t/32 + g from bytes[32g ..)
// -- bytes_to_chunks_kernel (upload.cuh) with a destination offset; kept separate so that the aligned bulk of a
// feed can go through either mover of the upload queue.

// ---- capture().errors: the Myers score at the current column (myers_matcher_restorable.hpp:57-59 exposes the
// whole pattern state, of which `errors` is the one value a caller can interpret).  D[m][j] <= m and the best
// alignment ending at j spans at most 2m symbols, so a scan of the last min(tail, 2m) symbols from the initial
// state gives it exactly.  One thread per needle; runs only when a state is captured.
struct StreamScoreArgs {
    const uint2 *tail;
    uint32_t tail_len;
    const uint32_t *eq;
    const MyersPatMeta *meta;
    uint32_t n_pat;
    uint32_t *out;
};

template <int NW>
__global__ void stream_score_kernel(StreamScoreArgs a)
{
    constexpr int NWP = (NW <= 4) ? 4 : 8;
    const uint32_t pid = blockIdx.x * blockDim.x + threadIdx.x;
    if (pid >= a.n_pat) return;
    const MyersPatMeta pm = a.meta[pid];
    if (pm.m == 0) { a.out[pm.pattern_id] = 0; return; }
    const uint32_t *eqp = a.eq + (size_t)pid * 4 * NWP;
    uint32_t vp[NW], vn[NW];
#pragma unroll
    for (int w = 0; w < NW; ++w) { vp[w] = 0xffffffffu; vn[w] = 0; }
    int score = (int)pm.m;
    const uint32_t span = min(a.tail_len, 2u * pm.m);
    for (uint32_t t = a.tail_len - span; t < a.tail_len; ++t) {
        const uint2 wv = a.tail[t >> 5];
        const uint32_t j = t & 31u;
        const uint32_t word = j < 16 ? wv.x : wv.y;
        const uint32_t c = (word >> (30 - 2 * (j & 15u))) & 3u;
        uint32_t eq[NW], hp[NW], hn[NW];
        pg_load_eq<NW>(eqp, c, eq);
        pg_myers_col<NW, false>(vp, vn, eq, hp, hn);
        score += pg_score_delta<NW>(hp, hn, pm.m);
    }
    a.out[pm.pattern_id] = (uint32_t)score; // the caller's pattern id (children of a bucketed set: spmb_capi.cu)
}

// the same for needles longer than 256 symbols: one warp per needle, words chained across the lanes (myers_chain.cuh)
template <int WPL>
__global__ void __launch_bounds__(128) stream_score_chain_kernel(StreamScoreArgs a)
{
    constexpr uint32_t NWT = 32 * WPL;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t ci = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (ci >= a.n_pat) return; // warp-uniform
    const MyersPatMeta pm = a.meta[ci];
    const uint32_t *eqp = a.eq + (size_t)ci * 4 * NWT + lane * WPL;
    const uint32_t top = (pm.m - 1) >> 5;
    const bool owner = (top / WPL) == lane;
    const uint32_t ow = top % WPL, ob = (pm.m - 1) & 31u;
    uint32_t vp[WPL], vn[WPL];
#pragma unroll
    for (int w = 0; w < WPL; ++w) { vp[w] = 0xffffffffu; vn[w] = 0; }
    int score = (int)pm.m;
    const uint32_t span = min(a.tail_len, 2u * pm.m);
    for (uint32_t t = a.tail_len - span; t < a.tail_len; ++t) {
        const uint2 wv = a.tail[t >> 5];
        const uint32_t j = t & 31u;
        const uint32_t word = j < 16 ? wv.x : wv.y;
        const uint32_t c = (word >> (30 - 2 * (j & 15u))) & 3u;
        uint32_t eq[WPL], hp[WPL], hn[WPL];
#pragma unroll
        for (int w = 0; w < WPL; ++w) eq[w] = __ldg(eqp + c * NWT + w);
        myers_chain_step<WPL>(vp, vn, eq, lane, hp, hn);
        uint32_t hpw = hp[0], hnw = hn[0];
#pragma unroll
        for (int w = 1; w < WPL; ++w)
            if ((uint32_t)w == ow) { hpw = hp[w]; hnw = hn[w]; }
        score += (int)((hpw >> ob) & 1u) - (int)((hnw >> ob) & 1u);
    }
    if (owner) a.out[pm.pattern_id] = (uint32_t)score;
}
