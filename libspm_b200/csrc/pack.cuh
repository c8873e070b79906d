// pack.cuh -- kernels that build / read the packed store.
//   pack_kernel<ByteSource>     1-byte dna4 ranks (a std::vector<spm::dna4> image)  -> strided bit planes
//   journal_derive_kernel       base store + journal events (a14)                   -> linear lane-chunk words of the derived sequence
//   pack_chunks_kernel<Source>  lane-chunk words (upload, journal, JST windows)     -> planes (+ lane chunks)
//   unpack_kernel               planes -> 1-byte ranks (round trips, tests)
// One CTA (32 warps) produces one 1024-column block (8 KB); warp w owns thread-row w of the
// block (32 columns).  Lane b reads the 32 symbols of its row for those columns, the warp
// transposes them with ballots, the block is staged in shared memory in the permuted order and
// written out as one coalesced 8 KB burst.
#pragma once
#include "layout.cuh"
#include "../../include/spmb.h"

// symbol at absolute position p of a packed plain store
__device__ __forceinline__ uint32_t spmb_symbol_at(const uint2 *__restrict__ planes, const SpmbLayout &lay,
                                                   uint64_t p)
{
    uint64_t R = spmb_div_rl(lay, p);
    uint32_t col = (uint32_t)(p - R * lay.row_len);
    uint64_t g = (R >> 5) * lay.row_len + col;
    uint2 w = __ldg(planes + spmb_perm(g));
    uint32_t bit = (uint32_t)R & 31u;
    return (((w.x >> bit) & 1u) << 1) | ((w.y >> bit) & 1u);
}

struct ByteSource {
    const uint8_t *__restrict__ in;
    uint64_t n;
    uint32_t shift; // which two bits of a symbol byte become {H,L}: 0 (dna4; low pair of a 4-bit alphabet) or 2 (its high pair)
    // 32 symbols of one row starting at absolute position p0 (p0 % 32 == 0), packed 4 per word
    __device__ __forceinline__ void load32(uint64_t p0, uint32_t (&w)[8]) const
    {
        if (p0 + 32 <= n && ((reinterpret_cast<uintptr_t>(in) & 15) == 0)) {
            const uint4 *q = reinterpret_cast<const uint4 *>(in + p0);
            uint4 a = __ldg(q), b = __ldg(q + 1);
            w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w;
            w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                uint32_t v = 0;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    uint64_t p = p0 + 4 * i + j;
                    uint32_t s = (p < n) ? (uint32_t)in[p] : 0u;
                    v |= (s & 15u) << (8 * j);
                }
                w[i] = v;
            }
        }
    }
};

// ---- lane-chunk helpers: one uint2 = 32 consecutive symbols of a row, 2 bits each, first symbol in the
// top bits of .x (the layout lanes_kernel derives from the planes; the Myers and filter kernels read it) ----
__device__ __forceinline__ uint2 spmb_chunk(const uint2 *__restrict__ lanes, uint32_t rl32, uint64_t R, uint32_t c)
{
    return __ldg(lanes + ((R >> 5) * rl32 + c) * 32 + ((uint32_t)R & 31u));
}

// 32 symbols starting o symbols into the 64-symbol window (w0, w1), 0 <= o < 32
__device__ __forceinline__ uint2 spmb_chunk_shift(const uint2 w0, const uint2 w1, uint32_t o)
{
    const uint32_t a = o < 16 ? w0.x : w0.y;
    const uint32_t b = o < 16 ? w0.y : w1.x;
    const uint32_t c = o < 16 ? w1.x : w1.y;
    const uint32_t sh = 2 * (o & 15u);
    return make_uint2(__funnelshift_l(b, a, sh), __funnelshift_l(c, b, sh));
}

// the 32 symbols of a PLAIN store that start at absolute position p (any alignment; runs on into the
// next row; positions past the end read the zero padding)
__device__ __forceinline__ uint2 spmb_fetch32(const uint2 *__restrict__ lanes, const SpmbLayout &lay, uint64_t p)
{
    const uint32_t rl32 = lay.row_len >> 5;
    const uint64_t R = spmb_div_rl(lay, p);
    const uint32_t col = (uint32_t)(p - R * lay.row_len);
    const uint32_t c = col >> 5, o = col & 31u;
    const uint2 w0 = spmb_chunk(lanes, rl32, R, c);
    if (o == 0) return w0;
    const uint2 w1 = (c + 1 < rl32) ? spmb_chunk(lanes, rl32, R, c + 1) : spmb_chunk(lanes, rl32, R + 1, 0);
    return spmb_chunk_shift(w0, w1, o);
}

// journaled sequence: derived position d -> base symbol or inserted symbol.
// ev_der[i] = derived position where event i's inserted segment begins (device prefix sum);
// blk_idx[b] = index of the last event with ev_der <= b*1024 (or -1).
//
// Materialisation is two kernels (round 2, second form): journal_derive_kernel writes the derived sequence as LINEAR
// lane-chunk words (out[d/32] = the 32 symbols at d, the intermediate the upload path also produces), and the plain
// pack_chunks_kernel<ChunkSource> turns those into planes + lane chunks.  History: round 1 fetched inside the pack
// kernel with one binary search per 32-symbol chunk (2.9 ms for 3.2 Gbp + 32 M events); the first round-2 form
// shared one event batch per warp but still sat in the pack kernel's 1024-thread CTA -- 60 registers = ONE CTA per
// SM, every warp a chain of 5-6 dependent memory round trips (block index -> batch -> one base fetch after the
// other -> pool bytes) and a block-wide barrier behind the slowest: 5.1 ms under ncu.  Here a warp owns one
// 1024-symbol block of the derived sequence on its own (no barrier, 256-thread CTAs, ~5 CTAs per SM):
//   1. blk_idx[b], blk_idx[b+1]                     -> the block's events are [first, first + nb)        (one trip)
//   2. the warp loads those records coalesced into shared memory                                         (one trip)
//   3. the warp loads the 48 base lane-chunk words the block can read into shared memory                   (one trip)
//   4. every lane walks the runs of its chunk in shared memory alone: a base run is one unaligned read of the block's
//      base words, short insertions (<= 16 symbols) sit in the event record; one loop trip per event in the chunk.
//      Trips 1-3 of the NEXT blocks run under step 4 of this one (software pipeline).  Chunks that do not fit this
//      shape (a run outside the window, a long insertion, more than 32 events in the block) are derived by the
//      general walk from global memory (journal_chunk_dense).
// Coordinates inside a block are 32-bit offsets from its first derived position D0 (rel), and an event is kept as
// what a run needs: where its inserted segment begins / ends (rel, ins_end) and the two constants that turn a derived
// position into a source position -- base position = d + shift behind the event, pool position = d + pshift inside
// its insertion.
struct JournalRec { // the general walk (journal_chunk_dense): built from the event table on the fly
    int32_t rel, ins_end;
    long long shift, pshift;
};
// what the fast walk needs of an event, block-independent (journal_rec_kernel writes one per event) ...
struct JournalEvRec {
    uint64_t der;       // derived position where the inserted segment begins
    long long shift;    // base position = derived position + shift behind the event
    uint32_t ins_len;
    uint32_t sym;       // the first 16 inserted symbols, first one in the top two bits
};
// ... and in a block's coordinates (shared memory)
struct JournalFast {
    int32_t rel, ins_end;
    long long shift;
    uint32_t sym, pad;
};

struct JournalDeriveArgs {
    const uint2 *__restrict__ base_lanes;
    SpmbLayout base;
    const spmb_journal_event *__restrict__ ev;
    const uint64_t *__restrict__ ev_der;
    const int64_t *__restrict__ blk_idx;
    const uint2 *__restrict__ pool_words; // the insertion pool, 32 symbols per word pair, one spare pair behind the end
    const JournalEvRec *__restrict__ ev_rec; // per event: journal_rec_kernel
    uint64_t n_ev;
    uint64_t n;                           // derived length
    uint2 *__restrict__ out;              // [ceil(n / 32)]
};

enum { JOURNAL_BATCH = 32, JOURNAL_NBASE = 48 };

__device__ __forceinline__ int32_t journal_clamp(long long x)
{
    return (int32_t)max(-(1ll << 30), min(1ll << 30, x));
}

__device__ __forceinline__ JournalRec journal_rec(const spmb_journal_event &x, uint64_t der, uint64_t D0)
{
    JournalRec r;
    r.rel = journal_clamp((long long)der - (long long)D0);
    r.ins_end = journal_clamp((long long)der + (long long)x.ins_len - (long long)D0);
    r.shift = (long long)(x.ref_pos + x.del_len) - (long long)der - (long long)x.ins_len;
    r.pshift = (long long)x.ins_off - (long long)der;
    return r;
}

// source words of a run: the two lane-chunk words that hold the 32 symbols from `src` on, and where they begin
__device__ __forceinline__ void journal_issue(const JournalDeriveArgs &a, bool pool, uint64_t src, uint2 &w0, uint2 &w1, uint32_t &o)
{
    if (pool) {
        w0 = __ldg(a.pool_words + (src >> 5));
        w1 = __ldg(a.pool_words + (src >> 5) + 1);
        o = (uint32_t)src & 31u;
    } else {
        const uint32_t rl32 = a.base.row_len >> 5;
        const uint64_t R = spmb_div_rl(a.base, src);
        const uint32_t col = (uint32_t)(src - R * a.base.row_len);
        const uint32_t c = col >> 5;
        w0 = spmb_chunk(a.base_lanes, rl32, R, c);
        w1 = (c + 1 < rl32) ? spmb_chunk(a.base_lanes, rl32, R, c + 1) : spmb_chunk(a.base_lanes, rl32, R + 1, 0);
        o = col & 31u;
    }
}

__device__ __forceinline__ unsigned long long journal_place(const uint2 w0, const uint2 w1, uint32_t o, uint32_t t, uint32_t len)
{
    const uint2 w = spmb_chunk_shift(w0, w1, o);
    unsigned long long W = ((unsigned long long)w.x << 32) | w.y;
    if (len < 32) W &= ~(~0ull >> (2 * len)); // keep the first `len` symbols
    return W >> (2 * t);
}

// One chunk from shared memory alone: sm = events [0, nb) of the block (rel ascending), e = index of the last one
// with rel <= lane * 32 (or -1), bw = the JOURNAL_NBASE base lane-chunk words from linear chunk qb on.  A loop trip is
// one base run + the event behind it + that event's inserted symbols (from the record); a lane with no event in its
// chunk makes one trip.  false = something does not fit (a run outside the base window, an insertion longer than the
// 16 embedded symbols): the caller derives that chunk from global memory instead.
__device__ __forceinline__ bool journal_chunk(const JournalFast *sm, int nb, int e, const uint2 *bw, long long qb, uint64_t D0, int32_t n_rel,
                                              uint32_t lane, uint2 &res)
{
    const int32_t c0 = (int32_t)lane * 32;
    int32_t cur = c0;
    const int32_t end = min(c0 + 32, n_rel);
    long long shift = 0;
    int32_t ins_end = INT32_MIN, irel = 0, next_rel = INT32_MAX;
    uint32_t sym = 0;
    if (e >= 0) { const JournalFast r = sm[e]; shift = r.shift; ins_end = r.ins_end; irel = r.rel; sym = r.sym; }
    if (e + 1 < nb) next_rel = sm[e + 1].rel;
    unsigned long long out = 0;
    bool ok = true;
    for (;;) {
        if (cur < ins_end) { // inside the inserted segment of the event in effect
            const int32_t stop = min(end, ins_end);
            if (stop - irel > 16 || cur < irel) ok = false;
            else {
                unsigned long long W = (unsigned long long)(sym << (2 * (cur - irel))) << 32;
                W &= ~(~0ull >> (2 * (stop - cur)));
                out |= W >> (2 * (cur - c0));
            }
            cur = stop;
        }
        const int32_t stop = min(end, next_rel);
        if (stop > cur) { // a run copied from the base
            const long long pb = (long long)D0 + cur + shift;
            const long long qi = (pb >> 5) - qb;
            if (qi < 0 || qi >= JOURNAL_NBASE - 1) ok = false;
            else {
                const uint2 w = spmb_chunk_shift(bw[qi], bw[qi + 1], (uint32_t)pb & 31u);
                unsigned long long W = ((unsigned long long)w.x << 32) | w.y;
                const int32_t len = stop - cur;
                if (len < 32) W &= ~(~0ull >> (2 * len));
                out |= W >> (2 * (cur - c0));
            }
            cur = stop;
        }
        if (cur >= end) break;
        ++e; // the event that begins here
        const JournalFast r = sm[e];
        shift = r.shift; ins_end = r.ins_end; irel = r.rel; sym = r.sym;
        next_rel = (e + 1 < nb) ? sm[e + 1].rel : INT32_MAX;
    }
    res = make_uint2((uint32_t)(out >> 32), (uint32_t)out);
    return ok;
}

// The same without a batch (a block with more events than JOURNAL_BATCH: every record comes from global memory).
__device__ __noinline__ uint2 journal_chunk_dense(const JournalDeriveArgs &a, int64_t first, int64_t last, uint64_t D0, int32_t n_rel,
                                                  uint32_t lane)
{
    const uint64_t d0 = D0 + lane * 32;
    // last event with der <= d0 among [first, last] (der[first] <= D0 unless first == 0)
    int64_t lo = first - 1, hi = last;
    while (lo < hi) {
        const int64_t mid = (lo + hi + 1) >> 1;
        if (a.ev_der[mid] <= d0) lo = mid; else hi = mid - 1;
    }
    int64_t e = lo;
    const int32_t c0 = (int32_t)lane * 32;
    int32_t cur = c0;
    const int32_t end = min(c0 + 32, n_rel);
    long long shift = 0, pshift = 0;
    int32_t ins_end = INT32_MIN, next_rel = INT32_MAX;
    if (e >= 0) { const JournalRec r = journal_rec(a.ev[e], a.ev_der[e], D0); shift = r.shift; pshift = r.pshift; ins_end = r.ins_end; }
    if (e + 1 < (int64_t)a.n_ev) next_rel = journal_clamp((long long)a.ev_der[e + 1] - (long long)D0);
    unsigned long long out = 0;
    while (cur < end) {
        while (cur >= next_rel) {
            ++e;
            const JournalRec r = journal_rec(a.ev[e], a.ev_der[e], D0);
            shift = r.shift; pshift = r.pshift; ins_end = r.ins_end;
            next_rel = (e + 1 < (int64_t)a.n_ev) ? journal_clamp((long long)a.ev_der[e + 1] - (long long)D0) : INT32_MAX;
        }
        const bool pool = cur < ins_end;
        const int32_t stop = pool ? min(end, ins_end) : min(end, next_rel);
        uint2 w0, w1;
        uint32_t o;
        journal_issue(a, pool, (uint64_t)((long long)D0 + cur + (pool ? pshift : shift)), w0, w1, o);
        out |= journal_place(w0, w1, o, (uint32_t)(cur - c0), (uint32_t)(stop - cur));
        cur = stop;
    }
    return make_uint2((uint32_t)(out >> 32), (uint32_t)out);
}

// per event, once: what the fast walk needs, so that staging a block's events is one 24-byte load per event and a
// subtraction (the per-block straight-line code -- index, records, base words, e0 -- was 440 of the 660 instructions
// a block cost, most of it 64-bit address and clamp arithmetic on three arrays)
__global__ void journal_rec_kernel(const spmb_journal_event *__restrict__ ev, const uint64_t *__restrict__ der, uint64_t n_ev,
                                   const uint2 *__restrict__ pool_words, JournalEvRec *__restrict__ out)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_ev) return;
    const spmb_journal_event x = ev[i];
    JournalEvRec r;
    r.der = der[i];
    r.shift = (long long)(x.ref_pos + x.del_len) - (long long)r.der - (long long)x.ins_len;
    r.ins_len = x.ins_len;
    r.sym = 0;
    if (x.ins_len) {
        const uint2 w = spmb_chunk_shift(__ldg(pool_words + (x.ins_off >> 5)), __ldg(pool_words + (x.ins_off >> 5) + 1), (uint32_t)x.ins_off & 31u);
        r.sym = w.x;
        if (x.ins_len < 16) r.sym &= ~(0xffffffffu >> (2 * x.ins_len));
    }
    out[i] = r;
}

// One warp = one 1024-symbol block of the derived sequence at a time, software-pipelined over the warp's blocks:
// while block j is derived from shared memory alone (its event records and the 48 base lane-chunk words it can read),
// the base words of block j+1, the event records of block j+2 and the block index of block j+3 are on their way into
// registers -- no memory round trip is left on a block's critical path.
__global__ void __launch_bounds__(256, 3) journal_derive_kernel(JournalDeriveArgs a)
{
    __shared__ JournalFast s_rec[8][2][JOURNAL_BATCH];
    __shared__ uint2 s_base[8][2][JOURNAL_NBASE];
    __shared__ int s_last[8][32];
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int *last_of = s_last[warp];
    const uint64_t n_blk = (a.n + 1023) >> 10;
    const uint64_t stride = (uint64_t)gridDim.x * 8;
    const uint64_t blk0 = (uint64_t)blockIdx.x * 8 + warp;
    if (blk0 >= n_blk) return; // no block-wide synchronisation anywhere
    const uint32_t rl32 = a.base.row_len >> 5;

    // events [first, first + want) matter to a block: the one in effect at its start, those inside, one behind
    auto span = [&](int64_t i_lo, int64_t i_hi, int64_t &first, uint32_t &want) {
        first = i_lo < 0 ? 0 : i_lo;
        uint64_t w = (uint64_t)(i_hi + 2 - first);
        if (w > a.n_ev - (uint64_t)first) w = a.n_ev - (uint64_t)first;
        want = (uint32_t)min(w, (uint64_t)0x7fffffff);
    };
    int64_t f_tmp;
    struct Idx { int64_t lo, hi; };
    auto load_idx = [&](uint64_t b) -> Idx {
        Idx x{-1, -1};
        if (b < n_blk) { x.lo = a.blk_idx[b]; x.hi = a.blk_idx[b + 1]; }
        return x;
    };
    // one event record per lane in flight (blocks with more than JOURNAL_BATCH events take the general walk)
    JournalEvRec xr;
    xr.der = 0; xr.shift = 0; xr.ins_len = 0; xr.sym = 0;
    auto load_events = [&](int64_t first, uint32_t want) {
        if (lane < want && want <= JOURNAL_BATCH) xr = a.ev_rec[first + lane];
    };
    auto stage_events = [&](JournalFast *sm, uint32_t want, uint64_t D0) {
        if (lane < want && want <= JOURNAL_BATCH) {
            const long long d = (long long)xr.der - (long long)D0;
            JournalFast r;
            r.rel = journal_clamp(d);
            r.ins_end = journal_clamp(d + (long long)xr.ins_len);
            r.shift = xr.shift; r.sym = xr.sym; r.pad = 0;
            sm[lane] = r;
        }
    };
    // base words in flight: linear chunks qb + lane and qb + 32 + lane (lanes 0..15)
    uint2 xb0 = make_uint2(0, 0), xb1 = make_uint2(0, 0);
    auto load_base = [&](long long qb) {
        // row / chunk of qb once, then steps inside the row (rows shorter than the window: a division per chunk)
        const uint64_t Rb = spmb_div_rl(a.base, (uint64_t)qb << 5);
        const uint32_t cb = (uint32_t)((uint64_t)qb - Rb * rl32);
        if (rl32 >= 2 * JOURNAL_NBASE) {
            uint32_t c = cb + lane;
            uint64_t R = Rb;
            if (c >= rl32) { c -= rl32; ++R; }
            xb0 = spmb_chunk(a.base_lanes, rl32, R, c);
            if (lane < JOURNAL_NBASE - 32) {
                c = cb + 32 + lane; R = Rb;
                if (c >= rl32) { c -= rl32; ++R; }
                xb1 = spmb_chunk(a.base_lanes, rl32, R, c);
            }
        } else {
            uint64_t R = spmb_div_rl(a.base, (uint64_t)(qb + lane) << 5);
            xb0 = spmb_chunk(a.base_lanes, rl32, R, (uint32_t)((uint64_t)(qb + lane) - R * rl32));
            if (lane < JOURNAL_NBASE - 32) {
                R = spmb_div_rl(a.base, (uint64_t)(qb + 32 + lane) << 5);
                xb1 = spmb_chunk(a.base_lanes, rl32, R, (uint32_t)((uint64_t)(qb + 32 + lane) - R * rl32));
            }
        }
    };
    // first base chunk a block may need: four chunks of slack before the position its first symbol comes from
    auto base_start = [&](const JournalFast *sm, uint32_t want, uint64_t D0) -> long long {
        long long s0 = 0;
        if (want && want <= JOURNAL_BATCH) { const JournalFast r = sm[0]; if (r.rel <= 0) s0 = r.shift; }
        long long q = (((long long)D0 + s0) >> 5) - 4;
        return q < 0 ? 0 : q;
    };

    // ---- prologue: records of block 0 staged, base words of block 0 and records of block 1 in flight ----
    uint64_t blk = blk0;
    Idx i0 = load_idx(blk), i1 = load_idx(blk + stride), i2 = load_idx(blk + 2 * stride);
    uint32_t want, want1 = 0;
    span(i0.lo, i0.hi, f_tmp, want);
    load_events(f_tmp, want);
    stage_events(s_rec[warp][0], want, blk << 10);
    __syncwarp();
    long long qb = base_start(s_rec[warp][0], want, blk << 10), qb1 = 0;
    load_base(qb);
    if (blk + stride < n_blk) { span(i1.lo, i1.hi, f_tmp, want1); load_events(f_tmp, want1); }
    int par = 0;
    for (;;) {
        const uint64_t D0 = blk << 10;
        const bool has_next = blk + stride < n_blk;
        JournalFast *sm = s_rec[warp][par];
        uint2 *bw = s_base[warp][par];
        // ---- stage what arrived: base words of this block, records of the next ----
        bw[lane] = xb0;
        if (lane < JOURNAL_NBASE - 32) bw[32 + lane] = xb1;
        if (has_next) stage_events(s_rec[warp][par ^ 1], want1, (blk + stride) << 10);
        last_of[lane] = -1;
        __syncwarp();
        // ---- issue the next loads: base of block j+1, records of block j+2, index of block j+3 ----
        uint32_t want2 = 0;
        if (has_next) {
            qb1 = base_start(s_rec[warp][par ^ 1], want1, (blk + stride) << 10);
            load_base(qb1);
            if (blk + 2 * stride < n_blk) { span(i2.lo, i2.hi, f_tmp, want2); load_events(f_tmp, want2); }
            i2 = load_idx(blk + 3 * stride);
        }
        // ---- this block: every lane's e0 = last event with rel <= lane * 32 ----
        const bool fast = want <= JOURNAL_BATCH;
        if (fast && lane < want) {
            const int32_t r = sm[lane].rel;
            const int32_t c = r <= 0 ? 0 : (r + 31) >> 5; // first chunk whose start lies at or behind the event
            int32_t cn = 64;
            if (lane + 1 < want) { const int32_t rn = sm[lane + 1].rel; cn = rn <= 0 ? 0 : (rn + 31) >> 5; }
            if (c < 32 && cn != c) last_of[c] = (int)lane;
        }
        __syncwarp();
        int e0 = last_of[lane];
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int up = __shfl_up_sync(0xffffffffu, e0, d);
            if ((int)lane >= d) e0 = max(e0, up);
        }
        const int32_t n_rel = (int32_t)min((uint64_t)1024, a.n - D0);
        const bool active = (int32_t)lane * 32 < n_rel;
        uint2 o = make_uint2(0, 0);
        bool ok = fast;
        if (active && ok) ok = journal_chunk(sm, (int)want, e0, bw, qb, D0, n_rel, lane, o);
        if (__any_sync(0xffffffffu, active && !ok)) { // dense blocks, long insertions, runs far from the window
            if (active && !ok) { // the block's event range again, from the index (this path is rare)
                const Idx ix = load_idx(blk);
                int64_t first;
                uint32_t w2;
                span(ix.lo, ix.hi, first, w2);
                o = journal_chunk_dense(a, first, first + (int64_t)w2 - 1, D0, n_rel, lane);
            }
        }
        if (active) a.out[(D0 >> 5) + lane] = o;
        if (!has_next) break;
        __syncwarp();
        blk += stride;
        want = want1; qb = qb1;
        want1 = want2;
        par ^= 1;
    }
}

// journal preparation on the device: validation + (ins - del) deltas, then an exclusive scan
// (CUB) gives every event's shift; ev_der = ref_pos + shift.
__global__ void journal_delta_kernel(const spmb_journal_event *__restrict__ ev, uint64_t n_ev, uint64_t n_ref,
                                     uint64_t pool_len, long long *__restrict__ delta, int *__restrict__ bad)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_ev) return;
    const spmb_journal_event x = ev[i];
    delta[i] = (long long)x.ins_len - (long long)x.del_len;
    bool ok = (x.ref_pos + x.del_len <= n_ref) && (x.ins_off + x.ins_len <= pool_len);
    if (i > 0) {
        const spmb_journal_event y = ev[i - 1];
        ok = ok && (x.ref_pos >= y.ref_pos + y.del_len);
    }
    if (!ok) atomicExch(bad, 1);
}

// events given in the coordinates of a longer sequence of which the base store holds [bias, ...): make them local
__global__ void journal_bias_kernel(spmb_journal_event *__restrict__ ev, uint64_t n_ev, uint64_t bias, int *__restrict__ bad)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_ev) return;
    const uint64_t p = ev[i].ref_pos;
    if (p < bias) atomicExch(bad, 1);
    ev[i].ref_pos = p - bias;
}

// shift[i] (exclusive prefix of the deltas) -> derived position of event i, in place
__global__ void journal_der_kernel(const spmb_journal_event *__restrict__ ev, uint64_t n_ev, long long *shift_der,
                                   const long long *__restrict__ delta, long long *__restrict__ total)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_ev) return;
    const long long sh = shift_der[i];
    if (i == n_ev - 1) *total = sh + delta[i];
    shift_der[i] = (long long)ev[i].ref_pos + sh;
}

__global__ void journal_blkidx_kernel(const uint64_t *__restrict__ der, uint64_t n_ev, uint64_t n_blk,
                                      int64_t *__restrict__ blk_idx)
{
    uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n_blk) return;
    const uint64_t d = b << 10;
    int64_t lo = -1, hi = (int64_t)n_ev - 1;
    while (lo < hi) {
        int64_t mid = (lo + hi + 1) >> 1;
        if (der[mid] <= d) lo = mid; else hi = mid - 1;
    }
    blk_idx[b] = lo;
}

// ---- 32x32 bit-matrix transpose across a warp: lane r holds row r (bit c = element [r][c]); afterwards
// lane c holds column c (bit r = element [r][c]).  Five butterfly steps, one shuffle each: step j swaps the
// off-diagonal j x j blocks between lanes r and r^j.  Replaces 32 ballots + 32 predicated moves per word.
__device__ __forceinline__ uint32_t spmb_transpose32(uint32_t x, uint32_t lane)
{
#pragma unroll
    for (int j = 16; j >= 1; j >>= 1) {
        // m_lo: bits c with (c & j) == 0
        const uint32_t m_lo = j == 16 ? 0x0000ffffu : j == 8 ? 0x00ff00ffu : j == 4 ? 0x0f0f0f0fu : j == 2 ? 0x33333333u : 0x55555555u;
        const uint32_t y = __shfl_xor_sync(0xffffffffu, x, j);
        const bool up = (lane & j) != 0;
        const uint32_t sh = up ? (y >> j) : (y << j);
        const uint32_t keep = up ? ~m_lo : m_lo;
        x = (x & keep) | (sh & ~keep);
    }
    return x;
}

template <class Source>
__global__ void __launch_bounds__(1024) pack_kernel(Source src, SpmbLayout lay, uint2 *__restrict__ planes)
{
    __shared__ uint2 tile[1024 + 16]; // one pad entry per 64: see pack_chunks_kernel
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint64_t g0 = (uint64_t)blockIdx.x * 1024 + warp * 32; // first column of this thread-row
    uint32_t w[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) w[i] = 0;
    if (g0 < lay.n_cols) {
        uint64_t st = spmb_div_rl(lay, g0);
        uint32_t col0 = (uint32_t)(g0 - st * lay.row_len);
        uint64_t p0 = (st * 32 + lane) * (uint64_t)lay.row_len + col0;
        if (p0 < lay.n) src.load32(p0, w);
    }
    uint32_t myH = 0, myL = 0;
#pragma unroll
    for (int c = 0; c < 32; ++c) {
        uint32_t s = w[c >> 2] >> (8 * (c & 3) + src.shift);
        uint32_t h = __ballot_sync(0xffffffffu, (s >> 1) & 1u);
        uint32_t l = __ballot_sync(0xffffffffu, s & 1u);
        if (lane == c) { myH = h; myL = l; }
    }
    // lane c holds column g0 + c  ->  permuted slot inside the block
    const uint32_t ti = ((lane >> 1) << 6) | (warp << 1) | (lane & 1);
    tile[ti + (ti >> 6)] = make_uint2(myH, myL);
    __syncthreads();
    planes[(uint64_t)blockIdx.x * 1024 + threadIdx.x] = tile[threadIdx.x + (threadIdx.x >> 6)];
}

// same block shape as pack_kernel for sources that deliver a lane chunk (32 symbols of a row, packed).
// Fetch phase: warp r = row r of the super-tile, lane c = the c-th 32-column group of the block, so a warp
// walks 1024 CONSECUTIVE symbols of one row: a linear source is read in 256-byte runs.  The chunks change hands
// through shared memory; then warp c
// holds group c with lane r = row r, the 2 x 32 x 32 bit transposition gives the {H,L} column words, and the
// chunk itself -- the store's lane-chunk layout -- is written as a by-product when `lanes_out` is given.
template <class Source>
__global__ void __launch_bounds__(1024) pack_chunks_kernel(Source src, SpmbLayout lay, uint2 *__restrict__ planes,
                                                           uint2 *__restrict__ lanes_out)
{
    // the tile is the block in the store's permuted column order, so the final copy is one coalesced 8 KB burst; one
    // pad entry per 64 keeps the scattered writes of a warp (stride 64 entries) off a single bank pair
    __shared__ uint2 tile[1024 + 16];
    __shared__ uint2 xchg[32][33];
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint64_t n_blocks = lay.n_cols_alloc / 1024;
    // a CTA walks its blocks with the NEXT block's chunk already on its way: the kernel has two barriers per block and
    // two CTAs per SM, so a load issued and awaited inside one block left DRAM idle most of the time (ncu: 33 % DRAM,
    // 56 % issue-active for one block per CTA)
    auto fetch = [&](uint64_t blk) -> uint2 {
        const uint64_t gf = blk * 1024 + lane * 32; // first column of group `lane`
        uint2 f = make_uint2(0, 0);
        if (gf < lay.n_cols) {
            const uint64_t st = spmb_div_rl(lay, gf);
            const uint32_t col0 = (uint32_t)(gf - st * lay.row_len);
            const uint64_t p0 = (st * 32 + warp) * (uint64_t)lay.row_len + col0;
            if (p0 < lay.n) f = src.load_chunk(p0);
        }
        return f;
    };
    uint2 f_next = fetch(blockIdx.x);
    for (uint64_t blk = blockIdx.x; blk < n_blocks; blk += gridDim.x) {
        xchg[warp][lane] = f_next;
        if (blk + gridDim.x < n_blocks) f_next = fetch(blk + gridDim.x);
        __syncthreads();
        const uint2 w = xchg[lane][warp];                 // row `lane`, group `warp`
        const uint64_t g0 = blk * 1024 + warp * 32;       // first column of this thread-row
        if (lanes_out) lanes_out[(g0 >> 5) * 32 + lane] = w; // group G = g0/32 (every group of the allocation is written)
        // bit j of a chunk word is the high (j odd) or low (j even) bit of symbol (31 - j) / 2, so the 32 x 32 bit
        // transposes of the two words ARE the plane words: lane j ends up with one plane of column s = (31 - j) / 2
        // (from .x) and of column 16 + s (from .y), bit r = row r -- no separate de-interleaving of the 2-bit codes
        const uint32_t tx = spmb_transpose32(w.x, lane), ty = spmb_transpose32(w.y, lane);
        const uint32_t s = (31u - lane) >> 1, plane = (lane & 1u) ^ 1u; // .x of a column entry = H, .y = L
        uint32_t *tile32 = reinterpret_cast<uint32_t *>(tile);
        const uint32_t i0 = ((s >> 1) << 6) | (warp << 1) | (s & 1u), i1 = (((16u + s) >> 1) << 6) | (warp << 1) | (s & 1u);
        tile32[2 * (i0 + (i0 >> 6)) + plane] = tx;
        tile32[2 * (i1 + (i1 >> 6)) + plane] = ty;
        __syncthreads();
        planes[blk * 1024 + threadIdx.x] = tile[threadIdx.x + (threadIdx.x >> 6)];
    }
}

__global__ void unpack_kernel(const uint2 *__restrict__ planes, SpmbLayout lay, uint64_t begin, uint64_t n,
                              uint8_t *__restrict__ out)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = (uint8_t)spmb_symbol_at(planes, lay, begin + i);
}

// planes -> "lane chunks", the layout the Myers kernels read: for the 32-column
// group G (global columns 32G .. 32G+31) lane r of a super-tile owns ONE uint2 = the 32 symbols of
// its row in those columns, 2-bit codes, first column in the top bits (.x = columns 0..15,
// .y = columns 16..31).  A Myers lane then pops one symbol per column with a shift by 30 and an
// IMAD by 4, and a warp's loads are one contiguous 256 B.  One warp transposes one group: lane j
// loads column j's {H,L}, 64 ballots hand lane r the words of row r.
__device__ __forceinline__ uint32_t spmb_spread16(uint32_t x)
{
    x &= 0xffffu;
    x = (x | (x << 8)) & 0x00ff00ffu;
    x = (x | (x << 4)) & 0x0f0f0f0fu;
    x = (x | (x << 2)) & 0x33333333u;
    x = (x | (x << 1)) & 0x55555555u;
    return x;
}

__global__ void __launch_bounds__(256) lanes_kernel(const uint2 *__restrict__ planes, uint64_t n_groups,
                                                    uint2 *__restrict__ lanes)
{
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t G = (uint64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (G >= n_groups) return; // warp-uniform
    const uint2 c = __ldg(planes + spmb_perm(G * 32 + lane));
    const uint32_t myH = spmb_transpose32(c.x, lane), myL = spmb_transpose32(c.y, lane); // row `lane`: bit j = column j
    const uint32_t bh = __brev(myH), bl = __brev(myL); // column j at bit 31-j
    uint2 o;
    o.x = (spmb_spread16(bh >> 16) << 1) | spmb_spread16(bl >> 16);
    o.y = (spmb_spread16(bh) << 1) | spmb_spread16(bl);
    lanes[G * 32 + lane] = o;
}

__global__ void zero_kernel(uint4 *p, uint64_t n16)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n16) p[i] = make_uint4(0, 0, 0, 0);
}
