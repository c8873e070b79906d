// pack.cuh -- kernels that build / read the packed store.
//   pack_kernel<ByteSource>     1-byte dna4 ranks (a std::vector<spm::dna4> image)  -> strided bit planes
//   pack_chunks_kernel<JournalSource>  base store + journal events (a14)            -> planes (+ lane chunks) of the derived sequence
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
// A lane produces the 32 symbols of its row chunk as ONE packed word pair: every maximal run copied
// from the base is one unaligned 32-symbol fetch from the base's lane chunks (two 8-byte loads), shifted
// into place -- the base is read at ~2x its packed size instead of one 8-byte load per SYMBOL.
//
// The event lookup is WARP-COOPERATIVE (round 2): the 32 lanes of a fetch warp cover up to 1024 consecutive
// derived symbols of one row, i.e. ~10 events at a 1 % journal.  Round 1 let every lane binary-search the event
// table on its own: ~10 dependent L2 round trips per 32 symbols (blk_idx x2, 4-5 search steps, the event, the
// next event's position, the base words), 2.9 ms for 3.2 Gbp where the bytes moved take 0.4 ms.  Now the warp
// reads blk_idx once, loads ONE batch of 32 consecutive event records with a coalesced access into shared
// memory, and every lane finds and walks its events there (5 LDS steps); the dependent chain is
// blk_idx -> batch -> base words.  Lanes whose events run past the batch fall back to global reads.
struct JournalSource {
    const uint2 *__restrict__ base_lanes;
    SpmbLayout base;
    const spmb_journal_event *__restrict__ ev;
    const uint64_t *__restrict__ ev_der;
    const int64_t *__restrict__ blk_idx;
    const uint8_t *__restrict__ pool;
    uint64_t n_ev;
    uint64_t n; // derived length

    enum { BATCH = 32 };
    struct Rec { uint64_t der, ref_end, ins_off; uint32_t ins_len, pad; };
    enum { WARP_SMEM = BATCH * 32 };

    __device__ __forceinline__ Rec fetch(int64_t i, const Rec *sm, int64_t b, uint32_t nb) const
    {
        const int64_t k = i - b;
        if (k >= 0 && k < (int64_t)nb) return sm[k];
        const spmb_journal_event x = ev[i];
        Rec r;
        r.der = ev_der[i]; r.ref_end = x.ref_pos + x.del_len; r.ins_off = x.ins_off; r.ins_len = x.ins_len; r.pad = 0;
        return r;
    }

    // the 32 symbols at derived position d0; batch = events [b, b + nb) in sm (der[b] <= d0 or b == 0)
    __device__ __forceinline__ uint2 walk(uint64_t d0, const Rec *sm, int64_t b, uint32_t nb) const
    {
        // number of batch events with der <= d0 (sorted): branch-free binary search over <= 32 entries
        uint32_t cnt = 0;
#pragma unroll
        for (uint32_t step = 16; step >= 1; step >>= 1)
            if (cnt + step <= nb && sm[cnt + step - 1].der <= d0) cnt += step;
        if (cnt < nb && sm[cnt].der <= d0) ++cnt; // nb == 32: the last entry
        int64_t e = b + (int64_t)cnt - 1;
        if (cnt == nb) // the lane's event may lie past the batch (dense journal): finish the search in global memory
            while (e + 1 < (int64_t)n_ev && ev_der[e + 1] <= d0) ++e;
        uint64_t seg_der = 0, seg_ref = 0, seg_off = 0;
        uint32_t seg_ins = 0;
        if (e >= 0) {
            const Rec r = fetch(e, sm, b, nb);
            seg_der = r.der; seg_ins = r.ins_len; seg_ref = r.ref_end; seg_off = r.ins_off;
        }
        Rec nx;
        nx.der = ~0ull;
        if (e + 1 < (int64_t)n_ev) nx = fetch(e + 1, sm, b, nb);
        unsigned long long out = 0;
        uint32_t t = 0;
        const uint32_t t_end = (uint32_t)min((uint64_t)32, n - d0);
        while (t < t_end) {
            const uint64_t d = d0 + t;
            while (d >= nx.der) { // enter the next event (several may start at the same d)
                ++e;
                seg_der = nx.der; seg_ins = nx.ins_len; seg_ref = nx.ref_end; seg_off = nx.ins_off;
                nx.der = ~0ull;
                if (e + 1 < (int64_t)n_ev) nx = fetch(e + 1, sm, b, nb);
            }
            const uint64_t o = d - seg_der;
            if (e >= 0 && o < seg_ins) { // inside the inserted segment: one pool symbol
                out |= (unsigned long long)(pool[seg_off + o] & 3u) << (62 - 2 * t);
                ++t;
                continue;
            }
            // a run copied from the base: up to the next event / the end of this chunk
            const uint64_t bp = (e >= 0) ? seg_ref + (o - seg_ins) : d;
            uint32_t run = t_end - t;
            if (nx.der - d < run) run = (uint32_t)(nx.der - d);
            const uint2 w = spmb_fetch32(base_lanes, base, bp);
            unsigned long long W = ((unsigned long long)w.x << 32) | w.y;
            if (run < 32) W &= ~(~0ull >> (2 * run)); // keep the first `run` symbols
            out |= W >> (2 * t);
            t += run;
        }
        return make_uint2((uint32_t)(out >> 32), (uint32_t)out);
    }

    // all 32 lanes of a warp call this together; lane l wants the chunk at derived position d0 (active) of the
    // row segment seg_key (lanes of one segment hold ascending, consecutive chunks).  sm = this warp's batch area.
    __device__ __forceinline__ uint2 load_chunk_warp(bool active, uint64_t d0, uint64_t seg_key, void *smem, uint32_t lane) const
    {
        Rec *sm = reinterpret_cast<Rec *>(smem);
        uint2 out = make_uint2(0, 0);
        uint32_t todo = __ballot_sync(0xffffffffu, active && d0 < n);
        while (todo) {
            const int leader = __ffs(todo) - 1;
            const uint64_t key = __shfl_sync(0xffffffffu, seg_key, leader);
            const bool mine = ((todo >> lane) & 1u) && seg_key == key;
            const uint32_t grp = __ballot_sync(0xffffffffu, mine);
            const uint64_t Dlo = __shfl_sync(0xffffffffu, d0, leader);
            int64_t b = blk_idx[Dlo >> 10];
            if (b < 0) b = 0;
            uint32_t nb;
            for (;;) {
                nb = (uint32_t)min((uint64_t)BATCH, n_ev - (uint64_t)b);
                if (lane < nb) {
                    const spmb_journal_event x = ev[b + lane];
                    Rec r;
                    r.der = ev_der[b + lane]; r.ref_end = x.ref_pos + x.del_len; r.ins_off = x.ins_off; r.ins_len = x.ins_len; r.pad = 0;
                    sm[lane] = r;
                }
                __syncwarp();
                const uint32_t le = __ballot_sync(0xffffffffu, lane < nb && sm[lane].der <= Dlo);
                if (le == 0xffffffffu && (uint64_t)b + BATCH < n_ev) { // every event of the batch lies before the segment: slide
                    b += BATCH - 1;
                    __syncwarp();
                    continue;
                }
                break;
            }
            if (mine) out = walk(d0, sm, b, nb);
            __syncwarp();
            todo &= ~grp;
        }
        return out;
    }
};

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

// 16 two-bit codes (symbol j at bits 31-2j, 30-2j) -> the 16 high bits and the 16 low bits, symbol j at bit 15-j
__device__ __forceinline__ uint32_t spmb_compress_even(uint32_t t) // even bit positions -> low 16 bits
{
    t = (t | (t >> 1)) & 0x33333333u;
    t = (t | (t >> 2)) & 0x0f0f0f0fu;
    t = (t | (t >> 4)) & 0x00ff00ffu;
    t = (t | (t >> 8)) & 0x0000ffffu;
    return t;
}

// lane-chunk word (32 symbols of one row) -> H / L words with symbol c at bit c
__device__ __forceinline__ void spmb_chunk_to_bits(const uint2 w, uint32_t &H, uint32_t &L)
{
    const uint32_t hx = spmb_compress_even((w.x >> 1) & 0x55555555u), lx = spmb_compress_even(w.x & 0x55555555u);
    const uint32_t hy = spmb_compress_even((w.y >> 1) & 0x55555555u), ly = spmb_compress_even(w.y & 0x55555555u);
    H = __brev((hx << 16) | hy);
    L = __brev((lx << 16) | ly);
}

template <class Source>
__global__ void __launch_bounds__(1024) pack_kernel(Source src, SpmbLayout lay, uint2 *__restrict__ planes)
{
    __shared__ uint2 tile[1024];
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
    tile[((lane >> 1) << 6) | (warp << 1) | (lane & 1)] = make_uint2(myH, myL);
    __syncthreads();
    planes[(uint64_t)blockIdx.x * 1024 + threadIdx.x] = tile[threadIdx.x];
}

// same block shape as pack_kernel for sources that deliver a lane chunk (32 symbols of a row, packed).
// Fetch phase: warp r = row r of the super-tile, lane c = the c-th 32-column group of the block, so a warp
// walks 1024 CONSECUTIVE symbols of one row: a linear source is read in 256-byte runs, and the lanes of a
// journal lookup share their block index, their binary-search path and their event records (L1 hits instead
// of a chain of DRAM round trips per lane).  The chunks change hands through shared memory; then warp c
// holds group c with lane r = row r, the 2 x 32 x 32 bit transposition gives the {H,L} column words, and the
// chunk itself -- the store's lane-chunk layout -- is written as a by-product when `lanes_out` is given.
template <class Source, class = void>
struct SpmbWarpSmem { enum { value = 0 }; };
template <class Source>
struct SpmbWarpSmem<Source, decltype((void)Source::WARP_SMEM)> { enum { value = Source::WARP_SMEM }; };

template <class Source>
__global__ void __launch_bounds__(1024) pack_chunks_kernel(Source src, SpmbLayout lay, uint2 *__restrict__ planes,
                                                           uint2 *__restrict__ lanes_out)
{
    // the fetch phase's per-warp event batches and the store phase's tile share one block of shared memory
    constexpr uint32_t BATCH_BYTES = 32u * SpmbWarpSmem<Source>::value;
    constexpr uint32_t UNION_BYTES = BATCH_BYTES > 8192u ? BATCH_BYTES : 8192u;
    __shared__ __align__(16) unsigned char s_union[UNION_BYTES];
    __shared__ uint2 xchg[32][33];
    uint2 *tile = reinterpret_cast<uint2 *>(s_union);
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    {
        const uint64_t gf = (uint64_t)blockIdx.x * 1024 + lane * 32; // first column of group `lane`
        uint2 f = make_uint2(0, 0);
        bool active = false;
        uint64_t p0 = 0, st = 0;
        if (gf < lay.n_cols) {
            st = spmb_div_rl(lay, gf);
            const uint32_t col0 = (uint32_t)(gf - st * lay.row_len);
            p0 = (st * 32 + warp) * (uint64_t)lay.row_len + col0;
            active = p0 < lay.n;
        }
        if constexpr (SpmbWarpSmem<Source>::value != 0) {
            f = src.load_chunk_warp(active, p0, st, s_union + warp * SpmbWarpSmem<Source>::value, lane);
        } else {
            if (active) f = src.load_chunk(p0);
        }
        xchg[warp][lane] = f;
    }
    __syncthreads();
    const uint2 w = xchg[lane][warp];                                // row `lane`, group `warp`
    const uint64_t g0 = (uint64_t)blockIdx.x * 1024 + warp * 32;     // first column of this thread-row
    if (lanes_out) lanes_out[(g0 >> 5) * 32 + lane] = w; // group G = g0/32 (every group of the allocation is written)
    uint32_t myH, myL;
    spmb_chunk_to_bits(w, myH, myL);          // row `lane`: bit c = column g0 + c
    myH = spmb_transpose32(myH, lane);        // column g0 + lane: bit r = row r
    myL = spmb_transpose32(myL, lane);
    tile[((lane >> 1) << 6) | (warp << 1) | (lane & 1)] = make_uint2(myH, myL);
    __syncthreads();
    planes[(uint64_t)blockIdx.x * 1024 + threadIdx.x] = tile[threadIdx.x];
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
