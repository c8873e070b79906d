// pigeon.cuh -- pigeonhole q-gram seed filter + anchored verification.
//
// Reference anchors:
//   spm::pigeonhole_matcher{needle(s), error_rate}   libspm/libspm/matcher/pigeonhole_matcher.hpp:129-243
//     (SeqAn2 Pattern<Index<StringSet, IndexQGram<SimpleShape, OpenAddressing>>, Pigeonhole<SeedOnly>>:
//      every needle is cut into q-grams at offsets 0, q, 2q, ..., the haystack is scanned once and every
//      position whose q-gram equals one of them is reported as a seed hit {needle index, needle offset, q},
//      pigeonhole_matcher.hpp:41-44,118-124; golden vectors pigeonhole_matcher_test.cpp:30-33)
//   spm::restorable_myers_prefix_matcher{needle, k}  libspm/libspm/matcher/myers_prefix_matcher_restorable.hpp:47-61,117-160
//     (SeqAn2 MyersUkkonenGlobal: the needle is aligned to PREFIXES of the haystack -- D[0][j] = j --
//      and the scan is clamped to m + k + 1 symbols, :47-50)
//
// What is built on top (no reference counterpart; the reference leaves verification to its callers):
//   SPMB_ALG_MYERS_FILTERED = the hit list of SPMB_ALG_MYERS, bit-exact, computed as
//     filter  : every k-error occurrence of needle i contains one of its k+1 disjoint pieces of length
//               q = floor(m_i / (k+1)) unchanged (pigeonhole), so only text positions whose q-gram is
//               one of the (k+1)*n_pat seeds can belong to a hit;
//     verify  : for a seed (needle i, offset o) found at text position p every occurrence through it
//               lies inside [p-o-k, p-o+m_i+k): one Myers scan of that window from the initial state.
//               A window score is an upper bound of the true score, and the window of the seed inside
//               the optimal alignment reproduces it exactly, so the minimum over windows is exact
//               (DESIGN.md 4b); duplicates across seeds are removed after the sort.
//
// Device side: ONE pass over the lane-chunk layout (pack.cuh lanes_kernel: one uint2 = 32 consecutive
// symbols of a lane's row).  Per text position: the q-gram (<= 32 symbols = 64 bits, top-aligned) falls out
// of two funnel shifts, a multiplicative hash (IMADs -- FMA pipe) indexes a bitmap held in shared
// memory (<= 128 KB per CTA), the pass bits of 32 positions are collected in one register and only
// survivors (a few %) probe the exact open-addressing table in L2.
#pragma once
#include <algorithm>
#include <vector>

#include "layout.cuh"
#include "hits.cuh"
#include "myers.cuh"
#include "pack.cuh"

// ------------------------------------------------------------------------------------------
// shared host/device definitions
// ------------------------------------------------------------------------------------------
// q-gram key: symbol j of the q-gram at bits [63-2j, 62-2j] of (hi:lo), zero below (kq <= 32 symbols)
// First-level bitmap (shared memory): hashed on the key's first min(kq, 16) symbols only -- `hi`, already in a
// register, one IMAD; a q-gram longer than 16 symbols differs from a seed behind them once in 4^16 positions, and
// the exact lookup behind the bitmap compares the whole key anyway.  The bitmap holds 2^B bits (B <= 20: 128 KB):
// x = h >> (32 - B) picks the word (x >> 5) and one bit (x & 31), h & 31 a second one for wide keys.  A bitmap of
// any size through a multiply-high of the hash (192 KB would halve the false positives) was measured and dropped:
// IMAD.HI made the scan 36 % slower, whatever the size.  So was a second table of 64 KB behind the first (one more
// bit per key from its own hash: 1.9 % false positives instead of 7.3 %): the five ALU instructions per position it
// costs outweigh the survivors it removes (C5 1.183 -> 1.224 ms) -- the scan is ALU-pipe-bound, the survivors' exact
// look-ups are latency that other warps cover.
SPMB_HD uint32_t pg_h1(uint32_t hi) { return hi * 0x9E3779B1u; }
SPMB_HD uint32_t pg_h3(uint32_t hi, uint32_t lo) // second-level bitmap
{
    uint32_t h = (hi ^ 0x7F4A7C15u) * 0x85EBCA6Bu + (lo ^ 0x165667B1u) * 0xC2B2AE35u;
    h ^= h >> 13;
    return h * 0x27D4EB2Fu;
}
SPMB_HD uint32_t pg_h2(uint32_t hi, uint32_t lo)
{
    uint32_t h = (hi * 0xC2B2AE3Du) ^ (lo * 0x27D4EB2Fu);
    h ^= h >> 15;
    return h * 0x2C1B3C6Du;
}

static const uint32_t PG_EMPTY = 0xffffffffu;

struct PigeonDev {
    const uint32_t *bitmap;            // 2^B bits; bit for index x is bit 31-(x&31) of word x>>5
    const uint32_t *bitmap2;           // second level, 2^B2 bits in L2: one bit per q-gram, asked by bitmap survivors only
    const ulonglong2 *tab;             // 2^T slots {key, value}, linear probing; value = index of the unique
                                       // q-gram, PG_EMPTY = free slot
    const uint32_t *ufirst;            // [n_unique+1] -> range of seeds carrying that q-gram
    const uint32_t *seed_pat;          // needle index of each seed
    const uint32_t *seed_off;          // offset of the seed inside its needle
    uint32_t q, kq, B, T, B2;
    uint32_t himask, lomask;           // the top 2*kq bits of (hi:lo)
};

struct PigeonCand {
    uint64_t pos; // plain: text position of the seed's first symbol; windows: row*row_len + column
    uint32_t u;   // unique q-gram index
    uint32_t pad;
};

// ------------------------------------------------------------------------------------------
// host: seed table
// ------------------------------------------------------------------------------------------
struct PigeonHost {
    uint32_t q = 0, kq = 0, B = 0, T = 0, B2 = 0;
    std::vector<uint32_t> bitmap, bitmap2;
    std::vector<unsigned long long> tab; // {key, value} pairs
    std::vector<uint32_t> ufirst, seed_pat, seed_off;
    uint64_t n_unique = 0;

    static unsigned long long key_of(const uint8_t *s, uint32_t kq)
    {
        unsigned long long k = 0;
        for (uint32_t j = 0; j < kq; ++j) k |= (unsigned long long)(s[j] & 3u) << (62 - 2 * j);
        return k;
    }

    // seeds = (needle, offset) pairs; ranks/off/len describe the concatenated needles
    void build(const std::vector<uint8_t> &ranks, const std::vector<uint32_t> &off,
               const std::vector<std::pair<uint32_t, uint32_t>> &seeds, uint32_t q_)
    {
        q = q_;
        kq = q < 32 ? q : 32;
        struct S { unsigned long long key; uint32_t pat, o; };
        std::vector<S> v;
        v.reserve(seeds.size());
        for (auto &s : seeds) v.push_back({key_of(ranks.data() + off[s.first] + s.second, kq), s.first, s.second});
        std::sort(v.begin(), v.end(), [](const S &a, const S &b) {
            if (a.key != b.key) return a.key < b.key;
            if (a.pat != b.pat) return a.pat < b.pat;
            return a.o < b.o;
        });
        seed_pat.clear(); seed_off.clear(); ufirst.clear();
        std::vector<unsigned long long> ukeys;
        for (size_t i = 0; i < v.size(); ++i) {
            if (i == 0 || v[i].key != v[i - 1].key) { ufirst.push_back((uint32_t)i); ukeys.push_back(v[i].key); }
            seed_pat.push_back(v[i].pat);
            seed_off.push_back(v[i].o);
        }
        ufirst.push_back((uint32_t)v.size());
        n_unique = ukeys.size();
        // bitmap: ~32 bits per unique q-gram, 2^15 .. 2^20 bits (4 KB .. 128 KB of shared memory)
        B = 15;
        while (B < 20 && (1ull << B) < n_unique * 32) ++B;
        bitmap.assign((size_t)1 << (B - 5), 0);
        T = 4;
        while ((1ull << T) < n_unique * 2 + 1) ++T;
        tab.assign((size_t)2 << T, 0);
        for (size_t i = 0; i < ((size_t)1 << T); ++i) tab[2 * i + 1] = PG_EMPTY;
        // second-level bitmap: <= 1/512 full, so that almost no first-level false positive reaches the table
        B2 = 16;
        while (B2 < 28 && (1ull << B2) < n_unique * 512) ++B2;
        bitmap2.assign((size_t)1 << (B2 - 5), 0);
        const uint32_t tmask = (1u << T) - 1;
        for (size_t u = 0; u < ukeys.size(); ++u) {
            const uint32_t hi = (uint32_t)(ukeys[u] >> 32), lo = (uint32_t)ukeys[u];
            const uint32_t h = pg_h1(hi);
            const uint32_t x = h >> (32 - B);
            bitmap[x >> 5] |= 0x80000000u >> (x & 31);
            if (kq > 16) bitmap[x >> 5] |= 0x80000000u >> (h & 31); // second bit, same word (wide keys)
            const uint32_t x2 = pg_h3(hi, lo) >> (32 - B2);
            bitmap2[x2 >> 5] |= 1u << (x2 & 31);
            uint32_t slot = pg_h2(hi, lo) >> (32 - T);
            while (tab[2 * slot + 1] != PG_EMPTY) slot = (slot + 1) & tmask;
            tab[2 * slot] = ukeys[u];
            tab[2 * slot + 1] = (uint32_t)u;
        }
    }

    uint32_t himask() const { return kq >= 16 ? 0xffffffffu : (kq == 0 ? 0u : ~0u << (32 - 2 * kq)); }
    uint32_t lomask() const { return kq <= 16 ? 0u : (kq >= 32 ? 0xffffffffu : ~0u << (64 - 2 * kq)); }
};

// ------------------------------------------------------------------------------------------
// device: text access through the lane-chunk layout
// ------------------------------------------------------------------------------------------
// symbol at (row R, column col) -- a sequential reader keeps the chunk of its current 32 columns
struct PgReader {
    const uint2 *__restrict__ lanes;
    uint32_t rl32;
    uint64_t R = ~0ull;
    uint32_t c = ~0u;
    uint2 w = make_uint2(0, 0);
    __device__ __forceinline__ uint32_t at(uint64_t row, uint32_t col)
    {
        const uint32_t cc = col >> 5;
        if (row != R || cc != c) { R = row; c = cc; w = spmb_chunk(lanes, rl32, row, cc); }
        const uint32_t j = col & 31u;
        const uint32_t word = j < 16 ? w.x : w.y;
        return (word >> (30 - 2 * (j & 15u))) & 3u;
    }
};

// ------------------------------------------------------------------------------------------
// filter kernel
// ------------------------------------------------------------------------------------------
struct PigeonFilterArgs {
    const uint2 *lanes;
    SpmbLayout lay;
    PigeonDev pg;
    const uint32_t *row_len; // windows mode
    uint64_t st0, st1;       // super-tiles scanned by this launch
    PigeonCand *cand;
    unsigned long long *cand_count;
    unsigned long long cand_cap;
};

// (hi:lo) = the 32 symbols starting at symbol j of the 64-symbol window (cur, nxt)
__device__ __forceinline__ void pg_window(const uint2 cur, const uint2 nxt, uint32_t j, uint32_t &hi, uint32_t &lo)
{
    const uint32_t a = j < 16 ? cur.x : cur.y;
    const uint32_t b = j < 16 ? cur.y : nxt.x;
    const uint32_t c = j < 16 ? nxt.x : nxt.y;
    const uint32_t s = 2 * (j & 15u);
    hi = __funnelshift_l(b, a, s);
    lo = __funnelshift_l(c, b, s);
}

// exact lookup of one bitmap survivor: (row R, chunk c, symbol j) -> candidate if its q-gram is a seed
template <int MODE>
__device__ __forceinline__ void pg_survivor(const PigeonFilterArgs &a, uint64_t st, uint32_t ln, uint32_t c, uint32_t j)
{
    const uint32_t RL = a.lay.row_len, rl32 = RL >> 5;
    const uint64_t R = st * 32 + ln;
    const uint32_t col = c * 32 + j;
    bool ok;
    if (MODE == SPMB_MODE_PLAIN) ok = (R * RL + col + a.pg.q <= a.lay.n);
    else ok = (R * RL < a.lay.n) && (col + a.pg.q <= a.row_len[R]);
    if (!ok) return;
    const uint2 *__restrict__ lp = a.lanes;
    const uint2 cur = __ldg(lp + (st * rl32 + c) * 32 + ln);
    uint2 nxt;
    if (c + 1 < rl32) nxt = __ldg(lp + (st * rl32 + c + 1) * 32 + ln);
    else if (MODE == SPMB_MODE_PLAIN)
        nxt = (ln < 31) ? __ldg(lp + (st * rl32) * 32 + ln + 1) : __ldg(lp + ((st + 1) * rl32) * 32);
    else nxt = make_uint2(0, 0);
    uint32_t hi, lo;
    pg_window(cur, nxt, j, hi, lo);
    hi &= a.pg.himask; lo &= a.pg.lomask;
    const uint32_t x2 = pg_h3(hi, lo) >> (32 - a.pg.B2);
    if (!((__ldg(a.pg.bitmap2 + (x2 >> 5)) >> (x2 & 31u)) & 1u)) return;
    const unsigned long long key = ((unsigned long long)hi << 32) | lo;
    const uint32_t tmask = (1u << a.pg.T) - 1;
    uint32_t slot = pg_h2(hi, lo) >> (32 - a.pg.T);
    for (;;) {
        const ulonglong2 ent = __ldg(a.pg.tab + slot);
        const uint32_t v = (uint32_t)ent.y;
        if (v == PG_EMPTY) return;
        if (ent.x == key) {
            const unsigned long long idx = spmb_append_slot(a.cand_count);
            if (idx < a.cand_cap) {
                PigeonCand cd;
                cd.pos = R * RL + col; cd.u = v; cd.pad = 0;
                a.cand[idx] = cd;
            }
            return;
        }
        slot = (slot + 1) & tmask;
    }
}

// WIDE: kq > 16 (the key has a low word; the bitmap then holds two bits per q-gram, both in one word).
// Work item = one chunk-column: 32 rows x 32 symbols of one super-tile.  Every warp of the (persistent, one CTA
// per SM) grid takes an EQUAL CONTIGUOUS share of the launch's chunk-columns -- round 1 handed out units of 32
// chunk-columns round-robin, which at 1/8 of the C5 text left 2.4 units per warp: three rounds where 2.4 would do
// (+25 % kernel time, the bulk of the poor strong scaling).  Bitmap survivors (a few per 1024 positions) are not
// looked up by the lane that found them -- that loop would run as long as the unluckiest lane of the warp -- but
// parked in a per-warp queue in shared memory and drained 32 at a time, one per lane.
static const uint32_t PG_QCAP = 256; // queue entries per warp (uint32: chunk-column in the warp's share | lane | symbol)

// STEP: only positions = 0 (mod STEP) of a row are looked at.  The filtered Myers path then indexes, for
// every pigeonhole piece of length L, the STEP sub-seeds of length L-STEP+1 at offsets 0..STEP-1 inside it:
// wherever the piece occurs, exactly one of them starts on a sampled position -- 1/STEP of the scan work
// for STEP times the seeds.
template <int MODE, bool WIDE, int STEP>
__global__ void __launch_bounds__(1024, 1) pigeon_filter_kernel(PigeonFilterArgs a)
{
    extern __shared__ uint32_t s_bits[];
    const uint32_t n_words = 1u << (a.pg.B - 5);
    for (uint32_t i = threadIdx.x; i < n_words; i += blockDim.x) s_bits[i] = a.pg.bitmap[i];
    __syncthreads();

    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t *s_q = s_bits + n_words + warp * PG_QCAP;
    const uint32_t RL = a.lay.row_len, rl32 = RL >> 5;
    const uint32_t shift = 32 - a.pg.B;
    const uint32_t himask = a.pg.himask;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const uint2 *__restrict__ lp = a.lanes;

    // this warp's share [t0, t1) of the chunk-columns (st - st0) * rl32 + c
    const uint64_t T = (a.st1 - a.st0) * rl32;
    const uint64_t n_warps = (uint64_t)gridDim.x * 32, gw = (uint64_t)blockIdx.x * 32 + warp;
    const uint64_t t0 = T / n_warps * gw + min(gw, T % n_warps);
    const uint64_t t1 = t0 + T / n_warps + (gw < T % n_warps ? 1 : 0);
    if (t0 >= t1) return; // no block-wide synchronisation below

    uint32_t qn = 0; // queue fill, warp-uniform
    const uint64_t st_first = a.st0 + t0 / rl32;
    const uint32_t c_first = (uint32_t)(t0 % rl32);
    auto drain = [&]() {
        __syncwarp();
        for (uint32_t b = 0; b < qn; b += 32) {
            if (b + lane < qn) {
                const uint32_t e = s_q[b + lane];
                // chunk-column e >> 10 of the warp's share -> (super-tile, chunk): 32-bit arithmetic (a share is far
                // below 2^22 chunk-columns; the 64-bit division here showed in the profile at 7 % survivors)
                const uint32_t cc = c_first + (e >> 10);
                const uint32_t ds = cc / rl32;
                pg_survivor<MODE>(a, st_first + ds, (e >> 5) & 31u, cc - ds * rl32, e & 31u);
            }
        }
        __syncwarp();
        qn = 0;
    };
    uint64_t st = a.st0 + t0 / rl32;
    uint32_t c = (uint32_t)(t0 % rl32);
    const uint2 *rowp = lp + (st * rl32) * 32 + lane;
    uint2 cur = __ldg(rowp + (uint64_t)c * 32);
    for (uint64_t t = t0; t < t1; ++t) {
        uint2 nxt;
        if (c + 1 < rl32) nxt = __ldg(rowp + (uint64_t)(c + 1) * 32);
        else if (MODE == SPMB_MODE_PLAIN)
            nxt = (lane < 31) ? __ldg(lp + (st * rl32) * 32 + lane + 1) : __ldg(lp + ((st + 1) * rl32) * 32);
        else nxt = make_uint2(0, 0);
        uint32_t acc = 0;
#pragma unroll
        for (int j = 0; j < 32; j += STEP) {
            // the 16 symbols from symbol j on (compile-time j: one funnel shift)
            const uint32_t wa = j < 16 ? cur.x : cur.y, wb = j < 16 ? cur.y : nxt.x;
            const uint32_t hi = __funnelshift_l(wb, wa, 2 * (j & 15));
            const uint32_t h = pg_h1(WIDE ? hi : (hi & himask));
            const uint32_t x = h >> shift;
            const uint32_t word = s_bits[x >> 5];
            // bit b of the word is its bit 31-b: shifted to the top (a funnel shift takes its amount mod 32)
            uint32_t tt = __funnelshift_l(0u, word, x);
            if (WIDE) tt &= __funnelshift_l(0u, word, h);
            acc = __funnelshift_l(tt, acc, 1); // sampled position k = j/STEP ends at bit 32/STEP-1-k
        }
        // ---- park the survivors of this chunk: one warp prefix sum hands every lane its queue slots ----
        const uint32_t cnt = __popc(acc);
        uint32_t inc = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t up = __shfl_up_sync(0xffffffffu, inc, d);
            if ((int)lane >= d) inc += up;
        }
        const uint32_t total = __shfl_sync(0xffffffffu, inc, 31);
        if (total) {
            const uint32_t tag = ((uint32_t)(t - t0) << 10) | (lane << 5);
            if (qn + total > PG_QCAP) drain();
            if (total <= PG_QCAP) {
                uint32_t slot = qn + inc - cnt;
                while (acc) {
                    const uint32_t b = __clz(acc); // 32 - 32/STEP + k
                    acc &= ~(0x80000000u >> b);
                    s_q[slot++] = tag | ((b - (32 - 32 / STEP)) * STEP);
                }
                qn += total;
            } else {
                // a chunk denser than the queue (low-complexity text): rounds of one survivor per lane
                uint32_t m = __ballot_sync(0xffffffffu, acc != 0);
                while (m) {
                    if (qn + 32 > PG_QCAP) drain();
                    if (acc) {
                        const uint32_t b = __clz(acc);
                        acc &= ~(0x80000000u >> b);
                        s_q[qn + __popc(m & lt_mask)] = tag | ((b - (32 - 32 / STEP)) * STEP);
                    }
                    qn += __popc(m);
                    m = __ballot_sync(0xffffffffu, acc != 0);
                }
            }
        }
        if (++c == rl32) { // next super-tile: its rows are other sequences' rows, nothing carries over
            c = 0;
            ++st;
            rowp = lp + (st * rl32) * 32 + lane;
            if (t + 1 < t1) cur = __ldg(rowp);
        } else cur = nxt;
    }
    drain();
}

// ------------------------------------------------------------------------------------------
// candidates -> hits
// ------------------------------------------------------------------------------------------
struct PigeonVerifyArgs {
    const uint2 *lanes;
    SpmbLayout lay;
    PigeonDev pg;
    const PigeonCand *cand;
    unsigned long long *counters; // [0] hits, [1] candidates of this launch, [2] max candidates of any launch
    unsigned long long cand_cap;
    const uint8_t *pat_ranks;     // concatenated needles
    const uint32_t *pat_off;      // [n_pat]
    const uint32_t *eq;           // [n_pat][4][NWP]
    const MyersPatMeta *meta;     // [n_pat]
    const uint32_t *row_len, *row_seq, *row_own; // windows mode
    uint32_t seq_id;
    uint64_t own_lo, own_hi;      // plain: owned range of (last symbol index of a hit | seed begin)
    HitSink sink;
    // duplicate suppression of the filtered Myers path (pg_emit_min); dd_keys == null: off
    unsigned long long *dd_keys;  // [2^dd_bits] 0 = free, else key + 1 (cleared before every launch)
    uint32_t *dd_err;             // [2^dd_bits] 256 - smallest error count reported for the key (0 = none yet)
    uint32_t dd_bits, dd_end_bits;
    uint64_t dd_nseq;
};

__device__ __forceinline__ void pg_emit(const HitSink &sink, uint64_t end, uint32_t seq, uint32_t pid, uint32_t err, uint32_t aux)
{
    const unsigned long long idx = spmb_append_slot(sink.count);
    if (idx < sink.cap) {
        spmb_hit h;
        h.end = end; h.seq_id = seq; h.pattern_id = pid; h.errors = err; h.reserved = aux;
        sink.hits[idx] = h;
    }
}

// SPMB_ALG_MYERS_FILTERED: the k+1 pieces of a needle (and neighbouring diagonals) verify overlapping windows, so the
// same (needle, sequence, end) is found up to ~(k+1) times, each time with an upper bound of the true score whose
// minimum is exact.  Sorting the raw list to drop them cost more than the search at small shards (C5 at 1/8 of the
// text: 13 ordering kernels, 0.12 ms, behind 0.2 ms of search).  So every finder of a hit goes through a hash table:
// the first one claims a slot (CAS on the key) and appends the record, with the slot number in `reserved`; ALL of
// them fold their error count into the slot's minimum (one atomicMax, fire and forget); when the kernel is done,
// pigeon_fixup_kernel writes each record's minimum back and clears `reserved`.  Nobody waits for anybody.  A table
// miss (4 probes) or a key that does not fit falls back to a plain append: the ordering downstream still removes
// those few duplicates, this table only removes the volume.
//
// A window's hits (2(k - e) + 1 neighbouring ends, typically) are collected first and reported TOGETHER behind the
// scan (PgHitBatch): the claims of all of them are in flight at once and one atomicAdd reserves their records --
// two memory round trips per candidate.  Reported one by one from inside the scan, every hit cost its warp a chain
// of three to four dependent atomics (profile of the first form: 60 % of the kernel's stall samples).
struct PgHitBatch {
    enum { CAP = 8 };
    unsigned long long ent[CAP]; // (end << 8) | errors
    int n = 0;
};

__device__ __forceinline__ void pg_batch_flush(const PigeonVerifyArgs &a, PgHitBatch &b, uint32_t seq, uint32_t pid)
{
    if (b.n == 0) return;
    const uint32_t mask = (1u << a.dd_bits) - 1u;
    unsigned long long key[PgHitBatch::CAP], seen[PgHitBatch::CAP];
    uint32_t slot[PgHitBatch::CAP];
    // ---- claims, all in flight together ----
#pragma unroll
    for (int j = 0; j < PgHitBatch::CAP; ++j) {
        seen[j] = ~0ull; key[j] = 0; slot[j] = 0;
        if (j < b.n && a.dd_keys) {
            key[j] = ((((unsigned long long)pid * a.dd_nseq + seq) << a.dd_end_bits) | (b.ent[j] >> 8)) + 1ull;
            slot[j] = pg_h2((uint32_t)(key[j] >> 32), (uint32_t)key[j]) >> (32 - a.dd_bits);
            seen[j] = atomicCAS(a.dd_keys + slot[j], 0ull, key[j]);
        }
    }
    // ---- who appends a record: the claimers, and whoever found no slot ----
    uint32_t rec = 0; // bit j: hit j gets a record
#pragma unroll
    for (int j = 0; j < PgHitBatch::CAP; ++j) {
        if (j < b.n) {
            bool mine = false, dup = false;
            if (a.dd_keys) {
                mine = seen[j] == 0ull;
                dup = seen[j] == key[j];
                for (int probe = 1; probe < 4 && !mine && !dup; ++probe) { // the slot holds another key: next ones
                    slot[j] = (slot[j] + 1) & mask;
                    const unsigned long long sn = atomicCAS(a.dd_keys + slot[j], 0ull, key[j]);
                    mine = sn == 0ull;
                    dup = sn == key[j];
                }
                if (mine || dup) atomicMax(a.dd_err + slot[j], 256u - (uint32_t)(b.ent[j] & 255ull));
                else slot[j] = 0xffffffffu; // no slot: a plain record
            } else slot[j] = 0xffffffffu;
            if (!dup) rec |= 1u << j;
        }
    }
    if (rec) {
        const unsigned long long base = atomicAdd(a.sink.count, (unsigned long long)__popc(rec));
        uint32_t o = 0;
#pragma unroll
        for (int j = 0; j < PgHitBatch::CAP; ++j) {
            if ((rec >> j) & 1u) {
                const unsigned long long idx = base + o++;
                if (idx < a.sink.cap) {
                    spmb_hit h;
                    h.end = b.ent[j] >> 8; h.seq_id = seq; h.pattern_id = pid; h.errors = (uint32_t)(b.ent[j] & 255ull);
                    h.reserved = slot[j] + 1u; // 0 for a plain record
                    a.sink.hits[idx] = h;
                }
            }
        }
    }
    b.n = 0;
}

// behind pigeon_verify_kernel: the records that claimed a slot take the slot's minimum
__global__ void __launch_bounds__(256) pigeon_fixup_kernel(spmb_hit *__restrict__ hits, const unsigned long long *__restrict__ count,
                                                           unsigned long long cap, const uint32_t *__restrict__ dd_err)
{
    const unsigned long long n = min(*count, cap);
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (unsigned long long)gridDim.x * blockDim.x) {
        const uint32_t r = hits[i].reserved;
        if (r) {
            hits[i].errors = 256u - dd_err[r - 1];
            hits[i].reserved = 0;
        }
    }
}

// symbol at offset d (>= 0) from (row R, column col); a plain text runs on into the next rows
template <int MODE>
__device__ __forceinline__ uint32_t pg_text(PgReader &rd, uint32_t RL, uint64_t R, uint64_t col)
{
    if (MODE == SPMB_MODE_PLAIN) {
        while (col >= RL) { col -= RL; ++R; }
    }
    return rd.at(R, (uint32_t)col);
}

// SPMB_ALG_PIGEONHOLE: every (candidate, seed) pair is a seed hit {end = begin + q, needle, offset}
// (pigeonhole_matcher.hpp:118-124); seeds longer than the 32-symbol key are compared in full.
template <int MODE>
__global__ void __launch_bounds__(128) pigeon_seed_kernel(PigeonVerifyArgs a)
{
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicMax(a.counters + 2, a.counters[1]);
    const unsigned long long n_cand = min(a.counters[1], a.cand_cap);
    const uint32_t RL = a.lay.row_len;
    PgReader rd{a.lanes, RL >> 5};
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_cand;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        const PigeonCand cd = a.cand[i];
        const uint64_t R = spmb_div_rl(a.lay, cd.pos);
        const uint32_t col = (uint32_t)(cd.pos - R * RL);
        uint64_t end;
        uint32_t seq = a.seq_id;
        if (MODE == SPMB_MODE_PLAIN) {
            end = cd.pos + a.pg.q;
            if (cd.pos < a.own_lo || cd.pos >= a.own_hi) continue;
        } else {
            end = (uint64_t)col + a.pg.q;
            if (a.row_own && end < a.row_own[R]) continue;
            if (a.row_seq) seq = a.row_seq[R];
        }
        for (uint32_t s = a.pg.ufirst[cd.u]; s < a.pg.ufirst[cd.u + 1]; ++s) {
            const uint32_t pid = a.pg.seed_pat[s], o = a.pg.seed_off[s];
            bool same = true;
            const uint8_t *pr = a.pat_ranks + a.pat_off[pid] + o;
            for (uint32_t j = a.pg.kq; j < a.pg.q && same; ++j)
                same = (pg_text<MODE>(rd, RL, R, (uint64_t)col + j) == pr[j]);
            if (same) pg_emit(a.sink, end, seq, pid, 0, o);
        }
    }
}

// one Myers column over NW words; hp/hn (unshifted) are handed back for the score row.
// GLOBAL: D[0][j] = j (MyersUkkonenGlobal, the anchored prefix matcher) -- a +1 enters at row 0.
template <int NW, bool GLOBAL>
__device__ __forceinline__ void pg_myers_col(uint32_t (&vp)[NW], uint32_t (&vn)[NW], const uint32_t (&eq)[NW],
                                             uint32_t (&hp)[NW], uint32_t (&hn)[NW])
{
    uint32_t x[NW], t[NW], s[NW];
#pragma unroll
    for (int i = 0; i < NW; ++i) {
        x[i] = eq[i] | vn[i];
        t[i] = x[i] & vp[i];
    }
    spmb_add_chain<NW>(s, t, vp);
    uint32_t hp_lo = GLOBAL ? 0x80000000u : 0u, hn_lo = 0;
#pragma unroll
    for (int i = 0; i < NW; ++i) {
        const uint32_t d0 = (s[i] ^ vp[i]) | x[i];
        hn[i] = vp[i] & d0;
        hp[i] = vn[i] | ~(vp[i] | d0);
        const uint32_t hps = __funnelshift_l(hp_lo, hp[i], 1);
        const uint32_t hns = __funnelshift_l(hn_lo, hn[i], 1);
        hp_lo = hp[i];
        hn_lo = hn[i];
        vn[i] = hps & d0;
        vp[i] = hns | ~(hps | d0);
    }
}

template <int NW>
__device__ __forceinline__ int pg_score_delta(const uint32_t (&hp)[NW], const uint32_t (&hn)[NW], uint32_t m)
{
    const uint32_t wi = (m - 1) >> 5, bi = (m - 1) & 31u;
    int d = 0;
#pragma unroll
    for (int i = 0; i < NW; ++i)
        if ((uint32_t)i == wi) d = (int)((hp[i] >> bi) & 1u) - (int)((hn[i] >> bi) & 1u);
    return d;
}

template <int NW>
__device__ __forceinline__ void pg_load_eq(const uint32_t *__restrict__ eqp, uint32_t c, uint32_t (&eq)[NW])
{
    constexpr int NWP = (NW <= 4) ? 4 : 8;
#pragma unroll
    for (int i = 0; i < NW; ++i) eq[i] = __ldg(eqp + c * NWP + i);
}

// SPMB_ALG_MYERS_FILTERED: Myers over the window [p-o-k, p-o+m+k) of every (candidate, seed) pair.
// One thread per candidate; the kernel is a chain of ~m+2k dependent column updates per thread, so what counts is
// the latency of one column, not the instruction count: round 1 fetched the needle's Eq row with four global loads
// per column and located the text symbol with a division per column -- 76 us for the C5 text, and just as long for
// 1/8 of it (a latency chain does not shrink with the shard: the bulk of C5's poor strong scaling).  Now the Eq rows
// of the thread's needle sit in shared memory (one LDS.128 per column), the text is walked chunk by chunk (32
// symbols per 8-byte load, the next chunk already on its way), and rows advance without a division.
template <int NW, int MODE>
__global__ void __launch_bounds__(128) pigeon_verify_kernel(PigeonVerifyArgs a)
{
    constexpr int NWP = (NW <= 4) ? 4 : 8;
    __shared__ __align__(16) uint32_t s_eq[128][4 * NWP + 4]; // +4 words: rows of neighbouring threads on different banks
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicMax(a.counters + 2, a.counters[1]);
    const unsigned long long n_cand = min(a.counters[1], a.cand_cap);
    const uint32_t RL = a.lay.row_len, rl32 = RL >> 5;
    uint32_t *my_eq = s_eq[threadIdx.x];
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_cand;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        const PigeonCand cd = a.cand[i];
        const uint64_t R0 = spmb_div_rl(a.lay, cd.pos);
        const uint32_t col = (uint32_t)(cd.pos - R0 * RL);
        // coordinates: plain = absolute text positions, windows = columns of row R0
        const int64_t p = (MODE == SPMB_MODE_PLAIN) ? (int64_t)cd.pos : (int64_t)col;
        const int64_t text_end = (MODE == SPMB_MODE_PLAIN) ? (int64_t)a.lay.n : (int64_t)a.row_len[R0];
        int64_t own_lo, own_hi; // of the last symbol index of a hit
        uint32_t seq = a.seq_id;
        if (MODE == SPMB_MODE_PLAIN) { own_lo = (int64_t)a.own_lo; own_hi = (int64_t)a.own_hi; }
        else {
            own_lo = a.row_own ? (int64_t)a.row_own[R0] - 1 : 0;
            own_hi = text_end;
            if (a.row_seq) seq = a.row_seq[R0];
        }
        for (uint32_t s = a.pg.ufirst[cd.u]; s < a.pg.ufirst[cd.u + 1]; ++s) {
            const uint32_t pid = a.pg.seed_pat[s], o = a.pg.seed_off[s];
            const MyersPatMeta pm = a.meta[pid];
            const int k = (int)pm.k;
            int64_t ws = p - (int64_t)o - k, we = p - (int64_t)o + (int64_t)pm.m + k;
            if (ws < 0) ws = 0;
            if (we > text_end) we = text_end;
            if (we <= own_lo || ws >= own_hi) continue;
            {   // this needle's Eq rows -> the thread's shared-memory row
                const uint4 *eg = reinterpret_cast<const uint4 *>(a.eq + (size_t)pid * 4 * NWP);
#pragma unroll
                for (int q = 0; q < NWP; ++q) reinterpret_cast<uint4 *>(my_eq)[q] = __ldg(eg + q);
            }
            // first chunk of the window: row / chunk index / offset inside the chunk
            uint64_t Rw = R0;
            uint32_t cw, ow;
            if (MODE == SPMB_MODE_PLAIN) {
                Rw = spmb_div_rl(a.lay, (uint64_t)ws);
                const uint32_t c0 = (uint32_t)((uint64_t)ws - Rw * RL);
                cw = c0 >> 5; ow = c0 & 31u;
            } else { cw = (uint32_t)ws >> 5; ow = (uint32_t)ws & 31u; }
            uint2 cur = spmb_chunk(a.lanes, rl32, Rw, cw);
            uint32_t vp[NW], vn[NW];
#pragma unroll
            for (int w = 0; w < NW; ++w) { vp[w] = 0xffffffffu; vn[w] = 0; }
            int score = (int)pm.m;
            int64_t t = ws;
            PgHitBatch hb;
            while (t < we) {
                // the chunk behind this one (plain text: rows follow one another; a window row simply ends)
                uint32_t cn = cw + 1;
                uint64_t Rn = Rw;
                if (cn == rl32 && MODE == SPMB_MODE_PLAIN) { cn = 0; ++Rn; }
                const int nb = (int)min((int64_t)(32 - ow), we - t);
                uint2 nxt = make_uint2(0, 0);
                if (t + nb < we) nxt = spmb_chunk(a.lanes, rl32, Rn, cn);
                unsigned long long W = (((unsigned long long)cur.x << 32) | cur.y) << (2 * ow);
                for (int j = 0; j < nb; ++j) {
                    const uint32_t c = (uint32_t)(W >> 62);
                    W <<= 2;
                    uint32_t eq[NW], hp[NW], hn[NW];
                    const uint4 e0 = *reinterpret_cast<const uint4 *>(my_eq + c * NWP);
                    eq[0] = e0.x;
                    if (NW > 1) eq[1 % NW] = e0.y;
                    if (NW > 2) eq[2 % NW] = e0.z;
                    if (NW > 3) eq[3 % NW] = e0.w;
                    if (NW > 4) {
                        const uint4 e1 = *reinterpret_cast<const uint4 *>(my_eq + c * NWP + 4);
                        eq[4 % NW] = e1.x;
                        if (NW > 5) eq[5 % NW] = e1.y;
                        if (NW > 6) eq[6 % NW] = e1.z;
                        if (NW > 7) eq[7 % NW] = e1.w;
                    }
                    pg_myers_col<NW, false>(vp, vn, eq, hp, hn);
                    score += pg_score_delta<NW>(hp, hn, pm.m);
                    const int64_t tt = t + j;
                    if (score <= k && tt >= own_lo && tt < own_hi) {
                        if (hb.n == PgHitBatch::CAP) pg_batch_flush(a, hb, seq, pid); // a hit-dense window
                        // constant indices only: the batch stays in registers
#pragma unroll
                        for (int q = 0; q < PgHitBatch::CAP; ++q)
                            if (q == hb.n) hb.ent[q] = ((unsigned long long)(tt + 1) << 8) | (unsigned long long)min(score, 255);
                        ++hb.n;
                    }
                }
                t += nb;
                ow = 0; cw = cn; Rw = Rn; cur = nxt;
            }
            pg_batch_flush(a, hb, seq, pid);
        }
    }
}

// SPMB_ALG_MYERS_PREFIX (restorable_myers_prefix_matcher): the needle against the prefixes of the text
// that starts at each anchor, scan clamped to m + k + 1 symbols (myers_prefix_matcher_restorable.hpp:47-50).
struct PrefixArgs {
    const uint2 *lanes;
    SpmbLayout lay;
    const uint64_t *anchors; // text positions; null = the single anchor 0
    uint64_t n_anchors;
    const uint32_t *eq;
    const MyersPatMeta *meta;
    uint32_t n_pat;
    uint32_t seq_id;
    HitSink sink;
};

template <int NW>
__global__ void __launch_bounds__(128) prefix_kernel(PrefixArgs a)
{
    constexpr int NWP = (NW <= 4) ? 4 : 8;
    const uint32_t RL = a.lay.row_len;
    PgReader rd{a.lanes, RL >> 5};
    const unsigned long long n_task = (unsigned long long)a.n_anchors * a.n_pat;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_task;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        const uint64_t ai = i / a.n_pat;
        const uint32_t pid = (uint32_t)(i % a.n_pat);
        const uint64_t p0 = a.anchors ? a.anchors[ai] : 0;
        const MyersPatMeta pm = a.meta[pid];
        if (pm.m == 0 || p0 >= a.lay.n) continue;
        const int k = (int)pm.k;
        uint64_t we = p0 + pm.m + pm.k + 1;
        if (we > a.lay.n) we = a.lay.n;
        const uint32_t *eqp = a.eq + (size_t)pid * 4 * NWP;
        uint32_t vp[NW], vn[NW];
#pragma unroll
        for (int w = 0; w < NW; ++w) { vp[w] = 0xffffffffu; vn[w] = 0; }
        int score = (int)pm.m;
        for (uint64_t t = p0; t < we; ++t) {
            const uint64_t Rt = spmb_div_rl(a.lay, t);
            const uint32_t c = rd.at(Rt, (uint32_t)(t - Rt * RL));
            uint32_t eq[NW], hp[NW], hn[NW];
            pg_load_eq<NW>(eqp, c, eq);
            pg_myers_col<NW, true>(vp, vn, eq, hp, hn);
            score += pg_score_delta<NW>(hp, hn, pm.m);
            if (score <= k) pg_emit(a.sink, t + 1, a.anchors ? (uint32_t)ai : a.seq_id, pid, (uint32_t)score, 0);
        }
    }
}
