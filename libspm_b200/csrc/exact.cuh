// exact.cuh -- exact search kernel (serves SPMB_ALG_HORSPOOL and SPMB_ALG_SHIFTOR).
//
// Replaces the inner loops behind spm::horspool_matcher / spm::shiftor_matcher
// (libspm/libspm/matcher/horspool_matcher.hpp:20-44, shiftor_matcher.hpp:20-44, both driven by
// seqan_pattern_base::operator(), seqan_pattern_base.hpp:40-52).  Semantics kept: EVERY
// occurrence, overlapping ones included, end = begin + m.  Control flow not kept: Horspool's
// data-dependent skip and Shift-Or's serial D-word recurrence are both hostile to SIMT, and the
// path is HBM-bound (0.25 B/bp), so the kernel is a bit-sliced shift-AND over ALIGNMENTS:
//
//   thread-row = 32 consecutive columns x 32 bit-rows = 1024 alignments.
//   mismatch[c] |= f_sym_j(H[c+j], L[c+j])   for the first J<=8 pattern symbols j
// where f_sym is ONE LOP3 whose truth table is picked by a warp-uniform switch on the pattern
// symbol (text position p+j lives in column c+j, same bit: no shifts at all).  After J=8 steps
// a random alignment survives with probability 4^-8; survivors (and all true hits) go to a
// warp-cooperative verification of the remaining m-J symbols and are appended to the hit buffer
// with one warp-aggregated atomic.
//
// Per thread-row: 20 coalesced LDG.128 + ~8.5 LOP3 per column  => ~0.3 instr / searched bp.
//
// Tried and measured against this shape on B200 (3.2 Gbp, 800 MB): a persistent TMA ring
// (cp.async.bulk + mbarriers, 3 x 72 KB and 5 x 40 KB stages, 512 consumer threads) 148 / 170 us, and
// private cp.async (LDGSTS) slots with 160 KB in flight per SM 155 us, against 141-146 us then and 126-131 us now; the
// kernel is co-limited by the ALU pipe (8 LOP3 per column = 70 us) and DRAM (114 us for a pure read,
// tools/ubench_read.cu), and independent load-then-compute warps overlap the two best.
#pragma once
#include "layout.cuh"
#include "hits.cuh"

struct ExactArgs {
    const uint4 *planes;     // packed store viewed as uint4 (two columns)
    SpmbLayout lay;
    uint64_t n_tiles;        // one past the last thread-row of this launch
    uint64_t tile0;          // first thread-row of this launch (shards cover disjoint super-tile ranges)
    const uint8_t *pat;      // m ranks (device)
    const uint32_t *row_len; // windows mode: valid symbols per row; else null
    const uint32_t *row_seq; // windows mode: seq id per row; else null
    const uint32_t *row_own; // windows mode: first owned end per row (hits ending earlier belong elsewhere); may be null
    uint32_t m;
    uint32_t J;              // filter depth = min(m, 8)
    uint32_t fo;             // offset of the filter window inside the needle: a candidate at column c is the alignment
                             // that begins at column c - fo.  The host picks the least repetitive 8-symbol window of
                             // the needle (pattern_build): on low-complexity text (poly-A, satellites) the FIRST eight
                             // symbols may match everywhere while a later window still rejects almost every alignment
    uint32_t filt_syms;      // pattern symbol fo + j at bits [2j, 2j+1], j < J
    // 4-bit alphabets (spm::dna5 / spm::dna15, alphabet.hpp:119-120): a second plane pair holds bits 3..2 of
    // every symbol, in the same geometry; the filter accumulates over both pairs
    const uint4 *planes_hi;  // null for dna4
    uint32_t filt_syms_hi;   // bits 3..2 of pattern symbol j at bits [2j, 2j+1]
    uint32_t seq_id;
    uint32_t pattern_id;
    HitSink sink;
};

// mismatch-accumulate for one filter step; SYM is a compile-time constant inside each case so
// every line below is a single LOP3 with an immediate truth table.
template <int SYM>
__device__ __forceinline__ uint32_t exact_acc(uint32_t acc, uint32_t h, uint32_t l)
{
    if (SYM == 0) return acc | h | l;
    if (SYM == 1) return acc | h | ~l;
    if (SYM == 2) return acc | ~h | l;
    return acc | ~h | ~l;
}

#define SPMB_EXACT_STEP(j)                                                                       \
    if (J > (j)) {                                                                               \
        switch ((fs >> (2 * (j))) & 3u) {                                                        \
        case 0:  _Pragma("unroll") for (int c = 0; c < 32; ++c) acc[c] = exact_acc<0>(acc[c], H[c + (j)], L[c + (j)]); break; \
        case 1:  _Pragma("unroll") for (int c = 0; c < 32; ++c) acc[c] = exact_acc<1>(acc[c], H[c + (j)], L[c + (j)]); break; \
        case 2:  _Pragma("unroll") for (int c = 0; c < 32; ++c) acc[c] = exact_acc<2>(acc[c], H[c + (j)], L[c + (j)]); break; \
        default: _Pragma("unroll") for (int c = 0; c < 32; ++c) acc[c] = exact_acc<3>(acc[c], H[c + (j)], L[c + (j)]); break; \
        }                                                                                        \
    }

// Warp-cooperative verification of the candidates in one column word.
// gc = global column of the candidates (= alignment begin + fo), cm = candidate bits (bit b = row b).  Lane i checks
// pattern offsets i, i+32, ...; returns the surviving candidate mask (same value in every lane).
// Needle symbol j of the alignment sits at column col - fo + j of the same row: past the row end it continues in
// the next row(s) (one bit up; row 31 continues in row 0 of the next super-tile), before the row start in the
// previous row(s) (one bit down).
template <int MODE, bool WIDE>
__device__ __forceinline__ uint32_t exact_verify(const ExactArgs &a, uint64_t gc, uint32_t cm, uint32_t lane)
{
    const uint2 *P2 = reinterpret_cast<const uint2 *>(a.planes);
    const uint2 *P2H = reinterpret_cast<const uint2 *>(a.planes_hi);
    const int64_t RL = (int64_t)a.lay.row_len;
    const uint64_t st = spmb_div_rl(a.lay, gc);
    const int64_t col = (int64_t)(gc - st * (uint64_t)RL);
    {   // alignments that begin before the first row of the text (plain) / before their own row (windows)
        const int64_t x0 = col - (int64_t)a.fo;
        if (x0 < 0) {
            if (MODE != SPMB_MODE_PLAIN) return 0;
            const uint32_t back = (uint32_t)((-x0 + RL - 1) / RL);
            if (st == 0) cm &= (back >= 32) ? 0u : ~((1u << back) - 1u);
        }
    }
    for (uint32_t j0 = 0; j0 < a.m && cm; j0 += 32) {
        const uint32_t j = j0 + lane;
        uint32_t mm = 0xffffffffu;
        if (j < a.m) {
            // column of needle symbol j relative to the row start: inside the row almost always (no division then);
            // |xi| < 32 rows of at most 65536 columns, so 32-bit arithmetic (a 64-bit division here cost the C2x
            // kernel 4 % through its 49 k filter survivors per pass)
            const int32_t rl = (int32_t)RL;
            const int32_t xi = (int32_t)col - (int32_t)a.fo + (int32_t)j;
            int32_t wraps = 0;
            if (xi >= rl) wraps = xi / rl;
            else if (xi < 0) wraps = -((-xi + rl - 1) / rl);
            const uint64_t x = (uint64_t)(xi - wraps * rl);
            const uint32_t s = a.pat[j];
            uint32_t h, l, hh = 0, lh = 0;
            if (wraps == 0) {
                const uint2 v = __ldg(P2 + spmb_perm(st * RL + x));
                h = v.x; l = v.y;
                if (WIDE) { const uint2 u = __ldg(P2H + spmb_perm(st * RL + x)); hh = u.x; lh = u.y; }
            } else if (MODE != SPMB_MODE_PLAIN) {
                h = l = 0;
                mm = 0; // windows do not continue: nothing can match outside the row
            } else if (wraps > 0) {
                const uint2 v0 = __ldg(P2 + spmb_perm(st * RL + x));
                const uint2 v1 = __ldg(P2 + spmb_perm((st + 1) * RL + x));
                h = __funnelshift_r(v0.x, v1.x, (uint32_t)wraps);
                l = __funnelshift_r(v0.y, v1.y, (uint32_t)wraps);
                if (WIDE) {
                    const uint2 u0 = __ldg(P2H + spmb_perm(st * RL + x));
                    const uint2 u1 = __ldg(P2H + spmb_perm((st + 1) * RL + x));
                    hh = __funnelshift_r(u0.x, u1.x, (uint32_t)wraps);
                    lh = __funnelshift_r(u0.y, u1.y, (uint32_t)wraps);
                }
            } else {
                const uint32_t w = (uint32_t)(-wraps);
                const uint2 v0 = __ldg(P2 + spmb_perm(st * RL + x));
                const uint2 vp = st ? __ldg(P2 + spmb_perm((st - 1) * RL + x)) : make_uint2(0, 0);
                h = __funnelshift_l(vp.x, v0.x, w);
                l = __funnelshift_l(vp.y, v0.y, w);
                if (WIDE) {
                    const uint2 u0 = __ldg(P2H + spmb_perm(st * RL + x));
                    const uint2 up = st ? __ldg(P2H + spmb_perm((st - 1) * RL + x)) : make_uint2(0, 0);
                    hh = __funnelshift_l(up.x, u0.x, w);
                    lh = __funnelshift_l(up.y, u0.y, w);
                }
            }
            if (mm) {
                mm = ~((h ^ (0u - ((s >> 1) & 1u))) | (l ^ (0u - (s & 1u))));
                if (WIDE) mm &= ~((hh ^ (0u - ((s >> 3) & 1u))) | (lh ^ (0u - ((s >> 2) & 1u))));
            }
        }
        cm &= __reduce_and_sync(0xffffffffu, mm);
    }
    return cm;
}

template <int MODE>
__device__ __forceinline__ void exact_emit(const ExactArgs &a, uint64_t gc, uint32_t cm, uint32_t lane)
{
    const uint32_t RL = a.lay.row_len;
    uint64_t st = spmb_div_rl(a.lay, gc);
    uint32_t col = (uint32_t)(gc - st * RL);
    uint64_t R = st * 32 + lane;
    bool ok = (cm >> lane) & 1u;
    uint64_t end;
    uint32_t seq = a.seq_id;
    if (MODE == SPMB_MODE_PLAIN) {
        const int64_t begin = (int64_t)(R * RL + col) - (int64_t)a.fo;
        end = (uint64_t)(begin + (int64_t)a.m);
        ok = ok && begin >= 0 && (end <= a.lay.n);
    } else {
        end = (uint64_t)col + a.m - a.fo;
        ok = ok && (col >= a.fo) && (R * RL < a.lay.n) && (end <= a.row_len[R]) && (!a.row_own || end >= a.row_own[R]);
        if (ok && a.row_seq) seq = a.row_seq[R];
    }
    uint32_t em = __ballot_sync(0xffffffffu, ok);
    if (em == 0) return;
    unsigned long long base = 0;
    if (lane == 0) base = atomicAdd(a.sink.count, (unsigned long long)__popc(em));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (ok) {
        unsigned long long idx = base + __popc(em & ((1u << lane) - 1u));
        if (idx < a.sink.cap) {
            spmb_hit hrec;
            hrec.end = end; hrec.seq_id = seq; hrec.pattern_id = a.pattern_id; hrec.errors = 0; hrec.reserved = 0;
            a.sink.hits[idx] = hrec;
        }
    }
}

// one filter pass over one plane pair: load the thread-row (+ 8 halo columns) and OR the mismatches of the
// first J pattern symbols (their two bits in this pair = fs) into acc
template <int MODE>
__device__ __forceinline__ void exact_filter_pass(const uint4 *__restrict__ planes, const SpmbLayout &lay, uint64_t tile,
                                                  uint32_t J, uint32_t fs, uint32_t (&acc)[32])
{
    const uint32_t RL = lay.row_len;
    uint32_t H[40], L[40];
    // own thread-row: chunk k (columns 2k, 2k+1) of lane l sits at uint4 index blk*512 + k*32 + l
    const uint4 *p = planes + (tile >> 5) * 512 + (tile & 31);
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        uint4 q = __ldg(p + k * 32);
        H[2 * k] = q.x; L[2 * k] = q.y; H[2 * k + 1] = q.z; L[2 * k + 1] = q.w;
    }
    // halo: first 8 columns of the next thread-row
    const uint64_t g0 = tile * 32;
    const uint64_t st = spmb_div_rl(lay, g0);
    const uint32_t col0 = (uint32_t)(g0 - st * RL);
    if (col0 + 32 != RL) {
        const uint4 *pn = planes + ((tile + 1) >> 5) * 512 + ((tile + 1) & 31);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            uint4 q = __ldg(pn + k * 32);
            H[32 + 2 * k] = q.x; L[32 + 2 * k] = q.y; H[33 + 2 * k] = q.z; L[33 + 2 * k] = q.w;
        }
    } else if (MODE == SPMB_MODE_PLAIN) {
        // last thread-row of a super-tile: row r continues in row r+1, i.e. the same
        // columns 0..7 one bit up; row 31 continues in row 0 of the next super-tile.
        const uint2 *P2 = reinterpret_cast<const uint2 *>(planes);
#pragma unroll
        for (int t = 0; t < 8; ++t) {
            uint2 v0 = __ldg(P2 + spmb_perm(st * RL + t));
            uint2 v1 = __ldg(P2 + spmb_perm((st + 1) * RL + t));
            H[32 + t] = (v0.x >> 1) | (v1.x << 31);
            L[32 + t] = (v0.y >> 1) | (v1.y << 31);
        }
    } else {
#pragma unroll
        for (int t = 0; t < 8; ++t) { H[32 + t] = 0; L[32 + t] = 0; }
    }
    SPMB_EXACT_STEP(0) SPMB_EXACT_STEP(1) SPMB_EXACT_STEP(2) SPMB_EXACT_STEP(3)
    SPMB_EXACT_STEP(4) SPMB_EXACT_STEP(5) SPMB_EXACT_STEP(6) SPMB_EXACT_STEP(7)
}

template <int MODE, bool WIDE = false>
__global__ void __launch_bounds__(256, 2) exact_kernel(ExactArgs a)
{
    __shared__ uint32_t sbuf[8][32][33];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // grid-stride over CTA tiles of 256 thread-rows (8 CTAs per SM in the grid, 2 resident)
    for (uint64_t ct = blockIdx.x; a.tile0 + ct * 256 < a.n_tiles; ct += gridDim.x) {
    const uint64_t tile = a.tile0 + ct * 256 + threadIdx.x;
    uint32_t acc[32];
    uint32_t all = 0xffffffffu;

    if (tile < a.n_tiles) {
#pragma unroll
        for (int c = 0; c < 32; ++c) acc[c] = 0;
        exact_filter_pass<MODE>(a.planes, a.lay, tile, a.J, a.filt_syms, acc);
        if (WIDE) exact_filter_pass<MODE>(a.planes_hi, a.lay, tile, a.J, a.filt_syms_hi, acc);
#pragma unroll
        for (int c = 0; c < 32; ++c) all &= acc[c];
    }

    const bool cand = (all != 0xffffffffu);
    uint32_t vote = __ballot_sync(0xffffffffu, cand);
    if (vote == 0) continue;

    // ---- rare path: a lane saw a candidate ----
    if (cand) {
#pragma unroll
        for (int c = 0; c < 32; ++c) sbuf[warp][lane][c] = acc[c];
    }
    __syncwarp();
    while (vote) {
        const int src = __ffs(vote) - 1;
        vote &= vote - 1;
        const uint64_t stile = __shfl_sync(0xffffffffu, tile, src);
        const uint32_t w = ~sbuf[warp][src][lane]; // lane i: candidate bits of column i
        uint32_t cvote = __ballot_sync(0xffffffffu, w != 0);
        while (cvote) {
            const int c = __ffs(cvote) - 1;
            cvote &= cvote - 1;
            uint32_t cm = __shfl_sync(0xffffffffu, w, c);
            const uint64_t gc = stile * 32 + c;
            cm = exact_verify<MODE, WIDE>(a, gc, cm, lane);
            if (cm) exact_emit<MODE>(a, gc, cm, lane);
        }
    }
    __syncwarp();
    }
}
