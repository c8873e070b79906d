// myers_cut.cuh -- Myers bit-vector k-error search with the Ukkonen cut-off (SPMB_ALG_MYERS, NW >= 2).
//
// The reference reaches SeqAn2's _findMyersLargePatterns for needles longer than one machine word
// (libspm/libspm/matcher/myers_matcher.hpp:47-49 -> seqan_pattern_base.hpp:60-65); that scan keeps a
// `lastBlock` and only updates the blocks that can still hold a cell <= k (Ukkonen cut-off; SURVEY.md
// 8(a) row a10, oracle/spm_oracle.c spmo_myers_large).  myers_kernel<NW> (myers.cuh) always runs all
// NW words; this kernel is its cut-off form: same lane-per-row geometry, same registers, same hit
// list, but a warp only updates the first A <= NW words of VP/VN.
//
// A is re-decided every C = 8 columns and is warp-uniform (max over the 32 rows of the warp):
//   invariant   at a decision column all rows >= 32*A hold a true value D > k;
//   diagonal    D[i][j] >= D[i-1][j-1], so a row r with D > k for all rows >= r at column j gives
//               D > k for all rows >= r + t at column j + t;
//   decision    the smallest a whose rows >= 32a - C are provably > k now (lower bounds from popcounts, word by
//               word: see myers_cut_next)
//               keeps rows >= 32a above k for the next C columns; if no a <= A qualifies, A grows by one
//               word, initialised VP = ~0, VN = 0 (an upper bound of the true column, as in SeqAn2).
// Computing words 0..A-1 alone is the exact recurrence of the needle prefix (carries only run upward),
// re-initialised words hold upper bounds that are > k where it matters, so every cell with D <= k is
// exact -- in particular the score row, which is only inspected while A == NW.
// ALU work per column: 10*A + 1 (+ ~1 for the decision) instead of 10*NW + 4; on random DNA with
// k <= 4 a warp sits at A = 1 most of the time.
#pragma once
#include "myers.cuh"

template <int A, int NW, bool TRACK>
__device__ __forceinline__ void myers_cut_step(uint32_t (&vp)[NW], uint32_t (&vn)[NW], const uint32_t (&eq)[A],
                                               uint32_t &hp_acc, uint32_t &hn_acc)
{
    uint32_t x[A], t[A], s[A], v[A];
#pragma unroll
    for (int i = 0; i < A; ++i) {
        v[i] = vp[i];
        x[i] = eq[i] | vn[i];
        t[i] = eq[i] & v[i]; // = x & vp: VP and VN are disjoint
    }
    spmb_add_chain<A>(s, t, v);
    uint32_t hp_lo = 0, hn_lo = 0;
#pragma unroll
    for (int i = 0; i < A; ++i) {
        uint32_t d0 = (s[i] ^ v[i]) | x[i];
        uint32_t hn = v[i] & d0;
        uint32_t hp = vn[i] | ~(v[i] | d0);
        uint32_t hps = (i == 0) ? (hp << 1) : __funnelshift_l(hp_lo, hp, 1);
        uint32_t hns = (i == 0) ? (hn << 1) : __funnelshift_l(hn_lo, hn, 1);
        hp_lo = hp;
        hn_lo = hn;
        vn[i] = hps & d0;
        vp[i] = hns | ~(hps | d0);
    }
    if (TRACK) {
        hp_acc = __funnelshift_l(hp_lo, hp_acc, 1);
        hn_acc = __funnelshift_l(hn_lo, hn_acc, 1);
    }
}

__device__ __forceinline__ uint32_t spmb_lds32(uint32_t addr)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ uint2 spmb_lds64(uint32_t addr)
{
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
    return v;
}
__device__ __forceinline__ uint4 spmb_lds128(uint32_t addr)
{
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}

// C = 8 columns on the first A words; w holds the symbols, next one in the top two bits; eq_base is the
// shared-window address of the Eq table (rows of eq_stride bytes, one per symbol)
template <int A, int NW, bool TRACK, bool WIDE>
__device__ __forceinline__ void myers_cut_chunk(uint32_t (&vp)[NW], uint32_t (&vn)[NW], uint32_t &w, uint32_t &wh, uint32_t eq_base,
                                                uint32_t mul4, uint32_t eq_stride, uint32_t &hp_acc, uint32_t &hn_acc)
{
    constexpr int NWP = (NW <= 4) ? 4 : 8;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        uint32_t c = w >> 30;
        w *= mul4; // IMAD (FMA pipe), see MyersArgs
        if (WIDE) { c = (wh >> 30) * mul4 + c; wh *= mul4; } // 4-bit alphabets: bits 3..2 come from the second chunk
        const uint32_t row = c * eq_stride + eq_base;
        uint32_t eq[A];
        if (A == 1) {
            eq[0] = spmb_lds32(row);
        } else if (A == 2) {
            const uint2 e = spmb_lds64(row);
            eq[0] = e.x; eq[1 % A] = e.y;
        } else {
            const uint4 e0 = spmb_lds128(row);
            eq[0] = e0.x; eq[1 % A] = e0.y; eq[2 % A] = e0.z;
            if (A > 3) eq[3 % A] = e0.w;
            if (A > 4) {
                const uint4 e1 = spmb_lds128(row + (NWP / 4 - 1) * 16);
                eq[4 % A] = e1.x;
                if (A > 5) eq[5 % A] = e1.y;
                if (A > 6) eq[6 % A] = e1.z;
                if (A > 7) eq[7 % A] = e1.w;
            }
        }
        myers_cut_step<A, NW, TRACK>(vp, vn, eq, hp_acc, hn_acc);
    }
}

// the A = 1 fast path for 2-bit alphabets: 8 columns on word 0, TWO symbols per table access.  eq2_base is the
// shared-window address of a 16-row table {Eq0[first], Eq0[second]} indexed by a symbol pair (4 text bits):
// one SHF + two IMADs + one LDS.64 per two columns instead of that per column.  A half-warp's 16 possible rows
// cover the 32 banks exactly once, so the access is conflict-free whatever the symbols are.
template <int NW>
__device__ __forceinline__ void myers_cut_chunk1_pairs(uint32_t (&vp)[NW], uint32_t (&vn)[NW], uint32_t &w, uint32_t eq2_base,
                                                       uint32_t mul16, uint32_t mul8)
{
    uint32_t d0 = 0, d1 = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const uint32_t c2 = w >> 28;
        w *= mul16; // IMAD (FMA pipe)
        const uint2 e = spmb_lds64(c2 * mul8 + eq2_base);
        uint32_t eq[1];
        eq[0] = e.x;
        myers_cut_step<1, NW, false>(vp, vn, eq, d0, d1);
        eq[0] = e.y;
        myers_cut_step<1, NW, false>(vp, vn, eq, d0, d1);
    }
}

// number of words this row needs for the next 8 columns, given words 0..A-1 are valid now.
// Lower bounds from popcounts (S = value at the bottom of the previous word, P/N = VP/VN bits of a word):
//   every row of word w            D >= S_{w-1} - N_w
//   rows 24..31 of word a-1        D >= S_{a-2} + popc(VP_{a-1} & rows 0..23) - N_{a-1}
// a words suffice when both kinds of rows from 32a - 8 upward are above k.  (Bounding the upper words one by one
// instead of subtracting all their VN bits at once lets a warp fall back to one word right after an excursion:
// with the coarse bound a return from 2 words to 1 was a coin toss at k = 4 and rare at k = 8.)
template <int A, int NW>
__device__ __forceinline__ int myers_cut_next(const uint32_t (&vp)[NW], const uint32_t (&vn)[NW], int k)
{
    int N[A], S[A + 1];
    S[0] = 0;
#pragma unroll
    for (int w = 0; w < A; ++w) { N[w] = __popc(vn[w]); S[w + 1] = S[w] + __popc(vp[w]) - N[w]; }
    int M[A + 1]; // M[w] = min over words v >= w of (S_{v-1} - N_v)
    M[A] = 0x3fffffff;
#pragma unroll
    for (int w = A - 1; w >= 1; --w) M[w] = min(M[w + 1], S[w] - N[w]);
    int res = (A < NW) ? A + 1 : NW;
    bool found = false;
#pragma unroll
    for (int a = 1; a <= A; ++a) {
        const int lb = min(S[a - 1] + __popc(vp[a - 1] & 0x00ffffffu) - N[a - 1], M[a]);
        if (!found && lb > k) { res = a; found = true; }
    }
    return res;
}

template <int A, int NW, bool WIDE>
struct MyersCutDispatch {
    static __device__ __forceinline__ int run(int Acur, uint32_t (&vp)[NW], uint32_t (&vn)[NW], uint32_t &w, uint32_t &wh, uint32_t s_eq,
                                              uint32_t mul4, uint32_t eq_stride, int k, uint32_t &hp_acc, uint32_t &hn_acc)
    {
        if (A == NW || Acur == A) {
            myers_cut_chunk<A, NW, A == NW, WIDE>(vp, vn, w, wh, s_eq, mul4, eq_stride, hp_acc, hn_acc);
            return myers_cut_next<A, NW>(vp, vn, k);
        }
        return MyersCutDispatch<(A < NW ? A + 1 : NW), NW, WIDE>::run(Acur, vp, vn, w, wh, s_eq, mul4, eq_stride, k, hp_acc, hn_acc);
    }
};

template <int NW, int MODE, bool WIDE = false>
__global__ void __launch_bounds__(128) myers_cut_kernel(MyersArgs a)
{
    constexpr int NWP = (NW <= 4) ? 4 : 8;
    constexpr int NSYM = WIDE ? 16 : 4;
    __shared__ __align__(16) uint32_t s_eq[NSYM * NWP];
    __shared__ __align__(8) uint32_t s_eq2[32]; // word 0 of Eq for a symbol PAIR: row (first << 2 | second) = {Eq0[first], Eq0[second]}
    const uint32_t pat = blockIdx.y;
    for (uint32_t t = threadIdx.x; t < NSYM * NWP; t += blockDim.x) s_eq[t] = a.eq[(size_t)pat * NSYM * NWP + t];
    if (!WIDE && threadIdx.x < 32) {
        const uint32_t pair = threadIdx.x >> 1, sym = (threadIdx.x & 1) ? (pair & 3u) : (pair >> 2);
        s_eq2[threadIdx.x] = a.eq[(size_t)pat * NSYM * NWP + sym * NWP];
    }
    __syncthreads();
    const MyersPatMeta pm = a.meta[pat];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // plain mode: a row is cut into a.split pieces (each with its own warm-up), one warp per piece, so
    // that the short dependency chain of the A = 1 update finds ~12 warps per scheduler instead of the 6
    // the store geometry gives (ncu: 72 % issue-active at 6 warps)
    const uint32_t S = (MODE == SPMB_MODE_PLAIN) ? a.split : 1u;
    const uint64_t unit = (uint64_t)blockIdx.x * 4 + warp;
    const uint64_t st = a.st0 + unit / S;
    const uint32_t piece = (uint32_t)(unit - (unit / S) * S);
    if (st >= a.st1) return; // warp-uniform
    const uint32_t RL = a.lay.row_len;
    const uint64_t R = st * 32 + lane;
    // rows past the end stay in the loop (the active-word count is a warp-wide maximum) and own nothing
    const bool valid = R * RL < a.lay.n;

    const int k = (int)pm.k;
    const int W = (int)pm.pad;

    uint32_t vp[NW], vn[NW];
#pragma unroll
    for (int i = 0; i < NW; ++i) { vp[i] = 0xffffffffu; vn[i] = 0; }

    int64_t p, end_p, own_lo, own_hi;
    int64_t cur_row;
    uint32_t cur_col;
    uint32_t seq = a.seq_id;
    if (MODE == SPMB_MODE_PLAIN) {
        const uint32_t b0 = ((RL >> 5) * piece / S) << 5, b1 = ((RL >> 5) * (piece + 1) / S) << 5;
        own_lo = (int64_t)(R * RL) + b0;
        const int64_t own_end = (int64_t)(R * RL) + b1;
        own_hi = own_end;
        if (own_hi > (int64_t)a.lay.n) own_hi = (int64_t)a.lay.n;
        if (!valid || own_hi < own_lo) own_hi = own_lo;
        p = own_lo - (int64_t)a.warm;
        if (p < 0) p = 0;
        cur_row = (int64_t)spmb_div_rl(a.lay, (uint64_t)p);
        cur_col = (uint32_t)(p - cur_row * RL);
        end_p = own_end + (W ? 32 : 0);
    } else {
        own_lo = 0; own_hi = 0;
        p = 0;
        cur_row = (int64_t)R;
        cur_col = 0;
        end_p = 0;
        if (valid) {
            own_lo = a.row_own ? (int64_t)a.row_own[R] - 1 : 0;
            own_hi = (int64_t)a.row_len[R];
            if (a.row_seq) seq = a.row_seq[R];
            end_p = ((own_hi + W + 31) / 32) * 32;
            if (end_p > (int64_t)RL) end_p = RL;
        }
    }
    // one trip count for the warp: rows that are done early run on (reads stay inside the allocation:
    // the lane-chunk array covers whole super-tiles plus a row of padding) and emit nothing, because
    // every column past end_p lies outside [own_lo, own_hi)
    const int my_blocks = (int)((end_p - p) >> 5);
    const int n_blocks = __reduce_max_sync(0xffffffffu, my_blocks);

    const uint32_t rl32 = RL >> 5;
    const uint32_t mul4 = a.mul4, eq_stride = a.eq_stride;
    const uint2 *__restrict__ lp = a.lanes;
    uint32_t eq_base; // pinned in a register: ptxas otherwise re-derives the shared window (S2R + LEA) per block
    asm volatile("mov.u32 %0, %1;" : "=r"(eq_base) : "r"((uint32_t)__cvta_generic_to_shared(s_eq)));
    uint32_t eq2_base;
    asm volatile("mov.u32 %0, %1;" : "=r"(eq2_base) : "r"((uint32_t)__cvta_generic_to_shared(s_eq2)));
    const uint32_t mul16 = mul4 * mul4, mul8 = mul4 + mul4; // run-time multipliers: the shifts stay IMADs (MyersArgs)
    int A = __reduce_max_sync(0xffffffffu, myers_cut_next<NW, NW>(vp, vn, k));
    int score = 32 * NW; // meaningful while A == NW

    const uint2 *__restrict__ lph = a.lanes_hi;
    uint64_t li = ((uint64_t)(cur_row >> 5) * rl32 + (cur_col >> 5)) * 32 + ((uint32_t)cur_row & 31u);
    uint2 nxt = __ldg(lp + li), nxt_hi = WIDE ? __ldg(lph + li) : make_uint2(0, 0);
    for (int blk = 0; blk < n_blocks; ++blk) {
        const uint2 cur = nxt, cur_hi = nxt_hi;
        cur_col += 32;
        if (cur_col == RL) { cur_col = 0; ++cur_row; }
        if (blk + 1 < n_blocks) {
            li = ((uint64_t)(cur_row >> 5) * rl32 + (cur_col >> 5)) * 32 + ((uint32_t)cur_row & 31u);
            nxt = __ldg(lp + li);
            if (WIDE) nxt_hi = __ldg(lph + li);
        }
        uint32_t w = cur.x, wh = cur_hi.x;
        int q = 0;
        if (A == 1) {
            // the common state on random text: one word, no score row in reach (NW >= 2), no bookkeeping
#pragma unroll
            for (int qq = 0; qq < 4; ++qq) {
                if (qq == 2) { w = cur.y; wh = cur_hi.y; }
                if (WIDE) {
                    uint32_t d0 = 0, d1 = 0;
                    myers_cut_chunk<1, NW, false, WIDE>(vp, vn, w, wh, eq_base, mul4, eq_stride, d0, d1);
                } else {
                    myers_cut_chunk1_pairs<NW>(vp, vn, w, eq2_base, mul16, mul8);
                }
                q = qq + 1;
                const bool grow = (int)__popc(vp[0] & 0x00ffffffu) - (int)__popc(vn[0]) <= k;
                if (__any_sync(0xffffffffu, grow)) {
                    A = 2;
                    vp[1 % NW] = 0xffffffffu; vn[1 % NW] = 0;
                    if (NW == 2) score = 32 + (int)__popc(vp[0]) - (int)__popc(vn[0]);
                    break;
                }
            }
            p += 8 * q;
        } else if (NW > 2 && A == 2) {
            // two words (k around 10, or the columns right after an excursion): same shape, the warp leaves as
            // soon as its word count changes
#pragma unroll
            for (int qq = 0; qq < 4; ++qq) {
                if (qq == 2) { w = cur.y; wh = cur_hi.y; }
                uint32_t d0 = 0, d1 = 0;
                myers_cut_chunk<2, NW, false, WIDE>(vp, vn, w, wh, eq_base, mul4, eq_stride, d0, d1);
                q = qq + 1;
                const int An = __reduce_max_sync(0xffffffffu, myers_cut_next<(NW > 2 ? 2 : 1), NW>(vp, vn, k));
                if (An != 2) {
                    if (An > 2) {
                        vp[2 % NW] = 0xffffffffu; vn[2 % NW] = 0;
                        if (NW == 3) {
                            score = 0;
#pragma unroll
                            for (int i = 0; i < NW; ++i) score += (int)__popc(vp[i]) - (int)__popc(vn[i]);
                        }
                    }
                    A = An;
                    break;
                }
            }
            p += 8 * q;
        }
#pragma unroll 1
        for (; q < 4; ++q) {
            if (q == 2) { w = cur.y; wh = cur_hi.y; }
            uint32_t hp_acc = 0, hn_acc = 0;
            int An = MyersCutDispatch<1, NW, WIDE>::run(A, vp, vn, w, wh, eq_base, mul4, eq_stride, k, hp_acc, hn_acc);
            if (A == NW) {
                if (score - (int)__popc(hn_acc) <= k) {
                    int s = score;
                    for (int t = 0; t < 8; ++t) {
                        const int b = 7 - t;
                        s += (int)((hp_acc >> b) & 1u) - (int)((hn_acc >> b) & 1u);
                        if (s <= k) {
                            const int64_t j = p + t - W;
                            if (j >= own_lo && j < own_hi) myers_emit(a.sink, (uint64_t)(j + 1), seq, pm.pattern_id, (uint32_t)s);
                        }
                    }
                }
                score += (int)__popc(hp_acc) - (int)__popc(hn_acc);
            }
            An = __reduce_max_sync(0xffffffffu, An);
            if (An > A) { // one more word: an upper bound of its true column (all vertical deltas +1)
#pragma unroll
                for (int i = 1; i < NW; ++i)
                    if (i == A) { vp[i] = 0xffffffffu; vn[i] = 0; }
                if (An == NW) {
                    score = 0;
#pragma unroll
                    for (int i = 0; i < NW; ++i) score += (int)__popc(vp[i]) - (int)__popc(vn[i]);
                }
            }
            A = An;
            p += 8;
        }
    }
}
