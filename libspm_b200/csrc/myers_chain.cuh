// myers_chain.cuh -- Myers bit-vector k-error search with the needle's words CHAINED ACROSS THE LANES of a warp
// (north_star item 3: "chain multi-word patterns across lanes with warp shuffles").
//
// myers_kernel / myers_cut_kernel keep all ceil(m/32) words of one row's state in ONE lane: carries cost nothing
// (IADD3.X, funnel shifts) and 32 rows advance per warp instruction -- the right shape while the words fit the
// register file, i.e. m <= 256.  The reference takes any needle (myers_matcher.hpp:37-43; SeqAn2's large-pattern path
// is unbounded), so beyond 256 symbols the state is spread over the warp instead: lane l holds words
// [l*WPL, (l+1)*WPL) of VP/VN (WPL = 1, 2, 4, 8: needles up to 1024 / 2048 / 4096 / 8192 symbols), ONE text column
// advances per step for the whole warp, and the three carries of the recurrence cross the lanes like this:
//   * the addition X&VP + VP: every lane adds its words with no carry-in and reports generate (carry out) and
//     propagate (all ones); two ballots turn the lane-level carry chain into ONE 32-bit addition
//     (A = G|P, B = G: the carry INTO lane l is bit l of (A + B) ^ A ^ B), and each lane adds its carry-in;
//   * HP << 1 and HN << 1: the top bit of the previous lane's last word arrives by __shfl_up.
// Per column and lane: 10*WPL recurrence instructions + ~14 for the ballots, the carry arithmetic and the two
// shuffles -- about 2.4x the instructions of the one-lane form per word, and 32x fewer rows in flight per warp.
// It is therefore used where the one-lane form cannot go:
//   myers_chain_verify_kernel  needles longer than 256 symbols: the first 256 symbols run through the one-lane
//                              kernels as a FILTER, and one warp per filter hit runs the full needle.
//                              Which hit a candidate stands for: let E be the (exclusive) end of an occurrence of
//                              the whole needle with <= k errors, and e* = E - (m - 256).  An optimal alignment
//                              crosses needle row 256 at some text column e0 with a errors before and b after it,
//                              a + b <= k, and |E - e0 - (m - 256)| <= b (the suffix's indels); the prefix score
//                              D[256][.] moves by at most 1 per column, so D[256][e*] <= a + |e* - e0| <= k: e* is
//                              ALWAYS a filter hit.  So the candidate with prefix end e checks exactly one end,
//                              E = e + (m - 256), over the window [E - m - k, E) that holds every occurrence with
//                              <= k errors ending there: every hit is found once, no duplicates, and the window's
//                              score is the true one whenever it is <= k;
//   myers_chain_scan_kernel    the same column step over whole rows (SPMB_MYERS_CHAIN=1): the A/B measurement of
//                              the two forms at m <= 256 (tools/myers_chain_ab.py).
#pragma once
#include "layout.cuh"
#include "hits.cuh"
#include "myers.cuh"
#include "pack.cuh"

struct ChainArgs {
    const uint2 *lanes;               // lane chunks of the store
    SpmbLayout lay;
    const uint32_t *eq;               // [n_needles][4][32*WPL]: Eq rows of the full needles, zero beyond ceil(m/32) words
    const MyersPatMeta *meta;         // [n_needles]: m, k, pattern_id of the caller's set (pad unused)
    // verify: candidates = hits of the prefix filter (end = exclusive end of a prefix occurrence, pattern_id = index
    // into eq / meta, seq_id = the filter's sequence id: windows mode, the window index)
    const spmb_hit *cand;
    unsigned long long *counters;     // [0] hits, [1] candidates of this launch, [2] max candidates of any launch
    unsigned long long cand_cap;
    uint32_t prefix_m;                // symbols of the needle the filter looked at (256)
    // windows mode: per-row length / first owned end; the row of a candidate is seq_id - seq_base
    const uint32_t *row_len, *row_own;
    uint32_t seq_base;
    // scan: every row of [st0, st1) x every needle
    uint64_t st0, st1;
    uint32_t n_needles;
    uint32_t warm;
    uint32_t seq_id;
    uint64_t own_lo, own_hi;          // verify, plain mode: owned range of the LAST symbol index of a hit
    HitSink sink;
};

// one text column for the whole warp; the lane's (hp, hn) words come back through the references for the score row
template <int WPL>
__device__ __forceinline__ void myers_chain_step(uint32_t (&vp)[WPL], uint32_t (&vn)[WPL], const uint32_t (&eq)[WPL], uint32_t lane,
                                                 uint32_t (&hp)[WPL], uint32_t (&hn)[WPL])
{
    uint32_t x[WPL], s[WPL];
    uint32_t carry = 0, all_ones = 0xffffffffu;
#pragma unroll
    for (int w = 0; w < WPL; ++w) {
        x[w] = eq[w] | vn[w];
        const unsigned long long sum = (unsigned long long)(x[w] & vp[w]) + vp[w] + carry;
        s[w] = (uint32_t)sum;
        carry = (uint32_t)(sum >> 32);
        all_ones &= s[w];
    }
    // lane-level carry chain as one addition of the ballot masks (generate and propagate exclude each other)
    const uint32_t G = __ballot_sync(0xffffffffu, carry != 0);
    const uint32_t P = __ballot_sync(0xffffffffu, all_ones == 0xffffffffu);
    const uint32_t A = G | P;
    uint32_t cin = (((A + G) ^ A ^ G) >> lane) & 1u;
#pragma unroll
    for (int w = 0; w < WPL; ++w) { // s += cin (a lane whose words are all ones wraps to zero: its carry out is in the chain)
        const uint32_t r = s[w] + cin;
        cin = (r < s[w]) ? 1u : 0u;
        s[w] = r;
    }
#pragma unroll
    for (int w = 0; w < WPL; ++w) {
        const uint32_t d0 = (s[w] ^ vp[w]) | x[w];
        hn[w] = vp[w] & d0;
        hp[w] = vn[w] | ~(vp[w] | d0);
        x[w] = d0; // keep d0
    }
    uint32_t hp_in = __shfl_up_sync(0xffffffffu, hp[WPL - 1], 1);
    uint32_t hn_in = __shfl_up_sync(0xffffffffu, hn[WPL - 1], 1);
    if (lane == 0) { hp_in = 0; hn_in = 0; }
#pragma unroll
    for (int w = 0; w < WPL; ++w) {
        const uint32_t hps = __funnelshift_l(w == 0 ? hp_in : hp[w - 1], hp[w], 1);
        const uint32_t hns = __funnelshift_l(w == 0 ? hn_in : hn[w - 1], hn[w], 1);
        vn[w] = hps & x[w];
        vp[w] = hns | ~(hps | x[w]);
    }
}

// needle `ci` over the symbols [ws, we) of the sequence that starts at store position `base`, from the initial state.
// LAST_ONLY: report the last column only (verification); else every column whose last symbol index lies in
// [own_lo, own_hi).  Reported ends are relative to `base`.
template <int WPL, bool LAST_ONLY>
__device__ __forceinline__ void myers_chain_run(const ChainArgs &a, uint32_t ci, uint64_t base, int64_t ws, int64_t we, int64_t own_lo,
                                                int64_t own_hi, uint32_t seq, uint32_t lane)
{
    constexpr uint32_t NWT = 32 * WPL;
    const MyersPatMeta pm = a.meta[ci];
    const uint32_t *eqp = a.eq + (size_t)ci * 4 * NWT + lane * WPL;
    const uint32_t top = (pm.m - 1) >> 5;                 // word that holds the score row
    const bool owner = (top / WPL) == lane;
    const uint32_t ow = top % WPL, ob = (pm.m - 1) & 31u;
    const int k = (int)pm.k;
    uint32_t vp[WPL], vn[WPL];
#pragma unroll
    for (int w = 0; w < WPL; ++w) { vp[w] = 0xffffffffu; vn[w] = 0; }
    int score = (int)pm.m;
    for (int64_t t = ws; t < we;) {
        const uint2 cw = spmb_fetch32(a.lanes, a.lay, base + (uint64_t)t); // the same 32 symbols in every lane (broadcast loads)
        unsigned long long W = ((unsigned long long)cw.x << 32) | cw.y;
        const int nb = (int)min((int64_t)32, we - t);
        for (int j = 0; j < nb; ++j) {
            const uint32_t c = (uint32_t)(W >> 62);
            W <<= 2;
            uint32_t eq[WPL], hp[WPL], hn[WPL];
#pragma unroll
            for (int w = 0; w < WPL; ++w) eq[w] = __ldg(eqp + c * NWT + w);
            myers_chain_step<WPL>(vp, vn, eq, lane, hp, hn);
            if (owner) {
                uint32_t hpw = hp[0], hnw = hn[0];
#pragma unroll
                for (int w = 1; w < WPL; ++w)
                    if ((uint32_t)w == ow) { hpw = hp[w]; hnw = hn[w]; }
                score += (int)((hpw >> ob) & 1u) - (int)((hnw >> ob) & 1u);
                const int64_t pos = t + j;
                if (!LAST_ONLY && score <= k && pos >= own_lo && pos < own_hi)
                    myers_emit(a.sink, (uint64_t)(pos + 1), seq, pm.pattern_id, (uint32_t)score);
            }
        }
        t += nb;
    }
    if (LAST_ONLY && owner && score <= k) myers_emit(a.sink, (uint64_t)we, seq, pm.pattern_id, (uint32_t)score);
}

template <int WPL, int MODE>
__global__ void __launch_bounds__(128) myers_chain_verify_kernel(ChainArgs a)
{
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicMax(a.counters + 2, a.counters[1]);
    const unsigned long long n_cand = min(a.counters[1], a.cand_cap);
    const uint32_t lane = threadIdx.x & 31;
    const unsigned long long warps = (unsigned long long)gridDim.x * (blockDim.x >> 5);
    for (unsigned long long i = (unsigned long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); i < n_cand; i += warps) {
        const spmb_hit cd = a.cand[i];
        const MyersPatMeta pm = a.meta[cd.pattern_id];
        // the one end this candidate stands for (header comment); everything below is warp-uniform
        const int64_t E = (int64_t)cd.end + (int64_t)(pm.m - a.prefix_m);
        uint64_t base = 0;
        int64_t limit, own_lo, own_hi;
        if (MODE == SPMB_MODE_PLAIN) {
            limit = (int64_t)a.lay.n;
            own_lo = (int64_t)a.own_lo; own_hi = (int64_t)a.own_hi;
        } else {
            const uint64_t R = cd.seq_id - a.seq_base;
            base = R * a.lay.row_len;
            limit = (int64_t)a.row_len[R];
            own_lo = a.row_own ? (int64_t)a.row_own[R] - 1 : 0;
            own_hi = limit;
        }
        if (E > limit || E - 1 < own_lo || E - 1 >= own_hi) continue;
        int64_t ws = E - (int64_t)pm.m - (int64_t)pm.k;
        if (ws < 0) ws = 0;
        myers_chain_run<WPL, true>(a, cd.pattern_id, base, ws, E, own_lo, own_hi, MODE == SPMB_MODE_PLAIN ? a.seq_id : cd.seq_id, lane);
    }
}

// whole rows of a PLAIN store, one warp per (row, needle): the lane-chained form as a searcher of its own
template <int WPL>
__global__ void __launch_bounds__(128) myers_chain_scan_kernel(ChainArgs a)
{
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t RL = a.lay.row_len;
    const uint64_t n_rows = (a.st1 - a.st0) * 32;
    const unsigned long long tasks = (unsigned long long)n_rows * a.n_needles;
    const unsigned long long warps = (unsigned long long)gridDim.x * (blockDim.x >> 5);
    for (unsigned long long i = (unsigned long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); i < tasks; i += warps) {
        const uint64_t R = a.st0 * 32 + i / a.n_needles;
        const uint32_t ci = (uint32_t)(i % a.n_needles);
        int64_t own_lo = (int64_t)(R * RL), own_hi = own_lo + (int64_t)RL;
        if (own_lo >= (int64_t)a.lay.n) continue;
        if (own_hi > (int64_t)a.lay.n) own_hi = (int64_t)a.lay.n;
        int64_t ws = own_lo - (int64_t)a.warm;
        if (ws < 0) ws = 0;
        myers_chain_run<WPL, false>(a, ci, 0, ws, own_hi, own_lo, own_hi, a.seq_id, lane);
    }
}
