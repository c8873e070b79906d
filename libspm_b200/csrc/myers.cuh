// myers.cuh -- Myers bit-vector k-error search kernel (SPMB_ALG_MYERS).
//
// Replaces the inner loops behind spm::myers_matcher (libspm/libspm/matcher/myers_matcher.hpp:19-60;
// SeqAn2's _findMyersSmallPatterns/_findMyersLargePatterns reached through
// seqan_pattern_base.hpp:60-65).  Semantics kept: EVERY end position whose best alignment has
// <= k errors is reported, end exclusive, plus the error count.
//
// Parallelisation (the reference is a single serial scan):
//   * one LANE owns one row of the packed store (row_len consecutive end positions) and runs
//     the full column recurrence over it; all NW = ceil(m/32) words of VP/VN live in that lane's
//     registers, the add carry runs through IADD3.X and the two shift carries through funnel
//     shifts -- no cross-lane traffic at all (the shuffle-chained variant of north_star costs
//     3 shuffles per step on the critical path; DESIGN.md 5 discusses the choice).
//   * a lane warms up on the window_size-1 = m+k-1 symbols before its row from the initial
//     state; every hit (<= k errors) it then reports is exact.
//   * the 32 lanes of a warp own the 32 rows of one super-tile and read the same {H,L} column
//     words (one broadcast LDG.128 per two columns, L1-resident).
//   * the pattern is padded with wildcard rows up to 32*NW, so that the score row is bit 31 of
//     the top word; a hit ending at j then shows at column j + pad with the same score
//     (horizontal deltas are in {-1,0,1}).  The +-1 score deltas are shifted into two 32-bit
//     accumulators (2 SHF per column) and only inspected every 32 columns.
//   * Eq[c] comes from a 64..128-byte shared-memory table with one LDS.128 per column.
// Algorithmic ALU work per (bp x pattern): 10 per 32-bit word + 4 = the figure of SURVEY.md 8(d).
#pragma once
#include "layout.cuh"
#include "hits.cuh"
#include "addchain.cuh"

struct MyersPatMeta {
    uint32_t m, k, pad, pattern_id;
};

struct MyersArgs {
    const uint2 *lanes;       // lane-chunk layout of the store (pack.cuh lanes_kernel)
    const uint2 *lanes_hi;    // 4-bit alphabets: lane chunks of bits 3..2 of every symbol (else null)
    SpmbLayout lay;
    const uint32_t *eq;       // [n_pat][4][NWP]  (4-bit alphabets: [n_pat][16][NWP])
    const MyersPatMeta *meta; // [n_pat]
    const uint32_t *row_len;  // windows mode
    const uint32_t *row_seq;  // windows mode
    const uint32_t *row_own;  // windows mode: first owned end index (local), may be null
    uint32_t warm;            // plain mode: warm-up columns, multiple of 32, >= max(m+k-1)
    uint32_t seq_id;
    uint64_t st0, st1;        // super-tile range of this launch
    // run-time multipliers (4 and the byte size of one Eq row): ptxas cannot turn a multiply by an
    // unknown into SHF/LEA, so the symbol pop and the table address are IMADs on the FMA pipe,
    // which runs in the shadow of the ALU pipe the recurrence saturates (tools/ubench_pipes.cu)
    uint32_t mul4, eq_stride;
    uint32_t split;           // myers_cut_kernel, plain mode: pieces per row (>= 1)
    HitSink sink;
};

template <int NW>
__device__ __forceinline__ void myers_step(uint32_t (&vp)[NW], uint32_t (&vn)[NW], const uint32_t (&eq)[NW],
                                           uint32_t &hp_acc, uint32_t &hn_acc)
{
    uint32_t x[NW], t[NW], s[NW];
#pragma unroll
    for (int i = 0; i < NW; ++i) {
        x[i] = eq[i] | vn[i];
        t[i] = x[i] & vp[i];
    }
    spmb_add_chain<NW>(s, t, vp);
    uint32_t hp_lo = 0, hn_lo = 0;
#pragma unroll
    for (int i = 0; i < NW; ++i) {
        uint32_t d0 = (s[i] ^ vp[i]) | x[i];
        uint32_t hn = vp[i] & d0;
        uint32_t hp = vn[i] | ~(vp[i] | d0);
        uint32_t hps = (i == 0) ? (hp << 1) : __funnelshift_l(hp_lo, hp, 1);
        uint32_t hns = (i == 0) ? (hn << 1) : __funnelshift_l(hn_lo, hn, 1);
        hp_lo = hp;
        hn_lo = hn;
        vn[i] = hps & d0;
        vp[i] = hns | ~(hps | d0);
    }
    hp_acc = __funnelshift_l(hp_lo, hp_acc, 1); // (acc << 1) | bit31(HP_top)
    hn_acc = __funnelshift_l(hn_lo, hn_acc, 1);
}

__device__ __forceinline__ void myers_emit(const HitSink &sink, uint64_t end, uint32_t seq, uint32_t pid, uint32_t err)
{
    const unsigned long long idx = spmb_append_slot(sink.count);
    if (idx < sink.cap) {
        spmb_hit h;
        h.end = end; h.seq_id = seq; h.pattern_id = pid; h.errors = err; h.reserved = 0;
        sink.hits[idx] = h;
    }
}

template <int NW, int MODE, bool WIDE = false>
__global__ void __launch_bounds__(128) myers_kernel(MyersArgs a)
{
    constexpr int NWP = (NW <= 4) ? 4 : 8;
    constexpr int NSYM = WIDE ? 16 : 4;
    __shared__ __align__(16) uint32_t s_eq[NSYM * NWP];
    const uint32_t pat = blockIdx.y;
    for (uint32_t t = threadIdx.x; t < NSYM * NWP; t += blockDim.x) s_eq[t] = a.eq[(size_t)pat * NSYM * NWP + t];
    __syncthreads();
    const MyersPatMeta pm = a.meta[pat];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint64_t st = a.st0 + (uint64_t)blockIdx.x * 4 + warp;
    if (st >= a.st1) return;
    const uint32_t RL = a.lay.row_len;
    const uint64_t R = st * 32 + lane;
    if (R * RL >= a.lay.n) return; // no warp-synchronous operation below this line

    const int k = (int)pm.k;
    const int W = (int)pm.pad;

    uint32_t vp[NW], vn[NW];
#pragma unroll
    for (int i = 0; i < NW; ++i) { vp[i] = 0xffffffffu; vn[i] = 0; }
    int score = 32 * NW;

    // positions are absolute (plain) or row-local (windows)
    int64_t p, end_p, own_lo, own_hi;
    int64_t cur_row;
    uint32_t cur_col;
    uint32_t seq = a.seq_id;
    if (MODE == SPMB_MODE_PLAIN) {
        own_lo = (int64_t)(R * RL);
        own_hi = own_lo + RL;
        if (own_hi > (int64_t)a.lay.n) own_hi = (int64_t)a.lay.n;
        const uint32_t rows_back = (a.warm + RL - 1) / RL;
        p = own_lo - (int64_t)a.warm;
        cur_row = (int64_t)R - rows_back;
        cur_col = rows_back * RL - a.warm;
        if (p < 0) { p = 0; cur_row = 0; cur_col = 0; }
        end_p = own_lo + RL + (W ? 32 : 0);
    } else {
        own_lo = a.row_own ? (int64_t)a.row_own[R] - 1 : 0; // row_own is an exclusive-end coordinate; j is the last symbol
        own_hi = (int64_t)a.row_len[R];
        if (a.row_seq) seq = a.row_seq[R];
        p = 0;
        cur_row = (int64_t)R;
        cur_col = 0;
        // stop as soon as this row's last end (+pad) has been seen, rounded up to a block
        end_p = ((own_hi + W + 31) / 32) * 32;
        if (end_p > (int64_t)RL) end_p = RL;
    }

    const uint32_t rl32 = RL >> 5;
    const uint32_t mul4 = a.mul4, eq_stride = a.eq_stride;
    const uint2 *__restrict__ lp = a.lanes;
    const uint2 *__restrict__ lph = a.lanes_hi;
    // one uint2 = this row's next 32 symbols; fetched one whole block (32 columns) ahead
    uint64_t li = ((uint64_t)(cur_row >> 5) * rl32 + (cur_col >> 5)) * 32 + ((uint32_t)cur_row & 31u);
    uint2 nxt = __ldg(lp + li), nxt_hi = WIDE ? __ldg(lph + li) : make_uint2(0, 0);
    while (p < end_p) {
        const uint2 cur = nxt, cur_hi = nxt_hi;
        cur_col += 32;
        if (cur_col == RL) { cur_col = 0; ++cur_row; }
        if (p + 32 < end_p) {
            li = ((uint64_t)(cur_row >> 5) * rl32 + (cur_col >> 5)) * 32 + ((uint32_t)cur_row & 31u);
            nxt = __ldg(lp + li);
            if (WIDE) nxt_hi = __ldg(lph + li);
        }
        uint32_t hp_acc = 0, hn_acc = 0;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            uint32_t w = half ? cur.y : cur.x;
            uint32_t wh = half ? cur_hi.y : cur_hi.x;
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                uint32_t c = w >> 30;       // the one ALU instruction of the symbol fetch
                w *= mul4;                  // IMAD
                if (WIDE) { c = (wh >> 30) * mul4 + c; wh *= mul4; }
                uint32_t eq[NW];
                const uint4 *e4 = reinterpret_cast<const uint4 *>(reinterpret_cast<const char *>(s_eq) + c * eq_stride);
                const uint4 e0 = e4[0];
                eq[0] = e0.x;
                if (NW > 1) eq[1 % NW] = e0.y;
                if (NW > 2) eq[2 % NW] = e0.z;
                if (NW > 3) eq[3 % NW] = e0.w;
                if (NW > 4) {
                    const uint4 e1 = e4[NWP / 4 - 1];
                    eq[4 % NW] = e1.x;
                    if (NW > 5) eq[5 % NW] = e1.y;
                    if (NW > 6) eq[6 % NW] = e1.z;
                    if (NW > 7) eq[7 % NW] = e1.w;
                }
                myers_step<NW>(vp, vn, eq, hp_acc, hn_acc);
            }
        }
        // ---- every 32 columns: can the score have dipped to <= k anywhere in this block? ----
        if (score - (int)__popc(hn_acc) <= k) {
            int s = score;
            for (int t = 0; t < 32; ++t) {
                const int b = 31 - t;
                s += (int)((hp_acc >> b) & 1u) - (int)((hn_acc >> b) & 1u);
                if (s <= k) {
                    const int64_t j = p + t - W; // index of the last symbol of the occurrence
                    if (j >= own_lo && j < own_hi) myers_emit(a.sink, (uint64_t)(j + 1), seq, pm.pattern_id, (uint32_t)s);
                }
            }
        }
        score += (int)__popc(hp_acc) - (int)__popc(hn_acc);
        p += 32;
    }
}

// ---------------- INT32 ALU-pipe peak probe (roofline denominator for the kernel above) ----------------
template <int OP>
__global__ void __launch_bounds__(256) int_peak_kernel(uint32_t *out, int iters, uint32_t k1)
{
    uint32_t a[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = threadIdx.x * 2654435761u + j * 40503u + blockIdx.x;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (OP == 0) // exactly one LOP3 (asm: the compiler must not merge rounds algebraically)
                    asm volatile("lop3.b32 %0, %0, %1, %2, 0x6a;" : "+r"(a[j]) : "r"(k1), "r"(a[(j + 1) & 7]));
                else         // exactly one IADD3
                    asm volatile("{ .reg .u32 t; add.u32 t, %0, %1; add.u32 %0, t, %2; }" : "+r"(a[j]) : "r"(a[(j + 1) & 7]), "r"(k1));
            }
        }
    }
    uint32_t r = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) r ^= a[j];
    if (r == 0x12345678u) out[0] = r; // practically never; keeps the chains alive
}
