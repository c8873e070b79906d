// layout.cuh -- geometry of the packed sequence store ("strided bit planes").
//
// A sequence of n symbols (2 bits each) is cut into rows of row_len (RL) symbols; 32 consecutive
// rows form a super-tile.  Column `col` of super-tile `st` is ONE pair of 32-bit words {H, L}:
// bit r of H / L is the high / low bit of the symbol at position (st*32 + r)*RL + col.
// So the 32 rows of a super-tile advance in lock step, one bit lane each:
//   * the exact kernel compares 32 alignments per LOP3 and never shifts (text position p+j is
//     simply column col+j, same bit),
//   * a Myers lane owns one row and reads its symbols column by column, all 32 lanes of a warp
//     reading the same words (broadcast).
// Columns are numbered globally g = st*RL + col and stored permuted inside blocks of 1024
// columns so that a thread owning 32 consecutive columns ("thread-row") finds its k-th 16-byte
// chunk next to its neighbour lanes' k-th chunks: a warp-wide LDG.128 is one contiguous 512 B.
//
// 2 bits/symbol => 0.25 B per searched bp, the algorithmic figure of SURVEY.md 8(d).
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define SPMB_HD __host__ __device__ __forceinline__
#else
#define SPMB_HD inline
#endif

struct SpmbLayout {
    uint64_t n;            // symbols
    uint64_t n_st;         // super-tiles
    uint64_t n_cols;       // n_st * row_len        (multiple of 32)
    uint64_t n_cols_alloc; // allocated columns     (multiple of 1024, >= n_cols + row_len + 1024)
    uint64_t rl_magic;     // ceil(2^64 / row_len)  (exact division for x < 2^42)
    uint32_t row_len;      // RL, multiple of 32
    uint32_t mode;         // 0 = plain text (row r+1 continues row r), 1 = windows (independent rows)
};

enum { SPMB_MODE_PLAIN = 0, SPMB_MODE_WINDOWS = 1 };

// index (in uint2 {H,L} units) of global column g
SPMB_HD uint64_t spmb_perm(uint64_t g)
{
    return (g & ~1023ull) | (((g & 31ull) >> 1) << 6) | (((g >> 5) & 31ull) << 1) | (g & 1ull);
}

// x / row_len for x < 2^42
SPMB_HD uint64_t spmb_div_rl(const SpmbLayout &l, uint64_t x)
{
#ifdef __CUDA_ARCH__
    return __umul64hi(x, l.rl_magic);
#else
    return (uint64_t)(((unsigned __int128)x * l.rl_magic) >> 64);
#endif
}

static inline uint32_t spmb_auto_row_len(uint64_t n)
{
    // aim at ~6 Myers warps per SM sub-partition on a 148-SM part (148*4*32*6 rows); keep the
    // Myers warm-up (m+k-1 symbols per row) a small fraction of a row.
    const uint64_t target_rows = 148ull * 4 * 32 * 6;
    uint64_t rl = (n + target_rows - 1) / target_rows;
    rl = (rl + 31) / 32 * 32;
    if (rl < 1024) {
        // short texts: fewer, longer rows -- but at least one full super-tile
        uint64_t one_tile = ((n + 31) / 32 + 31) / 32 * 32;
        rl = one_tile < 1024 ? one_tile : 1024;
        if (rl < 32) rl = 32;
    }
    if (rl > 8192) rl = 8192;
    return (uint32_t)rl;
}

static inline SpmbLayout spmb_make_layout(uint64_t n, uint32_t row_len, uint32_t mode)
{
    SpmbLayout l;
    l.n = n;
    l.row_len = row_len;
    l.mode = mode;
    uint64_t per_st = 32ull * row_len;
    l.n_st = (n + per_st - 1) / per_st;
    if (l.n_st == 0) l.n_st = 1;
    l.n_cols = l.n_st * row_len;
    uint64_t pad = ((uint64_t)row_len + 1023) / 1024 * 1024 + 1024;
    l.n_cols_alloc = (l.n_cols + 1023) / 1024 * 1024 + pad;
    l.rl_magic = (uint64_t)((((unsigned __int128)1) << 64) / row_len) + 1;
    return l;
}
