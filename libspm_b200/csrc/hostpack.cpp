// hostpack.cpp -- host half of the store upload (spmb_store_create with a HOST haystack).
//
// A haystack arrives as 1 byte per symbol (a std::vector<spm::dna4> image,
// libspm/libspm/seqan/alphabet.hpp:34-83): over PCIe that is 4x the bytes the device keeps.  The
// upload therefore runs two movers over one queue of chunks (upload.cuh): the copy engine takes raw
// chunks from the head (1 B/symbol, packed by a kernel), host threads take chunks from the tail, pack
// them to 2 bits/symbol here and send a quarter of the bytes.  This file is the packing loop only:
// format conversion, no search (the library has no CPU search path).
//
// Output format = the store's "lane chunk" word (pack.cuh): one uint2 per 32 symbols, 2-bit codes,
// first symbol in the top bits of .x, symbols 16..31 in .y; in memory x then y (little endian).
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#include "../../include/spmb.h"

#if defined(__x86_64__)
#include <immintrin.h>
#endif

#include <stdlib.h>

namespace {

// SPMB_HOST_NT=1 turns non-temporal stores on in the packing loops.  Off by default: measured on the B200 box
// (profiles/r01_e2e_host_nt.txt) the e2e step is 3-4 ms SLOWER with them (43.2 / 44.4 vs 39.6 / 41.6 ms) -- the
// copy engine reads a freshly packed staging buffer out of the last-level cache, a streamed one out of DRAM.
inline bool nt_enabled()
{
    static const bool on = [] { const char *e = getenv("SPMB_HOST_NT"); return e && *e == '1'; }();
    return on;
}

// 8 symbols (one per byte, first symbol in the LOWEST byte) -> 16 bits, first symbol on top
inline uint32_t pack8_swar(uint64_t q)
{
    q = __builtin_bswap64(q) & 0x0303030303030303ull;     // first symbol now in the top byte
    q = (q | (q >> 6)) & 0x000f000f000f000full;
    q = (q | (q >> 12)) & 0x000000ff000000ffull;
    q = (q | (q >> 24)) & 0xffffull;
    return (uint32_t)q;
}

inline uint64_t pack32_swar(const uint8_t *p)
{
    uint64_t a, b, c, d;
    memcpy(&a, p, 8); memcpy(&b, p + 8, 8); memcpy(&c, p + 16, 8); memcpy(&d, p + 24, 8);
    const uint64_t x = ((uint64_t)pack8_swar(a) << 16) | pack8_swar(b);
    const uint64_t y = ((uint64_t)pack8_swar(c) << 16) | pack8_swar(d);
    return x | (y << 32);
}

void pack_scalar(const uint8_t *in, uint64_t n_groups, uint64_t *out)
{
    for (uint64_t g = 0; g < n_groups; ++g) out[g] = pack32_swar(in + 32 * g);
}

#if defined(__x86_64__)
// 32 symbols per iteration: two multiply-adds fold 4 bytes into one (b0<<6|b1<<4|b2<<2|b3), a byte
// shuffle puts the four folded bytes of each 128-bit half into one dword (first symbol on top).
__attribute__((target("avx2"))) void pack_avx2(const uint8_t *in, uint64_t n_groups, uint64_t *out)
{
    const __m256i m3 = _mm256_set1_epi8(3);
    const __m256i k1 = _mm256_set1_epi16(0x0104);           // bytes [4, 1]: b0*4 + b1
    const __m256i k2 = _mm256_set1_epi32(0x00010010);       // words [16, 1]: t0*16 + t1
    const __m256i sh = _mm256_setr_epi8(12, 8, 4, 0, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1,
                                        12, 8, 4, 0, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1);
    const __m256i gather = _mm256_setr_epi32(0, 4, 0, 4, 0, 4, 0, 4);
    // optional non-temporal stores (see nt_enabled)
    const bool nt = ((uintptr_t)out & 31u) == 0 && nt_enabled();
    uint64_t g = 0;
    for (; g + 4 <= n_groups; g += 4) {
        __m256i r[4];
        for (int i = 0; i < 4; ++i) {
            __m256i v = _mm256_and_si256(_mm256_loadu_si256((const __m256i *)(in + 32 * (g + i))), m3);
            v = _mm256_maddubs_epi16(v, k1);
            v = _mm256_madd_epi16(v, k2);
            v = _mm256_shuffle_epi8(v, sh);                  // dword 0 = x, dword 4 = y
            r[i] = _mm256_permutevar8x32_epi32(v, gather);   // {x,y} repeated in every qword
        }
        __m256i o = _mm256_blend_epi32(_mm256_blend_epi32(r[0], r[1], 0x0c), _mm256_blend_epi32(r[2], r[3], 0xc0), 0xf0);
        if (nt) _mm256_stream_si256((__m256i *)(out + g), o);
        else _mm256_storeu_si256((__m256i *)(out + g), o);
    }
    if (nt) _mm_sfence();
    for (; g < n_groups; ++g) out[g] = pack32_swar(in + 32 * g);
}
#endif

} // namespace

// ranks[0, n) -> out[0, ceil(n/32)); symbols past n are zero.  Values are masked to 2 bits like every
// other entry point of the C ABI.
extern "C" void spmb_host_pack_chunks(const uint8_t *ranks, uint64_t n, uint64_t *out)
{
    const uint64_t full = n / 32;
#if defined(__x86_64__)
    static const bool avx2 = __builtin_cpu_supports("avx2");
    if (avx2) pack_avx2(ranks, full, out);
    else
#endif
        pack_scalar(ranks, full, out);
    if (n % 32) {
        uint8_t tmp[32] = {0};
        memcpy(tmp, ranks + 32 * full, n % 32);
        out[full] = pack32_swar(tmp);
    }
}

// ---- journal event table: 24-byte events -> 8-byte records (the short form of upload.cuh: per block of 1024
// events a header {base_pos, base_off}, per event {dpos:32, doff:16, del_len:8, ins_len:8} = one 64-bit word).
// false = some event does not fit the short form (the chunk then travels raw).  Host threads run this beside
// the haystack packing loop above; the scalar form is CPU-bound at ~5 ns per event, the AVX2 form (4 events per
// iteration: three 32-byte loads, lane blends + permutes to {pos x4, len x4, off x4}) at ~2 ns.
namespace {

enum { EV_BLOCK_HOST = 1024 };

inline uint64_t pack_events_tail(const uint64_t *w, uint32_t i, uint32_t cnt, uint64_t base_pos, uint64_t base_off,
                                 uint64_t pool_len, uint64_t *rec)
{
    uint64_t b = 0;
    for (; i < cnt; ++i) {
        const uint64_t p = w[3 * (size_t)i], l = w[3 * (size_t)i + 1], o = w[3 * (size_t)i + 2];
        const uint64_t dp = p - base_pos;
        const bool z = (l >> 32) == 0;                      // inserts nothing: the offset is only range-checked
        const uint64_t dof = z ? 0 : o - base_off;
        b |= (uint64_t)(z && o > pool_len);
        b |= (dp >> 32) | (dof >> 16) | (l & 0xffffff00ffffff00ull);
        rec[i] = (dp & 0xffffffffull) | (dof << 32) | ((l & 0xff) << 48) | ((l >> 32) << 56);
    }
    return b;
}

#if defined(__x86_64__)
__attribute__((target("avx2"))) uint64_t pack_events_block_avx2(const uint64_t *w, uint32_t cnt, uint64_t base_pos,
                                                                uint64_t base_off, uint64_t pool_len, uint64_t *rec)
{
    const __m256i bp = _mm256_set1_epi64x((long long)base_pos), bo = _mm256_set1_epi64x((long long)base_off);
    const __m256i pl = _mm256_set1_epi64x((long long)pool_len), zero = _mm256_setzero_si256();
    const __m256i lo32 = _mm256_set1_epi64x(0xffffffffll), ff = _mm256_set1_epi64x(0xff);
    const __m256i lenbad = _mm256_set1_epi64x((long long)0xffffff00ffffff00ull);
    __m256i bad = zero;
    const bool nt = ((uintptr_t)rec & 15u) == 0 && nt_enabled(); // staging buffer: non-temporal stores (see pack_avx2)
    uint32_t i = 0;
    for (; i + 4 <= cnt; i += 4) {
        // 64-bit lanes: v0 = [p0 l0 o0 p1], v1 = [l1 o1 p2 l2], v2 = [o2 p3 l3 o3]
        const __m256i v0 = _mm256_loadu_si256((const __m256i *)(w + 3 * (size_t)i));
        const __m256i v1 = _mm256_loadu_si256((const __m256i *)(w + 3 * (size_t)i + 4));
        const __m256i v2 = _mm256_loadu_si256((const __m256i *)(w + 3 * (size_t)i + 8));
        __m256i P = _mm256_blend_epi32(_mm256_blend_epi32(v0, v2, 0x0c), v1, 0x30); // [p0 p3 p2 p1]
        P = _mm256_permute4x64_epi64(P, 0x6c);
        __m256i L = _mm256_blend_epi32(_mm256_blend_epi32(v1, v0, 0x0c), v2, 0x30); // [l1 l0 l3 l2]
        L = _mm256_permute4x64_epi64(L, 0xb1);
        __m256i O = _mm256_blend_epi32(_mm256_blend_epi32(v2, v1, 0x0c), v0, 0x30); // [o2 o1 o0 o3]
        O = _mm256_permute4x64_epi64(O, 0xc6);
        const __m256i dp = _mm256_sub_epi64(P, bp);
        __m256i dof = _mm256_sub_epi64(O, bo);
        const __m256i ins = _mm256_srli_epi64(L, 32);
        const __m256i z = _mm256_cmpeq_epi64(ins, zero);
        // off > pool_len, unsigned: signed compare, or the top bit of off set
        const __m256i over = _mm256_or_si256(_mm256_cmpgt_epi64(O, pl), _mm256_srai_epi32(_mm256_shuffle_epi32(O, 0xf5), 31));
        bad = _mm256_or_si256(bad, _mm256_and_si256(z, over));
        dof = _mm256_andnot_si256(z, dof);
        bad = _mm256_or_si256(bad, _mm256_or_si256(_mm256_srli_epi64(dp, 32),
                                                   _mm256_or_si256(_mm256_srli_epi64(dof, 16), _mm256_and_si256(L, lenbad))));
        __m256i r = _mm256_or_si256(_mm256_and_si256(dp, lo32), _mm256_slli_epi64(dof, 32));
        r = _mm256_or_si256(r, _mm256_or_si256(_mm256_slli_epi64(_mm256_and_si256(L, ff), 48), _mm256_slli_epi64(ins, 56)));
        if (nt) {
            _mm_stream_si128((__m128i *)(rec + i), _mm256_castsi256_si128(r));
            _mm_stream_si128((__m128i *)(rec + i + 2), _mm256_extracti128_si256(r, 1));
        } else {
            _mm256_storeu_si256((__m256i *)(rec + i), r);
        }
    }
    if (nt) _mm_sfence();
    alignas(32) uint64_t t[4];
    _mm256_store_si256((__m256i *)t, bad);
    return t[0] | t[1] | t[2] | t[3] | pack_events_tail(w, i, cnt, base_pos, base_off, pool_len, rec);
}
#endif

} // namespace

static_assert(sizeof(spmb_journal_event) == 24, "event layout of the C ABI");

extern "C" int spmb_host_pack_events(const spmb_journal_event *ev, uint32_t n, uint64_t pool_len, uint8_t *out)
{
    const uint32_t n_blk = (n + EV_BLOCK_HOST - 1) / EV_BLOCK_HOST;
    uint64_t *hdr = reinterpret_cast<uint64_t *>(out);                       // {base_pos, base_off} per block
    uint64_t *rec = reinterpret_cast<uint64_t *>(out + (size_t)n_blk * 16);
    const uint64_t *w = reinterpret_cast<const uint64_t *>(ev);
#if defined(__x86_64__)
    static const bool avx2 = __builtin_cpu_supports("avx2");
#endif
    uint64_t bad = 0;
    for (uint32_t b = 0; b < n_blk; ++b) {
        const uint32_t i0 = b * EV_BLOCK_HOST, i1 = (i0 + EV_BLOCK_HOST < n) ? i0 + EV_BLOCK_HOST : n;
        const uint64_t base_pos = ev[i0].ref_pos;
        uint64_t base_off = 0;
        for (uint32_t i = i0; i < i1; ++i)
            if (ev[i].ins_len) { base_off = ev[i].ins_off; break; }
        if (base_off > pool_len) return 0;
        hdr[2 * (size_t)b] = base_pos;
        hdr[2 * (size_t)b + 1] = base_off;
#if defined(__x86_64__)
        if (avx2) bad |= pack_events_block_avx2(w + 3 * (size_t)i0, i1 - i0, base_pos, base_off, pool_len, rec + i0);
        else
#endif
            bad |= pack_events_tail(w + 3 * (size_t)i0, 0, i1 - i0, base_pos, base_off, pool_len, rec + i0);
    }
    return bad == 0;
}
