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

#if defined(__x86_64__)
#include <immintrin.h>
#endif

namespace {

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
        _mm256_storeu_si256((__m256i *)(out + g), o);
    }
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
