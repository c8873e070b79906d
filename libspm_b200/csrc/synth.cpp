// synth.cpp -- the one synthetic-input generator (include/spmb_synth.h): xoshiro256** seeded via splitmix64,
// counter-based per (seed, stream, block).  Harness code: built as libspmsynth.so, no CUDA, not linked into
// libspmb.so.  Compile with -ffp-contract=off: the few double operations below must round the same everywhere.
#include "../../include/spmb_synth.h"

#include <cstring>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif
#ifdef __BMI2__
#include <immintrin.h>
#endif

namespace {

inline uint64_t splitmix64(uint64_t &x)
{
    x += 0x9E3779B97F4A7C15ull;
    uint64_t z = x;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

inline uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }

struct Xoshiro {
    uint64_t s[4];
    Xoshiro(uint64_t seed, uint64_t stream, uint64_t block)
    {
        uint64_t x = seed;
        x = splitmix64(x) ^ (stream * 0xD1342543DE82EF95ull);
        x = splitmix64(x) ^ (block * 0xAF251AF3B0F025B5ull);
        for (int i = 0; i < 4; ++i) s[i] = splitmix64(x);
    }
    inline uint64_t next()
    {
        const uint64_t r = rotl(s[1] * 5, 7) * 9;
        const uint64_t t = s[1] << 17;
        s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3];
        s[2] ^= t;
        s[3] = rotl(s[3], 45);
        return r;
    }
    // uniform in [0, range): multiply-high (bias < range / 2^64)
    inline uint64_t below(uint64_t range) { return (uint64_t)(((unsigned __int128)next() * range) >> 64); }
    inline double unit() { return (double)(next() >> 11) * (1.0 / 9007199254740992.0); }
};

int g_threads = 0;

// log2(x), x >= 1, by repeated squaring: IEEE multiplications only
double det_log2(double x)
{
    double r = 0.0;
    while (x >= 2.0) { x *= 0.5; r += 1.0; }
    double b = 0.5;
    for (int i = 0; i < 52; ++i) {
        x = x * x;
        if (x >= 2.0) { r += b; x *= 0.5; }
        b *= 0.5;
    }
    return r;
}

// 2^f, 0 <= f < 1: Taylor series of exp(f ln 2)
double det_exp2_frac(double f)
{
    const double t = f * 0.6931471805599453094;
    double term = 1.0, sum = 1.0;
    for (int i = 1; i <= 22; ++i) { term = term * t / (double)i; sum += term; }
    return sum;
}

const uint64_t TEXT_BLOCK = 65536; // symbols per independently seeded block (2048 draws of 32 symbols)

inline void spread32(uint64_t x, uint8_t *out)
{
#ifdef __BMI2__
    for (int q = 0; q < 4; ++q) {
        const uint64_t b = _pdep_u64((x >> (16 * q)) & 0xffffu, 0x0303030303030303ull);
        memcpy(out + 8 * q, &b, 8);
    }
#else
    for (int j = 0; j < 32; ++j) out[j] = (uint8_t)((x >> (2 * j)) & 3u);
#endif
}

void text_block(uint64_t seed, uint64_t blk, uint64_t lo, uint64_t hi, uint8_t *out) // symbols [lo, hi) of block blk -> out
{
    Xoshiro g(seed, SPMS_STREAM_TEXT, blk);
    uint8_t tmp[32];
    uint64_t p = 0;
    for (; p + 32 <= lo; p += 32) g.next();
    while (p < hi) {
        const uint64_t x = g.next();
        if (p >= lo && p + 32 <= hi) spread32(x, out + (p - lo));
        else {
            spread32(x, tmp);
            for (uint64_t j = 0; j < 32; ++j)
                if (p + j >= lo && p + j < hi) out[p + j - lo] = tmp[j];
        }
        p += 32;
    }
}

} // namespace

extern "C" void spms_set_threads(int n) { g_threads = n; }

extern "C" uint64_t spms_u64(uint64_t seed, uint64_t stream, uint64_t index)
{
    Xoshiro g(seed, stream, index);
    return g.next();
}

extern "C" void spms_uniform(uint64_t seed, uint64_t stream, uint64_t index0, uint64_t n, uint64_t lo, uint64_t hi, uint64_t *out)
{
    const uint64_t range = hi > lo ? hi - lo : 1;
    for (uint64_t i = 0; i < n; ++i) {
        Xoshiro g(seed, stream, index0 + i);
        out[i] = lo + g.below(range);
    }
}

extern "C" void spms_text(uint64_t seed, uint64_t begin, uint64_t n, uint8_t *out)
{
    if (n == 0) return;
    const uint64_t b0 = begin / TEXT_BLOCK, b1 = (begin + n - 1) / TEXT_BLOCK;
#ifdef _OPENMP
    const int nt = g_threads > 0 ? g_threads : omp_get_num_procs();
#pragma omp parallel for schedule(dynamic, 16) num_threads(nt)
#endif
    for (long long b = (long long)b0; b <= (long long)b1; ++b) {
        const uint64_t base = (uint64_t)b * TEXT_BLOCK;
        const uint64_t lo = begin > base ? begin - base : 0;
        const uint64_t hi = (begin + n < base + TEXT_BLOCK) ? begin + n - base : TEXT_BLOCK;
        text_block(seed, (uint64_t)b, lo, hi, out + (base + lo - begin));
    }
}

extern "C" uint32_t spms_mutate(uint64_t seed, uint64_t stream, const uint8_t *pat, uint32_t m, uint32_t e, uint8_t *out)
{
    Xoshiro g(seed, stream, 0);
    std::vector<uint8_t> p(pat, pat + m);
    for (uint32_t t = 0; t < e && !p.empty(); ++t) {
        const uint64_t op = g.below(3);
        const size_t i = (size_t)g.below(p.size());
        if (op == 0) p[i] = (uint8_t)((p[i] + 1 + g.below(3)) & 3u);
        else if (op == 1) p.insert(p.begin() + i, (uint8_t)g.below(4));
        else p.erase(p.begin() + i);
    }
    if (!p.empty()) memcpy(out, p.data(), p.size());
    return (uint32_t)p.size();
}

namespace {
// shared by the journal and the JST: one variant stream
struct VariantWalk {
    Xoshiro g, gp;
    uint64_t pos = 0, lo, span, pool_len = 0;
    VariantWalk(uint64_t seed, uint64_t stream, double mean_gap, uint32_t max_indel)
        : g(seed, stream, 0), gp(seed, SPMS_STREAM_POOL, stream)
    {
        lo = (uint64_t)max_indel + 2;
        uint64_t hi = (uint64_t)(2.0 * mean_gap);
        hi = hi > 2 * lo ? hi - lo : lo + 1;   // gaps uniform in [lo, hi], mean = mean_gap
        span = hi - lo + 1;
    }
    // next variant; false when the reference is exhausted
    bool step(const uint8_t *ref, uint64_t n_ref, double p_snp, double p_ins, uint32_t max_indel, spms_event *e, uint8_t *kind,
              uint8_t *pool, uint64_t cap_pool)
    {
        pos += lo + g.below(span);
        if (pos + 2ull * max_indel + 2 >= n_ref) return false;
        const double u = g.unit();
        const uint32_t len = 1 + (uint32_t)g.below(max_indel);
        const uint8_t k = u < p_snp ? 0 : (u < p_snp + p_ins ? 1 : 2);
        e->ref_pos = pos;
        e->ins_off = pool_len;
        if (k == 0) {
            e->del_len = 1; e->ins_len = 1;
            const uint8_t alt = (uint8_t)((ref[pos] + 1 + gp.below(3)) & 3u);
            if (pool && pool_len < cap_pool) pool[pool_len] = alt;
            pool_len += 1;
        } else if (k == 1) {
            e->del_len = 0; e->ins_len = len;
            for (uint32_t j = 0; j < len; ++j) {
                const uint8_t s = (uint8_t)gp.below(4);
                if (pool && pool_len + j < cap_pool) pool[pool_len + j] = s;
            }
            pool_len += len;
        } else {
            e->del_len = len; e->ins_len = 0;
        }
        *kind = k;
        return true;
    }
};
} // namespace

extern "C" uint64_t spms_journal(uint64_t seed, const uint8_t *ref, uint64_t n_ref, double density, double p_snp, double p_ins,
                                 uint32_t max_indel, spms_event *ev, uint64_t cap_ev, uint8_t *pool, uint64_t cap_pool,
                                 uint64_t *pool_len, uint8_t *kinds)
{
    VariantWalk w(seed, SPMS_STREAM_JOURNAL, 1.0 / density, max_indel);
    uint64_t n = 0;
    spms_event e;
    uint8_t k;
    while (w.step(ref, n_ref, p_snp, p_ins, max_indel, &e, &k, ev ? pool : nullptr, cap_pool)) {
        if (ev) {
            if (n >= cap_ev) break;
            ev[n] = e;
            if (kinds) kinds[n] = k;
        }
        ++n;
    }
    if (pool_len) *pool_len = w.pool_len;
    return n;
}

extern "C" uint64_t spms_jst(uint64_t seed, const uint8_t *ref, uint64_t n_ref, uint64_t n_var, uint32_t n_hap, double p_snp,
                             double p_ins, uint32_t max_indel, double af_max, spms_event *var, uint8_t *pool, uint64_t cap_pool,
                             uint64_t *pool_len, uint8_t *kinds, uint64_t *cov)
{
    VariantWalk w(seed, SPMS_STREAM_JST, (double)n_ref / (double)(n_var + 1), max_indel);
    uint64_t n = 0;
    spms_event e;
    uint8_t k;
    while (n < n_var && w.step(ref, n_ref, p_snp, p_ins, max_indel, &e, &k, pool, cap_pool)) {
        if (var) var[n] = e;
        if (kinds) kinds[n] = k;
        ++n;
    }
    if (pool_len) *pool_len = w.pool_len;
    if (!cov) return n;
    const uint32_t wpr = (n_hap + 63) / 64;
    const double L = det_log2(af_max * (double)n_hap); // AF = 2^(u L) / n_hap: density ~ 1/x on [1/n_hap, af_max]
#ifdef _OPENMP
    const int nt = g_threads > 0 ? g_threads : omp_get_num_procs();
#pragma omp parallel for schedule(static, 1024) num_threads(nt)
#endif
    for (long long v = 0; v < (long long)n; ++v) {
        Xoshiro g(seed, SPMS_STREAM_COVERAGE, (uint64_t)v);
        const double y = g.unit() * L;
        const double yi = (double)(uint64_t)y;
        double af = det_exp2_frac(y - yi);
        for (uint64_t i = 0; i < (uint64_t)yi; ++i) af *= 2.0;
        af /= (double)n_hap;
        uint64_t thr = (uint64_t)(af * 65536.0);
        if (thr > 65535) thr = 65535;
        uint64_t *row = cov + (uint64_t)v * wpr;
        for (uint32_t q = 0; q < wpr; ++q) row[q] = 0;
        bool any = false;
        for (uint32_t h = 0; h < n_hap; h += 4) {
            const uint64_t x = g.next();
            for (uint32_t j = 0; j < 4 && h + j < n_hap; ++j)
                if (((x >> (16 * j)) & 0xffffu) < thr) { row[(h + j) >> 6] |= 1ull << ((h + j) & 63); any = true; }
        }
        if (!any) { const uint64_t h = g.below(n_hap); row[h >> 6] |= 1ull << (h & 63); }
    }
    return n;
}
