// hit_order.hpp -- host-side ordering of a short raw hit list (pure C++, no CUDA).
//
// The kernels append hits through an atomic counter, so a list arrives in no particular order; it is handed back
// sorted by (pattern, sequence, end) like the reference's callback sequence (seqan_pattern_base.hpp:40-52 runs the
// needles one after the other, each over ascending positions).  Lists below SPMB_ORDER_AT_LAUNCH_MIN records are
// ordered here, on the thread that called spmb_search_finish -- which makes this loop the host's share of a pass:
// measured on the C5 shards (13.6 k hits per pass at N = 2, 3.4 k at N = 8) the first form (LSD radix sort over
// separate key / index arrays, fresh vectors per call) took 0.53 ms / 0.16 ms per list, as long as the kernels of
// the pass (0.55 / 0.18 ms), and 0.12 ms of that was page faults of the fresh output vector.  This form:
//   * (key, record index) packed into ONE 64-bit word, so a pass moves 8 bytes per record;
//   * many needles (the seed-filter sets): ONE counting pass over the needle id, then the few records of a needle
//     are put in order by insertion sort -- two sweeps instead of six radix passes;
//   * few needles: LSD radix sort over the key bits only, all histograms from one sweep, constant digits skipped;
//   * every buffer lives in a HitOrderScratch that the hit list keeps between calls.
#pragma once
#include <algorithm>
#include <cstdint>
#include <vector>

#include "../../include/spmb.h"

struct HitOrderScratch {
    std::vector<uint64_t> a, b;
    std::vector<uint32_t> hist;
    std::vector<spmb_hit> out;
};

static inline int hit_order_bits(unsigned long long max_value)
{
    int b = 1;
    while (b < 64 && (max_value >> b)) ++b;
    return b;
}

// Orders `v` by (pattern_id, seq_id, end[, errors]); with `unique_keys` only the first record (smallest error count)
// of every (pattern_id, seq_id, end) is kept.  false = the packed key does not fit 64 bits (the caller falls back to
// a comparison sort).
static inline bool hits_order_packed(std::vector<spmb_hit> &v, bool with_errors, bool unique_keys, HitOrderScratch &sc)
{
    const size_t n = v.size();
    if (n == 0) return true;
    uint64_t max_end = 0;
    uint32_t max_seq = 0, max_pat = 0, max_err = 0;
    for (const spmb_hit &h : v) {
        max_end = std::max<uint64_t>(max_end, h.end); max_seq = std::max(max_seq, h.seq_id);
        max_pat = std::max(max_pat, h.pattern_id); max_err = std::max(max_err, h.errors);
    }
    const int b_err = with_errors ? hit_order_bits(max_err) : 0, b_end = hit_order_bits(max_end);
    const int b_seq = max_seq ? hit_order_bits(max_seq) : 0, b_pat = max_pat ? hit_order_bits(max_pat) : 0;
    const int b_idx = hit_order_bits(n);
    const int b_rest = b_seq + b_end + b_err;        // the key below the needle id
    if (b_pat + b_rest + b_idx > 64) return false;
    if (sc.a.size() < n) { sc.a.resize(n); sc.b.resize(n); }
    uint64_t *src = sc.a.data(), *dst = sc.b.data();
    const size_t P = (size_t)max_pat + 1;
    if (P >= 64 && n >= P / 4 && n <= 16 * P) {
        // ---- many needles: counting sort over the needle id, insertion sort inside a needle ----
        if (sc.hist.size() < P + 1) sc.hist.resize(P + 1);
        uint32_t *cnt = sc.hist.data();
        std::fill(cnt, cnt + P + 1, 0u);
        for (size_t i = 0; i < n; ++i) ++cnt[v[i].pattern_id + 1];
        for (size_t p = 0; p < P; ++p) cnt[p + 1] += cnt[p];
        // cnt[p] = first slot of needle p; the scatter advances it to the first slot of needle p + 1
        for (size_t i = 0; i < n; ++i) {
            const spmb_hit &h = v[i];
            uint64_t k = b_seq ? h.seq_id : 0;
            k = (k << b_end) | h.end;
            if (with_errors) k = (k << b_err) | h.errors;
            dst[cnt[h.pattern_id]++] = (k << b_idx) | i;
        }
        size_t lo = 0;
        for (size_t p = 0; p < P; ++p) {
            const size_t hi = cnt[p];
            if (hi - lo > 24) std::sort(dst + lo, dst + hi);
            else
                for (size_t i = lo + 1; i < hi; ++i) {
                    const uint64_t x = dst[i];
                    size_t j = i;
                    while (j > lo && dst[j - 1] > x) { dst[j] = dst[j - 1]; --j; }
                    dst[j] = x;
                }
            lo = hi;
        }
        std::swap(src, dst);
        // src = ordered (rest key | index); the needle id comes from the record itself below
        if (sc.out.size() < n) sc.out.resize(n);
        spmb_hit *out = sc.out.data();
        const uint64_t im = ((uint64_t)1 << b_idx) - 1;
        size_t o = 0;
        if (unique_keys) {
            uint64_t last_k = ~0ull;
            uint32_t last_p = 0xffffffffu;
            for (size_t i = 0; i < n; ++i) {
                const spmb_hit &h = v[src[i] & im];
                const uint64_t kk = src[i] >> (b_idx + b_err);
                if (kk != last_k || h.pattern_id != last_p) { out[o++] = h; last_k = kk; last_p = h.pattern_id; }
            }
        } else {
            for (size_t i = 0; i < n; ++i) out[o++] = v[src[i] & im];
        }
        v.swap(sc.out);
        v.resize(o);
        return true;
    }
    // ---- few needles: LSD radix sort over (pattern | sequence | end | errors), the index rides in the low bits ----
    const int total = b_pat + b_rest;
    for (size_t i = 0; i < n; ++i) {
        const spmb_hit &h = v[i];
        uint64_t k = h.pattern_id;
        k = b_seq ? (k << b_seq) | h.seq_id : k;
        k = (k << b_end) | h.end;
        if (with_errors) k = (k << b_err) | h.errors;
        src[i] = (k << b_idx) | i;
    }
    const int W = 8;
    const int passes = (total + W - 1) / W;
    const size_t R = (size_t)1 << W;
    if (sc.hist.size() < (size_t)passes * R) sc.hist.resize((size_t)passes * R);
    std::fill(sc.hist.begin(), sc.hist.begin() + (size_t)passes * R, 0u);
    for (size_t i = 0; i < n; ++i) {
        const uint64_t k = src[i] >> b_idx;
        for (int p = 0; p < passes; ++p) ++sc.hist[(size_t)p * R + ((k >> (p * W)) & (R - 1))];
    }
    for (int p = 0; p < passes; ++p) {
        uint32_t *hp = sc.hist.data() + (size_t)p * R;
        uint32_t sum = 0;
        bool constant = false;
        for (size_t d = 0; d < R; ++d) { const uint32_t c = hp[d]; constant |= (c == n); hp[d] = sum; sum += c; }
        if (constant) continue;
        const int sh = b_idx + p * W;
        for (size_t i = 0; i < n; ++i) { const uint64_t x = src[i]; dst[hp[(x >> sh) & (R - 1)]++] = x; }
        std::swap(src, dst);
    }
    if (sc.out.size() < n) sc.out.resize(n);
    spmb_hit *out = sc.out.data();
    const uint64_t im = ((uint64_t)1 << b_idx) - 1;
    size_t o = 0;
    if (unique_keys) {
        uint64_t last = ~0ull;
        for (size_t i = 0; i < n; ++i) {
            const uint64_t kk = src[i] >> (b_idx + b_err);
            if (kk != last) { out[o++] = v[src[i] & im]; last = kk; }
        }
    } else {
        for (size_t i = 0; i < n; ++i) out[o++] = v[src[i] & im];
    }
    v.swap(sc.out);
    v.resize(o);
    return true;
}
