// hit_order.hpp -- host-side ordering of a short raw hit list (pure C++, no CUDA).
//
// The kernels append hits through an atomic counter, so a list arrives in no particular order; it is handed back
// sorted by (pattern, sequence, end) like the reference's callback sequence (seqan_pattern_base.hpp:40-52 runs the
// needles one after the other, each over ascending positions).  Lists below SPMB_ORDER_AT_LAUNCH_MIN records are
// ordered here, on the thread that called spmb_search_finish -- which makes this loop the host's share of a pass:
// measured on the C5 shards (13.6 k hits per pass at N = 2, 3.4 k at N = 8) the first form (LSD radix sort over
// separate key / index arrays, six passes, fresh vectors per call) took 0.53 ms / 0.16 ms per list, as long as the
// kernels of the pass (0.55 / 0.18 ms), and 0.12 ms of that was page faults of the fresh output vector.  This form:
//   * (key, record index) packed into ONE 64-bit word;
//   * one scatter into ~n/2 buckets by the key's position inside [smallest key, largest key] (order-preserving), then
//     an insertion sort inside each bucket (a handful of records; std::sort when a bucket is crowded, so a clustered
//     list degrades to n log n, never worse) -- two sweeps over the words whatever the key width;
//   * every buffer lives in a HitOrderScratch that is kept between calls.
// Same host, same lists: 3.4 k hits 76 -> 35 us, 13.6 k hits 370 -> 210 us (tools/ lists of C5; g++ -O3, one thread).
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
// of every (pattern_id, seq_id, end) is kept.  Records with equal keys keep their order of arrival.  false = the
// packed key does not fit 64 bits (the caller falls back to a comparison sort).
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
    const int total = b_pat + b_seq + b_end + b_err;
    if (total + b_idx > 64) return false;
    if (sc.a.size() < n) { sc.a.resize(n); sc.b.resize(n); }
    uint64_t *src = sc.a.data(), *dst = sc.b.data();
    uint64_t kmin = ~0ull, kmax = 0;
    for (size_t i = 0; i < n; ++i) {
        const spmb_hit &h = v[i];
        uint64_t k = h.pattern_id;
        k = b_seq ? (k << b_seq) | h.seq_id : k;
        k = (k << b_end) | h.end;
        if (with_errors) k = (k << b_err) | h.errors;
        kmin = std::min(kmin, k); kmax = std::max(kmax, k);
        src[i] = (k << b_idx) | i;
    }
    int lb = 8;
    while (((size_t)1 << lb) < n / 2 && lb < 16) ++lb;
    const size_t B = (size_t)1 << lb;
    const int span_bits = hit_order_bits(kmax - kmin);
    const int sh = span_bits > lb ? span_bits - lb : 0;
    if (sc.hist.size() < B + 1) sc.hist.resize(B + 1);
    uint32_t *cnt = sc.hist.data();
    std::fill(cnt, cnt + B + 1, 0u);
    for (size_t i = 0; i < n; ++i) ++cnt[(((src[i] >> b_idx) - kmin) >> sh) + 1];
    for (size_t b = 0; b < B; ++b) cnt[b + 1] += cnt[b];
    for (size_t i = 0; i < n; ++i) { const uint64_t x = src[i]; dst[cnt[((x >> b_idx) - kmin) >> sh]++] = x; }
    size_t lo = 0;
    for (size_t b = 0; b < B; ++b) {
        const size_t hi = cnt[b];
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
    if (sc.out.size() < n) sc.out.resize(n);
    spmb_hit *out = sc.out.data();
    const uint64_t im = ((uint64_t)1 << b_idx) - 1;
    size_t o = 0;
    if (unique_keys) {
        uint64_t last = ~0ull;
        for (size_t i = 0; i < n; ++i) {
            const uint64_t kk = dst[i] >> (b_idx + b_err);
            if (kk != last) { out[o++] = v[dst[i] & im]; last = kk; }
        }
    } else {
        for (size_t i = 0; i < n; ++i) out[o++] = v[dst[i] & im];
    }
    v.swap(sc.out);
    v.resize(o);
    return true;
}
