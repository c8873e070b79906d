// host-side hit ordering (libspm_b200/csrc/hit_order.hpp) against std::sort + std::unique on the records:
// same lists byte for byte, over key distributions from one needle to more needles than hits, plain and
// windows-mode (seq_id) lists, clustered keys (every record in one bucket) and lists too wide for the packed key.
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <random>
#include <vector>

#include "../../libspm_b200/csrc/hit_order.hpp"

static bool key_less(const spmb_hit &a, const spmb_hit &b)
{
    if (a.pattern_id != b.pattern_id) return a.pattern_id < b.pattern_id;
    if (a.seq_id != b.seq_id) return a.seq_id < b.seq_id;
    return a.end < b.end;
}

static std::vector<spmb_hit> reference_order(std::vector<spmb_hit> v, bool filtered)
{
    if (filtered) {
        std::stable_sort(v.begin(), v.end(), [](const spmb_hit &a, const spmb_hit &b) {
            if (key_less(a, b)) return true;
            if (key_less(b, a)) return false;
            return a.errors < b.errors;
        });
        v.erase(std::unique(v.begin(), v.end(), [](const spmb_hit &a, const spmb_hit &b) {
                    return a.pattern_id == b.pattern_id && a.seq_id == b.seq_id && a.end == b.end;
                }), v.end());
    } else {
        std::stable_sort(v.begin(), v.end(), key_less);
    }
    return v;
}

int main()
{
    HitOrderScratch sc;
    int failures = 0, cases = 0, refused = 0;
    for (int npat : {1, 3, 100, 10000, 60000})
        for (size_t n : {size_t(1), size_t(2), size_t(255), size_t(1000), size_t(3400), size_t(16383), size_t(40000)})
            for (int nseq : {1, 2504})
                for (int shape = 0; shape < 3; ++shape)
                    for (int filtered = 0; filtered < 2; ++filtered) {
                        std::mt19937_64 g(n * 31 + (unsigned)npat + 7u * (unsigned)shape);
                        std::vector<spmb_hit> base(n);
                        for (size_t i = 0; i < n; ++i) {
                            spmb_hit &h = base[i];
                            const uint64_t span = shape == 1 ? 50ull : (nseq > 1 ? 70000ull : 3000000000ull); // shape 1: clustered ends
                            h.end = (shape == 2 ? (1ull << 40) : 0ull) + g() % span;
                            h.seq_id = (uint32_t)(g() % (uint64_t)nseq);
                            h.pattern_id = shape == 1 ? (uint32_t)(npat - 1) : (uint32_t)(g() % (uint64_t)npat);
                            h.errors = (uint32_t)(g() % 4);
                            h.reserved = (uint32_t)i; // tells equal keys apart: their order of arrival must be kept
                        }
                        for (size_t i = 0; i < n / 10; ++i) { // duplicates of a key with other error counts
                            const size_t j = g() % n, s = g() % n;
                            const uint32_t keep = base[j].reserved;
                            base[j] = base[s];
                            base[j].reserved = keep;
                            base[j].errors = (base[j].errors + 1 + (uint32_t)(g() % 3)) % 4;
                        }
                        const std::vector<spmb_hit> want = reference_order(base, filtered != 0);
                        std::vector<spmb_hit> got = base;
                        ++cases;
                        if (!hits_order_packed(got, filtered != 0, filtered != 0, sc)) { ++refused; continue; }
                        bool same = got.size() == want.size();
                        for (size_t i = 0; same && i < got.size(); ++i) {
                            const spmb_hit &a = got[i], &b = want[i];
                            same = a.end == b.end && a.seq_id == b.seq_id && a.pattern_id == b.pattern_id && a.errors == b.errors &&
                                   (filtered || a.reserved == b.reserved);
                        }
                        if (!same) {
                            ++failures;
                            std::printf("MISMATCH npat=%d n=%zu nseq=%d shape=%d filtered=%d (%zu vs %zu records)\n", npat, n, nseq, shape,
                                        filtered, got.size(), want.size());
                        }
                    }
    // a key that cannot be packed (pattern 2^31, end 2^40, seq 2^20): refused, the list untouched
    {
        std::vector<spmb_hit> v(300);
        for (size_t i = 0; i < v.size(); ++i) { v[i] = spmb_hit{(1ull << 40) + i, 1u << 20, 1u << 31, 3, 0}; }
        const std::vector<spmb_hit> before = v;
        ++cases;
        if (hits_order_packed(v, true, true, sc) || std::memcmp(v.data(), before.data(), v.size() * sizeof(spmb_hit)) != 0) {
            ++failures;
            std::printf("MISMATCH: an unpackable key was not refused\n");
        }
    }
    std::printf("%d cases, %d refused (key too wide), %d failures\n", cases, refused, failures);
    return failures ? 1 : 0;
}
