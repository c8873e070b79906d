// multi.cpp -- the multi-GPU driver of the C ABI (include/spmb.h, spmb_multi_*; north_star item 4).
//
// A libspm caller is ONE C++ process that owns one host haystack (seqan_pattern_base::operator(),
// libspm/libspm/matcher/seqan_pattern_base.hpp:40-52).  This driver lets that process use every GPU of the box:
//   * one spmb_context (one stream, one upload pipeline) per device;
//   * the haystack is cut into contiguous shards, one per device; shard i holds its owned symbols plus `overlap`
//     symbols of left context (a hit is owned by the shard that holds its END; its window_size-1 earlier symbols
//     are in the context -- the seam rule of myers_matcher.hpp:51-53) and `overlap` symbols behind its end;
//   * every shard crosses its own PCIe link (per-device host threads run the upload pipelines side by side);
//   * a journal's event table is split at the cuts -- slices of the caller's array, no copy -- and every device
//     materialises its part of the derived sequence;
//   * search = launch on every device, then per-device D2H of the hit lists and a host merge into one sorted list:
//     the "final gather of the small hit lists".  There is no NCCL and no collective anywhere.
// Everything here is host code on top of the single-device entry points; no kernel is launched from this file.
#include <sched.h>

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "../../include/spmb.h"

// internal entry points of spmb_capi.cu
int spmb_store_create_journaled_ex(spmb_context *ctx, const spmb_store *base, const spmb_journal_event *ev, uint64_t n_ev,
                                   const uint8_t *pool, uint64_t pool_len, uint32_t row_len, uint64_t ref_bias,
                                   uint64_t probe_index, uint64_t *probe_der, spmb_store **out);
int spmb_hits_set_host(spmb_hits *h, std::vector<spmb_hit> &&list);

struct spmb_multi {
    std::vector<spmb_context *> ctx;
    std::vector<spmb_hits *> hits;
    std::string err;
    double ms[4] = {0, 0, 0, 0}; // of the last calls: upload, journal, search (launch -> last hit list on the host), merge
};

struct MultiShard {
    spmb_store *store = nullptr;
    uint32_t dev = 0;
    uint64_t start = 0, cut_lo = 0, cut_hi = 0, stop = 0; // coordinates of the base sequence
    uint64_t own_lo = 0, own_hi = 0;                      // owned exclusive ends, local: (own_lo, own_hi]
    uint64_t global_base = 0;                             // global end = global_base + (local end - own_lo)
    uint64_t unusable = 0;                                // leading local symbols that are not the true sequence
};

struct spmb_multi_store {
    spmb_multi *m = nullptr;
    uint64_t n = 0;
    uint32_t overlap = 0;
    bool journaled = false;
    std::vector<MultiShard> sh;
};

struct spmb_multi_pattern {
    spmb_multi *m = nullptr;
    std::vector<spmb_pattern *> p;
    int alg = 0;
    uint32_t window = 0;
};

namespace {
using clk = std::chrono::steady_clock;
double ms_since(clk::time_point t0) { return std::chrono::duration<double, std::milli>(clk::now() - t0).count(); }

int host_cpus()
{
    cpu_set_t set;
    if (sched_getaffinity(0, sizeof(set), &set) == 0) {
        int c = CPU_COUNT(&set);
        if (c > 0) return c;
    }
    unsigned h = std::thread::hardware_concurrency();
    return h ? (int)h : 1;
}

// fn(i) for every i < n on its own host thread (the calling thread takes i = 0); first non-zero status wins
template <class F>
int parallel_devices(size_t n, F fn)
{
    std::vector<int> rc(n, SPMB_OK);
    std::vector<std::thread> th;
    for (size_t i = 1; i < n; ++i) th.emplace_back([&rc, &fn, i] { rc[i] = fn(i); });
    if (n) rc[0] = fn(0);
    for (auto &t : th) t.join();
    for (size_t i = 0; i < n; ++i)
        if (rc[i]) return rc[i];
    return SPMB_OK;
}

int fail(spmb_multi *m, int code, const std::string &what)
{
    if (m) m->err = what;
    return code;
}

std::mutex g_err_mu;

int fail_dev(spmb_multi *m, int code, size_t dev, const char *what)
{
    std::lock_guard<std::mutex> lk(g_err_mu);
    if (m) {
        m->err = std::string(what) + " (device shard " + std::to_string(dev) + "): " + spmb_status_string(code);
        const char *d = dev < m->ctx.size() ? spmb_last_error(m->ctx[dev]) : nullptr;
        if (d && *d) { m->err += " -- "; m->err += d; }
    }
    return code;
}

uint64_t first_event_at_or_after(const spmb_journal_event *ev, uint64_t n_ev, uint64_t pos)
{
    return (uint64_t)(std::lower_bound(ev, ev + n_ev, pos, [](const spmb_journal_event &e, uint64_t p) { return e.ref_pos < p; }) - ev);
}
} // namespace

extern "C" int spmb_multi_create(const int *devices, uint32_t n_devices, spmb_multi **out)
{
    if (!out) return SPMB_ERR_ARG;
    *out = nullptr;
    const int avail = spmb_device_count();
    if (avail <= 0) return SPMB_ERR_NO_DEVICE;
    std::vector<int> devs;
    if (n_devices == 0) for (int d = 0; d < avail; ++d) devs.push_back(d);
    else for (uint32_t i = 0; i < n_devices; ++i) devs.push_back(devices ? devices[i] : (int)i);
    for (int d : devs) if (d < 0 || d >= avail) return SPMB_ERR_ARG;
    spmb_multi *m = new (std::nothrow) spmb_multi();
    if (!m) return SPMB_ERR_OOM;
    // the devices' upload pipelines share the host's cores: one feeding thread per device (the caller's thread
    // of parallel_devices) plus an equal share of packing threads.  With many links the copy engines move raw
    // bytes faster than the host could pack them, with few the packing threads quarter the PCIe bytes -- both
    // claim from one chunk queue per device, so the split follows the machine (upload.cuh).
    int per_dev = (host_cpus() - (int)devs.size()) / (int)devs.size();
    if (per_dev < 0) per_dev = 0;
    if (const char *e = getenv("SPMB_MULTI_HOST_THREADS")) per_dev = atoi(e);
    for (int d : devs) {
        spmb_context *c = nullptr;
        int rc = spmb_context_create(d, nullptr, &c);
        spmb_hits *h = nullptr;
        if (rc == SPMB_OK) { spmb_context_set_upload_threads(c, per_dev); rc = spmb_hits_create(c, 1 << 16, &h); }
        if (rc) {
            if (c) spmb_context_destroy(c);
            spmb_multi_destroy(m);
            return rc;
        }
        m->ctx.push_back(c);
        m->hits.push_back(h);
    }
    *out = m;
    return SPMB_OK;
}

extern "C" int spmb_multi_destroy(spmb_multi *m)
{
    if (!m) return SPMB_OK;
    for (auto *h : m->hits) spmb_hits_destroy(h);
    for (auto *c : m->ctx) spmb_context_destroy(c);
    delete m;
    return SPMB_OK;
}

extern "C" uint32_t spmb_multi_device_count(const spmb_multi *m) { return m ? (uint32_t)m->ctx.size() : 0; }
extern "C" spmb_context *spmb_multi_context(spmb_multi *m, uint32_t i) { return (m && i < m->ctx.size()) ? m->ctx[i] : nullptr; }
extern "C" const char *spmb_multi_last_error(const spmb_multi *m) { return m ? m->err.c_str() : "null driver"; }

extern "C" int spmb_multi_last_stats(const spmb_multi *m, double ms[4])
{
    if (!m || !ms) return SPMB_ERR_ARG;
    for (int i = 0; i < 4; ++i) ms[i] = m->ms[i];
    return SPMB_OK;
}

extern "C" int spmb_multi_store_create(spmb_multi *m, const uint8_t *host_ranks, uint64_t n, uint32_t overlap,
                                       spmb_multi_store **out)
{
    if (!m || !out || (!host_ranks && n)) return SPMB_ERR_ARG;
    *out = nullptr;
    const auto t0 = clk::now();
    spmb_multi_store *s = new (std::nothrow) spmb_multi_store();
    if (!s) return SPMB_ERR_OOM;
    s->m = m;
    s->n = n;
    s->overlap = overlap ? overlap : 4096;
    // shards of at least 64 Ki owned symbols (a tiny haystack stays on one device), cuts on multiples of 64
    size_t n_sh = m->ctx.size();
    while (n_sh > 1 && n / n_sh < (1u << 16)) --n_sh;
    for (size_t i = 0; i < n_sh; ++i) {
        MultiShard sh;
        sh.dev = (uint32_t)i;
        sh.cut_lo = i == 0 ? 0 : (n / n_sh * i) / 64 * 64;
        sh.cut_hi = i + 1 == n_sh ? n : (n / n_sh * (i + 1)) / 64 * 64;
        sh.start = sh.cut_lo > s->overlap ? (sh.cut_lo - s->overlap) / 64 * 64 : 0;
        sh.stop = std::min<uint64_t>(n, sh.cut_hi + s->overlap);
        sh.own_lo = sh.cut_lo - sh.start;
        sh.own_hi = sh.cut_hi - sh.start;
        sh.global_base = sh.cut_lo;
        s->sh.push_back(sh);
    }
    int rc = parallel_devices(n_sh, [&](size_t i) -> int {
        MultiShard &sh = s->sh[i];
        int r = spmb_store_create(m->ctx[sh.dev], host_ranks + sh.start, sh.stop - sh.start, 0, &sh.store);
        return r ? fail_dev(m, r, i, "spmb_store_create") : SPMB_OK;
    });
    m->ms[0] = ms_since(t0);
    if (rc) { spmb_multi_store_destroy(s); return rc; }
    *out = s;
    return SPMB_OK;
}

extern "C" int spmb_multi_store_create_journaled(spmb_multi *m, const spmb_multi_store *base, const spmb_journal_event *ev,
                                                 uint64_t n_ev, const uint8_t *pool, uint64_t pool_len, spmb_multi_store **out)
{
    if (!m || !base || !out || base->m != m || (n_ev && !ev) || (pool_len && !pool)) return SPMB_ERR_ARG;
    if (base->journaled) return fail(m, SPMB_ERR_ARG, "the base of a journaled store must be a plain store");
    *out = nullptr;
    const auto t0 = clk::now();
    spmb_multi_store *s = new (std::nothrow) spmb_multi_store();
    if (!s) return SPMB_ERR_OOM;
    s->m = m;
    s->overlap = base->overlap;
    s->journaled = true;
    const size_t n_sh = base->sh.size();
    s->sh = base->sh;
    for (auto &sh : s->sh) sh.store = nullptr;
    // cuts in the coordinates of the base: moved behind an event that deletes across them, so that every event
    // belongs to exactly one owned range; the shards' margins (overlap symbols on both sides) absorb the move
    std::vector<uint64_t> ea(n_sh), ec(n_sh), eb(n_sh);
    for (size_t i = 0; i < n_sh; ++i) {
        MultiShard &sh = s->sh[i];
        if (i > 0) {
            uint64_t c = base->sh[i].cut_lo;
            const uint64_t e = first_event_at_or_after(ev, n_ev, c);
            if (e > 0 && ev[e - 1].ref_pos + ev[e - 1].del_len > c) c = ev[e - 1].ref_pos + ev[e - 1].del_len;
            if (c > base->sh[i - 1].stop || c > base->sh[i].cut_hi) {
                spmb_multi_store_destroy(s);
                return fail(m, SPMB_ERR_UNSUPPORTED, "a journal event deletes more symbols than the shards' overlap");
            }
            sh.cut_lo = c;
            s->sh[i - 1].cut_hi = c;
        }
    }
    for (size_t i = 0; i < n_sh; ++i) {
        MultiShard &sh = s->sh[i];
        ea[i] = first_event_at_or_after(ev, n_ev, sh.start);
        ec[i] = first_event_at_or_after(ev, n_ev, sh.cut_lo);
        eb[i] = first_event_at_or_after(ev, n_ev, sh.cut_hi);
        // an event that begins before the shard's first symbol and deletes into it is not applied here: the
        // symbols it would have removed stay at the front of the left context, where no owned hit may reach
        sh.unusable = 0;
        if (ea[i] > 0 && ev[ea[i] - 1].ref_pos + ev[ea[i] - 1].del_len > sh.start)
            sh.unusable = ev[ea[i] - 1].ref_pos + ev[ea[i] - 1].del_len - sh.start;
    }
    std::vector<uint64_t> L(n_sh), R(n_sh);
    int rc = parallel_devices(n_sh, [&](size_t i) -> int {
        MultiShard &sh = s->sh[i];
        uint64_t probe = 0;
        int r = spmb_store_create_journaled_ex(m->ctx[sh.dev], base->sh[i].store, ev + ea[i], eb[i] - ea[i], pool, pool_len, 0,
                                               sh.start, ec[i] - ea[i], &probe, &sh.store);
        if (r) return fail_dev(m, r, i, "spmb_store_create_journaled");
        const uint64_t n_der = spmb_store_length(sh.store);
        // derived position of the lower cut: up to the first owned event the base maps 1:1
        L[i] = ec[i] < eb[i] ? probe - (ev[ec[i]].ref_pos - sh.cut_lo) : n_der - (sh.stop - sh.cut_lo);
        R[i] = n_der - (sh.stop - sh.cut_hi); // no event of this slice lies behind the upper cut
        return SPMB_OK;
    });
    m->ms[1] = ms_since(t0);
    if (rc) { spmb_multi_store_destroy(s); return rc; }
    uint64_t g = 0;
    for (size_t i = 0; i < n_sh; ++i) {
        s->sh[i].own_lo = L[i];
        s->sh[i].own_hi = R[i];
        s->sh[i].global_base = g;
        g += R[i] - L[i];
    }
    s->n = g;
    *out = s;
    return SPMB_OK;
}

extern "C" int spmb_multi_store_destroy(spmb_multi_store *s)
{
    if (!s) return SPMB_OK;
    for (auto &sh : s->sh) spmb_store_destroy(sh.store);
    delete s;
    return SPMB_OK;
}

extern "C" uint64_t spmb_multi_store_length(const spmb_multi_store *s) { return s ? s->n : 0; }
extern "C" uint32_t spmb_multi_store_shards(const spmb_multi_store *s) { return s ? (uint32_t)s->sh.size() : 0; }

extern "C" int spmb_multi_pattern_create(spmb_multi *m, int alg, const uint8_t *ranks, const uint32_t *lengths,
                                         uint32_t n_patterns, uint32_t k, spmb_multi_pattern **out)
{
    if (!m || !out || !lengths || n_patterns == 0) return SPMB_ERR_ARG;
    *out = nullptr;
    spmb_multi_pattern *p = new (std::nothrow) spmb_multi_pattern();
    if (!p) return SPMB_ERR_OOM;
    p->m = m;
    p->alg = alg;
    p->p.assign(m->ctx.size(), nullptr);
    int rc = parallel_devices(m->ctx.size(), [&](size_t i) -> int {
        int r = spmb_pattern_set_create(m->ctx[i], alg, ranks, lengths, n_patterns, k, &p->p[i]);
        return r ? fail_dev(m, r, i, "spmb_pattern_set_create") : SPMB_OK;
    });
    if (rc) { spmb_multi_pattern_destroy(p); return rc; }
    p->window = spmb_pattern_window_size(p->p[0]);
    *out = p;
    return SPMB_OK;
}

extern "C" int spmb_multi_pattern_destroy(spmb_multi_pattern *p)
{
    if (!p) return SPMB_OK;
    for (auto *x : p->p) spmb_pattern_destroy(x);
    delete p;
    return SPMB_OK;
}

extern "C" int spmb_multi_search(spmb_multi *m, const spmb_multi_store *s, const spmb_multi_pattern *p, spmb_hits *merged)
{
    if (!m || !s || !p || !merged || s->m != m || p->m != m) return SPMB_ERR_ARG;
    if (p->alg == SPMB_ALG_MYERS_PREFIX) return fail(m, SPMB_ERR_UNSUPPORTED, "the anchored matcher has one anchor: use a single device");
    const auto t0 = clk::now();
    const size_t n_sh = s->sh.size();
    const uint64_t need = p->window ? p->window - 1 : 0;
    for (size_t i = 1; i < n_sh; ++i)
        if (s->sh[i].own_lo < need + s->sh[i].unusable)
            return fail(m, SPMB_ERR_UNSUPPORTED, "window_size - 1 exceeds the left context of the shards (create the store with a larger overlap)");
    for (size_t i = 0; i < n_sh; ++i) {
        const MultiShard &sh = s->sh[i];
        int rc = spmb_search_launch(m->ctx[sh.dev], sh.store, p->p[sh.dev], m->hits[sh.dev]);
        if (rc) return fail_dev(m, rc, i, "spmb_search_launch");
    }
    // per-device hit lists -> host (each finish waits for its own device only)
    std::vector<const spmb_hit *> data(n_sh);
    std::vector<uint64_t> cnt(n_sh);
    for (size_t i = 0; i < n_sh; ++i) {
        const MultiShard &sh = s->sh[i];
        int rc = spmb_search_finish(m->ctx[sh.dev], m->hits[sh.dev]);
        if (rc) return fail_dev(m, rc, i, "spmb_search_finish");
        data[i] = spmb_hits_data(m->hits[sh.dev]);
        cnt[i] = spmb_hits_count(m->hits[sh.dev]);
    }
    m->ms[2] = ms_since(t0);
    const auto t1 = clk::now();
    // merge: every list is sorted by (pattern, end) (seed matcher: by end first); the shards are in text order and
    // own disjoint ranges of ends, so per pattern the lists simply follow one another
    std::vector<spmb_hit> out;
    uint64_t total = 0;
    for (size_t i = 0; i < n_sh; ++i) total += cnt[i];
    out.reserve(total);
    auto take = [&](size_t i, const spmb_hit &h) {
        const MultiShard &sh = s->sh[i];
        if (h.end <= sh.own_lo || h.end > sh.own_hi) return; // ends inside the context / the margin belong to a neighbour
        spmb_hit x = h;
        x.end = sh.global_base + (h.end - sh.own_lo);
        out.push_back(x);
    };
    if (p->alg == SPMB_ALG_PIGEONHOLE) {
        for (size_t i = 0; i < n_sh; ++i)
            for (uint64_t j = 0; j < cnt[i]; ++j) take(i, data[i][j]);
    } else {
        std::vector<uint64_t> cur(n_sh, 0);
        for (;;) {
            uint32_t pid = 0xffffffffu;
            bool any = false;
            for (size_t i = 0; i < n_sh; ++i)
                if (cur[i] < cnt[i]) { pid = std::min(pid, data[i][cur[i]].pattern_id); any = true; }
            if (!any) break;
            for (size_t i = 0; i < n_sh; ++i)
                while (cur[i] < cnt[i] && data[i][cur[i]].pattern_id == pid) take(i, data[i][cur[i]++]);
        }
    }
    int rc = spmb_hits_set_host(merged, std::move(out));
    m->ms[3] = ms_since(t1);
    return rc;
}

extern "C" int spmb_multi_search_host(spmb_multi *m, const uint8_t *host_ranks, uint64_t n, const spmb_multi_pattern *p,
                                      spmb_hits *merged)
{
    if (!m || !p || !merged || (!host_ranks && n)) return SPMB_ERR_ARG;
    spmb_multi_store *s = nullptr;
    uint32_t overlap = (p->window + 63) / 64 * 64;
    if (overlap < 1024) overlap = 1024;
    int rc = spmb_multi_store_create(m, host_ranks, n, overlap, &s);
    if (rc) return rc;
    rc = spmb_multi_search(m, s, p, merged);
    spmb_multi_store_destroy(s);
    return rc;
}
