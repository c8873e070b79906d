// jst.cuh -- journaled string tree (JST) search: host-side flattening into branch-context
// windows + the device source that materialises them (north_star items 1, 2).
//
// The reference has no JST code (only the matcher hooks window_size/capture/restore that a JST
// traverser calls, libspm/libspm/matcher/concept.hpp:27-102, and a UML sketch,
// specs/journaled_sequence_class_diagram.drawio); this file defines it.
//
// Model: reference R, variants sorted by ref_pos (same record as a journal event), coverage
// bit matrix (bit h of row v = haplotype h carries v).  Haplotype h = R with its variants applied.
// Let W = window_size of the pattern (m for exact, m+k for Myers): whether position E of a
// haplotype is the end of a hit depends only on its last W symbols.  Hence
//   * an end E is CLEAN for h if those W symbols are an unmodified copy of a reference
//     interval: its hits are exactly the reference's hits there -> ONE scan of the reference,
//     each reference hit expanded to the haplotypes with no variant in its window;
//   * all other ends are DIRTY: they lie within W of a carried variant.  Consecutive carried
//     variants closer than W chain into one branch context.  A context is identified by the
//     ordered set of variants applied; its text and its owned ends are the same for every
//     haplotype carrying exactly that set in that span.  The DFS below enumerates every context
//     realised by at least one haplotype (coverage intersections), each exactly once ->
//     a flat array of windows, searched by the windows-mode kernels.
// Every (haplotype, end) is owned by exactly one of the two, so the union is duplicate-free.
#pragma once
#include <algorithm>
#include <atomic>
#include <chrono>
#include <thread>
#include <vector>

#include "layout.cuh"
#include "pack.cuh"

struct JstWindow {
    uint64_t ref_start; // reference position of the first symbol of the window text
    uint32_t len;       // symbols in the window text
    uint32_t own_lo;    // first owned end (exclusive-end coordinate, window-local)
    uint32_t var_off;   // into the window-variant id list
    uint32_t n_var;
};

// materialises window texts (one row of a windows-mode store = one window) as lane-chunk words, 32 symbols
// of a row per call, for pack_chunks_kernel: every maximal run copied from the reference is ONE unaligned
// 32-symbol fetch from the reference's lane chunks (two 8-byte loads); inserted symbols come from the pool
// one by one.
struct WindowChunkSource {
    const uint2 *__restrict__ base_lanes;
    SpmbLayout base;
    const spmb_journal_event *__restrict__ var;
    const uint8_t *__restrict__ pool;
    const JstWindow *__restrict__ win; // this bucket's windows
    const uint32_t *__restrict__ win_vars;
    uint64_t n_win;
    SpmbLayout lay; // of the windows store

    __device__ __forceinline__ uint2 load_chunk(uint64_t p0) const
    {
        const uint64_t R = spmb_div_rl(lay, p0);
        const uint32_t col0 = (uint32_t)(p0 - R * lay.row_len);
        if (R >= n_win) return make_uint2(0, 0);
        const JstWindow wd = win[R];
        if (col0 >= wd.len) return make_uint2(0, 0);
        unsigned long long out = 0;
        uint32_t vi = 0;
        uint32_t seg_local = 0;          // local position where the current reference segment starts
        uint64_t seg_ref = wd.ref_start; // reference position at seg_local
        const uint32_t t_end = min(32u, wd.len - col0);
        uint32_t t = 0;
        while (t < t_end) {
            const uint32_t x = col0 + t;
            uint32_t lim = col0 + t_end;  // local end (exclusive) of the reference run that starts at x
            bool inserted = false;
            while (vi < wd.n_var) {
                const spmb_journal_event v = var[win_vars[wd.var_off + vi]];
                const uint32_t lv = seg_local + (uint32_t)(v.ref_pos - seg_ref); // alt begins here
                if (x < lv) { lim = min(lim, lv); break; }
                if (x < lv + v.ins_len) {
                    out |= (unsigned long long)(pool[v.ins_off + (x - lv)] & 3u) << (62 - 2 * t);
                    ++t;
                    inserted = true;
                    break;
                }
                seg_local = lv + v.ins_len;
                seg_ref = v.ref_pos + v.del_len;
                ++vi;
            }
            if (inserted) continue;
            const uint32_t run = lim - x;
            const uint2 w = spmb_fetch32(base_lanes, base, seg_ref + (x - seg_local));
            unsigned long long W = ((unsigned long long)w.x << 32) | w.y;
            if (run < 32) W &= ~(~0ull >> (2 * run)); // keep the first `run` symbols
            out |= W >> (2 * t);
            t += run;
        }
        return make_uint2((uint32_t)(out >> 32), (uint32_t)out);
    }
};

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
struct JstHost {
    std::vector<spmb_journal_event> var;
    std::vector<uint8_t> pool;
    std::vector<uint64_t> cov; // n_var x wpr
    uint32_t n_hap = 0, wpr = 0;
    uint64_t n_ref = 0;
    uint32_t max_del = 0;
    // per-haplotype coordinate shift at every CK-th variant: ckpt[(i/CK)*n_hap + h] =
    // sum of (ins-del) over carried variants with index < (i/CK)*CK
    static constexpr uint32_t CK = 256;
    std::vector<int64_t> ckpt;

    const uint64_t *row(uint64_t v) const { return cov.data() + v * wpr; }

    void build_ckpt()
    {
        const uint64_t nv = var.size();
        const uint64_t nb = nv / CK + 1;
        ckpt.assign(nb * n_hap, 0);
        std::vector<int64_t> run(n_hap, 0);
        for (uint64_t v = 0; v < nv; ++v) {
            if (v % CK == 0) std::copy(run.begin(), run.end(), ckpt.begin() + (v / CK) * n_hap);
            const int64_t d = (int64_t)var[v].ins_len - (int64_t)var[v].del_len;
            if (d == 0) continue;
            const uint64_t *r = row(v);
            for (uint32_t w = 0; w < wpr; ++w) {
                uint64_t bits = r[w];
                while (bits) {
                    const uint32_t h = w * 64 + (uint32_t)__builtin_ctzll(bits);
                    bits &= bits - 1;
                    if (h < n_hap) run[h] += d;
                }
            }
        }
        if (nv % CK == 0) std::copy(run.begin(), run.end(), ckpt.begin() + (nv / CK) * n_hap);
    }

    // shift of every haplotype from carried variants with index < i0
    void shifts_before(uint64_t i0, std::vector<int64_t> &out) const
    {
        const uint64_t b = i0 / CK;
        out.assign(ckpt.begin() + b * n_hap, ckpt.begin() + (b + 1) * n_hap);
        for (uint64_t v = b * CK; v < i0; ++v) {
            const int64_t d = (int64_t)var[v].ins_len - (int64_t)var[v].del_len;
            if (d == 0) continue;
            const uint64_t *r = row(v);
            for (uint32_t w = 0; w < wpr; ++w) {
                uint64_t bits = r[w];
                while (bits) {
                    const uint32_t h = w * 64 + (uint32_t)__builtin_ctzll(bits);
                    bits &= bits - 1;
                    if (h < n_hap) out[h] += d;
                }
            }
        }
    }

    // coverage of haplotypes that start a context at variant v1: carry v1 and no earlier
    // variant closer than W
    bool start_cov(uint64_t v1, uint32_t W, uint64_t *c) const
    {
        const uint64_t *r = row(v1);
        uint64_t any = 0;
        for (uint32_t w = 0; w < wpr; ++w) { c[w] = r[w]; any |= c[w]; }
        if (!any) return false;
        const uint64_t p1 = var[v1].ref_pos;
        for (uint64_t u = v1; u-- > 0;) {
            if (var[u].ref_pos + max_del + W <= p1) break; // no earlier variant can reach any more
            if (var[u].ref_pos + var[u].del_len + W > p1) { // gap < W
                const uint64_t *ru = row(u);
                any = 0;
                for (uint32_t w = 0; w < wpr; ++w) { c[w] &= ~ru[w]; any |= c[w]; }
                if (!any) return false;
            }
        }
        return true;
    }

    // haplotypes realising a window (same walk as the DFS, along the window's own variant list)
    void window_cov(const uint32_t *ids, uint32_t n, uint32_t W, std::vector<uint64_t> &c) const
    {
        c.assign(wpr, 0);
        if (!start_cov(ids[0], W, c.data())) { std::fill(c.begin(), c.end(), 0); return; }
        uint64_t last_end = var[ids[0]].ref_pos + var[ids[0]].del_len;
        uint32_t next = 1;
        for (uint64_t u = (uint64_t)ids[0] + 1; u < var.size(); ++u) {
            if (var[u].ref_pos >= last_end + W) break;
            const uint64_t *ru = row(u);
            if (next < n && ids[next] == u) {
                for (uint32_t w = 0; w < wpr; ++w) c[w] &= ru[w];
                last_end = var[u].ref_pos + var[u].del_len;
                ++next;
            } else if (var[u].ref_pos >= last_end) {
                for (uint32_t w = 0; w < wpr; ++w) c[w] &= ~ru[w];
            }
        }
    }

    // haplotypes for which a REFERENCE hit ending at E (exclusive) is valid: none of their
    // variants touches [E-W, E)
    void clean_cov(uint64_t E, uint32_t W, std::vector<uint64_t> &c) const
    {
        c.assign(wpr, ~0ull);
        const int64_t a = (int64_t)E - (int64_t)W;
        // first variant with ref_pos >= E
        uint64_t hi = std::lower_bound(var.begin(), var.end(), E,
                                       [](const spmb_journal_event &x, uint64_t e) { return x.ref_pos < e; }) - var.begin();
        for (uint64_t u = hi; u-- > 0;) {
            const int64_t p = (int64_t)var[u].ref_pos;
            if (p + (int64_t)max_del + 1 <= a) break;
            const bool conflict = var[u].del_len ? (p + (int64_t)var[u].del_len > a) : (p > a);
            if (conflict) {
                const uint64_t *ru = row(u);
                for (uint32_t w = 0; w < wpr; ++w) c[w] &= ~ru[w];
            }
        }
        // mask the unused high bits of the last word
        if (n_hap % 64) c[wpr - 1] &= (1ull << (n_hap % 64)) - 1;
    }
};

struct JstFlatHost {
    uint32_t W = 0;
    bool overflow = false; // a window longer than 2^32-1 symbols or more than 2^32-1 window-variant ids: not representable
    std::vector<JstWindow> wins;
    std::vector<uint32_t> win_vars;
    uint64_t total_symbols = 0;
    double flatten_ms = 0;
};

// One branch context = the chain of variants in `ids`, realised by the haplotypes in the
// coverage vector at arena[depth*wpr ..].  Candidates u (reference order) within reach (gap < W
// from the end of the last applied variant) split the haplotypes: those carrying u continue
// with u applied (child), the others go on to the next candidate.  When no candidate is left in
// reach, whoever remains realises exactly this chain: emit one window.
struct JstDfs {
    const JstHost &J;
    uint32_t W;
    std::vector<JstWindow> &wins;
    std::vector<uint32_t> &ids_out;
    std::vector<uint64_t> arena;
    std::vector<uint32_t> ids;
    bool overflow = false;

    void emit()
    {
        const spmb_journal_event &first = J.var[ids[0]];
        const uint64_t rs = first.ref_pos + 1 >= W ? first.ref_pos + 1 - W : 0;
        uint64_t len = 0;
        uint64_t prev_end = rs;
        for (uint32_t id : ids) {
            const spmb_journal_event &x = J.var[id];
            len += (x.ref_pos - prev_end) + x.ins_len;
            prev_end = x.ref_pos + x.del_len;
        }
        len += std::min<uint64_t>(W - 1, J.n_ref - prev_end);
        if (len > 0xffffffffull || ids_out.size() + ids.size() > 0xffffffffull) { overflow = true; return; }
        JstWindow wd;
        wd.ref_start = rs;
        wd.len = (uint32_t)len;
        wd.own_lo = (uint32_t)(first.ref_pos - rs) + 1;
        wd.var_off = (uint32_t)ids_out.size();
        wd.n_var = (uint32_t)ids.size();
        ids_out.insert(ids_out.end(), ids.begin(), ids.end());
        wins.push_back(wd);
    }

    // Depth-first walk with an explicit stack (a chain of densely called variants is as deep as it is long:
    // no recursion, so no stack limit of the flattening threads applies).  Frame d owns the coverage vector at
    // arena[d*wpr ..) and ids[0..d]; arena[0 ..) holds the starting coverage of ids[0] on entry.
    struct Frame { uint64_t u, last_end; bool stop; };
    std::vector<Frame> stack;

    void walk(uint64_t v1)
    {
        const uint32_t wpr = J.wpr;
        const uint64_t nv = J.var.size();
        ids.clear();
        ids.push_back((uint32_t)v1);
        stack.clear();
        stack.push_back({v1 + 1, J.var[v1].ref_pos + J.var[v1].del_len, false});
        while (!stack.empty()) {
            if (stack.back().stop) { stack.pop_back(); ids.pop_back(); continue; } // nobody was left without the child's variant
            const size_t depth = stack.size() - 1;
            if (arena.size() < (depth + 2) * wpr) arena.resize((depth + 2) * wpr);
            bool descended = false;
            while (stack[depth].u < nv && J.var[stack[depth].u].ref_pos < stack[depth].last_end + W) {
                const uint64_t u = stack[depth].u++;
                if (J.var[u].ref_pos < stack[depth].last_end) continue; // overlaps the chain: cannot be carried with it
                const uint64_t *ru = J.row(u);
                uint64_t *c = arena.data() + depth * wpr;
                uint64_t *c1 = arena.data() + (depth + 1) * wpr;
                uint64_t any1 = 0, any0 = 0;
                for (uint32_t w = 0; w < wpr; ++w) {
                    c1[w] = c[w] & ru[w];
                    any1 |= c1[w];
                    c[w] &= ~ru[w];
                    any0 |= c[w];
                }
                if (!any0) stack[depth].stop = true; // after the child: nobody is left without u
                if (any1) {
                    ids.push_back((uint32_t)u);
                    stack.push_back({u + 1, J.var[u].ref_pos + J.var[u].del_len, false});
                    descended = true;
                    break;
                }
                if (!any0) break;
            }
            if (descended) continue;
            if (!stack[depth].stop) emit();
            stack.pop_back();
            ids.pop_back();
        }
    }
};

static bool jst_flatten_range(const JstHost &J, uint32_t W, uint64_t v_begin, uint64_t v_end,
                              std::vector<JstWindow> &wins, std::vector<uint32_t> &ids_out)
{
    JstDfs d{J, W, wins, ids_out, {}, {}};
    d.arena.resize(64 * J.wpr);
    for (uint64_t v1 = v_begin; v1 < v_end; ++v1) {
        if (!J.start_cov(v1, W, d.arena.data())) continue;
        d.walk(v1);
    }
    return !d.overflow;
}

static void jst_flatten(const JstHost &J, uint32_t W, JstFlatHost &out)
{
    const auto t0 = std::chrono::steady_clock::now();
    out.W = W;
    out.wins.clear();
    out.win_vars.clear();
    const uint64_t nv = J.var.size();
    unsigned nt = std::thread::hardware_concurrency();
    if (nt == 0) nt = 4;
    if (nt > 64) nt = 64;
    const uint64_t chunk = 4096;
    const uint64_t n_chunks = (nv + chunk - 1) / chunk;
    if (n_chunks < nt) nt = (unsigned)std::max<uint64_t>(1, n_chunks);
    std::vector<std::vector<JstWindow>> pw(n_chunks);
    std::vector<std::vector<uint32_t>> pi(n_chunks);
    std::atomic<uint64_t> next{0};
    std::atomic<int> bad{0};
    auto work = [&]() {
        for (;;) {
            const uint64_t c = next.fetch_add(1);
            if (c >= n_chunks) break;
            if (!jst_flatten_range(J, W, c * chunk, std::min(nv, (c + 1) * chunk), pw[c], pi[c])) bad.store(1);
        }
    };
    std::vector<std::thread> th;
    for (unsigned t = 1; t < nt; ++t) th.emplace_back(work);
    work();
    for (auto &t : th) t.join();
    uint64_t total_w = 0, total_i = 0;
    for (uint64_t c = 0; c < n_chunks; ++c) { total_w += pw[c].size(); total_i += pi[c].size(); }
    out.overflow = bad.load() != 0 || total_i > 0xffffffffull;
    if (out.overflow) return;
    out.wins.reserve(total_w);
    out.win_vars.reserve(total_i);
    out.total_symbols = 0;
    for (uint64_t c = 0; c < n_chunks; ++c) {
        const uint32_t base = (uint32_t)out.win_vars.size();
        for (JstWindow w : pw[c]) { w.var_off += base; out.total_symbols += w.len; out.wins.push_back(w); }
        out.win_vars.insert(out.win_vars.end(), pi[c].begin(), pi[c].end());
    }
    out.flatten_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}
