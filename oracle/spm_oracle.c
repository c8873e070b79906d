/*
 * spm_oracle.c -- CPU ORACLE (test infrastructure; see spm_oracle.h for the rules).
 *
 * Restates, in plain C, the algorithms the reference reaches through
 *   spm::seqan_pattern_base::operator()  libspm/libspm/matcher/seqan_pattern_base.hpp:40-52
 *   -> seqan2::find(finder, pattern[, minScore])                           ...:60-65
 * for the three searchers of the hot path:
 *   spm::horspool_matcher  libspm/libspm/matcher/horspool_matcher.hpp:20-44   (a3/a8)
 *   spm::shiftor_matcher   libspm/libspm/matcher/shiftor_matcher.hpp:20-44    (a4/a9)
 *   spm::myers_matcher     libspm/libspm/matcher/myers_matcher.hpp:19-60      (a5/a10)
 * and the resumable variants
 *   restorable_myers_matcher   libspm/libspm/matcher/myers_matcher_restorable.hpp:33-82 (a6)
 *   restorable_shiftor_matcher libspm/libspm/matcher/shiftor_matcher_restorable.hpp:29-67 (a7)
 * The inner loops themselves are SeqAn2's (third-party, absent from /root/reference);
 * they are restated here from the published algorithms and validated against the
 * definitional oracles in this file plus the reference's golden vectors.
 */
#include "spm_oracle.h"
#include <stdlib.h>
#include <string.h>
#include <math.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ------------------------------------------------------------------ */
/* definitional oracles                                                */
/* ------------------------------------------------------------------ */

uint64_t orc_exact_bruteforce(const uint8_t *text, uint64_t n, const uint8_t *pat, uint32_t m,
                              uint64_t *ends, uint64_t cap)
{
    uint64_t cnt = 0;
    if (m == 0 || n < m) return 0;
    for (uint64_t i = 0; i + m <= n; ++i) {
        uint32_t j = 0;
        while (j < m && text[i + j] == pat[j]) ++j;
        if (j == m) {
            if (ends && cnt < cap) ends[cnt] = i + m;
            ++cnt;
        }
    }
    return cnt;
}

uint64_t orc_sellers(const uint8_t *text, uint64_t n, const uint8_t *pat, uint32_t m, uint32_t k,
                     uint64_t *ends, uint32_t *errs, uint64_t cap)
{
    /* column-wise DP, D[0][j] = 0 (free start in the text), D[i][-1] = i */
    uint64_t cnt = 0;
    if (m == 0) return 0;
    uint32_t *col = (uint32_t *)malloc((size_t)(m + 1) * sizeof(uint32_t));
    for (uint32_t i = 0; i <= m; ++i) col[i] = i;
    for (uint64_t j = 0; j < n; ++j) {
        uint32_t diag = col[0]; /* D[i-1][j-1] */
        col[0] = 0;
        uint8_t c = text[j];
        for (uint32_t i = 1; i <= m; ++i) {
            uint32_t up = col[i - 1];   /* D[i-1][j]  (already updated) */
            uint32_t left = col[i];     /* D[i][j-1] */
            uint32_t v = diag + (pat[i - 1] != c);
            if (up + 1 < v) v = up + 1;
            if (left + 1 < v) v = left + 1;
            diag = left;
            col[i] = v;
        }
        if (col[m] <= k) {
            if (cnt < cap) {
                if (ends) ends[cnt] = j + 1;
                if (errs) errs[cnt] = col[m];
            }
            ++cnt;
        }
    }
    free(col);
    return cnt;
}

/* ------------------------------------------------------------------ */
/* Horspool (a8)                                                       */
/* ------------------------------------------------------------------ */

uint64_t orc_horspool(const uint8_t *text, uint64_t n, const uint8_t *pat, uint32_t m,
                      uint64_t *ends, uint64_t cap)
{
    uint64_t cnt = 0;
    if (m == 0 || n < m) return 0;
    /* shift[c] = distance from the last occurrence of c in pat[0..m-2] to the end; m if absent */
    uint32_t shift[256];
    for (int c = 0; c < 256; ++c) shift[c] = m;
    for (uint32_t i = 0; i + 1 < m; ++i) shift[pat[i]] = m - 1 - i;
    uint64_t pos = 0;
    const uint64_t last = n - m;
    while (pos <= last) {
        /* compare the window right to left */
        int64_t j = (int64_t)m - 1;
        while (j >= 0 && text[pos + (uint64_t)j] == pat[j]) --j;
        if (j < 0) {
            if (ends && cnt < cap) ends[cnt] = pos + m;
            ++cnt;
        }
        /* on a mismatch AND after a hit: shift by the table entry of the window's last symbol */
        pos += shift[text[pos + m - 1]];
    }
    return cnt;
}

/* ------------------------------------------------------------------ */
/* Shift-Or, 32-bit blocks (a9 / a7)                                   */
/* ------------------------------------------------------------------ */

int orc_shiftor_init(orc_shiftor_t *s, const uint8_t *pat, uint32_t m)
{
    if (m == 0 || m > 32u * 32u) return -1;
    memset(s, 0, sizeof(*s));
    s->m = m;
    s->blocks = (m + 31) / 32;
    for (int c = 0; c < ORC_SIGMA; ++c)
        for (uint32_t b = 0; b < s->blocks; ++b) s->mask[c][b] = 0xffffffffu;
    for (uint32_t i = 0; i < m; ++i) s->mask[pat[i] & (ORC_SIGMA - 1)][i / 32] &= ~(1u << (i % 32));
    for (uint32_t b = 0; b < s->blocks; ++b) s->d[b] = 0xffffffffu;
    return 0;
}

uint64_t orc_shiftor_run(orc_shiftor_t *s, const uint8_t *text, uint64_t n, uint64_t base,
                         uint64_t *ends, uint64_t cap)
{
    uint64_t cnt = 0;
    const uint32_t nb = s->blocks;
    const uint32_t hit_bit = 1u << ((s->m - 1) % 32);
    if (nb == 1) { /* small needle */
        uint32_t d = s->d[0];
        const uint32_t *b0 = &s->mask[0][0];
        for (uint64_t j = 0; j < n; ++j) {
            d = (d << 1) | b0[(text[j] & (ORC_SIGMA - 1)) * 32];
            if (!(d & hit_bit)) {
                if (ends && cnt < cap) ends[cnt] = base + j + 1;
                ++cnt;
            }
        }
        s->d[0] = d;
        return cnt;
    }
    /* large needle: shift with carry across blocks (a 0 shifts into block 0) */
    for (uint64_t j = 0; j < n; ++j) {
        const uint32_t *mk = s->mask[text[j] & (ORC_SIGMA - 1)];
        uint32_t carry = 0;
        for (uint32_t b = 0; b < nb; ++b) {
            uint32_t v = s->d[b];
            uint32_t nc = v >> 31;
            s->d[b] = ((v << 1) | carry) | mk[b];
            carry = nc;
        }
        if (!(s->d[nb - 1] & hit_bit)) {
            if (ends && cnt < cap) ends[cnt] = base + j + 1;
            ++cnt;
        }
    }
    return cnt;
}

uint64_t orc_shiftor(const uint8_t *text, uint64_t n, const uint8_t *pat, uint32_t m,
                     uint64_t *ends, uint64_t cap)
{
    orc_shiftor_t s;
    if (orc_shiftor_init(&s, pat, m)) return 0;
    return orc_shiftor_run(&s, text, n, 0, ends, cap);
}

/* ------------------------------------------------------------------ */
/* Myers bit-vector (a10 / a6)                                         */
/* ------------------------------------------------------------------ */

int orc_myers_init(orc_myers_t *s, const uint8_t *pat, uint32_t m, uint32_t k)
{
    if (m == 0 || m > 64u * ORC_MAX_BLOCKS) return -1;
    memset(s, 0, sizeof(*s));
    s->m = m;
    s->k = k;
    s->blocks = (m + 63) / 64;
    s->last_rows = m - 64 * (s->blocks - 1);
    for (uint32_t i = 0; i < m; ++i) s->peq[pat[i] & (ORC_SIGMA - 1)][i / 64] |= 1ull << (i % 64);
    /* _patternInit: VP = all ones, VN = 0, score = rows above */
    for (uint32_t b = 0; b < s->blocks; ++b) {
        s->vp[b] = ~0ull;
        s->vn[b] = 0;
        s->score[b] = (b + 1 < s->blocks) ? 64ll * (b + 1) : (int64_t)m;
    }
    /* Ukkonen: only blocks that can hold a cell <= k are active */
    uint32_t need = (k + 1 + 63) / 64;
    if (need > s->blocks) need = s->blocks;
    if (need == 0) need = 1;
    s->last_block = (int32_t)need - 1;
    return 0;
}

/* One block, one text symbol.  hin/hout in {-1,0,+1}: horizontal delta entering at the top
 * row / leaving at row 'out_bit'.  (Myers 1999, block formulation.) */
static inline int myers_block(uint64_t *vp, uint64_t *vn, uint64_t eq, int hin, uint64_t out_bit)
{
    uint64_t pv = *vp, mv = *vn;
    uint64_t hneg = (hin < 0) ? 1ull : 0ull;
    uint64_t xv = eq | mv;
    eq |= hneg;
    uint64_t xh = (((eq & pv) + pv) ^ pv) | eq;
    uint64_t ph = mv | ~(xh | pv);
    uint64_t mh = pv & xh;
    int hout = 0;
    if (ph & out_bit) hout = 1;
    else if (mh & out_bit) hout = -1;
    ph <<= 1;
    mh <<= 1;
    mh |= hneg;
    if (hin > 0) ph |= 1ull;
    *vp = mh | ~(xv | ph);
    *vn = ph & xv;
    return hout;
}

uint64_t orc_myers_run(orc_myers_t *s, const uint8_t *text, uint64_t n, uint64_t base,
                       uint64_t *ends, uint32_t *errs, uint64_t cap)
{
    uint64_t cnt = 0;
    const uint32_t nb = s->blocks;
    const int64_t k = (int64_t)s->k;
    const uint64_t last_bit = 1ull << (s->last_rows - 1);

    if (nb == 1) { /* small pattern: one word, no cut-off */
        uint64_t vp = s->vp[0], vn = s->vn[0];
        int64_t score = s->score[0];
        for (uint64_t j = 0; j < n; ++j) {
            uint64_t eq = s->peq[text[j] & (ORC_SIGMA - 1)][0];
            uint64_t x = eq | vn;
            uint64_t d0 = (((x & vp) + vp) ^ vp) | x;
            uint64_t hn = vp & d0;
            uint64_t hp = vn | ~(vp | d0);
            if (hp & last_bit) ++score;
            else if (hn & last_bit) --score;
            uint64_t xs = hp << 1;
            vn = xs & d0;
            vp = (hn << 1) | ~(xs | d0);
            if (score <= k) {
                if (cnt < cap) {
                    if (ends) ends[cnt] = base + j + 1;
                    if (errs) errs[cnt] = (uint32_t)score;
                }
                ++cnt;
            }
        }
        s->vp[0] = vp;
        s->vn[0] = vn;
        s->score[0] = score;
        return cnt;
    }

    int32_t lb = s->last_block;
    for (uint64_t j = 0; j < n; ++j) {
        const uint64_t *eq = s->peq[text[j] & (ORC_SIGMA - 1)];
        int hin = 0, hout = 0;
        for (int32_t b = 0; b <= lb; ++b) {
            uint64_t ob = ((uint32_t)b + 1 == nb) ? last_bit : (1ull << 63);
            hout = myers_block(&s->vp[b], &s->vn[b], eq[b], hin, ob);
            s->score[b] += hout;
            hin = hout;
        }
        /* grow the active band by one block if the cell diagonally above its top can be <= k */
        if ((uint32_t)lb + 1 < nb && s->score[lb] - hout <= k && ((eq[lb + 1] & 1ull) || hout < 0)) {
            ++lb;
            s->vp[lb] = ~0ull;
            s->vn[lb] = 0;
            int64_t rows = ((uint32_t)lb + 1 == nb) ? (int64_t)s->last_rows : 64;
            uint64_t ob = ((uint32_t)lb + 1 == nb) ? last_bit : (1ull << 63);
            int nh = myers_block(&s->vp[lb], &s->vn[lb], eq[lb], hout, ob);
            s->score[lb] = s->score[lb - 1] - hout + rows + nh;
            hout = nh;
        }
        /* shrink it while the whole last block is out of reach */
        for (;;) {
            int64_t rows = ((uint32_t)lb + 1 == nb) ? (int64_t)s->last_rows : 64;
            if (lb > 0 && s->score[lb] >= k + rows) --lb;
            else break;
        }
        if ((uint32_t)lb + 1 == nb && s->score[lb] <= k) {
            if (cnt < cap) {
                if (ends) ends[cnt] = base + j + 1;
                if (errs) errs[cnt] = (uint32_t)s->score[lb];
            }
            ++cnt;
        }
    }
    s->last_block = lb;
    return cnt;
}

uint64_t orc_myers(const uint8_t *text, uint64_t n, const uint8_t *pat, uint32_t m, uint32_t k,
                   uint64_t *ends, uint32_t *errs, uint64_t cap)
{
    orc_myers_t *s = (orc_myers_t *)malloc(sizeof(orc_myers_t));
    if (orc_myers_init(s, pat, m, k)) { free(s); return 0; }
    uint64_t c = orc_myers_run(s, text, n, 0, ends, errs, cap);
    free(s);
    return c;
}

/* ------------------------------------------------------------------ */
/* multi-threaded driver                                               */
/* ------------------------------------------------------------------ */

void orc_set_threads(int n)
{
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

int orc_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

typedef struct { uint64_t *e; uint32_t *r; uint64_t n, cap; } hitvec_t;
static void hv_push(hitvec_t *v, uint64_t e, uint32_t r)
{
    if (v->n == v->cap) {
        v->cap = v->cap ? v->cap * 2 : 1024;
        v->e = (uint64_t *)realloc(v->e, v->cap * sizeof(uint64_t));
        v->r = (uint32_t *)realloc(v->r, v->cap * sizeof(uint32_t));
    }
    v->e[v->n] = e;
    v->r[v->n] = r;
    ++v->n;
}

/* Each thread owns the hits whose last symbol index j lies in [a,b); it starts scanning
 * window_size-1 symbols earlier from the initial state and drops hits ending before a
 * ("long-context" row). */
uint64_t orc_search_mt(int alg, const uint8_t *text, uint64_t n, const uint8_t *pat, uint32_t m,
                       uint32_t k, int threads, uint64_t *ends, uint32_t *errs, uint64_t cap)
{
    if (threads < 1) threads = 1;
    if (m == 0 || n == 0) return 0;
    const uint64_t overlap = (alg == 3) ? (uint64_t)m + k - 1 : (uint64_t)m - 1;
    const uint64_t chunk = (n + (uint64_t)threads - 1) / (uint64_t)threads;
    hitvec_t *per = (hitvec_t *)calloc((size_t)threads, sizeof(hitvec_t));
    uint64_t *counts = (uint64_t *)calloc((size_t)threads, sizeof(uint64_t));
    const int store = (ends != NULL);

#ifdef _OPENMP
#pragma omp parallel for num_threads(threads) schedule(static, 1)
#endif
    for (int t = 0; t < threads; ++t) {
        uint64_t a = (uint64_t)t * chunk;
        uint64_t b = a + chunk;
        if (a >= n) continue;
        if (b > n) b = n;
        uint64_t s0 = (a > overlap) ? a - overlap : 0;
        uint64_t len = b - s0;
        /* chunk-local scan in steps of 2^20 symbols; the local buffer holds one hit per symbol of a step, so a
         * hit-dense text (every position a hit) is stored completely */
        const uint64_t LB = 1 << 20;
        uint64_t *le = (uint64_t *)malloc(LB * sizeof(uint64_t));
        uint32_t *lr = (uint32_t *)malloc(LB * sizeof(uint32_t));
        uint64_t cnt = 0;
        if (alg == 3) {
            orc_myers_t *ms = (orc_myers_t *)malloc(sizeof(orc_myers_t));
            orc_myers_init(ms, pat, m, k);
            uint64_t off = 0;
            while (off < len) {
                uint64_t step = len - off < (1u << 20) ? len - off : (1u << 20);
                uint64_t c = orc_myers_run(ms, text + s0 + off, step, s0 + off, le, lr, LB);
                uint64_t lim = c < LB ? c : LB;
                for (uint64_t i = 0; i < lim; ++i)
                    if (le[i] > a) { ++cnt; if (store) hv_push(&per[t], le[i], lr[i]); }
                if (c > LB) cnt += c - LB; /* cannot happen: at most one hit per symbol */
                off += step;
            }
            free(ms);
        } else if (alg == 2) {
            orc_shiftor_t ss;
            orc_shiftor_init(&ss, pat, m);
            uint64_t off = 0;
            while (off < len) {
                uint64_t step = len - off < (1u << 20) ? len - off : (1u << 20);
                uint64_t c = orc_shiftor_run(&ss, text + s0 + off, step, s0 + off, le, LB);
                uint64_t lim = c < LB ? c : LB;
                for (uint64_t i = 0; i < lim; ++i)
                    if (le[i] > a) { ++cnt; if (store) hv_push(&per[t], le[i], 0); }
                if (c > LB) cnt += c - LB;
                off += step;
            }
        } else {
            /* Horspool is stateless: windows starting in [s0, b-m], taken in steps of 2^20 start positions */
            uint64_t off = 0;
            while (off + m <= len) {
                uint64_t span = len - off < (1u << 20) + (uint64_t)m - 1 ? len - off : (1u << 20) + (uint64_t)m - 1;
                uint64_t c = orc_horspool(text + s0 + off, span, pat, m, le, LB);
                uint64_t lim = c < LB ? c : LB;
                for (uint64_t i = 0; i < lim; ++i)
                    if (le[i] + s0 + off > a) { ++cnt; if (store) hv_push(&per[t], le[i] + s0 + off, 0); }
                if (c > LB) cnt += c - LB;
                off += 1u << 20;
            }
        }
        free(le);
        free(lr);
        counts[t] = cnt;
    }
    uint64_t total = 0;
    for (int t = 0; t < threads; ++t) {
        if (store)
            for (uint64_t i = 0; i < per[t].n; ++i) {
                if (total + i < cap) {
                    ends[total + i] = per[t].e[i];
                    if (errs) errs[total + i] = per[t].r[i];
                }
            }
        total += store ? per[t].n : counts[t];
        free(per[t].e);
        free(per[t].r);
    }
    free(per);
    free(counts);
    return total;
}

/* ------------------------------------------------------------------ */
/* journaled sequence                                                  */
/* ------------------------------------------------------------------ */

uint64_t orc_journal_apply(const uint8_t *ref, uint64_t n, const orc_event_t *ev, uint64_t n_ev,
                           const uint8_t *pool, uint8_t *out, uint64_t out_cap)
{
    /* pass 1 (serial, cheap): validate and find the derived start of every event */
    uint64_t o = 0, r = 0;
    uint64_t *der = out ? (uint64_t *)malloc((n_ev + 1) * sizeof(uint64_t)) : NULL;
    for (uint64_t i = 0; i < n_ev; ++i) {
        uint64_t p = ev[i].ref_pos;
        if (p < r || p > n || p + ev[i].del_len > n) { free(der); return (uint64_t)-1; } /* unsorted / overlapping */
        o += p - r;
        if (der) der[i] = o;
        o += ev[i].ins_len;
        r = p + ev[i].del_len;
    }
    uint64_t total = o + (n - r);
    if (!out) return total;
    if (total > out_cap) { free(der); return total; }
    /* pass 2 (parallel): copy the reference slice before each event and its insertion */
#ifdef _OPENMP
#pragma omp parallel for schedule(static)
#endif
    for (int64_t i = 0; i <= (int64_t)n_ev; ++i) {
        uint64_t prev_end = (i == 0) ? 0 : ev[i - 1].ref_pos + ev[i - 1].del_len;
        uint64_t prev_out = (i == 0) ? 0 : der[i - 1] + ev[i - 1].ins_len;
        uint64_t upto = (i == (int64_t)n_ev) ? n : ev[i].ref_pos;
        memcpy(out + prev_out, ref + prev_end, upto - prev_end);
        if (i < (int64_t)n_ev && ev[i].ins_len) memcpy(out + der[i], pool + ev[i].ins_off, ev[i].ins_len);
    }
    free(der);
    return total;
}

/* ------------------------------------------------------------------ */
/* journaled string tree: per-haplotype materialise + search           */
/* ------------------------------------------------------------------ */

static uint64_t jst_collect(const orc_event_t *var, uint64_t n_var, const uint64_t *cov, uint32_t wpr,
                            uint32_t hap, orc_event_t *out)
{
    uint64_t c = 0;
    const uint64_t bit = 1ull << (hap & 63);
    const uint32_t w = hap >> 6;
    for (uint64_t v = 0; v < n_var; ++v)
        if (cov[v * wpr + w] & bit) { if (out) out[c] = var[v]; ++c; }
    return c;
}

uint64_t orc_jst_haplotype(const uint8_t *ref, uint64_t n, const orc_event_t *var, uint64_t n_var,
                           const uint8_t *pool, const uint64_t *cov, uint32_t wpr, uint32_t hap,
                           uint8_t *out, uint64_t out_cap)
{
    uint64_t c = jst_collect(var, n_var, cov, wpr, hap, NULL);
    orc_event_t *ev = (orc_event_t *)malloc((c + 1) * sizeof(orc_event_t));
    jst_collect(var, n_var, cov, wpr, hap, ev);
    uint64_t len = orc_journal_apply(ref, n, ev, c, pool, out, out_cap);
    free(ev);
    return len;
}

static int jst_hit_cmp(const void *a, const void *b)
{
    const orc_jst_hit_t *x = (const orc_jst_hit_t *)a, *y = (const orc_jst_hit_t *)b;
    if (x->hap != y->hap) return x->hap < y->hap ? -1 : 1;
    if (x->end != y->end) return x->end < y->end ? -1 : 1;
    return 0;
}

uint64_t orc_jst_search(int alg, const uint8_t *ref, uint64_t n, const orc_event_t *var, uint64_t n_var,
                        const uint8_t *pool, const uint64_t *cov, uint32_t n_hap, uint32_t wpr,
                        const uint32_t *haps, uint32_t n_list, const uint8_t *pat, uint32_t m, uint32_t k,
                        int threads, orc_jst_hit_t *out, uint64_t cap)
{
    const uint32_t count = haps ? n_list : n_hap;
    uint64_t total = 0;
    if (threads < 1) threads = 1;
#ifdef _OPENMP
#pragma omp parallel for num_threads(threads) schedule(dynamic, 1)
#endif
    for (int64_t i = 0; i < (int64_t)count; ++i) {
        const uint32_t h = haps ? haps[i] : (uint32_t)i;
        uint64_t len = orc_jst_haplotype(ref, n, var, n_var, pool, cov, wpr, h, NULL, 0);
        if (len == (uint64_t)-1) continue;
        uint8_t *seq = (uint8_t *)malloc(len + 1);
        orc_jst_haplotype(ref, n, var, n_var, pool, cov, wpr, h, seq, len);
        uint64_t hc = 1024, c;
        uint64_t *ends = NULL;
        uint32_t *errs = NULL;
        for (;;) {
            ends = (uint64_t *)realloc(ends, hc * sizeof(uint64_t));
            errs = (uint32_t *)realloc(errs, hc * sizeof(uint32_t));
            if (alg == 3) c = orc_myers(seq, len, pat, m, k, ends, errs, hc);
            else if (alg == 2) c = orc_shiftor(seq, len, pat, m, ends, hc);
            else c = orc_horspool(seq, len, pat, m, ends, hc);
            if (c <= hc) break;
            hc = c;
        }
#ifdef _OPENMP
#pragma omp critical
#endif
        {
            for (uint64_t j = 0; j < c; ++j) {
                if (total < cap && out) {
                    out[total].end = ends[j];
                    out[total].hap = h;
                    out[total].err = (alg == 3) ? errs[j] : 0;
                }
                ++total;
            }
        }
        free(ends); free(errs); free(seq);
    }
    if (out) qsort(out, total < cap ? total : cap, sizeof(orc_jst_hit_t), jst_hit_cmp);
    return total;
}

/* ------------------------------------------------------------------ */
/* pigeonhole q-gram seed matcher + anchored (prefix) Myers            */
/*                                          */
/* ------------------------------------------------------------------ */

/* q of spm::pigeonhole_matcher{needles, error_rate} (pigeonhole_matcher.hpp:157-180 -> SeqAn2
 * _patternInit(Pattern<TIndex, Pigeonhole<TSpec>>, errorRate), UNVERIFIED-UPSTREAM: per needle
 * delta = len / (floor(rate*len) + 1), deltas < 3 ignored, q = smallest delta, step = q).
 * Pinned by pigeonhole_matcher_test.cpp:50-53 (window_size == 5 for needle GCACG, rate 0) and the
 * multi-needle vector :31-33 (needles of 5 and 10 symbols -> q = 5, seeds at needle offsets 0 and 5). */
uint32_t orc_pigeonhole_q(const uint32_t *lengths, uint32_t n_needles, double error_rate)
{
    uint32_t q = 0xffffffffu;
    for (uint32_t i = 0; i < n_needles; ++i) {
        uint32_t len = lengths[i];
        if (len == 0) continue;
        uint32_t errors = (uint32_t)floor(error_rate * (double)len);
        uint32_t delta = len / (errors + 1);
        if (delta < 3) continue;
        if (delta < q) q = delta;
    }
    return q == 0xffffffffu ? 0 : q;
}

/* definitional: every haystack position whose q-gram equals the q-gram at offset 0, q, 2q, ... of a
 * needle; hit = (begin, needle index, needle offset), in haystack order, ties by (needle, offset)
 * (pigeonhole_matcher.hpp:118-124 position() = {index, offset, count = q}). */
uint64_t orc_pigeonhole(const uint8_t *text, uint64_t n, const uint8_t *needles, const uint32_t *lengths,
                        uint32_t n_needles, double error_rate, uint64_t *begins, uint32_t *needle_idx,
                        uint32_t *needle_off, uint64_t cap)
{
    const uint32_t q = orc_pigeonhole_q(lengths, n_needles, error_rate);
    uint64_t cnt = 0;
    if (q == 0 || n < q) return 0;
    for (uint64_t p = 0; p + q <= n; ++p) {
        uint64_t base = 0;
        for (uint32_t i = 0; i < n_needles; ++i) {
            for (uint32_t o = 0; o + q <= lengths[i]; o += q) {
                if (memcmp(text + p, needles + base + o, q) == 0) {
                    if (cnt < cap) {
                        if (begins) begins[cnt] = p;
                        if (needle_idx) needle_idx[cnt] = i;
                        if (needle_off) needle_off[cnt] = o;
                    }
                    ++cnt;
                }
            }
            base += lengths[i];
        }
    }
    return cnt;
}

/* definitional: restorable_myers_prefix_matcher (myers_prefix_matcher_restorable.hpp:47-61,117-160;
 * SeqAn2 MyersUkkonenGlobal): edit distance of the needle to every PREFIX text[0..j) -- D[0][j] = j,
 * D[i][0] = i -- scanning at most m + k + 1 symbols (:47-50); end = j, errors = D[m][j] <= k. */
uint64_t orc_myers_prefix(const uint8_t *text, uint64_t n, const uint8_t *pat, uint32_t m, uint32_t k,
                          uint64_t *ends, uint32_t *errs, uint64_t cap)
{
    uint64_t cnt = 0;
    if (m == 0) return 0;
    uint64_t lim = (uint64_t)m + k + 1;
    if (lim > n) lim = n;
    uint32_t *col = (uint32_t *)malloc((size_t)(m + 1) * sizeof(uint32_t));
    for (uint32_t i = 0; i <= m; ++i) col[i] = i;
    for (uint64_t j = 0; j < lim; ++j) {
        uint32_t diag = col[0];
        col[0] = (uint32_t)(j + 1);
        uint8_t c = text[j];
        for (uint32_t i = 1; i <= m; ++i) {
            uint32_t up = col[i - 1], left = col[i];
            uint32_t v = diag + (pat[i - 1] != c);
            if (up + 1 < v) v = up + 1;
            if (left + 1 < v) v = left + 1;
            diag = left;
            col[i] = v;
        }
        if (col[m] <= k) {
            if (cnt < cap) {
                if (ends) ends[cnt] = j + 1;
                if (errs) errs[cnt] = col[m];
            }
            ++cnt;
        }
    }
    free(col);
    return cnt;
}
