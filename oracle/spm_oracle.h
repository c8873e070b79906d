/*
 * spm_oracle.h -- CPU ORACLE for the libspm online pattern-matching hot path.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load it.  The product
 * (libspm_b200/) never links, imports or calls anything in oracle/.
 *
 * Parity status: PINNED (thinly) for the single-word paths by the reference's own
 * golden vectors (test/api/libspm/matcher/{horspool,shiftor,myers,myers_matcher_restorable}_test.cpp),
 * which tests/test_oracle_golden.py replays; UNPINNED by any reference test for the
 * multi-word paths, overlapping exact hits, error values and journaled/JST inputs --
 * those are closed by the definitional oracles below (brute force / Sellers DP), see DESIGN.md.
 *
 * The arithmetic of the reference lives in SeqAn2 (rrahn/seqan @ 7a8ef3ce..., not vendored,
 * not on this machine: cmake/package-lock.cmake:20-28).  The functions here restate the
 * published algorithms (Horspool 1980; Baeza-Yates/Gonnet Shift-Or 1992; Myers 1999 with
 * Hyyro's formulation and the Ukkonen block cut-off) behind the call shape of
 * libspm/libspm/matcher/seqan_pattern_base.hpp:40-52: scan the haystack left to right,
 * report every hit in ascending order.
 *
 * Symbols are 1 byte each holding the dna4 rank (A=0,C=1,G=2,T=3), exactly like a
 * std::vector<spm::dna4> (libspm/libspm/seqan/alphabet.hpp:34-83,118).
 * Positions: "end" is EXCLUSIVE (seqan2::endPosition); exact begin = end - m.
 */
#ifndef SPM_ORACLE_H
#define SPM_ORACLE_H
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_MAX_BLOCKS 128 /* 64-bit blocks: patterns up to 8192 symbols */

/* ---- definitional oracles (authoritative semantics) ---- */
/* every i with text[i..i+m) == pat; writes end=i+m.  Returns hit count (may exceed cap). */
uint64_t orc_exact_bruteforce(const uint8_t *text, uint64_t n, const uint8_t *pat, uint32_t m,
                              uint64_t *ends, uint64_t cap);
/* Sellers DP with free text start: every j with min_s ed(pat, text[s..j]) <= k; end=j+1. */
uint64_t orc_sellers(const uint8_t *text, uint64_t n, const uint8_t *pat, uint32_t m, uint32_t k,
                     uint64_t *ends, uint32_t *errs, uint64_t cap);

/* ---- algorithmic restatements (these are also the timed CPU baseline) ---- */
/* a8: Horspool, all (overlapping) occurrences. */
uint64_t orc_horspool(const uint8_t *text, uint64_t n, const uint8_t *pat, uint32_t m,
                      uint64_t *ends, uint64_t cap);

/* a9: Shift-Or, 32-bit blocks like SeqAn2 (blockCount = ceil(m/32)); resumable state =
 * prefSufMatch (libspm/libspm/matcher/shiftor_matcher_restorable.hpp:29,44-50). */
/* symbols are 1-byte ranks; 16 covers spm::dna4 (4), spm::dna5 (5) and spm::dna15 (15)
 * (libspm/libspm/seqan/alphabet.hpp:118-120) */
#define ORC_SIGMA 16
typedef struct {
    uint32_t m, blocks;
    uint32_t mask[ORC_SIGMA][32]; /* B[c][block] : bit clear where pat[i]==c */
    uint32_t d[32];             /* state D (prefSufMatch) */
} orc_shiftor_t;
int      orc_shiftor_init(orc_shiftor_t *s, const uint8_t *pat, uint32_t m);
/* scans a chunk, continuing from the state; reported ends are chunk-local + base. */
uint64_t orc_shiftor_run(orc_shiftor_t *s, const uint8_t *text, uint64_t n, uint64_t base,
                         uint64_t *ends, uint64_t cap);
uint64_t orc_shiftor(const uint8_t *text, uint64_t n, const uint8_t *pat, uint32_t m,
                     uint64_t *ends, uint64_t cap);

/* a10/a6: Myers bit-vector k-error search, 64-bit blocks, small (m<=64) and blocked large
 * path with Ukkonen cut-off; resumable state = whole pattern state
 * (libspm/libspm/matcher/myers_matcher_restorable.hpp:57-63). */
typedef struct {
    uint32_t m, k, blocks;
    uint32_t last_rows;                    /* rows in the last block (1..64) */
    uint64_t peq[ORC_SIGMA][ORC_MAX_BLOCKS];
    /* --- state (capture/restore copies from here on) --- */
    uint64_t vp[ORC_MAX_BLOCKS], vn[ORC_MAX_BLOCKS];
    int64_t  score[ORC_MAX_BLOCKS];        /* D at the last row of each block */
    int32_t  last_block;                   /* Ukkonen: last active block */
} orc_myers_t;
int      orc_myers_init(orc_myers_t *s, const uint8_t *pat, uint32_t m, uint32_t k);
uint64_t orc_myers_run(orc_myers_t *s, const uint8_t *text, uint64_t n, uint64_t base,
                       uint64_t *ends, uint32_t *errs, uint64_t cap);
uint64_t orc_myers(const uint8_t *text, uint64_t n, const uint8_t *pat, uint32_t m, uint32_t k,
                   uint64_t *ends, uint32_t *errs, uint64_t cap);

/* ---- multi-threaded drivers (chunks overlapped by window_size-1) ---- */
/* alg: 1 horspool, 2 shiftor, 3 myers.  Returns hit count; hits are NOT stored (timing only)
 * unless ends != NULL (then sorted ascending). */
uint64_t orc_search_mt(int alg, const uint8_t *text, uint64_t n, const uint8_t *pat, uint32_t m,
                       uint32_t k, int threads, uint64_t *ends, uint32_t *errs, uint64_t cap);
int orc_max_threads(void);
void orc_set_threads(int n); /* launchers like torchrun export OMP_NUM_THREADS=1 */

/* ---- journaled sequence (a14; semantics of specs/journaled_sequence_class_diagram.drawio:196,444) ---- */
/* events sorted by ref_pos, non-overlapping: replace ref[ref_pos, ref_pos+del_len) by
 * pool[ins_off, ins_off+ins_len).  Returns derived length; writes it if out != NULL. */
typedef struct { uint64_t ref_pos; uint32_t del_len; uint32_t ins_len; uint64_t ins_off; } orc_event_t;
uint64_t orc_journal_apply(const uint8_t *ref, uint64_t n, const orc_event_t *ev, uint64_t n_ev,
                           const uint8_t *pool, uint8_t *out, uint64_t out_cap);

/* ---- journaled string tree (a14): oracle = materialise every haplotype, then plain search ----
 * var[]: n_var variants sorted by ref_pos (same record as a journal event); cov: n_var rows of
 * words_per_row 64-bit words, bit h of row v set = haplotype h carries variant v.  A haplotype's
 * carried variants must not overlap.  haps == NULL means all n_hap haplotypes, else the listed ones.
 * Hits (haplotype, end in that haplotype's coordinates, errors), sorted by (hap, end). */
typedef struct { uint64_t end; uint32_t hap; uint32_t err; } orc_jst_hit_t;
uint64_t orc_jst_search(int alg, const uint8_t *ref, uint64_t n, const orc_event_t *var, uint64_t n_var,
                        const uint8_t *pool, const uint64_t *cov, uint32_t n_hap, uint32_t words_per_row,
                        const uint32_t *haps, uint32_t n_list, const uint8_t *pat, uint32_t m, uint32_t k,
                        int threads, orc_jst_hit_t *out, uint64_t cap);
/* one haplotype's sequence; returns its length (out may be NULL for a size query) */
uint64_t orc_jst_haplotype(const uint8_t *ref, uint64_t n, const orc_event_t *var, uint64_t n_var,
                           const uint8_t *pool, const uint64_t *cov, uint32_t words_per_row, uint32_t hap,
                           uint8_t *out, uint64_t out_cap);

/* ---- "next" rows of SURVEY.md 8(f): pigeonhole seed matcher, anchored (prefix) Myers ---- */
uint32_t orc_pigeonhole_q(const uint32_t *lengths, uint32_t n_needles, double error_rate);
uint64_t orc_pigeonhole(const uint8_t *text, uint64_t n, const uint8_t *needles, const uint32_t *lengths,
                        uint32_t n_needles, double error_rate, uint64_t *begins, uint32_t *needle_idx,
                        uint32_t *needle_off, uint64_t cap);
uint64_t orc_myers_prefix(const uint8_t *text, uint64_t n, const uint8_t *pat, uint32_t m, uint32_t k,
                          uint64_t *ends, uint32_t *errs, uint64_t cap);

#ifdef __cplusplus
}
#endif
#endif
