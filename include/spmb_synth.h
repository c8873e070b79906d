/*
 * spmb_synth.h -- the ONE synthetic-input generator of this repository ("xoshiro256**
 * seeded via splitmix64 ... one generator implementation in C++ shared by oracle and GPU harness").
 *
 * Harness code, not product code: built as libspm_b200/lib/libspmsynth.so (plain C++, no CUDA), used by
 * bench.py, tools/ and tests/ for BOTH sides of every comparison -- the CUDA path and the CPU oracle are
 * always handed the same bytes.  Nothing in libspmb.so depends on it.
 *
 * Everything is counter-based on top of xoshiro256**: a generator state is derived with splitmix64 from
 * (seed, stream, block), so any range of any stream can be produced independently (in parallel, on any
 * rank, on the GPU box or in the CPU container) and is bit-identical everywhere.  Only integer arithmetic
 * and IEEE double +,-,* with contraction off: no libm call can make two machines disagree.
 *
 * Symbols are dna4 ranks 0..3, one byte each (libspm/libspm/seqan/alphabet.hpp:34-83,118).
 */
#ifndef SPMB_SYNTH_H
#define SPMB_SYNTH_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* same record as spmb_journal_event (include/spmb.h) */
typedef struct spms_event {
    uint64_t ref_pos;
    uint32_t del_len;
    uint32_t ins_len;
    uint64_t ins_off;
} spms_event;

/* streams of one seed (any other value is free for callers) */
enum {
    SPMS_STREAM_TEXT = 0,     /* spms_text */
    SPMS_STREAM_JOURNAL = 1,  /* spms_journal */
    SPMS_STREAM_JST = 2,      /* spms_jst: positions / kinds / lengths */
    SPMS_STREAM_COVERAGE = 3, /* spms_jst: one block per variant */
    SPMS_STREAM_POOL = 4,     /* inserted symbols */
    SPMS_STREAM_USER = 16     /* first stream id for harness draws (offsets, edits) */
};

/* stateless draw number `index` of (seed, stream) */
uint64_t spms_u64(uint64_t seed, uint64_t stream, uint64_t index);
/* out[i] = lo + draw(index0 + i) mod-free mapped to [lo, hi)   (hi > lo) */
void spms_uniform(uint64_t seed, uint64_t stream, uint64_t index0, uint64_t n, uint64_t lo, uint64_t hi, uint64_t *out);

/* iid uniform ranks of text positions [begin, begin + n) of stream `seed` (any range, same bytes) */
void spms_text(uint64_t seed, uint64_t begin, uint64_t n, uint8_t *out);

/* e random edits (substitution to a different symbol / insertion / deletion, equiprobable) of pat[0, m);
 * out has room for m + e symbols; returns the new length.  Draws come from (seed, stream). */
uint32_t spms_mutate(uint64_t seed, uint64_t stream, const uint8_t *pat, uint32_t m, uint32_t e, uint8_t *out);

/* journal of configs[1]: events at sorted, non-overlapping positions with mean gap
 * 1/density, p_snp SNPs (alt != ref[pos]), p_ins insertions U[1, max_indel], the rest deletions
 * U[1, max_indel].  ref = the n_ref reference ranks.  Returns the number of events (<= cap_ev), writes the
 * pool length; kinds[i] = 0 SNP, 1 insertion, 2 deletion (may be NULL).  Call with ev == NULL to size. */
uint64_t spms_journal(uint64_t seed, const uint8_t *ref, uint64_t n_ref, double density, double p_snp, double p_ins,
                      uint32_t max_indel, spms_event *ev, uint64_t cap_ev, uint8_t *pool, uint64_t cap_pool,
                      uint64_t *pool_len, uint8_t *kinds);

/* journaled string tree of configs[3] (C4): n_var variants at sorted non-overlapping positions (p_snp SNP,
 * p_ins insertion, rest deletion, lengths U[1, max_indel]); allele frequency ~ truncated 1/x on
 * [1/n_hap, af_max]; membership Bernoulli(AF) per haplotype, every variant carried at least once.
 * cov: n_var rows of ceil(n_hap/64) words.  Returns the number of variants (<= n_var). */
uint64_t spms_jst(uint64_t seed, const uint8_t *ref, uint64_t n_ref, uint64_t n_var, uint32_t n_hap, double p_snp,
                  double p_ins, uint32_t max_indel, double af_max, spms_event *var, uint8_t *pool, uint64_t cap_pool,
                  uint64_t *pool_len, uint8_t *kinds, uint64_t *cov);

/* threads used by the bulk generators (0 = all) */
void spms_set_threads(int n);

#ifdef __cplusplus
}
#endif
#endif
