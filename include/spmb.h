/*
 * spmb.h -- C ABI of the B200-native online pattern-matching path (libspmb.so).
 *
 * The reference (rrahn/libspm) is a header-only C++20 template library with NO plugin/FFI
 * boundary of its own; this header is the boundary a maintainer binds:
 * every entry point names the reference interface it stands in for.  The C++ facade
 * include/libspm_b200/spm.hpp re-creates the reference's spm:: names on top of it.
 *
 * Conventions: plain pointers and sizes only; symbols are 1 byte each holding the dna4
 * rank 0..3 (libspm/libspm/seqan/alphabet.hpp:34-83,118 -- one spm::dna4 is one byte whose
 * integral value is the rank); positions are 0-based, "end" is EXCLUSIVE like
 * seqan2::endPosition(finder).  All functions return SPMB_OK (0) or an error code; the
 * reference's entry points are noexcept and have no error channel
 * (seqan_pattern_base.hpp:41,54,61,68), so any non-zero code is new information.
 * A context is bound to one device and one CUDA stream and is used from one host thread at
 * a time (a reference matcher is mutable state and not shareable either: concept.hpp:126-130).
 * There is NO CPU fallback: without a CUDA device every call fails with SPMB_ERR_NO_DEVICE.
 */
#ifndef SPMB_H
#define SPMB_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SPMB_VERSION 100

enum {
    SPMB_OK = 0,
    SPMB_ERR_ARG = 1,         /* null pointer, bad size, unsorted journal ... */
    SPMB_ERR_CUDA = 2,        /* a CUDA runtime call failed: see spmb_last_error() */
    SPMB_ERR_OOM = 3,
    SPMB_ERR_UNSUPPORTED = 4, /* pattern too long for the algorithm / store geometry */
    SPMB_ERR_NO_DEVICE = 5
};

/* algorithms: the three searchers of the hot path */
enum {
    SPMB_ALG_HORSPOOL = 1, /* spm::horspool_matcher  libspm/libspm/matcher/horspool_matcher.hpp:20-44 */
    SPMB_ALG_SHIFTOR = 2,  /* spm::shiftor_matcher   libspm/libspm/matcher/shiftor_matcher.hpp:20-44  */
    SPMB_ALG_MYERS = 3,    /* spm::myers_matcher     libspm/libspm/matcher/myers_matcher.hpp:19-60     */
    /* SURVEY.md 8(f) "next" rows */
    SPMB_ALG_PIGEONHOLE = 4,     /* spm::pigeonhole_matcher  libspm/libspm/matcher/pigeonhole_matcher.hpp:129-243
                                    (create with spmb_pigeonhole_create) */
    SPMB_ALG_MYERS_FILTERED = 5, /* the hit list of SPMB_ALG_MYERS, computed as pigeonhole seed filter +
                                    windowed Myers verification (no reference counterpart: the reference
                                    leaves verification to the callers of pigeonhole_matcher) */
    SPMB_ALG_MYERS_PREFIX = 6    /* spm::restorable_myers_prefix_matcher
                                    libspm/libspm/matcher/myers_prefix_matcher_restorable.hpp:117-160 */
};

typedef struct spmb_context spmb_context;
typedef struct spmb_store spmb_store;
typedef struct spmb_pattern spmb_pattern;
typedef struct spmb_hits spmb_hits;
typedef struct spmb_stream spmb_stream;

/* One hit = what the reference's callback reads off the finder, plus what it hides:
 * seqan2::endPosition(finder) -> end; seqan2::beginPosition(finder) -> end - m (exact only);
 * errors is the edit distance of the best alignment ending there (the reference exposes it
 * only through capture().errors, myers_matcher_restorable.hpp:57-59).
 * Sorted by (pattern_id, seq_id, end) when handed back. */
typedef struct spmb_hit {
    uint64_t end;        /* exclusive end, in the coordinates of sequence seq_id */
    uint32_t seq_id;     /* sequence / haplotype id (0 for a plain store) */
    uint32_t pattern_id; /* index inside a pattern set (0 for a single pattern) */
    uint32_t errors;     /* 0 for the exact searchers */
    uint32_t reserved;   /* SPMB_ALG_PIGEONHOLE: offset of the seed inside its needle
                            (seqan2::PigeonholeSeedOnlyPosition::offset, pigeonhole_matcher.hpp:41-44); else 0 */
} spmb_hit;

/* journal event (a14; specs/journaled_sequence_class_diagram.drawio:196 record_sequence_edit):
 * replace ref[ref_pos, ref_pos+del_len) by pool[ins_off, ins_off+ins_len).
 * Events sorted by ref_pos, non-overlapping.  SNP = (p,1,1,off). */
typedef struct spmb_journal_event {
    uint64_t ref_pos;
    uint32_t del_len;
    uint32_t ins_len;
    uint64_t ins_off;
} spmb_journal_event;

/* ---------------- library / context ---------------- */
int spmb_version(void);
const char *spmb_status_string(int status);
/* stream: a cudaStream_t to enqueue on (NULL = the context creates its own). */
int spmb_context_create(int device, void *stream, spmb_context **out);
int spmb_context_destroy(spmb_context *ctx);
void *spmb_context_stream(const spmb_context *ctx);
const char *spmb_last_error(const spmb_context *ctx);
int spmb_context_sync(spmb_context *ctx);
/* give back what the context caches between calls (journal scratch block, upload threads and their pinned
 * buffers, the stream-ordered pool's free memory); everything is re-created on demand. */
int spmb_context_trim(spmb_context *ctx);
/* Multi-GPU sharding: every later search on this context covers only shard
 * `shard` of `n_shards` of each store it is given (contiguous ranges of super-tiles; a hit
 * belongs to the shard that holds the first symbol of its filter window (exact: a fixed offset from
 * its begin) / its end (Myers).  The shards of one
 * store partition its hit list: ranks search disjoint shards with no exchange, the driver
 * concatenates the small hit lists afterwards.  Default (0, 1) = everything. */
int spmb_context_set_shard(spmb_context *ctx, uint32_t shard, uint32_t n_shards);
/* host threads that pack chunks of a long host haystack for this context's upload pipeline;
 * -1 = default (SPMB_HOST_THREADS, else available cpus - 1).  A process that drives several devices gives
 * every context its share of the cores (the multi-GPU driver below does). */
int spmb_context_set_upload_threads(spmb_context *ctx, int n_threads);
int spmb_device_count(void); /* CUDA devices visible to this process (0: none -- nothing here works without one) */

/* SPMB_ALG_MYERS with needles longer than 32 symbols runs the Ukkonen cut-off form of the column update
 * (what SeqAn2's _findMyersLargePatterns does with `lastBlock`; reached from myers_matcher.hpp:47-49 for
 * such needles): only the 32-row words that can still hold a cell <= k are updated.  Same hit list.
 * enabled = 0 selects the full recurrence over all ceil(m/32) words (the kernel the INT-pipe roofline of
 * DESIGN.md 4 is stated on); default 1, or 0 when the environment has SPMB_MYERS_FULL=1. */
int spmb_context_set_myers_cutoff(spmb_context *ctx, int enabled);

/* ---------------- sequence store (north_star item 1) ----------------
 * Replaces the haystack argument of seqan_pattern_base::operator()
 * (seqan_pattern_base.hpp:40-47, wrapped by spm::seqan_container_adapter,
 * libspm/libspm/seqan/container_adapter.hpp:23-100): instead of viewing host memory per call
 * the sequence is packed once into 2 bits/symbol in HBM ("strided bit planes").
 * row_len = 0 lets the library choose the geometry. */
int spmb_store_create(spmb_context *ctx, const uint8_t *host_ranks, uint64_t n, uint32_t row_len,
                      spmb_store **out);
int spmb_store_create_from_device(spmb_context *ctx, const uint8_t *dev_ranks, uint64_t n,
                                  uint32_t row_len, spmb_store **out);
/* spmb_store_create with a long host haystack (>= SPMB_UPLOAD_MIN symbols, default 32 Mi) uploads in
 * chunks through two movers at once: the copy engine takes raw chunks (1 B/symbol), host threads
 * (SPMB_HOST_THREADS, default: available cpus - 1) pack chunks to 2 bits/symbol and send a quarter of the
 * bytes; both claim from one queue, so the split follows the machine.
 * The event table of spmb_store_create_journaled takes the same queue once the context has uploaded a long
 * haystack: host threads rewrite 24-byte events as 8-byte records where they fit, the device expands them.
 * stats[0] raw chunks, [1] host-packed chunks of the last such upload, [2] symbols per chunk, [3] host threads,
 * [4] raw / [5] host-packed event chunks of the last journal, [6] events per chunk, [7] 0. */
int spmb_upload_last_stats(const spmb_context *ctx, uint64_t stats[8]);
/* the host packing loop itself (no device needed): ranks[0,n) -> out[0, ceil(n/32)), one 64-bit word per
 * 32 symbols = the store's lane-chunk word (uint2 {x,y} in memory: 2-bit codes, first symbol in the top
 * bits of x, symbols 16..31 in y); ranks are masked to 2 bits, symbols past n are 0. */
void spmb_host_pack_chunks(const uint8_t *ranks, uint64_t n, uint64_t *out);
/* the host loop that rewrites a chunk of n 24-byte events as 8-byte records (no device needed; AVX2 where the CPU has
 * it): out = ceil(n/1024) headers {uint64 base_pos, uint64 base_off} followed by n words
 * {dpos:32 | doff:16 | del_len:8 | ins_len:8} (ref_pos - base_pos, ins_off - base_off of the block; an event that
 * inserts nothing stores doff 0 and its offset is only checked against pool_len).  Returns 1, or 0 when some event
 * does not fit this form (the chunk then travels as it is). */
int spmb_host_pack_events(const spmb_journal_event *events, uint32_t n, uint64_t pool_len, uint8_t *out);
/* journaled sequence: base + events, materialised on the device into a new packed store whose
 * coordinates are those of the derived sequence. */
int spmb_store_create_journaled(spmb_context *ctx, const spmb_store *base,
                                const spmb_journal_event *events, uint64_t n_events,
                                const uint8_t *ins_pool, uint64_t pool_len, uint32_t row_len,
                                spmb_store **out);
/* 4-bit alphabets: spm::dna5 (ranks 0..4) and spm::dna15 (ranks 0..14), libspm/libspm/seqan/alphabet.hpp:119-120
 * (seqan2::BitsPerValue = 3 / 4, :107-112).  One symbol = one byte holding the rank, as for dna4; on the device
 * the sequence is TWO plane pairs in the dna4 geometry (bits 1..0 and bits 3..2 of every rank): 0.5 B per bp.
 * bits = 3 or 4 (both stored as 4; 2 forwards to the dna4 entry point).  Such a store is searched with
 * patterns from spmb_pattern_create_wide (SPMB_ALG_HORSPOOL / SHIFTOR / MYERS); mixing widths is SPMB_ERR_ARG;
 * journals, the JST, streams, the seed filter and the prefix matcher stay dna4-only (SPMB_ERR_UNSUPPORTED). */
int spmb_store_create_wide(spmb_context *ctx, const uint8_t *host_ranks, uint64_t n, uint32_t bits,
                           uint32_t row_len, spmb_store **out);
int spmb_store_create_wide_from_device(spmb_context *ctx, const uint8_t *dev_ranks, uint64_t n, uint32_t bits,
                                       uint32_t row_len, spmb_store **out);
uint32_t spmb_store_symbol_bits(const spmb_store *s); /* 2 or 4 */
int spmb_store_destroy(spmb_store *s);
uint64_t spmb_store_length(const spmb_store *s);
uint32_t spmb_store_row_len(const spmb_store *s);
uint64_t spmb_store_device_bytes(const spmb_store *s);
/* unpack [begin, begin+n) back to 1-byte ranks in host memory (round-trip checks). */
int spmb_store_read(spmb_context *ctx, const spmb_store *s, uint64_t begin, uint64_t n,
                    uint8_t *host_out);

/* ---------------- patterns ----------------
 * spmb_pattern_create == the matcher constructors
 *   horspool_matcher{needle} / shiftor_matcher{needle} / myers_matcher{needle, k}
 * (horspool_matcher.hpp:35-40, shiftor_matcher.hpp:35-40, myers_matcher.hpp:37-43).
 * The needle is copied (the reference only views it). */
int spmb_pattern_create(spmb_context *ctx, int alg, const uint8_t *ranks, uint32_t m, uint32_t k,
                        spmb_pattern **out);
/* the same constructors over a 4-bit alphabet (needle ranks 0..15; bits = 3 or 4) */
int spmb_pattern_create_wide(spmb_context *ctx, int alg, const uint8_t *ranks, uint32_t m, uint32_t k,
                             uint32_t bits, spmb_pattern **out);
uint32_t spmb_pattern_symbol_bits(const spmb_pattern *p); /* 2 or 4 */
/* a set of n_patterns needles searched in one pass (config C5); lengths[i] symbols each,
 * concatenated in ranks.  SPMB_ALG_MYERS / SPMB_ALG_MYERS_FILTERED take any mix of lengths up to 8192 symbols
 * (the reference takes any needle, myers_matcher.hpp:37-43): needles are grouped by ceil(len/32) and searched
 * one group per launch; needles longer than 256 symbols run their first 256 symbols through the same kernels
 * as a filter and the whole needle, its words chained across the lanes of a warp, over every candidate
 * (libspm_b200/csrc/myers_chain.cuh).  Hits carry the index of the needle in this call's order.
 * SPMB_ALG_MYERS_PREFIX and the 4-bit alphabets stay at 256 symbols (SPMB_ERR_UNSUPPORTED beyond). */
int spmb_pattern_set_create(spmb_context *ctx, int alg, const uint8_t *ranks,
                            const uint32_t *lengths, uint32_t n_patterns, uint32_t k,
                            spmb_pattern **out);
/* spm::pigeonhole_matcher{needles, error_rate} (pigeonhole_matcher.hpp:157-180): q-gram seed search.
 * q = min over needles of floor(len / (floor(error_rate*len) + 1)) (needles giving q < 3 are ignored, as
 * SeqAn2's _patternInit does); every needle contributes the q-grams at offsets 0, q, 2q, ...; a hit is
 * {end = begin + q, pattern_id = needle index, reserved = needle offset}, sorted by
 * (seq_id, end, pattern_id, reserved) = haystack order like the reference's callback sequence
 * (pigeonhole_matcher_test.cpp:30-33).  spmb_pattern_window_size() returns q (:200-202).
 * Not reproduced: the Finder's masking of >= 1000-symbol period-1 repeats (:186-191). */
int spmb_pigeonhole_create(spmb_context *ctx, const uint8_t *ranks, const uint32_t *lengths,
                           uint32_t n_needles, double error_rate, spmb_pattern **out);
int spmb_pattern_destroy(spmb_pattern *p);
/* figures of a SPMB_ALG_PIGEONHOLE / SPMB_ALG_MYERS_FILTERED pattern: info[0] q, [1] seeds,
 * [2] distinct q-grams, [3] log2 bitmap bits, [4] 0: q too short, the direct Myers kernel runs instead;
 * else the sampling step of the scan (1, 2 or 4: every step-th text position is hashed against the
 * step sub-seeds of each pigeonhole piece, q = piece length - step + 1) */
int spmb_pattern_filter_info(const spmb_pattern *p, uint32_t info[5]);
/* spm::window_size(matcher): m for the exact searchers (seqan_pattern_base.hpp:97-99),
 * m + k for Myers (myers_matcher.hpp:51-53) and the prefix matcher
 * (myers_prefix_matcher_restorable.hpp:151-153); q for the pigeonhole matcher; 0 for an empty needle. */
uint32_t spmb_pattern_window_size(const spmb_pattern *p);
uint32_t spmb_pattern_count(const spmb_pattern *p);

/* ---------------- hits ---------------- */
int spmb_hits_create(spmb_context *ctx, uint64_t capacity_hint, spmb_hits **out);
int spmb_hits_destroy(spmb_hits *h);
uint64_t spmb_hits_count(const spmb_hits *h);
const spmb_hit *spmb_hits_data(const spmb_hits *h); /* host memory, valid until the next search */

/* ---------------- search ----------------
 * spmb_search == seqan_pattern_base::operator()(haystack, callback)
 * (seqan_pattern_base.hpp:40-52): all hits of the pattern (set) in the store, ascending.
 * launch enqueues the kernels on the context's stream and returns; finish waits, grows the
 * hit buffer and re-runs if it overflowed, copies the hits back and sorts them. */
int spmb_search_launch(spmb_context *ctx, const spmb_store *s, const spmb_pattern *p, spmb_hits *h);
int spmb_search_finish(spmb_context *ctx, spmb_hits *h);
int spmb_search(spmb_context *ctx, const spmb_store *s, const spmb_pattern *p, spmb_hits *h);
/* the reference's call shape with a HOST haystack: upload + pack + search + hits back. */
int spmb_search_host(spmb_context *ctx, const uint8_t *host_ranks, uint64_t n,
                     const spmb_pattern *p, spmb_hits *h);
/* the same with a host haystack over a 4-bit alphabet (bits = 3 or 4) and a spmb_pattern_create_wide pattern */
int spmb_search_host_wide(spmb_context *ctx, const uint8_t *host_ranks, uint64_t n, uint32_t bits,
                          const spmb_pattern *p, spmb_hits *h);
/* SPMB_ALG_MYERS_PREFIX with many anchors (what a seed-and-extend caller of
 * restorable_myers_prefix_matcher does, one matcher call per seed): the needle(s) are aligned to the
 * prefixes of the text starting at every anchors[i] (host array of text positions), scan clamped to
 * m + k + 1 symbols; hit = {end (absolute, exclusive), seq_id = anchor index, pattern_id, errors}.
 * spmb_search() with such a pattern is the single anchor 0 (the reference's call shape). */
int spmb_search_anchored(spmb_context *ctx, const spmb_store *s, const spmb_pattern *p,
                         const uint64_t *anchors, uint64_t n_anchors, spmb_hits *h);
/* number of kernel launches the last search enqueued (bench.py's gpu_launches). */
uint32_t spmb_last_launch_count(const spmb_context *ctx);

/* Allocation-free variant of spmb_search for SPMB_ALG_MYERS_FILTERED sets.  It has NO counterpart in the
 * reference: it was added for callers that cannot allocate device
 * memory where they search (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * spmb_search keeps the seed filter's candidate list, its counters and the duplicate-suppression table in
 * buffers of the spmb_hits object, allocated and grown on demand; here they live in ONE file-scope
 * `static __device__` array of 3.5 MiB + 256 B that every GPU loading libspmb.so holds until the process
 * exits.  The call makes no cudaMalloc / cudaFree / allocating library call; the caller passes the inputs
 * (a plain 2-bit store and a filtered Myers set, both made at set-up) and the outputs:
 *   dev_hits  DEVICE array of `capacity` records; receives the hit records UNORDERED (spmb_search sorts on
 *             the way back, this call does not touch the host list of any spmb_hits).  A (pattern_id, end)
 *             normally appears once, with its smallest error count; when the 2^17-slot table is crowded a few
 *             appear again -- keep the smallest `errors` (what spmb_search's ordering does);
 *   dev_count DEVICE counter, receives the number of records; *count_out the same number on the host.
 * Limits, refused with a clear spmb_last_error() before anything is launched: store longer than
 * SPMB_NOALLOC_MAX_SYMBOLS, capacity beyond SPMB_NOALLOC_MAX_HITS, a store without its lane-chunk copy (the
 * first Myers / filtered spmb_search on a store derives it, with an allocation: run one at set-up), another
 * algorithm, a set whose seed filter is off, windows-mode or 4-bit stores.  Data-dependent limits, reported
 * as SPMB_ERR_OOM after the run (dev_hits is then incomplete): more than SPMB_NOALLOC_MAX_CANDIDATES filter
 * survivors, or more records than `capacity` (*count_out then holds the number needed).
 * The call returns when the kernels are done (it synchronises the context's stream).  The first call on a
 * device looks the array up (cudaGetSymbolAddress, which also loads the module): make one call at set-up.
 * ALL calls of a process on one GPU share the array -- it keeps the last call's contents, the call clears what
 * must start at zero -- so two calls that overlap on the GPU (two contexts / streams / host threads on one
 * device) race on it: serialise them. */
#define SPMB_NOALLOC_MAX_CANDIDATES 131072u
#define SPMB_NOALLOC_MAX_SYMBOLS (1ull << 32)
#define SPMB_NOALLOC_MAX_HITS (1ull << 24)
int spmb_search_filtered_noalloc(spmb_context *ctx, const spmb_store *s, const spmb_pattern *p,
                                 spmb_hit *dev_hits, uint64_t capacity, unsigned long long *dev_count,
                                 uint64_t *count_out);

/* ---------------- resumable search (capture / restore) ----------------
 * restorable_myers_matcher / restorable_shiftor_matcher
 * (myers_matcher_restorable.hpp:57-63,72-82; shiftor_matcher_restorable.hpp:44-63):
 * feed the haystack in chunks of any size; the concatenated hit list equals the unchunked one.
 * The carried state is the last window_size-1 symbols of the stream (Myers: the last 2m, which also determine the
 * score at the current column) -- it determines every later hit exactly -- and it lives ON THE
 * DEVICE as packed words: a feed uploads only its new chunk (long chunks through both movers of the upload
 * queue), splices it behind the carried symbols, searches, and leaves the new tail on the device.
 * Pattern sets are allowed (hits carry pattern_id); dna4 only.
 * feed   = submit + collect: the hits of this chunk, ends relative to the chunk start like the reference's fresh
 *          Finder on each chunk (myers_matcher_restorable_test.cpp:66-69).
 * submit / collect: the asynchronous pair for streaming at full rate -- submit hands the chunk to the upload pipeline
 *          and returns at once; THE HOST BUFFER MUST STAY VALID UNTIL THAT CHUNK HAS BEEN COLLECTED.  collect hands
 *          back the hits of the OLDEST chunk in flight; at most three chunks may be in flight, so the packing threads
 *          and the copy engine run on from chunk i into chunk i+1 while chunk i is searched and its hits travel.
 * SPMB_ALG_MYERS_PREFIX streams (restorable_myers_prefix_matcher) are anchored at the START OF THE STREAM: the scan is
 * clamped to the first m + k + 1 symbols of the whole stream, whereas the reference clamps every call on its own
 * haystack (myers_prefix_matcher_restorable.hpp:53-54); a caller that restores a state and feeds a fresh context
 * should create a fresh stream instead.  Their state is that short prefix, kept on the host; feed only. */
int spmb_stream_create(spmb_context *ctx, const spmb_pattern *p, spmb_stream **out);
int spmb_stream_destroy(spmb_stream *st);
int spmb_stream_feed(spmb_context *ctx, spmb_stream *st, const uint8_t *host_ranks, uint64_t n,
                     spmb_hits *h);
int spmb_stream_submit(spmb_context *ctx, spmb_stream *st, const uint8_t *host_ranks, uint64_t n);
int spmb_stream_collect(spmb_context *ctx, spmb_stream *st, spmb_hits *h);
/* capture(): copies the state into buf (size query when buf==NULL); restore(): sets it.
 * blob = {u64 offset, u32 tail_len, u32 n_patterns, u32 errors[n_patterns], u8 tail[tail_len]}: `errors` is what
 * the reference's captured state shows a caller (capture().errors, myers_matcher_restorable.hpp:57-59) -- the
 * score of each needle at the current column, computed on the device from the carried symbols when the state is
 * captured (0 for the exact searchers).  No chunk may be in flight. */
int spmb_stream_capture(const spmb_stream *st, void *buf, size_t *size);
int spmb_stream_restore(spmb_stream *st, const void *buf, size_t size);
int spmb_stream_errors(spmb_stream *st, uint32_t pattern_index, uint32_t *errors);
uint64_t spmb_stream_offset(const spmb_stream *st); /* symbols consumed so far */

/* ---------------- journaled string tree (north_star items 1-2; a14) ----------------
 * What an external JST traverser does with the reference's matchers (it sizes branch contexts
 * with spm::window_size, concept.hpp:27-50, and runs matcher(context, callback) on each),
 * done here in one call: the tree is flattened on the host into an array of branch-context
 * windows (each carries window_size-1 symbols of context before its first variant), the
 * reference is scanned once, the windows are scanned by the windows-mode kernels, and every hit
 * is expanded to (haplotype id = seq_id, end in that haplotype's coordinates, errors).
 * variants: sorted by ref_pos (insertions first on ties); coverage: n_variants rows of
 * ceil(n_haplotypes/64) 64-bit words, bit h of row v set = haplotype h carries variant v; the
 * variants carried by one haplotype must form a valid journal (sorted, non-overlapping: every carried variant
 * starts at or behind the end of the previous one, which puts insertions first on ties); spmb_jst_create checks
 * this per haplotype and returns SPMB_ERR_ARG otherwise (spmb_vcf_read never produces such input: it clears the
 * later of two overlapping calls, spmb_vcf_dropped_calls).  The reference store must outlive the JST. */
typedef struct spmb_jst spmb_jst;
int spmb_jst_create(spmb_context *ctx, const spmb_store *reference, const spmb_journal_event *variants,
                    uint64_t n_variants, const uint8_t *ins_pool, uint64_t pool_len,
                    const uint64_t *coverage, uint32_t n_haplotypes, spmb_jst **out);
int spmb_jst_destroy(spmb_jst *j);
int spmb_jst_search(spmb_context *ctx, spmb_jst *j, const spmb_pattern *p, spmb_hits *h);
/* the same search in two halves, like spmb_search_launch / spmb_search_finish (a traverser that runs one matcher
 * call per context, concept.hpp:27-50, back to back): launch enqueues the kernels and returns, finish waits, brings
 * the device hits back and expands them to haplotypes on the host.  With two or three spmb_hits in rotation the
 * host half of a search runs under the kernels of the next one.  Searches in flight on one tree share the window
 * size (a launch with another one is refused with SPMB_ERR_ARG until the pending ones are finished: it would rebuild
 * the window stores they read); spmb_jst_destroy and the one-call spmb_jst_search want no search in flight.
 * spmb_jst_last_stats reports no kernel time for a split search. */
int spmb_jst_search_launch(spmb_context *ctx, spmb_jst *j, const spmb_pattern *p, spmb_hits *h);
int spmb_jst_search_finish(spmb_context *ctx, spmb_hits *h);
/* figures of the last search: stats[0] windows, [1] window symbols, [2] length buckets,
 * [3] raw device hits before expansion; ms[0] host flattening (0 when cached), [1] window
 * materialisation, [2] search kernels (reference + windows), [3] host expansion + sort */
int spmb_jst_last_stats(const spmb_jst *j, uint64_t stats[4], double ms[4]);

/* ---------------- multi-GPU driver (north_star item 4) ----------------
 * What a single-process libspm caller (seqan_pattern_base::operator(), seqan_pattern_base.hpp:40-52: one host
 * haystack, one callback sequence) needs to use every GPU of the box: one context, stream and upload pipeline per
 * device inside ONE process; a host haystack is cut into contiguous shards, one per device, each with `overlap`
 * symbols of left context (a hit belongs to the shard that holds its END; the window_size-1 symbols before it --
 * the seam rule of myers_matcher.hpp:51-53 -- are in the context), every shard crosses its own PCIe link; the
 * search runs on all devices at once and the per-device hit lists are copied back and merged on the host into one
 * list sorted like spmb_search's.  No NCCL, no collective: the only exchange is this final gather of the hit lists.
 * devices = NULL: devices 0..n_devices-1; n_devices = 0: every visible device.  overlap = 0: 4096 symbols.
 * Searching with window_size - 1 > overlap is SPMB_ERR_UNSUPPORTED.  `merged` is a hit list of device 0
 * (spmb_hits_create(spmb_multi_context(m, 0), ...)). */
typedef struct spmb_multi spmb_multi;
typedef struct spmb_multi_store spmb_multi_store;
typedef struct spmb_multi_pattern spmb_multi_pattern;
int spmb_multi_create(const int *devices, uint32_t n_devices, spmb_multi **out);
int spmb_multi_destroy(spmb_multi *m);
uint32_t spmb_multi_device_count(const spmb_multi *m);
spmb_context *spmb_multi_context(spmb_multi *m, uint32_t i);
const char *spmb_multi_last_error(const spmb_multi *m);
int spmb_multi_store_create(spmb_multi *m, const uint8_t *host_ranks, uint64_t n, uint32_t overlap,
                            spmb_multi_store **out);
/* journaled sequence over a sharded base: `events` (sorted, non-overlapping, coordinates of the whole base) is split
 * at the shard cuts -- every device reads its slice of the caller's array -- and every device materialises its part
 * of the derived sequence; hits come back in the coordinates of the whole derived sequence.  An event that deletes
 * more than `overlap` symbols across a cut is SPMB_ERR_UNSUPPORTED. */
int spmb_multi_store_create_journaled(spmb_multi *m, const spmb_multi_store *base, const spmb_journal_event *events,
                                      uint64_t n_events, const uint8_t *ins_pool, uint64_t pool_len,
                                      spmb_multi_store **out);
int spmb_multi_store_destroy(spmb_multi_store *s);
uint64_t spmb_multi_store_length(const spmb_multi_store *s);
uint32_t spmb_multi_store_shards(const spmb_multi_store *s);
/* the needle (set) on every device: spmb_pattern_set_create's arguments */
int spmb_multi_pattern_create(spmb_multi *m, int alg, const uint8_t *ranks, const uint32_t *lengths,
                              uint32_t n_patterns, uint32_t k, spmb_multi_pattern **out);
int spmb_multi_pattern_destroy(spmb_multi_pattern *p);
int spmb_multi_search(spmb_multi *m, const spmb_multi_store *s, const spmb_multi_pattern *p, spmb_hits *merged);
/* the reference's call shape over all devices: host haystack in, merged hits out */
int spmb_multi_search_host(spmb_multi *m, const uint8_t *host_ranks, uint64_t n, const spmb_multi_pattern *p,
                           spmb_hits *merged);
/* wall-clock of the last calls in ms: [0] store_create (all uploads), [1] store_create_journaled, [2] search
 * (first launch -> last device's hit list on the host), [3] host merge */
int spmb_multi_last_stats(const spmb_multi *m, double ms[4]);

/* ---------------- ingest: FASTA / VCF -> store inputs ----------------
 * Host-side readers for the formats of the reference's own test data (test/data/datasources.cmake:72-102:
 * sim_ref_10Kb.fasta.gz + sim_ref_10Kb_{SNPs,SNP_INDELs}.vcf, whose shipped haplotype FASTAs are the known
 * answer).  Plain or gzip files.  Characters outside ACGT(U) become rank 0 like seqan3::dna4 (counted).
 * They need no device and no context. */
typedef struct spmb_fasta spmb_fasta;
int spmb_fasta_read(const char *path, spmb_fasta **out);
/* the same reader for spm::dna5 (bits = 3: ranks ACGNT = 0..4) or spm::dna15 (bits = 4: ABCDGHKMNRSTVWY = 0..14),
 * seqan3's rank orders; characters outside the alphabet become N (counted); bits = 2 is spmb_fasta_read.
 * The ranks feed spmb_store_create_wide. */
int spmb_fasta_read_alphabet(const char *path, uint32_t bits, spmb_fasta **out);
int spmb_fasta_destroy(spmb_fasta *fa);
uint64_t spmb_fasta_count(const spmb_fasta *fa);
const char *spmb_fasta_name(const spmb_fasta *fa, uint64_t i);
const uint8_t *spmb_fasta_ranks(const spmb_fasta *fa, uint64_t i, uint64_t *len); /* 1-byte ranks of record i */
uint64_t spmb_fasta_unknown_symbols(const spmb_fasta *fa);
/* VCF records of one contig (NULL = the contig of the first record) -> the arrays spmb_jst_create takes:
 * one variant per ALT allele (common prefix with REF stripped: replace ref[pos, pos+del) by ins), sorted
 * by position with insertions first on ties; coverage bit h = sample h/ploidy, allele h%ploidy carries
 * it (GT a|b or a/b; '.' = reference).  Symbolic / breakend ALTs and records outside [0, ref_len) are
 * skipped and counted (ref_len = 0: no bound check). */
typedef struct spmb_vcf spmb_vcf;
int spmb_vcf_read(const char *path, const char *contig, uint64_t ref_len, spmb_vcf **out);
int spmb_vcf_destroy(spmb_vcf *v);
uint64_t spmb_vcf_variant_count(const spmb_vcf *v);
const spmb_journal_event *spmb_vcf_variants(const spmb_vcf *v);
const uint8_t *spmb_vcf_pool(const spmb_vcf *v, uint64_t *len);
const uint64_t *spmb_vcf_coverage(const spmb_vcf *v, uint32_t *words_per_row);
uint32_t spmb_vcf_haplotype_count(const spmb_vcf *v);
uint32_t spmb_vcf_sample_count(const spmb_vcf *v);
const char *spmb_vcf_sample_name(const spmb_vcf *v, uint32_t i);
int spmb_vcf_stats(const spmb_vcf *v, uint64_t stats[3]); /* [0] records of the contig, [1] skipped, [2] non-ACGT symbols */
/* (variant, haplotype) calls cleared because they overlap an earlier call of the same haplotype (a SNP inside a
 * called deletion ...): the first call wins, so that every haplotype's variants form a valid journal */
uint64_t spmb_vcf_dropped_calls(const spmb_vcf *v);

/* ---------------- measurement helpers ---------------- */
/* INT32 ALU-pipe peak (LOP3 / IADD3 chains on all SMs), in 1e9 lane-instructions per second:
 * the roofline denominator for the Myers kernels. */
int spmb_measure_int_peak(spmb_context *ctx, double *lop3_gops, double *iadd3_gops);
int spmb_device_info(spmb_context *ctx, int *sm_count, int *sm_clock_khz, uint64_t *mem_bytes);

#ifdef __cplusplus
}
#endif
#endif
