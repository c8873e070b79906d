// ingest.cpp -- sequence-store ingest: FASTA (+gzip) and VCF -> reference ranks, variants, insertion
// pool, haplotype coverage.  Host-side parsing only; no search logic.
//
// Reference anchors: the data the reference ships for its (absent) JST layer,
//   test/data/sim_ref_10Kb.fasta.gz + sim_ref_10Kb_{SNPs,SNP_INDELs}.vcf -> ..._haplotypes.fasta.gz
//   (test/data/datasources.cmake:72-102), and the journal semantics of
//   specs/journaled_sequence_class_diagram.drawio:196 (record_sequence_edit: replace [pos, pos+del) by ins).
// Symbols follow spm::dna4 (libspm/libspm/seqan/alphabet.hpp:34-83): one byte per symbol holding the
// rank A=0 C=1 G=2 T=3; like seqan3::dna4 every other character converts to A (counted and reported).
#include <zlib.h>

#include <algorithm>
#include <cstdint>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "../../include/spmb.h"

namespace {

inline int rank_of_char(char c, uint64_t &unknown)
{
    switch (c) {
    case 'A': case 'a': return 0;
    case 'C': case 'c': return 1;
    case 'G': case 'g': return 2;
    case 'T': case 't': case 'U': case 'u': return 3;
    default: ++unknown; return 0;
    }
}

// seqan3 rank orders (alphabetical): dna5 = ACGNT, dna15 = ABCDGHKMNRSTVWY; unknown characters convert to N
inline int rank_of_char_wide(char c, uint32_t bits, uint64_t &unknown)
{
    const char *letters = bits == 3 ? "ACGNT" : "ABCDGHKMNRSTVWY";
    if (c >= 'a' && c <= 'z') c = (char)(c - 'a' + 'A');
    if (c == 'U') c = 'T';
    int n_rank = 0;
    for (int i = 0; letters[i]; ++i) {
        if (letters[i] == c) return i;
        if (letters[i] == 'N') n_rank = i;
    }
    ++unknown;
    return n_rank;
}

// line reader over gzopen (reads plain files too)
struct LineReader {
    gzFile f = nullptr;
    std::vector<char> buf;
    explicit LineReader(const char *path) : buf(1 << 16) { f = gzopen(path, "rb"); if (f) gzbuffer(f, 1 << 20); }
    ~LineReader() { if (f) gzclose(f); }
    bool ok() const { return f != nullptr; }
    // next line without the trailing newline / carriage return; false at EOF
    bool next(std::string &line)
    {
        line.clear();
        for (;;) {
            if (!gzgets(f, buf.data(), (int)buf.size())) return !line.empty();
            const size_t n = strlen(buf.data());
            line.append(buf.data(), n);
            if (n && buf[n - 1] == '\n') break;
            if (gzeof(f)) break;
        }
        while (!line.empty() && (line.back() == '\n' || line.back() == '\r')) line.pop_back();
        return true;
    }
};

} // namespace

struct spmb_fasta {
    std::vector<std::string> names;
    std::vector<uint64_t> offset; // n_records + 1
    std::vector<uint8_t> ranks;
    uint64_t unknown = 0;
};

struct spmb_vcf {
    std::vector<spmb_journal_event> var;
    std::vector<uint8_t> pool;
    std::vector<uint64_t> cov; // n_var x wpr
    std::vector<std::string> samples;
    uint32_t n_hap = 0, ploidy = 0, wpr = 0;
    uint64_t records = 0, skipped = 0, unknown = 0;
    uint64_t dropped_calls = 0; // (variant, haplotype) calls cleared because they overlap an earlier call of that haplotype
    std::string contig;
};

extern "C" int spmb_fasta_read(const char *path, spmb_fasta **out) { return spmb_fasta_read_alphabet(path, 2, out); }

extern "C" int spmb_fasta_read_alphabet(const char *path, uint32_t bits, spmb_fasta **out)
{
    if (!path || !out || bits < 2 || bits > 4) return SPMB_ERR_ARG;
    *out = nullptr;
    LineReader rd(path);
    if (!rd.ok()) return SPMB_ERR_ARG;
    spmb_fasta *fa = new (std::nothrow) spmb_fasta();
    if (!fa) return SPMB_ERR_OOM;
    std::string line;
    bool have = false;
    while (rd.next(line)) {
        if (line.empty()) continue;
        if (line[0] == '>') {
            have = true;
            fa->offset.push_back(fa->ranks.size());
            // the record id is the header up to the first blank
            size_t e = line.find_first_of(" \t");
            fa->names.push_back(line.substr(1, e == std::string::npos ? std::string::npos : e - 1));
            continue;
        }
        if (line[0] == ';') continue;
        if (!have) { delete fa; return SPMB_ERR_ARG; } // sequence data before any header
        for (char c : line)
            if (c != ' ' && c != '\t')
                fa->ranks.push_back((uint8_t)(bits == 2 ? rank_of_char(c, fa->unknown) : rank_of_char_wide(c, bits, fa->unknown)));
    }
    fa->offset.push_back(fa->ranks.size());
    *out = fa;
    return SPMB_OK;
}

extern "C" int spmb_fasta_destroy(spmb_fasta *fa) { delete fa; return SPMB_OK; }
extern "C" uint64_t spmb_fasta_count(const spmb_fasta *fa) { return fa ? fa->names.size() : 0; }
extern "C" const char *spmb_fasta_name(const spmb_fasta *fa, uint64_t i) { return (fa && i < fa->names.size()) ? fa->names[i].c_str() : nullptr; }
extern "C" uint64_t spmb_fasta_unknown_symbols(const spmb_fasta *fa) { return fa ? fa->unknown : 0; }
extern "C" const uint8_t *spmb_fasta_ranks(const spmb_fasta *fa, uint64_t i, uint64_t *len)
{
    if (!fa || i >= fa->names.size()) { if (len) *len = 0; return nullptr; }
    if (len) *len = fa->offset[i + 1] - fa->offset[i];
    return fa->ranks.data() + fa->offset[i];
}

// one tab-separated field after another
static bool next_field(const std::string &s, size_t &pos, std::string &out)
{
    if (pos > s.size()) return false;
    size_t e = s.find('\t', pos);
    if (e == std::string::npos) e = s.size();
    out.assign(s, pos, e - pos);
    pos = e + 1;
    return true;
}

extern "C" int spmb_vcf_read(const char *path, const char *contig, uint64_t ref_len, spmb_vcf **out)
{
    if (!path || !out) return SPMB_ERR_ARG;
    *out = nullptr;
    LineReader rd(path);
    if (!rd.ok()) return SPMB_ERR_ARG;
    spmb_vcf *v = new (std::nothrow) spmb_vcf();
    if (!v) return SPMB_ERR_OOM;
    if (contig) v->contig = contig;
    struct Rec { uint64_t pos; uint32_t del; std::string ins; std::vector<uint64_t> cov; uint64_t order; };
    std::vector<Rec> recs;
    std::string line, f;
    std::vector<std::string> alts;
    std::vector<int> gt; // allele index per haplotype of this record
    bool header = false;
    while (rd.next(line)) {
        if (line.empty() || (line[0] == '#' && line.size() > 1 && line[1] == '#')) continue;
        size_t p = 0;
        if (line[0] == '#') { // #CHROM POS ID REF ALT QUAL FILTER INFO FORMAT samples...
            for (int i = 0; i < 9; ++i) if (!next_field(line, p, f)) break;
            while (p <= line.size() && next_field(line, p, f)) v->samples.push_back(f);
            header = true;
            continue;
        }
        if (!header) { delete v; return SPMB_ERR_ARG; }
        std::string chrom, ref, alt, fmt;
        uint64_t pos1 = 0;
        next_field(line, p, chrom);
        next_field(line, p, f); pos1 = strtoull(f.c_str(), nullptr, 10);
        next_field(line, p, f);           // ID
        next_field(line, p, ref);
        next_field(line, p, alt);
        next_field(line, p, f);           // QUAL
        next_field(line, p, f);           // FILTER
        next_field(line, p, f);           // INFO
        next_field(line, p, fmt);         // FORMAT (GT first, VCF 4.x 1.4.2)
        if (v->contig.empty()) v->contig = chrom;
        if (chrom != v->contig) continue;
        ++v->records;
        if (pos1 == 0 || (ref_len && pos1 - 1 + ref.size() > ref_len) || fmt.compare(0, 2, "GT") != 0) { ++v->skipped; continue; }
        alts.clear();
        for (size_t a = 0; a <= alt.size();) {
            size_t e = alt.find(',', a);
            if (e == std::string::npos) e = alt.size();
            alts.push_back(alt.substr(a, e - a));
            a = e + 1;
        }
        // genotypes: allele indices separated by | or /, '.' = reference
        gt.clear();
        uint32_t si = 0;
        while (p <= line.size() && next_field(line, p, f)) {
            size_t colon = f.find(':');
            if (colon != std::string::npos) f.resize(colon);
            uint32_t n_all = 0;
            for (size_t a = 0; a <= f.size();) {
                size_t e = f.find_first_of("|/", a);
                if (e == std::string::npos) e = f.size();
                gt.push_back((e > a && f[a] != '.') ? atoi(f.c_str() + a) : 0);
                ++n_all;
                a = e + 1;
            }
            if (v->ploidy == 0) v->ploidy = n_all;
            if (n_all != v->ploidy) { delete v; return SPMB_ERR_UNSUPPORTED; } // mixed ploidy
            ++si;
        }
        if (si != v->samples.size()) { delete v; return SPMB_ERR_ARG; }
        if (v->n_hap == 0) { v->n_hap = (uint32_t)gt.size(); v->wpr = (v->n_hap + 63) / 64; }
        for (size_t ai = 0; ai < alts.size(); ++ai) {
            const std::string &a = alts[ai];
            if (a.empty() || a[0] == '<' || a == "*" || a.find_first_of("[]") != std::string::npos) { ++v->skipped; continue; }
            size_t c = 0; // strip the common prefix: the event starts at the first differing symbol
            while (c < ref.size() && c < a.size() && toupper(ref[c]) == toupper(a[c])) ++c;
            Rec r;
            r.pos = pos1 - 1 + c;
            r.del = (uint32_t)(ref.size() - c);
            r.ins = a.substr(c);
            r.cov.assign(v->wpr, 0);
            for (uint32_t h = 0; h < v->n_hap; ++h)
                if (gt[h] == (int)ai + 1) r.cov[h >> 6] |= 1ull << (h & 63);
            r.order = recs.size();
            recs.push_back(std::move(r));
        }
    }
    // sorted by position, insertions before deletions/replacements on ties (spmb_jst_create's contract)
    std::stable_sort(recs.begin(), recs.end(), [](const Rec &a, const Rec &b) {
        if (a.pos != b.pos) return a.pos < b.pos;
        return (a.del != 0) < (b.del != 0);
    });
    if (v->n_hap == 0) { v->n_hap = (uint32_t)v->samples.size() * (v->ploidy ? v->ploidy : 2); v->wpr = (v->n_hap + 63) / 64; }
    // a haplotype's calls must form a valid journal (spmb_jst_create's contract): where one ALT overlaps an
    // earlier call of the same haplotype (a SNP inside a called deletion, two ALTs of one multi-allelic site
    // through different records), the first call wins and the later one is cleared for that haplotype
    {
        std::vector<uint64_t> last_end(v->n_hap, 0);
        for (Rec &r : recs)
            for (uint32_t w = 0; w < v->wpr; ++w) {
                uint64_t bits = r.cov[w];
                while (bits) {
                    const uint32_t h = w * 64 + (uint32_t)__builtin_ctzll(bits);
                    bits &= bits - 1;
                    if (h >= v->n_hap) continue;
                    if (r.pos < last_end[h]) { r.cov[w] &= ~(1ull << (h & 63)); ++v->dropped_calls; }
                    else last_end[h] = r.pos + r.del;
                }
            }
    }
    for (const Rec &r : recs) {
        spmb_journal_event e;
        e.ref_pos = r.pos; e.del_len = r.del; e.ins_len = (uint32_t)r.ins.size(); e.ins_off = v->pool.size();
        for (char c : r.ins) v->pool.push_back((uint8_t)rank_of_char(c, v->unknown));
        v->var.push_back(e);
        v->cov.insert(v->cov.end(), r.cov.begin(), r.cov.end());
    }
    *out = v;
    return SPMB_OK;
}

extern "C" int spmb_vcf_destroy(spmb_vcf *v) { delete v; return SPMB_OK; }
extern "C" uint64_t spmb_vcf_variant_count(const spmb_vcf *v) { return v ? v->var.size() : 0; }
extern "C" const spmb_journal_event *spmb_vcf_variants(const spmb_vcf *v) { return (v && !v->var.empty()) ? v->var.data() : nullptr; }
extern "C" const uint8_t *spmb_vcf_pool(const spmb_vcf *v, uint64_t *len)
{
    if (len) *len = v ? v->pool.size() : 0;
    return (v && !v->pool.empty()) ? v->pool.data() : nullptr;
}
extern "C" const uint64_t *spmb_vcf_coverage(const spmb_vcf *v, uint32_t *words_per_row)
{
    if (words_per_row) *words_per_row = v ? v->wpr : 0;
    return (v && !v->cov.empty()) ? v->cov.data() : nullptr;
}
extern "C" uint32_t spmb_vcf_haplotype_count(const spmb_vcf *v) { return v ? v->n_hap : 0; }
extern "C" uint32_t spmb_vcf_sample_count(const spmb_vcf *v) { return v ? (uint32_t)v->samples.size() : 0; }
extern "C" const char *spmb_vcf_sample_name(const spmb_vcf *v, uint32_t i) { return (v && i < v->samples.size()) ? v->samples[i].c_str() : nullptr; }
extern "C" uint64_t spmb_vcf_dropped_calls(const spmb_vcf *v) { return v ? v->dropped_calls : 0; }
extern "C" int spmb_vcf_stats(const spmb_vcf *v, uint64_t stats[3])
{
    if (!v || !stats) return SPMB_ERR_ARG;
    stats[0] = v->records; stats[1] = v->skipped; stats[2] = v->unknown;
    return SPMB_OK;
}
