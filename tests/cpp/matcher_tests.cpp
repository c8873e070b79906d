// matcher_tests.cpp -- the reference's matcher unit tests, replayed against the C++ facade.
// Same fixture, same calls, same expected vectors as
//   /root/reference/test/api/libspm/matcher/horspool_matcher_test.cpp:19-52
//   /root/reference/test/api/libspm/matcher/shiftor_matcher_test.cpp:19-52
//   /root/reference/test/api/libspm/matcher/myers_matcher_test.cpp:19-53
//   /root/reference/test/api/libspm/matcher/myers_matcher_restorable_test.cpp:19-74
//   /root/reference/test/api/libspm/matcher/pigeonhole_matcher_test.cpp:19-86
// (GoogleTest swapped for tests/cpp/mini_test.hpp).  Runs on the GPU box only.
#include <algorithm>
#include <cstddef>
#include <vector>

#include <libspm_b200/spm.hpp>

#include "mini_test.hpp"

using spm::operator""_dna4;
using sequence_t = std::vector<spm::dna4>;

static const sequence_t haystack = "ACGTGACTAGCACGTGACTAGCACGTGACTAGCACGTGACTAGC"_dna4;
static const sequence_t needle = "GCACG"_dna4;
static const std::vector<std::size_t> exact_begin{9, 20, 31};
static const std::vector<std::size_t> myers_end{13, 14, 15, 24, 25, 26, 35, 36, 37};

MINI_TEST(horspool_concept_and_window_size) {
    auto matcher = spm::horspool_matcher{needle};
    EXPECT_TRUE(spm::window_matcher<decltype(matcher)>);
    EXPECT_EQ(spm::window_size(matcher), std::ranges::size(needle));
}

MINI_TEST(horspool_dna4_pattern) {
    auto matcher = spm::horspool_matcher{needle};
    std::vector<std::size_t> actual{};
    matcher(haystack, [&](auto const &finder) { actual.push_back(seqan2::beginPosition(finder)); });
    EXPECT_TRUE(std::ranges::equal(actual, exact_begin));
}

MINI_TEST(shiftor_concept_and_window_size) {
    auto matcher = spm::shiftor_matcher{needle};
    EXPECT_TRUE(spm::window_matcher<decltype(matcher)>);
    EXPECT_EQ(spm::window_size(matcher), std::ranges::size(needle));
}

MINI_TEST(shiftor_dna4_pattern) {
    auto matcher = spm::shiftor_matcher{needle};
    std::vector<std::size_t> actual{};
    matcher(haystack, [&](auto const &finder) { actual.push_back(seqan2::beginPosition(finder)); });
    EXPECT_TRUE(std::ranges::equal(actual, exact_begin));
}

MINI_TEST(myers_concept_and_window_size) {
    std::size_t errors = 1;
    auto matcher = spm::myers_matcher{needle, errors};
    EXPECT_TRUE(spm::window_matcher<decltype(matcher)>);
    EXPECT_EQ(spm::window_size(matcher), std::ranges::size(needle) + errors);
}

MINI_TEST(myers_dna4_pattern) {
    auto matcher = spm::myers_matcher{needle, 1};
    std::vector<std::size_t> actual{};
    std::vector<unsigned> errs{};
    matcher(haystack, [&](auto const &finder) {
        actual.push_back(seqan2::endPosition(finder));
        errs.push_back(finder.errors());
    });
    EXPECT_TRUE(std::ranges::equal(actual, myers_end));
    EXPECT_TRUE(std::ranges::equal(errs, std::vector<unsigned>{1, 0, 1, 1, 0, 1, 1, 0, 1}));
}

MINI_TEST(restorable_myers_concept_and_window_size) {
    std::size_t errors = 1;
    auto matcher = spm::restorable_myers_matcher{needle, errors};
    EXPECT_TRUE(spm::window_matcher<decltype(matcher)>);
    EXPECT_TRUE(spm::restorable_matcher<decltype(matcher)>);
    EXPECT_EQ(spm::window_size(matcher), std::ranges::size(needle) + errors);
}

MINI_TEST(restorable_myers_dna4_pattern) {
    auto matcher = spm::restorable_myers_matcher{needle, 1u};
    std::vector<std::size_t> actual{};
    matcher(haystack, [&](auto const &finder) { actual.push_back(seqan2::endPosition(finder)); });
    EXPECT_TRUE(std::ranges::equal(actual, myers_end));
}

MINI_TEST(restorable_myers_dna4_pattern_captured) {
    std::size_t chunk_size{13};
    std::ptrdiff_t chunk_count = (haystack.size() + chunk_size - 1) / chunk_size;
    auto matcher = spm::restorable_myers_matcher{needle, 1u};
    auto state = matcher.capture();
    std::vector<std::size_t> actual{};
    for (std::ptrdiff_t chunk_idx = 0; chunk_idx < chunk_count; ++chunk_idx) {
        std::ptrdiff_t offset = chunk_idx * chunk_size;
        sequence_t chunk{haystack.begin() + offset,
                         haystack.begin() + std::min<std::ptrdiff_t>(offset + chunk_size, haystack.size())};
        matcher.restore(state);
        matcher(chunk, [&](auto const &finder) { actual.push_back(seqan2::endPosition(finder) + offset); });
        state = matcher.capture();
    }
    EXPECT_TRUE(std::ranges::equal(actual, myers_end));
}

// untested in the reference, covered here
MINI_TEST(restorable_shiftor_chunks) {
    auto matcher = spm::restorable_shiftor_matcher{needle};
    EXPECT_TRUE(spm::restorable_matcher<decltype(matcher)>);
    std::vector<std::size_t> actual{};
    for (std::size_t offset = 0; offset < haystack.size(); offset += 7) {
        sequence_t chunk{haystack.begin() + offset, haystack.begin() + std::min(offset + 7, haystack.size())};
        matcher(chunk, [&](auto const &finder) { actual.push_back(seqan2::beginPosition(finder) + offset); });
    }
    // beginPosition of a hit that started in an earlier chunk is clamped at 0 of the chunk; use ends
    std::vector<std::size_t> ends{};
    auto m2 = spm::restorable_shiftor_matcher{needle};
    for (std::size_t offset = 0; offset < haystack.size(); offset += 7) {
        sequence_t chunk{haystack.begin() + offset, haystack.begin() + std::min(offset + 7, haystack.size())};
        m2(chunk, [&](auto const &finder) { ends.push_back(seqan2::endPosition(finder) + offset); });
    }
    EXPECT_TRUE(std::ranges::equal(ends, std::vector<std::size_t>{14, 25, 36}));
}

MINI_TEST(device_resident_sequence_and_copies) {
    spm::device_sequence seq{haystack};
    EXPECT_EQ(seq.size(), haystack.size());
    auto matcher = spm::myers_matcher{needle, 1};
    auto copy = matcher; // matchers are copyable (concept.hpp:126-130)
    std::vector<std::size_t> a{}, b{};
    matcher(seq, [&](auto const &f) { a.push_back(seqan2::endPosition(f)); });
    copy(seq, [&](auto const &f) { b.push_back(seqan2::endPosition(f)); });
    EXPECT_TRUE(std::ranges::equal(a, myers_end));
    EXPECT_TRUE(std::ranges::equal(b, myers_end));
}

MINI_TEST(empty_needle_and_haystack) {
    sequence_t none{};
    auto matcher = spm::horspool_matcher{none};
    EXPECT_TRUE(matcher.empty());
    EXPECT_EQ(spm::window_size(matcher), 0u);
    int calls = 0;
    matcher(haystack, [&](auto const &) { ++calls; });
    auto m2 = spm::shiftor_matcher{needle};
    m2(none, [&](auto const &) { ++calls; });
    EXPECT_EQ(calls, 0);
}

// journaled sequence + journaled string tree behind the same matcher call: the hits of every haplotype
// equal the matcher run over that haplotype's materialised text
MINI_TEST(journaled_string_tree_equals_per_haplotype_search) {
    // variants over the shared fixture (ACGTGACTAGCACGTGACTAGC...): SNP G->T at 12, insertion "CA" at 20,
    // deletion of 2 at 30; pool = T, C, A
    const std::vector<spmb_journal_event> variants{{12, 1, 1, 0}, {20, 0, 2, 1}, {30, 2, 0, 0}};
    const std::vector<std::uint8_t> pool{3, 1, 0};
    // haplotype 0: none, 1: SNP, 2: insertion + deletion, 3: all three
    const std::vector<std::uint64_t> coverage{0b1010, 0b1100, 0b1100};
    spm::device_sequence reference{haystack};
    spm::journaled_string_tree tree{reference, variants, pool, coverage, 4};
    EXPECT_EQ(tree.haplotype_count(), 4u);
    auto apply = [&](unsigned h) {
        sequence_t out{};
        std::size_t r = 0;
        for (std::size_t v = 0; v < variants.size(); ++v) {
            if (!((coverage[v] >> h) & 1u)) continue;
            for (; r < variants[v].ref_pos; ++r) out.push_back(haystack[r]);
            for (std::uint32_t i = 0; i < variants[v].ins_len; ++i) { spm::dna4 s{}; s.assign_rank(pool[variants[v].ins_off + i]); out.push_back(s); }
            r += variants[v].del_len;
        }
        for (; r < haystack.size(); ++r) out.push_back(haystack[r]);
        return out;
    };
    for (int pass = 0; pass < 2; ++pass) {
        std::vector<std::vector<std::size_t>> got(4), want(4);
        if (pass == 0) {
            auto matcher = spm::shiftor_matcher{needle};
            matcher(tree, [&](auto const &f) { got[f.sequence_id()].push_back(seqan2::endPosition(f)); });
            for (unsigned h = 0; h < 4; ++h) matcher(apply(h), [&](auto const &f) { want[h].push_back(seqan2::endPosition(f)); });
        } else {
            auto matcher = spm::myers_matcher{needle, 1};
            matcher(tree, [&](auto const &f) { got[f.sequence_id()].push_back(seqan2::endPosition(f)); });
            for (unsigned h = 0; h < 4; ++h) matcher(apply(h), [&](auto const &f) { want[h].push_back(seqan2::endPosition(f)); });
        }
        for (unsigned h = 0; h < 4; ++h) EXPECT_TRUE(std::ranges::equal(got[h], want[h]));
        EXPECT_TRUE(!want[0].empty());
    }
    // the journaled sequence of haplotype 3 as a device-resident haystack
    auto h3 = reference.journaled(variants, pool);
    EXPECT_EQ(h3.size(), apply(3).size());
}

// ---- pigeonhole_matcher_test.cpp:19-86 ----
using needle_position_t = seqan2::PigeonholeSeedOnlyPosition;
static const sequence_t needle2 = "TGACTAGCAC"_dna4;
static const std::vector<sequence_t> multi_needle{needle, needle2};
static const std::vector<std::size_t> expected_multi_positions{3, 8, 9, 14, 19, 20, 25, 30, 31, 36};
static const std::vector<needle_position_t> expected_needle_positions{{1, 0, 5}, {1, 5, 5}, {0, 0, 5}, {1, 0, 5}, {1, 5, 5},
                                                                      {0, 0, 5}, {1, 0, 5}, {1, 5, 5}, {0, 0, 5}, {1, 0, 5}};

MINI_TEST(pigeonhole_concept_and_window_size) {
    double errors = 0.0;
    auto matcher = spm::pigeonhole_matcher{needle, errors};
    EXPECT_TRUE(spm::window_matcher<decltype(matcher)>);
    EXPECT_EQ(spm::window_size(matcher), std::ranges::size(needle));
}

MINI_TEST(pigeonhole_dna4_pattern) {
    auto matcher = spm::pigeonhole_matcher{needle, 0.0};
    std::vector<std::size_t> actual_positions{};
    matcher(haystack, [&](auto const &finder) { actual_positions.push_back(seqan2::beginPosition(finder)); });
    EXPECT_TRUE(std::ranges::equal(actual_positions, exact_begin));
}

MINI_TEST(pigeonhole_dna4_multi_pattern) {
    auto matcher = spm::pigeonhole_matcher{multi_needle, 0.0};
    std::vector<std::size_t> actual_positions{};
    matcher(haystack, [&](auto const &finder) { actual_positions.push_back(seqan2::beginPosition(finder)); });
    EXPECT_TRUE(std::ranges::equal(actual_positions, expected_multi_positions));
}

MINI_TEST(pigeonhole_dna4_multi_pattern_position) {
    auto matcher = spm::pigeonhole_matcher{multi_needle, 0.0};
    std::vector<needle_position_t> actual_needle_positions{};
    matcher(haystack, [&](auto const &) { actual_needle_positions.push_back(matcher.position()); });
    EXPECT_TRUE(actual_needle_positions == expected_needle_positions);
}

// the prefix matcher has no test in the reference: needle vs haystack prefixes
MINI_TEST(restorable_myers_prefix_matcher_anchored) {
    auto matcher = spm::restorable_myers_prefix_matcher{"ACGTGA"_dna4, 1u};
    EXPECT_TRUE(spm::restorable_matcher<decltype(matcher)>);
    EXPECT_EQ(spm::window_size(matcher), 7u);
    std::vector<std::size_t> ends{};
    std::vector<unsigned> errs{};
    matcher(haystack, [&](auto const &f) { ends.push_back(seqan2::endPosition(f)); errs.push_back(f.errors()); });
    EXPECT_TRUE(std::ranges::equal(ends, std::vector<std::size_t>{5, 6, 7}));
    EXPECT_TRUE(std::ranges::equal(errs, std::vector<unsigned>{1, 0, 1}));
    // anchored: the same needle does not match at haystack position 11 unless the haystack starts there
    auto m2 = spm::restorable_myers_prefix_matcher{"GCACG"_dna4, 0u};
    int calls = 0;
    m2(haystack, [&](auto const &) { ++calls; });
    EXPECT_EQ(calls, 0);
}

MINI_TEST(filtered_myers_matcher_equals_myers_matcher) {
    // 3 copies of a 48-symbol needle with 0, 1 and 2 edits inside filler
    sequence_t pat = "ACGTTGCAAGCTTAGGCTAACGGTACCATGGATCCAGTCAAGTCCGTA"_dna4;
    sequence_t text = "TTTTTTTTTT"_dna4;
    auto app = [&](sequence_t const &x) { text.insert(text.end(), x.begin(), x.end()); };
    app(pat); app("GGGGGGGGGGGG"_dna4);
    sequence_t one = pat; one[20] = "T"_dna4[0]; app(one); app("CCCCCCCCCCCCC"_dna4);
    sequence_t two = pat; two.erase(two.begin() + 7); two[30] = "A"_dna4[0]; app(two); app("AAAAAAAAAA"_dna4);
    std::vector<std::pair<std::size_t, unsigned>> a{}, b{};
    auto direct = spm::myers_matcher{pat, 2};
    direct(text, [&](auto const &f) { a.emplace_back(seqan2::endPosition(f), f.errors()); });
    auto filtered = spm::filtered_myers_matcher{std::vector<sequence_t>{pat}, 2};
    filtered(text, [&](auto const &f) { b.emplace_back(seqan2::endPosition(f), f.errors()); });
    EXPECT_TRUE(!a.empty());
    EXPECT_TRUE(a == b);
}

// ---- spm::dna5 / spm::dna15 (alphabet.hpp:119-120; typed over all three alphabets in the reference's
// test/api/libspm/seqan/alphabet_test.cpp:21-25): the same fixture with N / IUPAC letters mixed in ----
MINI_TEST(alphabet_ranks_and_chars) {
    using spm::operator""_dna5;
    using spm::operator""_dna15;
    auto s5 = "ACGNTx"_dna5;
    EXPECT_EQ(s5[3].to_rank(), 3u);
    EXPECT_EQ(s5[4].to_rank(), 4u);
    EXPECT_EQ(s5[5].to_char(), 'N'); // unknown -> N like seqan3::dna5
    auto s15 = "ABCDGHKMNRSTVWYu"_dna15;
    for (std::size_t i = 0; i < 15; ++i) EXPECT_EQ(static_cast<std::size_t>(s15[i].to_rank()), i);
    EXPECT_EQ(s15[15].to_char(), 'T');
    EXPECT_EQ(spm::dna5::alphabet_size, 5u);
    EXPECT_EQ(spm::dna15::alphabet_size, 15u);
}

MINI_TEST(dna5_patterns) {
    using spm::operator""_dna5;
    //                0         1         2         3         4
    //                01234567890123456789012345678901234567890123
    auto hay5    = "ACGTGACTAGCNCGTGACTAGCACGTGACTAGCNCGTGACTAGC"_dna5;
    auto needle5 = "GCNCG"_dna5;
    std::vector<std::size_t> begins{}, ends{};
    spm::horspool_matcher h{needle5};
    h(hay5, [&](auto const &f) { begins.push_back(seqan2::beginPosition(f)); });
    EXPECT_TRUE(std::ranges::equal(begins, std::vector<std::size_t>{9, 31}));
    begins.clear();
    spm::shiftor_matcher so{needle5};
    so(hay5, [&](auto const &f) { begins.push_back(seqan2::beginPosition(f)); });
    EXPECT_TRUE(std::ranges::equal(begins, std::vector<std::size_t>{9, 31}));
    // k = 1: the N occurrences within one error around 14 and 36, the GCACG at 20 by one substitution
    spm::myers_matcher my{needle5, 1};
    EXPECT_EQ(spm::window_size(my), 6u);
    my(hay5, [&](auto const &f) { ends.push_back(seqan2::endPosition(f)); });
    EXPECT_TRUE(std::ranges::equal(ends, std::vector<std::size_t>{13, 14, 15, 25, 35, 36, 37}));
    // the same haystack resident on the device
    spm::device_sequence dev{hay5};
    ends.clear();
    my(dev, [&](auto const &f) { ends.push_back(seqan2::endPosition(f)); });
    EXPECT_TRUE(std::ranges::equal(ends, std::vector<std::size_t>{13, 14, 15, 25, 35, 36, 37}));
}

MINI_TEST(dna15_patterns) {
    using spm::operator""_dna15;
    auto hay15    = "ACGTRYACKMSWGTBDHVNACGTRYAC"_dna15;
    auto needle15 = "RYAC"_dna15;
    std::vector<std::size_t> begins{};
    spm::shiftor_matcher so{needle15};
    so(hay15, [&](auto const &f) { begins.push_back(seqan2::beginPosition(f)); });
    EXPECT_TRUE(std::ranges::equal(begins, std::vector<std::size_t>{4, 23}));
    // letters are symbols, not ambiguity sets: R does not match A or G (the reference compares ranks too)
    auto needle_a = "AYAC"_dna15;
    begins.clear();
    spm::horspool_matcher h{needle_a};
    h(hay15, [&](auto const &f) { begins.push_back(seqan2::beginPosition(f)); });
    EXPECT_TRUE(begins.empty());
}

MINI_TEST(needle_and_haystack_of_different_alphabets_are_rejected) {
    using spm::operator""_dna5;
    auto needle5 = "GCNCG"_dna5;
    spm::shiftor_matcher so{needle5};
    bool threw = false;
    try {
        so(haystack, [&](auto const &) {});
    } catch (spm::device_error const &e) {
        threw = e.status == SPMB_ERR_ARG;
    }
    EXPECT_TRUE(threw);
}

// ---- round 2: what the captured state shows, and the matcher call over a group of devices ----
MINI_TEST(restorable_myers_state_exposes_errors) {
    // capture().errors of the reference's state (myers_matcher_restorable.hpp:57-59): the score at the current column
    auto matcher = spm::restorable_myers_matcher{needle, 1u};
    EXPECT_EQ(spm::capture(matcher).errors(), std::ranges::size(needle)); // nothing read yet
    sequence_t first(haystack.begin(), haystack.begin() + 13);           // ACGTGACTAGCAC: GCAC matches GCACG up to one deletion
    matcher(first, [](auto const &) {});
    EXPECT_EQ(spm::capture(matcher).errors(), 1u);
    EXPECT_EQ(spm::capture(matcher).offset(), 13u);
    sequence_t next(haystack.begin() + 13, haystack.begin() + 14);       // ...G: an exact occurrence ends here
    matcher(next, [](auto const &) {});
    EXPECT_EQ(spm::capture(matcher).errors(), 0u);
}

MINI_TEST(device_group_equals_single_device) {
    // a long haystack (the golden one repeated with a seeded twist) over three shards; two of them share device 0
    sequence_t big;
    std::uint32_t x = 12345;
    for (int r = 0; r < 20000; ++r) {
        for (auto c : haystack) big.push_back(c);
        x = x * 1664525u + 1013904223u;
        big.push_back(spm::dna4{}.assign_rank(static_cast<std::uint8_t>(x >> 30)));
    }
    spm::device_group gpus{std::vector<int>{0, 0, 0}};
    EXPECT_EQ(gpus.size(), 3u);
    {
        auto matcher = spm::horspool_matcher{needle};
        std::vector<std::size_t> one, many;
        matcher(big, [&](auto const &f) { one.push_back(seqan2::beginPosition(f)); });
        gpus.search(matcher, big, [&](auto const &f) { many.push_back(seqan2::beginPosition(f)); });
        EXPECT_TRUE(one.size() > 50000);
        EXPECT_TRUE(one == many);
    }
    {
        auto matcher = spm::myers_matcher{needle, 1};
        std::vector<std::size_t> one, many;
        matcher(big, [&](auto const &f) { one.push_back(seqan2::endPosition(f) * 4 + f.errors()); });
        gpus.search(matcher, big, [&](auto const &f) { many.push_back(seqan2::endPosition(f) * 4 + f.errors()); });
        EXPECT_TRUE(one.size() > 150000);
        EXPECT_TRUE(one == many);
    }
}

int main() { return mini::run_all(); }
