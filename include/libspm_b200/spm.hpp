// spm.hpp -- C++20 facade: the reference's matcher API surface on top of the C ABI (spmb.h).
//
// Mirrors, name for name, what callers of rrahn/libspm use on the hot path:
//   spm::horspool_matcher{needle}            libspm/libspm/matcher/horspool_matcher.hpp:20-44
//   spm::shiftor_matcher{needle}             libspm/libspm/matcher/shiftor_matcher.hpp:20-44
//   spm::myers_matcher{needle, k}            libspm/libspm/matcher/myers_matcher.hpp:19-60
//   spm::restorable_myers_matcher{needle,k}  libspm/libspm/matcher/myers_matcher_restorable.hpp:110-156
//   spm::restorable_shiftor_matcher{needle}  libspm/libspm/matcher/shiftor_matcher_restorable.hpp:89-129
//   spm::pigeonhole_matcher{needle(s), rate} libspm/libspm/matcher/pigeonhole_matcher.hpp:129-243
//   seqan2::PigeonholeSeedOnlyPosition       libspm/libspm/matcher/pigeonhole_matcher.hpp:41-57
//   spm::restorable_myers_prefix_matcher{needle,k}  libspm/libspm/matcher/myers_prefix_matcher_restorable.hpp:117-160
//   spm::filtered_myers_matcher{needles, k}  (new: myers_matcher's hit list through the pigeonhole seed filter)
//   matcher(haystack, callback)              libspm/libspm/matcher/seqan_pattern_base.hpp:40-52
//   spm::window_size / capture / restore     libspm/libspm/matcher/concept.hpp:27-102
//   spm::window_matcher / restorable_matcher libspm/libspm/matcher/concept.hpp:126-148
//   spm::dna4, operator""_dna4               libspm/libspm/seqan/alphabet.hpp:34-83,118,122-135
//   spm::dna5, spm::dna15, _dna5               libspm/libspm/seqan/alphabet.hpp:119-120,137-150
//   seqan2::beginPosition / endPosition      (what the reference's callbacks call on the finder)
// so that the reference's matcher tests read the same against this header.
//
// Differences a maintainer must know: the needle is copied at construction (the
// reference only views it); the callback receives a spm::hit_cursor instead of a seqan2::Finder
// (same free functions, plus errors()); hits are replayed on the host, ascending, after the
// kernels finished; a haystack may also be a device-resident spm::device_sequence.
// No CPU search path exists: every call needs a CUDA device and throws spm::device_error otherwise.
#pragma once

#include <concepts>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <ranges>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <utility>
#include <vector>

#include "../spmb.h"

namespace seqan2 {
// what pigeonhole_matcher::position() returns (pigeonhole_matcher.hpp:41-57)
struct PigeonholeSeedOnlyPosition {
    std::ptrdiff_t index{};  // needle index
    std::ptrdiff_t offset{}; // offset of the seed inside that needle
    std::ptrdiff_t count{};  // seed length q
    friend constexpr bool operator==(PigeonholeSeedOnlyPosition const &, PigeonholeSeedOnlyPosition const &) noexcept = default;
    template <typename stream_t>
    friend stream_t &operator<<(stream_t &stream, PigeonholeSeedOnlyPosition const &me) {
        stream << "<" << me.index << ", " << me.offset << ", " << me.count << ">";
        return stream;
    }
};
} // namespace seqan2

namespace spm {

// ------------------------------------------------------------------------------------------
// alphabet: one symbol = one byte holding the rank (A=0,C=1,G=2,T=3)
// ------------------------------------------------------------------------------------------
class dna4 {
    std::uint8_t _rank{0};
public:
    constexpr dna4() noexcept = default;
    constexpr dna4 &assign_rank(std::uint8_t r) noexcept { _rank = r & 3u; return *this; }
    constexpr dna4 &assign_char(char c) noexcept {
        switch (c) {
        case 'C': case 'c': _rank = 1; break;
        case 'G': case 'g': _rank = 2; break;
        case 'T': case 't': case 'U': case 'u': _rank = 3; break;
        default: _rank = 0; // like seqan3::dna4: unknown characters convert to 'A'
        }
        return *this;
    }
    constexpr std::uint8_t to_rank() const noexcept { return _rank; }
    constexpr char to_char() const noexcept { return "ACGT"[_rank]; }
    // alphabet_adaptor: implicit conversion to integral yields the rank (alphabet.hpp:68-72)
    template <std::integral T> requires (!std::same_as<T, char>)
    constexpr operator T() const noexcept { return static_cast<T>(_rank); }
    constexpr operator char() const noexcept { return to_char(); }
    friend constexpr bool operator==(dna4 a, dna4 b) noexcept { return a._rank == b._rank; }
    friend constexpr auto operator<=>(dna4 a, dna4 b) noexcept { return a._rank <=> b._rank; }
    static constexpr std::size_t alphabet_size = 4;
};
static_assert(sizeof(dna4) == 1);

inline std::vector<dna4> operator""_dna4(const char *s, std::size_t n) {
    std::vector<dna4> v(n);
    for (std::size_t i = 0; i < n; ++i) v[i].assign_char(s[i]);
    return v;
}

// spm::dna5 / spm::dna15 (alphabet.hpp:119-120 = alphabet_adaptor over seqan3::dna5 / seqan3::dna15): ranks in
// seqan3's order (alphabetical: ACGNT, ABCDGHKMNRSTVWY), unknown characters convert to 'N'.  On the device
// these travel in the 4-bit layout (two plane pairs; spmb_store_create_wide).
namespace detail {
template <std::size_t SIZE>
class wide_nucleotide {
    std::uint8_t _rank{0};
    static constexpr const char *letters() noexcept { return SIZE == 5 ? "ACGNT" : "ABCDGHKMNRSTVWY"; }
public:
    constexpr wide_nucleotide() noexcept = default;
    constexpr wide_nucleotide &assign_rank(std::uint8_t r) noexcept { _rank = r < SIZE ? r : std::uint8_t(SIZE - 1); return *this; }
    constexpr wide_nucleotide &assign_char(char c) noexcept {
        if (c >= 'a' && c <= 'z') c = static_cast<char>(c - 'a' + 'A');
        if (c == 'U') c = 'T';
        std::uint8_t n_rank = 0, found = 0xff;
        for (std::uint8_t i = 0; i < SIZE; ++i) {
            if (letters()[i] == 'N') n_rank = i;
            if (letters()[i] == c) found = i;
        }
        _rank = found != 0xff ? found : n_rank;
        return *this;
    }
    constexpr std::uint8_t to_rank() const noexcept { return _rank; }
    constexpr char to_char() const noexcept { return letters()[_rank]; }
    template <std::integral T> requires (!std::same_as<T, char>)
    constexpr operator T() const noexcept { return static_cast<T>(_rank); }
    constexpr operator char() const noexcept { return to_char(); }
    friend constexpr bool operator==(wide_nucleotide a, wide_nucleotide b) noexcept { return a._rank == b._rank; }
    friend constexpr auto operator<=>(wide_nucleotide a, wide_nucleotide b) noexcept { return a._rank <=> b._rank; }
    static constexpr std::size_t alphabet_size = SIZE;
};
} // namespace detail
using dna5 = detail::wide_nucleotide<5>;
using dna15 = detail::wide_nucleotide<15>;
static_assert(sizeof(dna5) == 1 && sizeof(dna15) == 1);

inline std::vector<dna5> operator""_dna5(const char *s, std::size_t n) {
    std::vector<dna5> v(n);
    for (std::size_t i = 0; i < n; ++i) v[i].assign_char(s[i]);
    return v;
}
inline std::vector<dna15> operator""_dna15(const char *s, std::size_t n) {
    std::vector<dna15> v(n);
    for (std::size_t i = 0; i < n; ++i) v[i].assign_char(s[i]);
    return v;
}

// ------------------------------------------------------------------------------------------
// errors + the process-wide device context
// ------------------------------------------------------------------------------------------
struct device_error : std::runtime_error {
    int status;
    device_error(int s, const std::string &what) : std::runtime_error(what), status(s) {}
};

namespace detail {

inline void check(int rc, spmb_context *ctx, const char *where) {
    if (rc == SPMB_OK) return;
    std::string msg = std::string(where) + ": " + spmb_status_string(rc);
    if (ctx) { msg += " ("; msg += spmb_last_error(ctx); msg += ")"; }
    throw device_error(rc, msg);
}

struct context_holder {
    spmb_context *ctx{nullptr};
    spmb_hits *hits{nullptr};
    context_holder() {
        check(spmb_context_create(0, nullptr, &ctx), nullptr, "spmb_context_create");
        check(spmb_hits_create(ctx, 0, &hits), ctx, "spmb_hits_create");
    }
    ~context_holder() { spmb_hits_destroy(hits); spmb_context_destroy(ctx); }
    context_holder(const context_holder &) = delete;
    context_holder &operator=(const context_holder &) = delete;
};

// one context per host thread: a context is single-threaded, like a reference matcher
inline context_holder &device() {
    thread_local context_holder h;
    return h;
}

// rank of one symbol of any supported value type
template <typename T>
constexpr std::uint8_t rank_of(const T &v) noexcept {
    if constexpr (requires { v.to_rank(); }) return static_cast<std::uint8_t>(v.to_rank());
    else return static_cast<std::uint8_t>(v) & 3u;
}

// symbol width of a range on the device: 2 bits for dna4 and plain integers, 4 for alphabets with more than
// four letters (seqan2::BitsPerValue<dna5> = 3, <dna15> = 4, alphabet.hpp:107-112; both travel as 4)
template <typename V>
constexpr std::uint32_t symbol_bits() noexcept {
    if constexpr (requires { V::alphabet_size; }) return V::alphabet_size > 4 ? 4u : 2u;
    else return 2u;
}
template <typename R>
constexpr std::uint32_t range_bits() noexcept { return symbol_bits<std::remove_cvref_t<std::ranges::range_value_t<R>>>(); }

// a view of a range as contiguous 1-byte ranks: zero-copy for contiguous ranges of 1-byte
// symbols (std::vector<spm::dna4>, std::span<uint8_t> ...), a packed temporary otherwise
template <std::ranges::input_range R>
struct rank_buffer {
    std::vector<std::uint8_t> tmp;
    const std::uint8_t *ptr{nullptr};
    std::size_t n{0};
    explicit rank_buffer(R &&r) {
        using V = std::ranges::range_value_t<R>;
        if constexpr (std::ranges::contiguous_range<R> && std::ranges::sized_range<R> && sizeof(V) == 1 &&
                      std::is_trivially_copyable_v<V>) {
            n = std::ranges::size(r);
            ptr = reinterpret_cast<const std::uint8_t *>(std::ranges::data(r));
        } else {
            for (auto &&v : r) tmp.push_back(rank_of(v));
            n = tmp.size();
            ptr = tmp.data();
        }
    }
};

struct pattern_deleter { void operator()(spmb_pattern *p) const noexcept { spmb_pattern_destroy(p); } };
struct store_deleter { void operator()(spmb_store *p) const noexcept { spmb_store_destroy(p); } };
struct stream_deleter { void operator()(spmb_stream *p) const noexcept { spmb_stream_destroy(p); } };

} // namespace detail

// ------------------------------------------------------------------------------------------
// what the callback receives (stands in for seqan2::Finder)
// ------------------------------------------------------------------------------------------
class hit_cursor {
    std::size_t _begin{}, _end{};
    std::uint32_t _errors{}, _seq{}, _pattern{};
public:
    hit_cursor() = default;
    hit_cursor(std::size_t b, std::size_t e, std::uint32_t err, std::uint32_t seq, std::uint32_t pat) noexcept
        : _begin{b}, _end{e}, _errors{err}, _seq{seq}, _pattern{pat} {}
    std::size_t begin_position() const noexcept { return _begin; } // exact: end - m; Myers: end - m (nominal)
    std::size_t end_position() const noexcept { return _end; }
    std::uint32_t errors() const noexcept { return _errors; }
    std::uint32_t sequence_id() const noexcept { return _seq; }
    std::uint32_t pattern_id() const noexcept { return _pattern; }
};

// a sequence packed once into HBM and searched many times (north_star item 1)
class device_sequence {
    std::shared_ptr<spmb_store> _store;
public:
    device_sequence() = default;
    template <std::ranges::input_range R>
    explicit device_sequence(R &&haystack) {
        auto &d = detail::device();
        constexpr std::uint32_t bits = detail::range_bits<R>();
        detail::rank_buffer<R> buf{std::forward<R>(haystack)};
        spmb_store *s = nullptr;
        detail::check(spmb_store_create_wide(d.ctx, buf.ptr, buf.n, bits, 0, &s), d.ctx, "spmb_store_create");
        _store.reset(s, detail::store_deleter{});
    }
    explicit device_sequence(spmb_store *adopt) : _store(adopt, detail::store_deleter{}) {}
    // journaled sequence: this + events -> derived sequence
    device_sequence journaled(const std::vector<spmb_journal_event> &events, const std::vector<std::uint8_t> &pool) const {
        auto &d = detail::device();
        spmb_store *s = nullptr;
        detail::check(spmb_store_create_journaled(d.ctx, _store.get(), events.data(), events.size(), pool.data(),
                                                  pool.size(), 0, &s), d.ctx, "spmb_store_create_journaled");
        return device_sequence{s};
    }
    std::size_t size() const noexcept { return _store ? spmb_store_length(_store.get()) : 0; }
    spmb_store *handle() const noexcept { return _store.get(); }
};

// journaled string tree (north_star items 1-2): a reference, its variants and the haplotype coverage.
// matcher(tree, callback) reports every hit of every haplotype, cursor.sequence_id() = haplotype,
// endPosition in that haplotype's coordinates -- what an external JST traverser computes by sizing branch
// contexts with spm::window_size and running the matcher on each (concept.hpp:27-50).
class journaled_string_tree {
    device_sequence _reference;
    std::shared_ptr<spmb_jst> _jst;
    std::uint32_t _n_haplotypes{0};
public:
    struct jst_deleter { void operator()(spmb_jst *j) const noexcept { spmb_jst_destroy(j); } };
    journaled_string_tree() = default;
    // variants sorted by ref_pos (insertions first on ties); coverage: one row of ceil(n_haplotypes/64)
    // 64-bit words per variant, bit h set = haplotype h carries the variant
    journaled_string_tree(device_sequence reference, const std::vector<spmb_journal_event> &variants,
                          const std::vector<std::uint8_t> &pool, const std::vector<std::uint64_t> &coverage,
                          std::uint32_t n_haplotypes)
        : _reference{std::move(reference)}, _n_haplotypes{n_haplotypes} {
        auto &d = detail::device();
        spmb_jst *j = nullptr;
        detail::check(spmb_jst_create(d.ctx, _reference.handle(), variants.data(), variants.size(), pool.data(), pool.size(),
                                      coverage.data(), n_haplotypes, &j), d.ctx, "spmb_jst_create");
        _jst.reset(j, jst_deleter{});
    }
    std::uint32_t haplotype_count() const noexcept { return _n_haplotypes; }
    const device_sequence &reference() const noexcept { return _reference; }
    spmb_jst *handle() const noexcept { return _jst.get(); }
};

// ------------------------------------------------------------------------------------------
// CPOs (concept.hpp:27-102): member-function dispatch is all the new matchers need
// ------------------------------------------------------------------------------------------
inline constexpr struct window_size_fn {
    template <typename M> requires requires (const M &m) { { m.window_size() } -> std::integral; }
    constexpr auto operator()(const M &m) const noexcept { return m.window_size(); }
} window_size{};

inline constexpr struct capture_fn {
    template <typename M> requires requires (const M &m) { m.capture(); }
    constexpr decltype(auto) operator()(const M &m) const { return m.capture(); }
} capture{};

inline constexpr struct restore_fn {
    template <typename M, typename S> requires requires (M &m, S &&s) { m.restore(std::forward<S>(s)); }
    constexpr void operator()(M &m, S &&s) const { m.restore(std::forward<S>(s)); }
} restore{};

template <typename M>
using matcher_state_t = std::remove_cvref_t<std::invoke_result_t<capture_fn, M>>;

template <typename M>
concept window_matcher = std::copyable<std::remove_cvref_t<M>> && requires (M &&m) {
    { spm::window_size(m) } -> std::integral;
};

template <typename M>
concept restorable_matcher = window_matcher<M> && requires (std::remove_cvref_t<M> &m) {
    typename matcher_state_t<M>;
    requires std::semiregular<matcher_state_t<M>>;
    { spm::capture(m) } -> std::convertible_to<matcher_state_t<M>>;
    { spm::restore(m, std::declval<matcher_state_t<M>>()) };
};

template <typename M, typename... Args>
concept online_matcher_for = window_matcher<M> && std::invocable<M, Args...>;

// ------------------------------------------------------------------------------------------
// matcher base: seqan_pattern_base's driver loop, device edition
// ------------------------------------------------------------------------------------------
template <typename derived_t, int ALG>
class device_pattern_base {
protected:
    std::shared_ptr<spmb_pattern> _pattern;
    std::size_t _needle_size{0};
    std::size_t _errors{0};
    std::uint32_t _bits{2};
    std::vector<std::uint8_t> _ranks; // the needle's ranks (device_group builds the pattern on every device from them)

    template <std::ranges::input_range N>
    void build(N &&needle, std::size_t k) {
        auto &d = detail::device();
        _bits = detail::range_bits<N>();
        detail::rank_buffer<N> buf{std::forward<N>(needle)};
        _needle_size = buf.n;
        _errors = k;
        _ranks.assign(buf.ptr, buf.ptr + buf.n);
        spmb_pattern *p = nullptr;
        if (_bits == 2)
            detail::check(spmb_pattern_create(d.ctx, ALG, buf.ptr, static_cast<std::uint32_t>(buf.n),
                                              static_cast<std::uint32_t>(k), &p), d.ctx, "spmb_pattern_create");
        else // spm::dna5 / spm::dna15 needles
            detail::check(spmb_pattern_create_wide(d.ctx, ALG, buf.ptr, static_cast<std::uint32_t>(buf.n),
                                                   static_cast<std::uint32_t>(k), _bits, &p), d.ctx, "spmb_pattern_create_wide");
        _pattern.reset(p, detail::pattern_deleter{});
    }

    template <typename callback_t>
    void replay(callback_t &&callback) const {
        auto &d = detail::device();
        const spmb_hit *h = spmb_hits_data(d.hits);
        const std::uint64_t n = spmb_hits_count(d.hits);
        for (std::uint64_t i = 0; i < n; ++i) {
            const std::size_t e = static_cast<std::size_t>(h[i].end);
            const std::size_t b = e >= _needle_size ? e - _needle_size : 0;
            callback(hit_cursor{b, e, h[i].errors, h[i].seq_id, h[i].pattern_id});
        }
    }

public:
    // matcher(haystack, callback): every hit, ascending, callback(cursor) per hit
    // (seqan_pattern_base.hpp:40-52).  Non-const like the reference.
    template <std::ranges::viewable_range haystack_t, typename callback_t>
        requires (!std::same_as<std::remove_cvref_t<haystack_t>, device_sequence> &&
                  !std::same_as<std::remove_cvref_t<haystack_t>, journaled_string_tree>)
    void operator()(haystack_t &&haystack, callback_t &&callback) {
        auto &d = detail::device();
        constexpr std::uint32_t bits = detail::range_bits<haystack_t>();
        detail::rank_buffer<haystack_t> buf{std::forward<haystack_t>(haystack)};
        if (bits == 2 && _bits == 2)
            detail::check(spmb_search_host(d.ctx, buf.ptr, buf.n, _pattern.get(), d.hits), d.ctx, "spmb_search_host");
        else // a needle and a haystack of different widths is the library's SPMB_ERR_ARG
            detail::check(spmb_search_host_wide(d.ctx, buf.ptr, buf.n, bits, _pattern.get(), d.hits), d.ctx, "spmb_search_host_wide");
        replay(std::forward<callback_t>(callback));
    }

    // same call over a device-resident sequence
    template <typename callback_t>
    void operator()(const device_sequence &haystack, callback_t &&callback) {
        auto &d = detail::device();
        detail::check(spmb_search(d.ctx, haystack.handle(), _pattern.get(), d.hits), d.ctx, "spmb_search");
        replay(std::forward<callback_t>(callback));
    }

    // same call over a journaled string tree: hits of all haplotypes, ordered by (haplotype, end)
    template <typename callback_t>
    void operator()(const journaled_string_tree &tree, callback_t &&callback) {
        auto &d = detail::device();
        detail::check(spmb_jst_search(d.ctx, tree.handle(), _pattern.get(), d.hits), d.ctx, "spmb_jst_search");
        replay(std::forward<callback_t>(callback));
    }

    bool empty() const noexcept { return _needle_size == 0; }
    static constexpr int algorithm = ALG;
    const std::vector<std::uint8_t> &needle_ranks() const noexcept { return _ranks; }
    std::size_t max_error_count() const noexcept { return _errors; }
    std::uint32_t symbol_bits() const noexcept { return _bits; }

    // seqan_pattern_base.hpp:97-99 (+ myers_matcher.hpp:51-53 through _errors)
    std::size_t window_size() const noexcept { return empty() ? 0 : _needle_size + _errors; }
};

template <std::ranges::random_access_range needle_t = std::vector<dna4>>
class horspool_matcher : public device_pattern_base<horspool_matcher<needle_t>, SPMB_ALG_HORSPOOL> {
public:
    horspool_matcher() = delete;
    template <std::ranges::viewable_range N> requires (!std::same_as<std::remove_cvref_t<N>, horspool_matcher>)
    explicit horspool_matcher(N &&needle) { this->build(std::forward<N>(needle), 0); }
};
template <std::ranges::viewable_range N> horspool_matcher(N &&) -> horspool_matcher<std::views::all_t<N>>;

template <std::ranges::random_access_range needle_t = std::vector<dna4>>
class shiftor_matcher : public device_pattern_base<shiftor_matcher<needle_t>, SPMB_ALG_SHIFTOR> {
public:
    shiftor_matcher() = delete;
    template <std::ranges::viewable_range N> requires (!std::same_as<std::remove_cvref_t<N>, shiftor_matcher>)
    explicit shiftor_matcher(N &&needle) { this->build(std::forward<N>(needle), 0); }
};
template <std::ranges::viewable_range N> shiftor_matcher(N &&) -> shiftor_matcher<std::views::all_t<N>>;

template <std::ranges::random_access_range needle_t = std::vector<dna4>>
class myers_matcher : public device_pattern_base<myers_matcher<needle_t>, SPMB_ALG_MYERS> {
public:
    myers_matcher() = delete;
    template <std::ranges::viewable_range N> requires (!std::same_as<std::remove_cvref_t<N>, myers_matcher>)
    explicit myers_matcher(N &&needle, std::size_t max_error_count = 0) { this->build(std::forward<N>(needle), max_error_count); }
};
template <std::ranges::viewable_range N> myers_matcher(N &&) -> myers_matcher<std::views::all_t<N>>;
template <std::ranges::viewable_range N> myers_matcher(N &&, std::size_t) -> myers_matcher<std::views::all_t<N>>;

// ------------------------------------------------------------------------------------------
// restorable matchers: capture()/restore() around chunked calls
// ------------------------------------------------------------------------------------------
// The state is the carried context (last window_size-1 symbols + offset).  It is a value type
// (semiregular) like the reference's TPatternState / prefSufMatch copies.
struct matcher_state {
    std::vector<std::uint8_t> blob; // spmb_stream_capture: {u64 offset, u32 tail_len, u32 n_patterns, u32 errors[], u8 tail[]}
    friend bool operator==(const matcher_state &, const matcher_state &) = default;
    // what the reference's captured state shows a caller: the score at the current column
    // (capture().errors, myers_matcher_restorable.hpp:57-59); 0 for the exact searchers
    std::uint32_t errors() const noexcept {
        std::uint32_t e = 0;
        if (blob.size() >= 20) std::memcpy(&e, blob.data() + 16, 4);
        return e;
    }
    std::uint64_t offset() const noexcept {
        std::uint64_t o = 0;
        if (blob.size() >= 8) std::memcpy(&o, blob.data(), 8);
        return o;
    }
};

template <typename derived_t, int ALG>
class restorable_base : public device_pattern_base<derived_t, ALG> {
    using base_t = device_pattern_base<derived_t, ALG>;
    std::shared_ptr<spmb_stream> _stream;
    mutable matcher_state _captured;

    void ensure_stream() {
        if (_stream) return;
        auto &d = detail::device();
        spmb_stream *s = nullptr;
        detail::check(spmb_stream_create(d.ctx, this->_pattern.get(), &s), d.ctx, "spmb_stream_create");
        _stream.reset(s, detail::stream_deleter{});
    }

protected:
    template <std::ranges::input_range N>
    void build_restorable(N &&needle, std::size_t k) {
        base_t::build(std::forward<N>(needle), k);
        ensure_stream(); // eager, so that capture() is valid before the first search
                         // (myers_matcher_restorable.hpp:35-40)
    }

public:
    using state_type = matcher_state;

    restorable_base() = default;
    // copies must not share the mutable stream: deep-copy its state
    restorable_base(const restorable_base &o) : base_t(o) {
        if (o._stream) { ensure_stream(); restore(o.capture()); }
    }
    restorable_base &operator=(const restorable_base &o) {
        if (this != &o) {
            base_t::operator=(o);
            _stream.reset();
            if (o._stream) { ensure_stream(); restore(o.capture()); }
        }
        return *this;
    }
    restorable_base(restorable_base &&) noexcept = default;
    restorable_base &operator=(restorable_base &&) noexcept = default;

    // each call sees a fresh "finder" on the chunk: positions are chunk-relative
    // (myers_matcher_restorable_test.cpp:66-69), matcher state flows across calls
    template <std::ranges::viewable_range haystack_t, typename callback_t>
    void operator()(haystack_t &&haystack, callback_t &&callback) {
        auto &d = detail::device();
        detail::rank_buffer<haystack_t> buf{std::forward<haystack_t>(haystack)};
        detail::check(spmb_stream_feed(d.ctx, _stream.get(), buf.ptr, buf.n, d.hits), d.ctx, "spmb_stream_feed");
        this->replay(std::forward<callback_t>(callback));
    }

    const state_type &capture() const {
        std::size_t sz = 0;
        spmb_stream_capture(_stream.get(), nullptr, &sz);
        _captured.blob.resize(sz);
        detail::check(spmb_stream_capture(_stream.get(), _captured.blob.data(), &sz), nullptr, "spmb_stream_capture");
        return _captured;
    }

    void restore(state_type state) {
        detail::check(spmb_stream_restore(_stream.get(), state.blob.data(), state.blob.size()), nullptr, "spmb_stream_restore");
    }
};

template <std::ranges::random_access_range needle_t = std::vector<dna4>>
class restorable_myers_matcher : public restorable_base<restorable_myers_matcher<needle_t>, SPMB_ALG_MYERS> {
public:
    restorable_myers_matcher() = delete;
    template <std::ranges::viewable_range N, std::unsigned_integral error_count_t>
        requires (!std::same_as<std::remove_cvref_t<N>, restorable_myers_matcher>)
    explicit restorable_myers_matcher(N &&needle, error_count_t const error_count) {
        this->build_restorable(std::forward<N>(needle), error_count);
    }
};
template <std::ranges::viewable_range N, std::unsigned_integral E>
restorable_myers_matcher(N &&, E) -> restorable_myers_matcher<std::views::all_t<N>>;

template <std::ranges::random_access_range needle_t = std::vector<dna4>>
class restorable_shiftor_matcher : public restorable_base<restorable_shiftor_matcher<needle_t>, SPMB_ALG_SHIFTOR> {
public:
    restorable_shiftor_matcher() = delete;
    template <std::ranges::viewable_range N> requires (!std::same_as<std::remove_cvref_t<N>, restorable_shiftor_matcher>)
    explicit restorable_shiftor_matcher(N &&needle) { this->build_restorable(std::forward<N>(needle), 0); }
};
template <std::ranges::viewable_range N>
restorable_shiftor_matcher(N &&) -> restorable_shiftor_matcher<std::views::all_t<N>>;

// anchored (prefix) Myers: the needle against prefixes of the haystack, scan clamped to m+k+1 symbols
// (myers_prefix_matcher_restorable.hpp:47-61); window_size = m + k (:151-153)
template <std::ranges::random_access_range needle_t = std::vector<dna4>>
class restorable_myers_prefix_matcher
    : public restorable_base<restorable_myers_prefix_matcher<needle_t>, SPMB_ALG_MYERS_PREFIX> {
public:
    restorable_myers_prefix_matcher() = delete;
    template <std::ranges::viewable_range N, std::unsigned_integral error_count_t>
        requires (!std::same_as<std::remove_cvref_t<N>, restorable_myers_prefix_matcher>)
    explicit restorable_myers_prefix_matcher(N &&needle, error_count_t const error_count) {
        this->build_restorable(std::forward<N>(needle), error_count);
    }
};
template <std::ranges::viewable_range N, std::unsigned_integral E>
restorable_myers_prefix_matcher(N &&, E) -> restorable_myers_prefix_matcher<std::views::all_t<N>>;

// ------------------------------------------------------------------------------------------
// needle sets: pigeonhole seed matcher and the filtered Myers search built on it
// ------------------------------------------------------------------------------------------
namespace detail {
template <typename R>
concept needle_set = std::ranges::input_range<R> && std::ranges::input_range<std::ranges::range_reference_t<R>>;

struct needle_pack {
    std::vector<std::uint8_t> ranks;
    std::vector<std::uint32_t> lengths;
    template <typename R>
    void add(R &&needle) {
        std::uint32_t n = 0;
        for (auto &&v : needle) { ranks.push_back(rank_of(v)); ++n; }
        lengths.push_back(n);
    }
};
} // namespace detail

template <std::ranges::random_access_range needle_t = std::vector<dna4>>
class pigeonhole_matcher {
    std::shared_ptr<spmb_pattern> _pattern;
    double _error_rate{};
    seqan2::PigeonholeSeedOnlyPosition _position{};

    void build(const detail::needle_pack &np) {
        auto &d = detail::device();
        spmb_pattern *p = nullptr;
        detail::check(spmb_pigeonhole_create(d.ctx, np.ranks.data(), np.lengths.data(),
                                             static_cast<std::uint32_t>(np.lengths.size()), _error_rate, &p),
                      d.ctx, "spmb_pigeonhole_create");
        _pattern.reset(p, detail::pattern_deleter{});
    }

    template <typename callback_t>
    void replay(callback_t &&callback) {
        auto &d = detail::device();
        const spmb_hit *h = spmb_hits_data(d.hits);
        const std::uint64_t n = spmb_hits_count(d.hits);
        const std::size_t q = window_size();
        for (std::uint64_t i = 0; i < n; ++i) {
            const std::size_t e = static_cast<std::size_t>(h[i].end);
            _position = {static_cast<std::ptrdiff_t>(h[i].pattern_id), static_cast<std::ptrdiff_t>(h[i].reserved),
                         static_cast<std::ptrdiff_t>(q)};
            callback(hit_cursor{e - q, e, 0, h[i].seq_id, h[i].pattern_id});
        }
    }

public:
    pigeonhole_matcher() = delete;
    // one needle (pigeonhole_matcher.hpp:157-167)
    template <std::ranges::viewable_range N>
        requires (!std::same_as<std::remove_cvref_t<N>, pigeonhole_matcher> && !detail::needle_set<N>)
    explicit pigeonhole_matcher(N &&needle, double error_rate = 0.0) : _error_rate{error_rate} {
        detail::needle_pack np;
        np.add(std::forward<N>(needle));
        build(np);
    }
    // several needles (pigeonhole_matcher.hpp:169-180)
    template <std::ranges::viewable_range M>
        requires (!std::same_as<std::remove_cvref_t<M>, pigeonhole_matcher> && detail::needle_set<M>)
    explicit pigeonhole_matcher(M &&multi_needle, double error_rate = 0.0) : _error_rate{error_rate} {
        detail::needle_pack np;
        for (auto &&needle : multi_needle) np.add(needle);
        build(np);
    }

    // {needle index, needle offset, q} of the seed the running callback is looking at (:182-184)
    constexpr auto position() const noexcept { return _position; }

    template <std::ranges::viewable_range haystack_t, typename callback_t>
        requires (!std::same_as<std::remove_cvref_t<haystack_t>, device_sequence>)
    void operator()(haystack_t &&haystack, callback_t &&callback) {
        auto &d = detail::device();
        detail::rank_buffer<haystack_t> buf{std::forward<haystack_t>(haystack)};
        detail::check(spmb_search_host(d.ctx, buf.ptr, buf.n, _pattern.get(), d.hits), d.ctx, "spmb_search_host");
        replay(std::forward<callback_t>(callback));
    }
    template <typename callback_t>
    void operator()(const device_sequence &haystack, callback_t &&callback) {
        auto &d = detail::device();
        detail::check(spmb_search(d.ctx, haystack.handle(), _pattern.get(), d.hits), d.ctx, "spmb_search");
        replay(std::forward<callback_t>(callback));
    }

    // the q-gram length (pigeonhole_matcher.hpp:200-202)
    std::size_t window_size() const noexcept { return spmb_pattern_window_size(_pattern.get()); }
};
template <std::ranges::viewable_range N> requires (!detail::needle_set<N>)
pigeonhole_matcher(N &&) -> pigeonhole_matcher<std::views::all_t<N>>;
template <std::ranges::viewable_range N> requires (!detail::needle_set<N>)
pigeonhole_matcher(N &&, double) -> pigeonhole_matcher<std::views::all_t<N>>;
template <std::ranges::viewable_range M> requires detail::needle_set<M>
pigeonhole_matcher(M &&) -> pigeonhole_matcher<std::views::all_t<std::ranges::range_reference_t<M>>>;
template <std::ranges::viewable_range M> requires detail::needle_set<M>
pigeonhole_matcher(M &&, double) -> pigeonhole_matcher<std::views::all_t<std::ranges::range_reference_t<M>>>;

// myers_matcher's hit list for a whole needle set in one pass: pigeonhole seed filter + windowed
// verification on the device (SPMB_ALG_MYERS_FILTERED).  callback(cursor): cursor.pattern_id() names the
// needle, hits ascend per needle.
class filtered_myers_matcher {
    std::shared_ptr<spmb_pattern> _pattern;
    std::vector<std::uint32_t> _lengths;
    std::size_t _errors{0};

    template <typename callback_t>
    void replay(callback_t &&callback) const {
        auto &d = detail::device();
        const spmb_hit *h = spmb_hits_data(d.hits);
        const std::uint64_t n = spmb_hits_count(d.hits);
        for (std::uint64_t i = 0; i < n; ++i) {
            const std::size_t e = static_cast<std::size_t>(h[i].end), m = _lengths[h[i].pattern_id];
            callback(hit_cursor{e >= m ? e - m : 0, e, h[i].errors, h[i].seq_id, h[i].pattern_id});
        }
    }

public:
    filtered_myers_matcher() = delete;
    template <std::ranges::viewable_range M> requires detail::needle_set<M>
    explicit filtered_myers_matcher(M &&multi_needle, std::size_t max_error_count = 0) : _errors{max_error_count} {
        detail::needle_pack np;
        for (auto &&needle : multi_needle) np.add(needle);
        _lengths = np.lengths;
        auto &d = detail::device();
        spmb_pattern *p = nullptr;
        detail::check(spmb_pattern_set_create(d.ctx, SPMB_ALG_MYERS_FILTERED, np.ranks.data(), np.lengths.data(),
                                              static_cast<std::uint32_t>(np.lengths.size()),
                                              static_cast<std::uint32_t>(max_error_count), &p),
                      d.ctx, "spmb_pattern_set_create");
        _pattern.reset(p, detail::pattern_deleter{});
    }
    template <std::ranges::viewable_range haystack_t, typename callback_t>
        requires (!std::same_as<std::remove_cvref_t<haystack_t>, device_sequence>)
    void operator()(haystack_t &&haystack, callback_t &&callback) {
        auto &d = detail::device();
        detail::rank_buffer<haystack_t> buf{std::forward<haystack_t>(haystack)};
        detail::check(spmb_search_host(d.ctx, buf.ptr, buf.n, _pattern.get(), d.hits), d.ctx, "spmb_search_host");
        replay(std::forward<callback_t>(callback));
    }
    template <typename callback_t>
    void operator()(const device_sequence &haystack, callback_t &&callback) {
        auto &d = detail::device();
        detail::check(spmb_search(d.ctx, haystack.handle(), _pattern.get(), d.hits), d.ctx, "spmb_search");
        replay(std::forward<callback_t>(callback));
    }
    std::size_t window_size() const noexcept { return spmb_pattern_window_size(_pattern.get()); }
};

// ------------------------------------------------------------------------------------------
// every GPU of the box behind one matcher call (north_star item 4; spmb_multi_*): ONE process, ONE host haystack.
//   spm::device_group gpus;                       // all visible devices
//   gpus.search(matcher, haystack, callback);     // == matcher(haystack, callback), hits merged and replayed ascending
// The haystack is cut into one shard per device (with window_size-1 symbols of left context), every shard crosses
// its own PCIe link, the per-device hit lists are merged on the host; no NCCL anywhere.
// ------------------------------------------------------------------------------------------
class device_group {
    spmb_multi *_m{nullptr};
    spmb_hits *_hits{nullptr};
    void check(int rc, const char *where) const {
        if (rc == SPMB_OK) return;
        throw device_error(rc, std::string(where) + ": " + spmb_status_string(rc) + " (" + spmb_multi_last_error(_m) + ")");
    }
public:
    explicit device_group(const std::vector<int> &devices = {}) {
        detail::check(spmb_multi_create(devices.empty() ? nullptr : devices.data(), static_cast<std::uint32_t>(devices.size()), &_m),
                      nullptr, "spmb_multi_create");
        int rc = spmb_hits_create(spmb_multi_context(_m, 0), 0, &_hits);
        if (rc) { spmb_multi_destroy(_m); detail::check(rc, nullptr, "spmb_hits_create"); }
    }
    ~device_group() { spmb_hits_destroy(_hits); spmb_multi_destroy(_m); }
    device_group(const device_group &) = delete;
    device_group &operator=(const device_group &) = delete;
    std::uint32_t size() const noexcept { return spmb_multi_device_count(_m); }
    spmb_multi *handle() const noexcept { return _m; }

    template <typename matcher_t, std::ranges::viewable_range haystack_t, typename callback_t>
    void search(const matcher_t &matcher, haystack_t &&haystack, callback_t &&callback) {
        if (matcher.symbol_bits() != 2) throw device_error(SPMB_ERR_UNSUPPORTED, "device_group: dna4 needles only");
        detail::rank_buffer<haystack_t> buf{std::forward<haystack_t>(haystack)};
        const std::vector<std::uint8_t> &nr = matcher.needle_ranks();
        const std::uint32_t len = static_cast<std::uint32_t>(nr.size());
        spmb_multi_pattern *p = nullptr;
        check(spmb_multi_pattern_create(_m, matcher_t::algorithm, nr.data(), &len, 1,
                                        static_cast<std::uint32_t>(matcher.max_error_count()), &p), "spmb_multi_pattern_create");
        int rc = spmb_multi_search_host(_m, buf.ptr, buf.n, p, _hits);
        spmb_multi_pattern_destroy(p);
        check(rc, "spmb_multi_search_host");
        const spmb_hit *h = spmb_hits_data(_hits);
        const std::uint64_t n = spmb_hits_count(_hits);
        for (std::uint64_t i = 0; i < n; ++i) {
            const std::size_t e = static_cast<std::size_t>(h[i].end);
            callback(hit_cursor{e >= nr.size() ? e - nr.size() : 0, e, h[i].errors, h[i].seq_id, h[i].pattern_id});
        }
    }
};

} // namespace spm

// ------------------------------------------------------------------------------------------
// the free functions the reference's callbacks call on the finder
// ------------------------------------------------------------------------------------------
namespace seqan2 {
inline std::size_t beginPosition(const spm::hit_cursor &c) noexcept { return c.begin_position(); }
inline std::size_t endPosition(const spm::hit_cursor &c) noexcept { return c.end_position(); }
inline std::size_t position(const spm::hit_cursor &c) noexcept { return c.begin_position(); }
} // namespace seqan2
