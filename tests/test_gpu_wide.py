"""GPU parity for the 4-bit alphabets (spm::dna5 / spm::dna15, libspm/libspm/seqan/alphabet.hpp:119-120) and for
the two forms of the Myers kernel (Ukkonen cut-off = default, full recurrence): hit lists bit-exact against the
CPU oracle (definitional: all-occurrence compare / Sellers DP; restated Horspool / Shift-Or / Myers)."""
import numpy as np
import pytest

import libspm_b200 as sp
import oracle as orc
from tests.util import rand_text

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = sp.Context(0)
    yield c
    c.close()


def _mutate(rng, pat, e, sigma):
    p = list(pat)
    for _ in range(e):
        op = rng.integers(0, 3)
        i = int(rng.integers(0, len(p)))
        if op == 0:
            p[i] = (p[i] + 1 + int(rng.integers(0, sigma - 1))) % sigma
        elif op == 1:
            p.insert(i, int(rng.integers(0, sigma)))
        elif len(p) > 1:
            del p[i]
    return np.array(p, dtype=np.uint8)


def _plant(rng, text, pat, copies, max_edits, sigma):
    n = len(text)
    for _ in range(copies):
        q = _mutate(rng, pat, int(rng.integers(0, max_edits + 1)), sigma)
        s = int(rng.integers(0, n - len(q)))
        text[s:s + len(q)] = q


@pytest.mark.parametrize("sigma,bits", [(5, 3), (15, 4), (16, 4)])
@pytest.mark.parametrize("n,rl", [(37, 0), (5000, 32), (70001, 64), (400_003, 0)])
def test_wide_round_trip_and_exact(ctx, sigma, bits, n, rl):
    rng = np.random.default_rng(1000 * sigma + n)
    text = rand_text(rng, n, sigma)
    for m in (1, 5, 8, 9, 33, 70):
        if m >= n:
            continue
        s0 = int(rng.integers(0, n - m))
        pat = text[s0:s0 + m].copy()
        if m >= 5:
            _plant(rng, text, pat, 5, 0, sigma)
    store = sp.Store.from_host_wide(ctx, text, bits, rl)
    assert store.symbol_bits == 4 and len(store) == n
    assert np.array_equal(store.read(), text)
    for m in (1, 5, 8, 9, 33, 70):
        if m >= n:
            continue
        s0 = int(rng.integers(0, n - m))
        pat = text[s0:s0 + m].copy()
        want = orc.exact_bruteforce(text, pat)
        assert len(want) >= 1
        # the restated searchers agree with the definition on this alphabet too
        assert np.array_equal(orc.horspool(text, pat), want)
        assert np.array_equal(orc.shiftor(text, pat), want)
        for alg in (sp.ALG_HORSPOOL, sp.ALG_SHIFTOR):
            p = sp.Pattern(ctx, alg, pat, bits=bits)
            assert p.window_size == m
            got = sp.search(ctx, store, p)
            assert np.array_equal(got["end"], want), (sigma, n, rl, m, alg)
            assert not got["errors"].any()
            p.close()
    store.close()


def test_wide_symbols_that_differ_only_in_the_high_bits(ctx):
    """ranks c and c+4 / c+8 / c+12 share their two low bits: the second plane pair must separate them."""
    rng = np.random.default_rng(7)
    n = 20000
    low = rand_text(rng, n, 4)
    text = (low + 4 * rand_text(rng, n, 4)).astype(np.uint8)
    pat = text[1234:1234 + 24].copy()
    twin = (pat ^ 4).astype(np.uint8)          # same low bits, other high bits
    text[5000:5024] = twin
    store = sp.Store.from_host_wide(ctx, text, 4)
    p = sp.Pattern(ctx, sp.ALG_SHIFTOR, pat, bits=4)
    got = sp.search(ctx, store, p)
    assert np.array_equal(got["end"], orc.exact_bruteforce(text, pat))
    assert 5024 not in got["end"].tolist()
    pm = sp.Pattern(ctx, sp.ALG_MYERS, pat, 2, bits=4)
    gm = sp.search(ctx, store, pm)
    we, wr = orc.sellers(text, pat, 2)
    assert np.array_equal(gm["end"], we) and np.array_equal(gm["errors"], wr)
    p.close(); pm.close(); store.close()


@pytest.mark.parametrize("sigma,bits", [(5, 3), (15, 4)])
@pytest.mark.parametrize("m,k", [(7, 1), (32, 3), (33, 2), (64, 5), (100, 4), (150, 2), (200, 12), (256, 6)])
def test_wide_myers(ctx, sigma, bits, m, k):
    rng = np.random.default_rng(31 * m + k + sigma)
    n = 150_000
    text = rand_text(rng, n, sigma)
    pat = rand_text(rng, m, sigma)
    _plant(rng, text, pat, 40, k + 2, sigma)
    store = sp.Store.from_host_wide(ctx, text, bits, 96 if m < 60 else 0)
    _, we, wr = orc.search_mt(3, text, pat, k)
    if m <= 64:   # the restated Myers against the definition (Sellers DP) on a slice
        se, sr = orc.sellers(text[:15000], pat, k)
        keep = we <= 15000
        assert np.array_equal(we[keep], se) and np.array_equal(wr[keep], sr)
    assert len(we) > 0
    p = sp.Pattern(ctx, sp.ALG_MYERS, pat, k, bits=bits)
    assert p.window_size == m + k
    got = sp.search(ctx, store, p)
    assert np.array_equal(got["end"], we), (sigma, m, k)
    assert np.array_equal(got["errors"], wr)
    # the reference's call shape: host haystack in, hits out
    gh = sp.search_host(ctx, text, p, bits=bits)
    assert np.array_equal(gh["end"], we) and np.array_equal(gh["errors"], wr)
    p.close(); store.close()


def test_width_mismatch_and_unsupported_paths_fail_loudly(ctx):
    rng = np.random.default_rng(3)
    text = rand_text(rng, 5000, 5)
    wide = sp.Store.from_host_wide(ctx, text, 3)
    narrow = sp.Store.from_host(ctx, text & 3)
    p2 = sp.Pattern(ctx, sp.ALG_MYERS, text[10:50] & 3, 1)
    p4 = sp.Pattern(ctx, sp.ALG_MYERS, text[10:50], 1, bits=3)
    for st, p in ((wide, p2), (narrow, p4)):
        with pytest.raises(sp.SpmbError) as ei:
            sp.search(ctx, st, p)
        assert ei.value.code == 1
    with pytest.raises(sp.SpmbError) as ei:
        sp.Pattern(ctx, sp.ALG_MYERS_FILTERED, text[10:110], 1, bits=4)
    assert ei.value.code == 4
    js = wide.journaled(np.zeros(0, sp.EVENT_DTYPE), np.zeros(0, np.uint8))     # journals take 4-bit stores (round 2)
    assert js.symbol_bits == 4 and np.array_equal(js.read(), text)
    js.close()
    with pytest.raises(sp.SpmbError):
        sp.Stream(ctx, p4)
    for x in (p2, p4, wide, narrow):
        x.close()


# ---------------- the two forms of the Myers kernel ----------------
def _adversarial_texts(rng, n):
    yield "random", rand_text(rng, n)
    yield "polyA", np.zeros(n, dtype=np.uint8)
    per = rand_text(rng, 700)
    yield "period-700", np.tile(per, n // 700 + 1)[:n].copy()
    two = rand_text(rng, n, 2)
    yield "binary", two


@pytest.mark.parametrize("m,k", [(33, 0), (40, 3), (64, 8), (65, 1), (100, 4), (100, 29), (130, 10), (200, 3), (256, 20)])
def test_myers_cutoff_equals_full_recurrence_and_oracle(ctx, m, k):
    rng = np.random.default_rng(17 * m + k)
    n = 120_000
    for name, text in _adversarial_texts(rng, n):
        s0 = int(rng.integers(0, n - m))
        pat = text[s0:s0 + m].copy()
        # needles a few substitutions off the text keep the active region deep without flooding the hit list
        for i in rng.integers(0, m, size=max(1, k // 2 + 1)):
            pat[i] = (pat[i] + 1) % 4
        for rl in (0, 64):
            store = sp.Store.from_host(ctx, text, rl)
            p = sp.Pattern(ctx, sp.ALG_MYERS, pat, k)
            ctx.set_myers_cutoff(True)
            a = sp.search(ctx, store, p).copy()
            ctx.set_myers_cutoff(False)
            b = sp.search(ctx, store, p).copy()
            ctx.set_myers_cutoff(True)
            _, we, wr = orc.search_mt(3, text, pat, k)
            assert np.array_equal(a["end"], b["end"]) and np.array_equal(a["errors"], b["errors"]), (name, m, k, rl)
            assert np.array_equal(a["end"], we) and np.array_equal(a["errors"], wr), (name, m, k, rl)
            p.close(); store.close()


def test_wide_empty_haystack_and_empty_needle(ctx):
    """like the dna4 edge cases: nothing to find, no error (window_size 0 for an empty needle)."""
    text = np.array([4, 0, 1, 4, 2], dtype=np.uint8)
    empty = sp.Store.from_host_wide(ctx, np.zeros(0, np.uint8), 3)
    store = sp.Store.from_host_wide(ctx, text, 3)
    assert len(empty) == 0 and empty.symbol_bits == 4
    for alg, k in ((sp.ALG_HORSPOOL, 0), (sp.ALG_SHIFTOR, 0), (sp.ALG_MYERS, 1)):
        p = sp.Pattern(ctx, alg, text[1:4], k, bits=3)
        assert len(sp.search(ctx, empty, p)) == 0
        p.close()
        p0 = sp.Pattern(ctx, alg, np.zeros(0, np.uint8), k, bits=3)
        assert p0.window_size == 0
        assert len(sp.search(ctx, store, p0)) == 0
        p0.close()
    # needle longer than the haystack
    p = sp.Pattern(ctx, sp.ALG_SHIFTOR, np.array([4, 0, 1, 4, 2, 2], dtype=np.uint8), bits=3)
    assert len(sp.search(ctx, store, p)) == 0
    p.close(); empty.close(); store.close()


def test_myers_row_split_gives_the_same_hits(ctx, monkeypatch):
    """SPMB_MYERS_SPLIT cuts every row into pieces with their own warm-up (one warp per piece): the hit list must
    not change, including hits that straddle a piece seam."""
    rng = np.random.default_rng(99)
    n = 400_000
    text = rand_text(rng, n)
    pat = rand_text(rng, 90)
    _plant(rng, text, pat, 300, 5, 4)
    store = sp.Store.from_host(ctx, text, 2048)       # long rows: up to 3 pieces of >= 4 x (warm-up + 32) columns
    p = sp.Pattern(ctx, sp.ALG_MYERS, pat, 4)
    monkeypatch.delenv("SPMB_MYERS_SPLIT", raising=False)
    base = sp.search(ctx, store, p).copy()
    _, we, wr = orc.search_mt(3, text, pat, 4)
    assert np.array_equal(base["end"], we) and np.array_equal(base["errors"], wr) and len(we) > 100
    for split in ("2", "3", "8"):
        monkeypatch.setenv("SPMB_MYERS_SPLIT", split)
        got = sp.search(ctx, store, p)
        assert np.array_equal(got["end"], we) and np.array_equal(got["errors"], wr), split
    p.close(); store.close()


def test_long_hit_lists_ordered_by_the_launch(ctx):
    """A hit list that was long (>= 16384 raw hits) once is sorted / deduplicated by the next launch itself
    (spmb_search_launch enqueues the ordering behind the kernels): same lists as the first pass and the oracle,
    through a reused Hits object, pipelined launches, and a later short list on the same object."""
    rng = np.random.default_rng(2024)
    n = 300_000
    text = rand_text(rng, n, 2)                       # binary text: a short needle hits often
    store = sp.Store.from_host(ctx, text)
    hits = sp.Hits(ctx)
    pat = text[1000:1012].copy()
    want_exact = orc.exact_bruteforce(text, pat)
    assert len(want_exact) >= 4096 // 64
    pm = text[5000:5040].copy()                       # two words: the cut-off kernel, ~37 k hits
    _, we, wr = orc.search_mt(3, text, pm, 10)
    assert len(we) >= 16384
    p = sp.Pattern(ctx, sp.ALG_MYERS, pm, 10)
    for rep in range(4):                              # pass 0 arms the path, passes 1.. use it
        got = sp.search(ctx, store, p, hits)
        assert np.array_equal(got["end"], we) and np.array_equal(got["errors"], wr), rep
    # two lists in flight
    hits_b = sp.Hits(ctx)
    sp.search(ctx, store, p, hits_b)
    sp.search_launch(ctx, store, p, hits)
    sp.search_launch(ctx, store, p, hits_b)
    a = sp.search_finish(ctx, hits).copy()
    b = sp.search_finish(ctx, hits_b).copy()
    assert np.array_equal(a["end"], we) and np.array_equal(b["end"], we) and np.array_equal(b["errors"], wr)
    # the same Hits object, now a short list (exact) and a long one again, then a pattern set through the filter
    pe = sp.Pattern(ctx, sp.ALG_SHIFTOR, pat)
    for rep in range(2):
        got = sp.search(ctx, store, pe, hits)
        assert np.array_equal(got["end"], want_exact)
    got = sp.search(ctx, store, p, hits)
    assert np.array_equal(got["end"], we) and np.array_equal(got["errors"], wr)
    n4 = 2_000_000
    t4 = rand_text(rng, n4)
    needles = []
    for i in range(40):
        q = t4[5000 * i + 77:5000 * i + 77 + 100].copy()
        needles.append(q)
        for c in range(500):
            s0 = int(rng.integers(0, n4 - 200))
            t4[s0:s0 + 100] = q
    st4 = sp.Store.from_host(ctx, t4)
    pf = sp.Pattern(ctx, sp.ALG_MYERS_FILTERED, needles, 3)
    pd = sp.Pattern(ctx, sp.ALG_MYERS, needles, 3)
    direct = sp.search(ctx, st4, pd).copy()
    assert len(direct) >= 16384                       # long enough for the launch to order the list itself
    for rep in range(3):
        got = sp.search(ctx, st4, pf, hits)
        assert len(got) == len(direct), (rep, len(got), len(direct))
        for f in ("pattern_id", "end", "errors"):
            assert np.array_equal(got[f], direct[f]), (rep, f)
    for x in (p, pe, pf, pd, hits, hits_b, store, st4):
        x.close()


@pytest.mark.parametrize("sigma,bits", [(5, 3), (15, 4)])
@pytest.mark.parametrize("rl", [0, 96])
def test_wide_journaled_store(ctx, sigma, bits, rl):
    """a journal over a 4-bit store (spm::dna5 / spm::dna15): both plane pairs take the same events, the pool gives
    bits 1..0 and 3..2; derived text = the oracle's journal apply, searches over it = the oracle's."""
    rng = np.random.default_rng(31 * sigma + rl)
    n = 300_000
    ref = rand_text(rng, n, sigma)
    # events: SNPs, insertions, deletions, sorted and non-overlapping
    pos = np.sort(rng.choice(np.arange(10, n - 50, 40), size=3000, replace=False))
    ev = np.zeros(len(pos), dtype=sp.EVENT_DTYPE)
    pool = []
    for i, p0 in enumerate(pos):
        kind = int(rng.integers(0, 3))
        dl = 1 if kind == 0 else (0 if kind == 1 else int(rng.integers(1, 9)))
        il = 1 if kind == 0 else (int(rng.integers(1, 9)) if kind == 1 else 0)
        ev[i] = (int(p0), dl, il, len(pool))
        pool.extend(int(x) for x in rng.integers(0, sigma, size=il))
    pool = np.array(pool, dtype=np.uint8)
    derived = orc.journal_apply(ref, ev.astype(orc.EVENT_DTYPE), pool)
    base = sp.Store.from_host_wide(ctx, ref, bits, rl)
    js = base.journaled(ev, pool, rl)
    assert js.symbol_bits == 4 and len(js) == len(derived)
    assert np.array_equal(js.read(), derived)
    s0 = int(pos[len(pos) // 2]) - 20
    pat = derived[s0:s0 + 64].copy()
    for alg in (sp.ALG_HORSPOOL, sp.ALG_SHIFTOR):
        p = sp.Pattern(ctx, alg, pat, 0, bits=bits)
        got = sp.search(ctx, js, p)
        assert np.array_equal(got["end"], orc.exact_bruteforce(derived, pat))
        p.close()
    pm = _mutate(rng, derived[s0 + 100:s0 + 200].copy(), 2, sigma)
    p = sp.Pattern(ctx, sp.ALG_MYERS, pm, 4, bits=bits)
    got = sp.search(ctx, js, p)
    we, wr = orc.sellers(derived, pm, 4)
    assert len(we) >= 1
    assert np.array_equal(got["end"], we) and np.array_equal(got["errors"], wr)
    p.close(); js.close(); base.close()
