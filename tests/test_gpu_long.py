"""GPU parity tests for Myers needles the one-lane kernels cannot take in one launch: needles longer than 256 symbols
(256-symbol prefix filter + lane-chained verification, libspm_b200/csrc/myers_chain.cuh) and sets that mix word
counts.  The reference takes any needle (libspm/libspm/matcher/myers_matcher.hpp:37-43); every list is compared with
the oracle's blocked Myers (oracle/spm_oracle.c) on the same seeded input, ends and error counts."""
import numpy as np
import pytest

import libspm_b200 as sp
import oracle as orc
from tests.util import mutate, rand_text

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = sp.Context(0)
    yield c
    c.close()


def _same(got, want, what):
    got, want = np.asarray(got), np.asarray(want)
    if got.shape == want.shape and np.array_equal(got, want):
        return
    gs, ws = set(got.tolist()), set(want.tolist())
    raise AssertionError(f"{what}: got {len(got)} want {len(want)}; missing {sorted(ws - gs)[:8]} extra {sorted(gs - ws)[:8]}")


def planted(rng, n, pat, k, copies):
    """random text with mutated copies of pat (0..k+2 edits each, so some fall outside k)."""
    text = rand_text(rng, n)
    m = len(pat)
    for s in rng.integers(0, n - 2 * m - 64, size=copies):
        mp = mutate(rng, pat, int(rng.integers(0, k + 3)))
        text[int(s):int(s) + len(mp)] = mp
    return text


@pytest.mark.parametrize("m", [257, 300, 512, 1000, 1025, 2500])
@pytest.mark.parametrize("k", [0, 5, 25])
def test_long_needles_equal_the_oracle(ctx, m, k):
    rng = np.random.default_rng(m * 131 + k)
    n = 400_000
    pat = rand_text(rng, m)
    text = planted(rng, n, pat, k, 25)
    # an occurrence at the very start and one cut off by the end of the text
    text[:m] = pat
    text[n - m + 7:] = pat[:m - 7]
    store = sp.Store.from_host(ctx, text)
    we, wr = orc.myers(text, pat, k)
    assert len(we) > (20 if k else 1)
    p = sp.Pattern(ctx, sp.ALG_MYERS, pat, k)
    assert p.window_size == m + k
    got = sp.search(ctx, store, p)
    _same(got["end"], we, f"ends m={m} k={k}")
    _same(got["errors"], wr, f"errors m={m} k={k}")
    assert np.all(got["pattern_id"] == 0)
    p.close(); store.close()


@pytest.mark.parametrize("rl", [32, 64, 1024])
def test_long_needle_across_row_and_tile_seams(ctx, rl):
    """short rows: the verification window of a candidate runs over many rows and super-tiles."""
    rng = np.random.default_rng(rl)
    m, k, n = 700, 6, 120_000
    pat = rand_text(rng, m)
    text = planted(rng, n, pat, k, 12)
    store = sp.Store.from_host(ctx, text, row_len=rl)
    we, wr = orc.myers(text, pat, k)
    p = sp.Pattern(ctx, sp.ALG_MYERS, pat, k)
    got = sp.search(ctx, store, p)
    _same(got["end"], we, f"ends rl={rl}")
    _same(got["errors"], wr, f"errors rl={rl}")
    p.close(); store.close()


def test_long_needle_hit_dense_low_complexity(ctx):
    """poly-A text and a needle one substitution off poly-A: every column is a candidate AND a hit."""
    n, m, k = 60_000, 400, 3
    text = np.zeros(n, dtype=np.uint8)
    text[20_000:30_000] = rand_text(np.random.default_rng(3), 10_000)
    pat = np.zeros(m, dtype=np.uint8)
    pat[m // 2] = 2
    store = sp.Store.from_host(ctx, text)
    we, wr = orc.myers(text, pat, k)
    assert len(we) > n // 2
    p = sp.Pattern(ctx, sp.ALG_MYERS, pat, k)
    got = sp.search(ctx, store, p)
    _same(got["end"], we, "dense ends")
    _same(got["errors"], wr, "dense errors")
    p.close(); store.close()


def test_mixed_word_counts_and_long_needles_in_one_set(ctx):
    """one set with needles of 1..8 words, two long ones and an empty one: hits carry the caller's pattern ids."""
    rng = np.random.default_rng(77)
    n, k = 500_000, 3
    lens = [20, 33, 64, 0, 100, 150, 31, 256, 257, 200, 900, 65]
    pats = [rand_text(rng, m) for m in lens]
    text = rand_text(rng, n)
    for i, pat in enumerate(pats):
        if len(pat) == 0:
            continue
        for s in rng.integers(0, n - 2000, size=10):
            mp = mutate(rng, pat, int(rng.integers(0, k + 2)))
            text[int(s):int(s) + len(mp)] = mp
    store = sp.Store.from_host(ctx, text)
    p = sp.Pattern(ctx, sp.ALG_MYERS, pats, k)
    assert p.window_size == max(lens) + k
    got = sp.search(ctx, store, p)
    for alg in (sp.ALG_MYERS_FILTERED,):     # same list through the other entry point (a long needle disables the filter)
        p2 = sp.Pattern(ctx, alg, pats, k)
        got2 = sp.search(ctx, store, p2)
        assert np.array_equal(got2, got)
        p2.close()
    assert np.all(np.diff(got["pattern_id"].astype(np.int64)) >= 0), "sorted by pattern first"
    for i, pat in enumerate(pats):
        sel = got[got["pattern_id"] == i]
        if len(pat) == 0:
            assert len(sel) == 0
            continue
        we, wr = orc.myers(text, pat, k)
        assert len(we) >= 5
        _same(sel["end"], we, f"ends needle {i} (m={len(pat)})")
        _same(sel["errors"], wr, f"errors needle {i} (m={len(pat)})")
    p.close(); store.close()


def test_long_needle_shards_partition_the_hit_list(ctx):
    rng = np.random.default_rng(5)
    m, k, n = 600, 8, 3_000_000
    pat = rand_text(rng, m)
    text = planted(rng, n, pat, k, 200)
    store = sp.Store.from_host(ctx, text, row_len=256)   # many super-tiles: cuts fall inside occurrences
    we, wr = orc.myers(text, pat, k)
    p = sp.Pattern(ctx, sp.ALG_MYERS, pat, k)
    parts = []
    for s in range(7):
        ctx.set_shard(s, 7)
        parts.append(sp.search(ctx, store, p))
    ctx.set_shard(0, 1)
    merged = np.concatenate(parts)
    assert len(merged) == len(we), "shards must not share or lose a hit"
    merged = merged[np.argsort(merged["end"], kind="stable")]
    _same(merged["end"], we, "sharded ends")
    _same(merged["errors"], wr, "sharded errors")
    p.close(); store.close()


def test_long_needle_stream_chunks_and_errors(ctx):
    """restorable matcher semantics with a long needle: chunked feeds give the one-shot list; errors() is the score
    at the current column (Sellers' last-row value)."""
    rng = np.random.default_rng(9)
    m, k, n = 300, 4, 40_000
    pat = rand_text(rng, m)
    text = planted(rng, n, pat, k, 10)
    we, wr = orc.myers(text, pat, k)
    p = sp.Pattern(ctx, sp.ALG_MYERS, pat, k)
    for size in (m - 1, m + k, 4096, 13_000):
        st = sp.Stream(ctx, p)
        ends, errs = [], []
        off = 0
        while off < n:
            h = st.feed(text[off:off + size])
            ends.extend((h["end"] + off).tolist()); errs.extend(h["errors"].tolist())
            off += size
        _same(np.array(ends, dtype=np.uint64), we, f"stream ends chunk={size}")
        _same(np.array(errs, dtype=np.uint32), wr, f"stream errors chunk={size}")
        # score at the last column by definition (Sellers, free start)
        tail = text[-2 * m:]
        d = np.arange(m + 1)
        for c in tail:
            nd = np.empty_like(d)
            nd[0] = 0
            sub = d[:-1] + (pat != c)
            nd[1:] = np.minimum(sub, d[1:] + 1)
            for i in range(1, m + 1):
                if nd[i - 1] + 1 < nd[i]:
                    nd[i] = nd[i - 1] + 1
            d = nd
        assert st.errors(0) == int(d[m])
        st.close()
    p.close()


def test_jst_search_with_a_long_needle(ctx):
    """branch-context windows of length m+k-1 with m > 256: windows-mode prefix filter + chained verification against
    the per-haplotype oracle; also a set that mixes a long needle with short ones."""
    from libspm_b200 import synth
    rng = np.random.default_rng(21)
    n_ref, n_hap, n_var = 60_000, 24, 400
    ref = rand_text(rng, n_ref)
    var, pool, kinds, cov = synth.jst(ref, n_var, n_hap, 21, p_snp=0.6, p_ins=0.2, max_indel=8, af_max=0.7)
    ev = var.astype(orc.EVENT_DTYPE)
    store = sp.Store.from_host(ctx, ref)
    jst = sp.Jst(ctx, store, var, pool, cov, n_hap)
    pats = []
    for (m, k, src_h) in [(300, 3, 0), (420, 5, n_hap - 1), (700, 2, None), (1030, 8, 3)]:
        seq = ref if src_h is None else orc.jst_haplotype(ref, ev, pool, cov, src_h)
        s = int(rng.integers(0, len(seq) - m))
        pat = mutate(rng, seq[s:s + m].copy(), 1)
        pats.append(pat)
        p = sp.Pattern(ctx, sp.ALG_MYERS, pat, k)
        got = jst.search(p)
        want = orc.jst_search(sp.ALG_MYERS, ref, ev, pool, cov, n_hap, pat, k)
        assert len(want) >= 1
        g = np.stack([got["seq_id"].astype(np.int64), got["end"].astype(np.int64), got["errors"].astype(np.int64)], 1)
        w = np.stack([want["hap"].astype(np.int64), want["end"].astype(np.int64), want["err"].astype(np.int64)], 1)
        assert g.shape == w.shape and np.array_equal(g, w), (m, k, len(g), len(w))
        p.close()
    # mixed set: a long needle beside 2- and 5-word ones
    k = 3
    mixed = [pats[0], pats[0][:150].copy(), pats[1][:40].copy()]
    p = sp.Pattern(ctx, sp.ALG_MYERS, mixed, k)
    got = jst.search(p)
    for i, pat in enumerate(mixed):
        want = orc.jst_search(sp.ALG_MYERS, ref, ev, pool, cov, n_hap, pat, k)
        sel = got[got["pattern_id"] == i]
        g = np.stack([sel["seq_id"].astype(np.int64), sel["end"].astype(np.int64), sel["errors"].astype(np.int64)], 1)
        w = np.stack([want["hap"].astype(np.int64), want["end"].astype(np.int64), want["err"].astype(np.int64)], 1)
        assert g.shape == w.shape and np.array_equal(g, w), (i, len(g), len(w))
    p.close(); jst.close(); store.close()
