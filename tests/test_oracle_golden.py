"""Pins the CPU oracle: (1) the reference's own golden vectors, (2) the algorithmic
restatements (Horspool / Shift-Or / Myers incl. multi-word + Ukkonen cut-off, resumable state,
multi-threaded chunking) against the definitional oracles on seeded random inputs."""
import numpy as np
import pytest

import oracle as orc
from tests.util import (GOLD_EXACT_BEGIN, GOLD_HAYSTACK, GOLD_MYERS_K1_END, GOLD_NEEDLE, dna4,
                        mutate, rand_text)


def test_golden_exact_all_algorithms():
    m = len(GOLD_NEEDLE)
    for fn in (orc.exact_bruteforce, orc.horspool, orc.shiftor):
        ends = fn(GOLD_HAYSTACK, GOLD_NEEDLE)
        assert [int(e) - m for e in ends] == GOLD_EXACT_BEGIN, fn.__name__


def test_golden_myers_k1():
    for fn in (orc.sellers, orc.myers):
        ends, errs = fn(GOLD_HAYSTACK, GOLD_NEEDLE, 1)
        assert [int(e) for e in ends] == GOLD_MYERS_K1_END
        assert [int(e) for e in errs] == [1, 0, 1] * 3   # re-derived error values


def test_golden_myers_restorable_chunks_of_13():
    # myers_matcher_restorable_test.cpp:55-74: capture -> per chunk restore/run/capture
    st = orc.MyersState(GOLD_NEEDLE, 1)
    state = st.capture()
    got = []
    for off in range(0, len(GOLD_HAYSTACK), 13):
        chunk = GOLD_HAYSTACK[off:off + 13]
        st.restore(state)
        ends, _ = st.run(chunk, base=0)
        got += [int(e) + off for e in ends]
        state = st.capture()
    assert got == GOLD_MYERS_K1_END


def test_shiftor_restorable_chunks():
    st = orc.ShiftOrState(GOLD_NEEDLE)
    state = st.capture()
    got = []
    for off in range(0, len(GOLD_HAYSTACK), 7):
        st.restore(state)
        got += [int(e) + off - len(GOLD_NEEDLE) for e in st.run(GOLD_HAYSTACK[off:off + 7])]
        state = st.capture()
    assert got == GOLD_EXACT_BEGIN


def test_overlapping_exact_hits():
    # period-4 run, pattern ACGTx8 inside ACGTx16 -> 9 overlapping hits
    text = np.concatenate([dna4("TTGCA"), np.tile(dna4("ACGT"), 16), dna4("GGAT")])
    pat = np.tile(dna4("ACGT"), 8)
    want = orc.exact_bruteforce(text, pat)
    assert len(want) == 9
    assert np.array_equal(orc.horspool(text, pat), want)
    assert np.array_equal(orc.shiftor(text, pat), want)


@pytest.mark.parametrize("m", [1, 2, 5, 31, 32, 33, 64, 65, 100, 150])
def test_exact_random_vs_bruteforce(m):
    rng = np.random.default_rng(1000 + m)
    for sigma in (2, 4):
        text = rand_text(rng, 20000, sigma)
        start = int(rng.integers(0, len(text) - m))
        pat = text[start:start + m].copy()
        # plant a few more copies, some overlapping for low sigma
        for _ in range(5):
            p = int(rng.integers(0, len(text) - m))
            text[p:p + m] = pat
        want = orc.exact_bruteforce(text, pat)
        assert len(want) >= 1
        assert np.array_equal(orc.horspool(text, pat), want)
        assert np.array_equal(orc.shiftor(text, pat), want)


@pytest.mark.parametrize("m,k", [(5, 1), (8, 0), (20, 3), (63, 4), (64, 2), (65, 2), (100, 4),
                                 (128, 6), (129, 3), (150, 2), (200, 40), (300, 10), (100, 99)])
def test_myers_random_vs_sellers(m, k):
    rng = np.random.default_rng(77 * m + k)
    for sigma in (2, 4):
        text = rand_text(rng, 6000, sigma)
        pat = rand_text(rng, m, sigma)
        for e in range(0, min(k + 3, m)):
            p = int(rng.integers(0, len(text) - 2 * m))
            mp = mutate(rng, pat, e)
            text[p:p + len(mp)] = mp
        we, wr = orc.sellers(text, pat, k)
        ge, gr = orc.myers(text, pat, k)
        assert np.array_equal(ge, we)
        assert np.array_equal(gr, wr)
        assert len(we) > 0


def test_myers_state_carry_equals_unchunked():
    rng = np.random.default_rng(5)
    text = rand_text(rng, 5000)
    pat = text[1234:1234 + 130].copy()
    mp = mutate(rng, pat, 3)
    text[3000:3000 + len(mp)] = mp
    we, wr = orc.myers(text, pat, 5)
    for chunk in (1, 13, 129, 130, 135, 1000):
        st = orc.MyersState(pat, 5)
        ge, gr = [], []
        for off in range(0, len(text), chunk):
            e, r = st.run(text[off:off + chunk], base=off)
            ge += list(e)
            gr += list(r)
        assert np.array_equal(np.array(ge, dtype=np.uint64), we), chunk
        assert np.array_equal(np.array(gr, dtype=np.uint32), wr), chunk


@pytest.mark.parametrize("alg,m,k", [(1, 32, 0), (2, 64, 0), (3, 100, 4), (3, 20, 2)])
def test_multithreaded_chunking_is_exact(alg, m, k):
    rng = np.random.default_rng(alg * 100 + m)
    text = rand_text(rng, 200000)
    pat = text[777:777 + m].copy()
    for i in range(20):
        p = int(rng.integers(0, len(text) - 2 * m))
        mp = mutate(rng, pat, int(rng.integers(0, k + 2))) if alg == 3 else pat
        text[p:p + len(mp)] = mp
    if alg == 3:
        we, wr = orc.sellers(text[:60000], pat, k)   # DP is slow: definitional check on a prefix
        ge, gr = orc.myers(text[:60000], pat, k)
        assert np.array_equal(ge, we) and np.array_equal(gr, wr)
        we, wr = orc.myers(text, pat, k)
    else:
        we, wr = orc.exact_bruteforce(text, pat), None
    for threads in (1, 3, 8):
        cnt, ge, gr = orc.search_mt(alg, text, pat, k, threads=threads)
        assert cnt == len(we)
        assert np.array_equal(ge, we)
        if wr is not None:
            assert np.array_equal(gr, wr)
        cnt2, _, _ = orc.search_mt(alg, text, pat, k, threads=threads, store=False)
        assert cnt2 == cnt


def test_journal_apply():
    ref = dna4("ACGTACGTAC")
    ev = np.array([(2, 1, 1, 0), (4, 0, 3, 1), (7, 2, 0, 0)], dtype=orc.EVENT_DTYPE)
    pool = dna4("TGGG")
    out = orc.journal_apply(ref, ev, pool)
    #  AC [G->T] T [+GGG] ACG [-TA] C
    assert "".join("ACGT"[x] for x in out) == "ACTTGGGACGC"


# ---- pigeonhole seed matcher -- the reference's own vectors ----
# /root/reference/test/api/libspm/matcher/pigeonhole_matcher_test.cpp:25-33
GOLD_NEEDLE2 = dna4("TGACTAGCAC")
GOLD_MULTI_POSITIONS = [3, 8, 9, 14, 19, 20, 25, 30, 31, 36]
GOLD_NEEDLE_POSITIONS = [(1, 0, 5), (1, 5, 5), (0, 0, 5), (1, 0, 5), (1, 5, 5),
                         (0, 0, 5), (1, 0, 5), (1, 5, 5), (0, 0, 5), (1, 0, 5)]


def test_pigeonhole_golden_single_needle():
    # pigeonhole_matcher_test.cpp:50-53 (window_size == needle size at error rate 0), :55-64
    assert orc.pigeonhole_q([len(GOLD_NEEDLE)], 0.0) == 5
    hits = orc.pigeonhole(GOLD_HAYSTACK, [GOLD_NEEDLE], 0.0)
    assert hits["begin"].tolist() == GOLD_EXACT_BEGIN


def test_pigeonhole_golden_multi_needle():
    # pigeonhole_matcher_test.cpp:66-86: begin positions and {index, offset, count} per callback
    q = orc.pigeonhole_q([5, 10], 0.0)
    assert q == 5
    hits = orc.pigeonhole(GOLD_HAYSTACK, [GOLD_NEEDLE, GOLD_NEEDLE2], 0.0)
    assert hits["begin"].tolist() == GOLD_MULTI_POSITIONS
    assert [(int(h["needle"]), int(h["offset"]), q) for h in hits] == GOLD_NEEDLE_POSITIONS


def test_pigeonhole_q_rule():
    # delta = len / (floor(rate*len)+1); deltas < 3 are ignored; q = the smallest
    assert orc.pigeonhole_q([100], 0.03) == 25
    assert orc.pigeonhole_q([100, 60], 0.05) == 15
    assert orc.pigeonhole_q([100, 4], 0.3) == 3      # 100/31 = 3; 4/2 = 2 is ignored
    assert orc.pigeonhole_q([2], 0.0) == 0


def test_myers_prefix_is_global_prefix_distance():
    # restorable_myers_prefix_matcher (myers_prefix_matcher_restorable.hpp:47-61): needle vs prefixes
    rng = np.random.default_rng(5)
    for _ in range(50):
        m = int(rng.integers(1, 40))
        k = int(rng.integers(0, 6))
        pat = rand_text(rng, m)
        text = np.concatenate([mutate(rng, pat, int(rng.integers(0, k + 2))), rand_text(rng, 30)])
        ends, errs = orc.myers_prefix(text, pat, k)
        # brute force: plain edit distance of pat against text[:j]
        for j in range(1, min(len(text), m + k + 1) + 1):
            d = list(range(m + 1))
            for c in text[:j]:
                nd = [d[0] + 1]
                for i in range(1, m + 1):
                    nd.append(min(d[i - 1] + (pat[i - 1] != c), d[i] + 1, nd[i - 1] + 1))
                d = nd
            if d[m] <= k:
                assert j in ends.tolist() and errs[ends.tolist().index(j)] == d[m]
            else:
                assert j not in ends.tolist()
