"""Device-resident resumable search (spmb_stream_*): chunked feeds of every awkward size
equal the unchunked oracle, for single needles and needle sets, through the short-chunk path and through the
upload queue; capture / restore round trips; the captured `errors` equals the Sellers score at the current column;
the asynchronous submit / collect pair delivers the same lists."""
import numpy as np
import pytest

import libspm_b200 as sp
import oracle as orc
from tests.util import GOLD_HAYSTACK, GOLD_MYERS_K1_END, GOLD_NEEDLE, mutate, rand_text

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = sp.Context(0)
    yield c
    c.close()


def sellers_last_column(text, pat):
    """min over start positions of ed(pat, text[s..]) at the last text symbol (m when the text is empty)."""
    m = len(pat)
    col = np.arange(m + 1)
    for c in text:
        new = np.empty_like(col)
        new[0] = 0
        for i in range(1, m + 1):
            new[i] = min(col[i - 1] + (0 if pat[i - 1] == c else 1), col[i] + 1, new[i - 1] + 1)
        col = new
    return int(col[m])


def feed_all(s, text, sizes):
    """feed `text` in pieces whose sizes cycle through `sizes`; returns (ends, errors, pattern_ids) in global coordinates."""
    ge, gr, gp = [], [], []
    off, i = 0, 0
    while off < len(text):
        ln = min(sizes[i % len(sizes)], len(text) - off)
        h = s.feed(text[off:off + ln])
        ge += [int(e) + off for e in h["end"]]
        gr += [int(e) for e in h["errors"]]
        gp += [int(e) for e in h["pattern_id"]]
        off += ln
        i += 1
    return ge, gr, gp


@pytest.mark.parametrize("m,k", [(5, 1), (31, 2), (64, 3), (100, 4), (150, 2)])
def test_chunk_size_sweep_myers(ctx, m, k):
    """SURVEY.md 7 hard part 5: chunk sizes {1, m-1, m, m+k, 13} and mixtures, against the unchunked oracle."""
    rng = np.random.default_rng(m)
    n = 3000 if m < 64 else 6000
    text = rand_text(rng, n)
    pat = rand_text(rng, m)
    for s0 in (0, 700, 1500, n - m - 3):
        mp = mutate(rng, pat, int(rng.integers(0, k + 1)))
        text[s0:s0 + len(mp)] = mp
    we, wr = orc.sellers(text, pat, k) if n <= 20000 else orc.myers(text, pat, k)
    assert len(we) > 0
    p = sp.Pattern(ctx, sp.ALG_MYERS, pat, k)
    for sizes in ([1], [m - 1], [m], [m + k], [13], [1, 500, 2, m + k, 31, 32, 33], [n]):
        if sizes == [1] and n > 3000:
            continue
        s = sp.Stream(ctx, p)
        ge, gr, _ = feed_all(s, text, sizes)
        assert ge == list(map(int, we)), (sizes, len(ge), len(we))
        assert gr == list(map(int, wr)), sizes
        assert s.offset == n
        s.close()
    p.close()


@pytest.mark.parametrize("alg", [sp.ALG_HORSPOOL, sp.ALG_SHIFTOR])
def test_chunk_size_sweep_exact(ctx, alg):
    rng = np.random.default_rng(9)
    n = 5000
    text = rand_text(rng, n, 2)                     # binary text: many overlapping occurrences
    for m in (1, 7, 33, 70):
        pat = text[100:100 + m].copy()
        want = [int(e) for e in orc.exact_bruteforce(text, pat)]
        p = sp.Pattern(ctx, alg, pat)
        for sizes in ([1], [max(1, m - 1)], [m], [13], [64, 1, 1, 31, 997]):
            if sizes == [1] and m > 7:
                continue
            s = sp.Stream(ctx, p)
            ge, _, _ = feed_all(s, text, sizes)
            assert ge == want, (m, sizes, len(ge), len(want))
            s.close()
        p.close()


def test_stream_pattern_sets(ctx):
    rng = np.random.default_rng(21)
    n = 40_000
    text = rand_text(rng, n)
    pats = []
    for i in range(6):
        s0 = int(rng.integers(0, n - 200))
        pats.append(mutate(rng, text[s0:s0 + 100].copy(), 2)[:128])
        if len(pats[-1]) < 97:
            pats[-1] = np.concatenate([pats[-1], text[s0:s0 + 97 - len(pats[-1])]])
    p = sp.Pattern(ctx, sp.ALG_MYERS, pats, 4)
    s = sp.Stream(ctx, p)
    got = {i: ([], []) for i in range(len(pats))}
    off = 0
    for ln in [1000, 77, 104, 5000, 1, 1, 33_000, 10_000]:
        ln = min(ln, n - off)
        if ln <= 0:
            break
        h = s.feed(text[off:off + ln])
        for x in h:
            got[int(x["pattern_id"])][0].append(int(x["end"]) + off)
            got[int(x["pattern_id"])][1].append(int(x["errors"]))
        off += ln
    assert off == n
    for i, q in enumerate(pats):
        we, wr = orc.myers(text, q, 4)
        assert got[i][0] == list(map(int, we)) and got[i][1] == list(map(int, wr)), i
    s.close(); p.close()


def test_capture_restore_and_errors(ctx):
    """the reference's own test shape (myers_matcher_restorable_test.cpp:55-74) plus the captured `errors`: the
    score at the current column, against a Sellers DP of the whole prefix."""
    p = sp.Pattern(ctx, sp.ALG_MYERS, GOLD_NEEDLE, 1)
    s = sp.Stream(ctx, p)
    state = s.capture()
    assert s.errors() == len(GOLD_NEEDLE)           # nothing read yet: D[m][-1] = m
    got = []
    for off in range(0, len(GOLD_HAYSTACK), 13):
        s2 = sp.Stream(ctx, p)                       # a different matcher object picks the state up
        s2.restore(state)
        h = s2.feed(GOLD_HAYSTACK[off:off + 13])
        got += [int(e) + off for e in h["end"]]
        state = s2.capture()
        seen = GOLD_HAYSTACK[:off + 13]
        assert s2.errors() == sellers_last_column(seen, GOLD_NEEDLE), off
        n_pat = np.frombuffer(state[12:16], dtype=np.uint32)[0]
        assert n_pat == 1 and np.frombuffer(state[16:20], dtype=np.uint32)[0] == s2.errors()
        s2.close()
    assert got == GOLD_MYERS_K1_END
    s.close(); p.close()
    # longer needle, errors far above k
    rng = np.random.default_rng(4)
    text = rand_text(rng, 700)
    pat = rand_text(rng, 40)
    p = sp.Pattern(ctx, sp.ALG_MYERS, pat, 3)
    s = sp.Stream(ctx, p)
    off = 0
    for ln in (1, 39, 40, 43, 200, 377):
        s.feed(text[off:off + ln])
        off += ln
        assert s.errors() == sellers_last_column(text[:off], pat), off
    s.close(); p.close()


def test_submit_collect_three_chunks_in_flight(ctx):
    rng = np.random.default_rng(77)
    n = 2_000_000
    text = rand_text(rng, n)
    pat = text[1000:1100].copy()
    for s0 in rng.integers(0, n - 300, size=40):
        mp = mutate(rng, pat, int(rng.integers(0, 6)))
        text[int(s0):int(s0) + len(mp)] = mp
    _, we, wr = orc.search_mt(3, text, pat, 4)
    p = sp.Pattern(ctx, sp.ALG_MYERS, pat, 4)
    s = sp.Stream(ctx, p)
    chunk = 300_001
    offs = list(range(0, n, chunk))
    ge, gr = [], []
    inflight = []
    for o in offs:
        if len(inflight) == 3:
            o0 = inflight.pop(0)
            h = s.collect()
            ge += [int(e) + o0 for e in h["end"]]; gr += [int(e) for e in h["errors"]]
        s.submit(text[o:o + chunk])
        inflight.append(o)
    assert len(inflight) == 3
    with pytest.raises(sp.SpmbError):
        s.submit(text[:10])                          # a fourth chunk: collect first
    while inflight:
        o0 = inflight.pop(0)
        h = s.collect()
        ge += [int(e) + o0 for e in h["end"]]; gr += [int(e) for e in h["errors"]]
    assert ge == list(map(int, we)) and gr == list(map(int, wr))
    with pytest.raises(sp.SpmbError):
        s.collect()
    s.close(); p.close()


@pytest.mark.parametrize("threads", [0, 3])
def test_long_chunks_through_the_upload_queue(monkeypatch, threads):
    """chunks above SPMB_UPLOAD_MIN take both movers of the upload queue, spliced behind a tail that ends anywhere
    inside a 32-symbol word; the exact searcher and Myers give the unchunked lists."""
    monkeypatch.setenv("SPMB_UPLOAD_MIN", "100000")
    monkeypatch.setenv("SPMB_UPLOAD_CHUNK", "65536")
    monkeypatch.setenv("SPMB_HOST_THREADS", str(threads))
    c = sp.Context(0)
    try:
        rng = np.random.default_rng(5 + threads)
        n = 1_700_000
        text = rand_text(rng, n)
        pat = text[500_000:500_064].copy()
        for s0 in (3, 199_990, 200_000, 200_030, 899_999, n - 64):
            text[s0:s0 + 64] = pat
        want = [int(e) for e in orc.exact_bruteforce(text, pat)]
        _, we, wr = orc.search_mt(3, text, pat, 2)
        for alg, k, ref in ((sp.ALG_SHIFTOR, 0, want), (sp.ALG_MYERS, 2, list(map(int, we)))):
            p = sp.Pattern(c, alg, pat, k)
            s = sp.Stream(c, p)
            ge, _, _ = feed_all(s, text, [200_000, 37, 300_003, 150_001, 700_000])
            assert ge == ref, (alg, len(ge), len(ref))
            s.close(); p.close()
        assert c.upload_last_stats()["host_threads"] == threads
    finally:
        c.close()


def test_prefix_stream_still_anchored_at_the_stream_start(ctx):
    rng = np.random.default_rng(2)
    text = rand_text(rng, 400)
    pat = mutate(rng, text[:50].copy(), 2)
    we, wr = orc.myers_prefix(text, pat, 3)
    p = sp.Pattern(ctx, sp.ALG_MYERS_PREFIX, pat, 3)
    s = sp.Stream(ctx, p)
    ge = []
    off = 0
    for ln in (10, 30, 5, 100, 255):
        h = s.feed(text[off:off + ln])
        ge += [int(e) + off for e in h["end"]]
        off += ln
    assert ge == list(map(int, we))
    s.close(); p.close()
