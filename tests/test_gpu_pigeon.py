"""GPU parity for SURVEY.md 8(f) ranks 2-3 (through the C ABI, bit-exact against the oracle):
pigeonhole q-gram seed matcher, the filtered Myers search built on it, anchored (prefix) Myers."""
import numpy as np
import pytest

import libspm_b200 as sp
import oracle as orc
from tests.test_oracle_golden import GOLD_MULTI_POSITIONS, GOLD_NEEDLE2, GOLD_NEEDLE_POSITIONS
from tests.util import GOLD_EXACT_BEGIN, GOLD_HAYSTACK, GOLD_NEEDLE, mutate, rand_text

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = sp.Context(0)
    yield c
    c.close()


def seed_tuples(h, q):
    return [(int(x["end"]) - q, int(x["pattern_id"]), int(x["reserved"])) for x in h]


def want_tuples(w):
    return [(int(x["begin"]), int(x["needle"]), int(x["offset"])) for x in w]


def test_pigeonhole_reference_golden_vectors(ctx):
    # pigeonhole_matcher_test.cpp:30-33,50-86
    st = sp.Store.from_host(ctx, GOLD_HAYSTACK)
    p = sp.Pigeonhole(ctx, GOLD_NEEDLE, 0.0)
    assert p.window_size == 5
    got = sp.search(ctx, st, p)
    assert [int(e) - 5 for e in got["end"]] == GOLD_EXACT_BEGIN
    p.close()
    p = sp.Pigeonhole(ctx, [GOLD_NEEDLE, GOLD_NEEDLE2], 0.0)
    assert p.window_size == 5
    got = sp.search(ctx, st, p)
    assert [int(e) - 5 for e in got["end"]] == GOLD_MULTI_POSITIONS
    assert [(int(x["pattern_id"]), int(x["reserved"]), 5) for x in got] == GOLD_NEEDLE_POSITIONS
    p.close(); st.close()


@pytest.mark.parametrize("n,rl,lens,rate", [
    (5000, 32, [12], 0.0), (5000, 64, [12, 30, 7], 0.1), (100_000, 0, [40, 55, 64, 33], 0.05),
    (100_000, 1056, [100] * 8, 0.03), (70_000, 96, [70, 200], 0.0), (300_000, 0, [17] * 50, 0.0),
])
def test_pigeonhole_vs_oracle(ctx, n, rl, lens, rate):
    rng = np.random.default_rng(n + len(lens))
    text = rand_text(rng, n)
    needles = []
    for L in lens:
        s = int(rng.integers(0, n - L))
        needles.append(text[s:s + L].copy())
    # plant pieces of the needles (also across row seams) and a shared q-gram in two needles
    q = orc.pigeonhole_q(lens, rate)
    if len(needles) > 1 and len(needles[1]) >= q and len(needles[0]) >= q:
        needles[1][:q] = needles[0][:q]
    for _ in range(30):
        nd = needles[int(rng.integers(0, len(needles)))]
        s = int(rng.integers(0, n - len(nd)))
        text[s:s + len(nd)] = nd
    text[n - q:] = needles[0][:q]            # a seed that ends exactly at the text end
    st = sp.Store.from_host(ctx, text, rl)
    p = sp.Pigeonhole(ctx, needles, rate)
    assert p.window_size == q
    got = sp.search(ctx, st, p)
    want = orc.pigeonhole(text, needles, rate)
    assert seed_tuples(got, q) == want_tuples(want), (p.filter_info, len(got), len(want))
    p.close(); st.close()


def test_pigeonhole_dense_hits_regrow_buffers(ctx):
    # poly-A: every position is a seed hit of every poly-A q-gram (candidate and hit buffers overflow)
    text = np.zeros(200_000, dtype=np.uint8)
    needles = [np.zeros(24, dtype=np.uint8)]
    st = sp.Store.from_host(ctx, text)
    p = sp.Pigeonhole(ctx, needles, 0.05)      # q = 12, seeds at 0 and 12
    got = sp.search(ctx, st, p, sp.Hits(ctx, 16))
    assert len(got) == 2 * (200_000 - 12 + 1)
    assert got["end"][0] == 12 and got["end"][-1] == 200_000
    p.close(); st.close()


def check_filtered(ctx, st, text, pats, k, tag=""):
    pf = sp.Pattern(ctx, sp.ALG_MYERS_FILTERED, pats, k)
    info = pf.filter_info
    got = sp.search(ctx, st, pf)
    pid, end, err = [], [], []
    for i, pat in enumerate(pats):
        if len(text) <= 20000:
            we, wr = orc.sellers(text, pat, k)
        else:
            _, we, wr = orc.search_mt(3, text, pat, k)
        pid += [i] * len(we); end += list(we); err += list(wr)
    assert got["pattern_id"].tolist() == pid, (tag, info, len(got), len(pid))
    assert got["end"].tolist() == [int(e) for e in end], (tag, info)
    assert got["errors"].tolist() == [int(e) for e in err], (tag, info)
    pf.close()
    return info


@pytest.mark.parametrize("n,rl,m,k,n_pat", [
    (20000, 32, 60, 2, 3), (20000, 64, 100, 4, 2), (150_000, 0, 100, 3, 6), (150_000, 1056, 150, 2, 4),
    (150_000, 96, 64, 1, 5), (400_000, 0, 250, 5, 2), (50_000, 0, 40, 1, 8),
])
def test_filtered_myers_equals_direct(ctx, n, rl, m, k, n_pat):
    rng = np.random.default_rng(n + m * 7 + k)
    text = rand_text(rng, n)
    pats = []
    for i in range(n_pat):
        s = int(rng.integers(0, n - m))
        pats.append(text[s:s + m].copy())
    for _ in range(60):      # mutated copies: 0..k+2 edits (hits with every error count, and decoys)
        pat = pats[int(rng.integers(0, n_pat))]
        mp = mutate(rng, pat, int(rng.integers(0, k + 3)))
        s = int(rng.integers(0, n - len(mp)))
        text[s:s + len(mp)] = mp
    text[:m // 2] = pats[0][m - m // 2:]              # a clipped occurrence at the text start
    text[n - m + 3:] = pats[0][:m - 3]                # and one cut off by the text end
    st = sp.Store.from_host(ctx, text, rl)
    info = check_filtered(ctx, st, text, pats, k, f"n={n} rl={rl} m={m} k={k}")
    assert info["filtered"] and info["q"] == m // (k + 1) - info["step"] + 1 and info["step"] in (1, 2, 4)
    st.close()


@pytest.mark.parametrize("step", [1, 2, 4])
def test_filtered_myers_every_sampling_step(ctx, step, monkeypatch):
    """the scan's sampling step is a pure performance knob: every setting gives the direct kernel's hits"""
    monkeypatch.setenv("SPMB_FILTER_STEP", str(step))
    rng = np.random.default_rng(41 + step)
    n, m, k = 90_000, 100, 3
    text = rand_text(rng, n)
    pats = [text[s:s + m].copy() for s in (100, 5000, 33333)]
    for i in range(90):      # occurrences at every alignment mod 4, with 0..k+1 edits
        pat = pats[i % 3]
        mp = mutate(rng, pat, int(rng.integers(0, k + 2)))
        s = 1000 + i * 900 + (i % 4)
        text[s:s + len(mp)] = mp
    st = sp.Store.from_host(ctx, text, 96)
    info = check_filtered(ctx, st, text, pats, k, f"step={step}")
    assert info["step"] == step and info["q"] == 25 - step + 1
    st.close()


def test_filtered_myers_mixed_lengths_and_repeats(ctx):
    rng = np.random.default_rng(99)
    n = 120_000
    text = rand_text(rng, n)
    # low-complexity stretch: many overlapping seed windows report the same ends (dedupe, min errors)
    text[5000:5400] = np.tile(np.array([0, 1], dtype=np.uint8), 200)
    pats = [np.tile(np.array([0, 1], dtype=np.uint8), 30), text[70_000:70_090].copy(), text[300:370].copy()]
    st = sp.Store.from_host(ctx, text, 0)
    info = check_filtered(ctx, st, text, pats, 3, "mixed")
    assert info["filtered"]
    st.close()


def test_filtered_falls_back_to_direct_kernel_for_short_q(ctx):
    rng = np.random.default_rng(3)
    text = rand_text(rng, 30000)
    pats = [text[100:130].copy()]
    st = sp.Store.from_host(ctx, text)
    info = check_filtered(ctx, st, text, pats, 4, "short q")        # q = 6 < 12: direct Myers kernel
    assert not info["filtered"]
    st.close()


def test_filtered_myers_sharded(ctx):
    rng = np.random.default_rng(17)
    n, m, k = 300_000, 80, 3
    text = rand_text(rng, n)
    pat = text[1234:1234 + m].copy()
    st = sp.Store.from_host(ctx, text, 64)             # many super-tiles
    # occurrences straddling shard seams (shards are contiguous super-tile ranges of 2048 symbols)
    for s in range(2048 * 7 - m, n - 2 * m, 2048 * 9 + 13):
        mp = mutate(rng, pat, int(rng.integers(0, k + 1)))
        text[s:s + len(mp)] = mp
    st.close()
    st = sp.Store.from_host(ctx, text, 64)
    pf = sp.Pattern(ctx, sp.ALG_MYERS_FILTERED, [pat], k)
    whole = sp.search(ctx, st, pf)
    _, we, wr = orc.search_mt(3, text, pat, k)
    assert whole["end"].tolist() == we.tolist() and whole["errors"].tolist() == wr.tolist()
    parts = []
    for sh in range(5):
        ctx.set_shard(sh, 5)
        parts.append(sp.search(ctx, st, pf))
    ctx.set_shard(0, 1)
    cat = np.concatenate(parts)
    cat = cat[np.argsort(cat["end"], kind="stable")]
    assert cat["end"].tolist() == we.tolist() and cat["errors"].tolist() == wr.tolist()
    pf.close(); st.close()


def test_prefix_matcher_vs_oracle(ctx):
    rng = np.random.default_rng(23)
    for m, k in [(5, 1), (31, 3), (32, 0), (33, 4), (100, 4), (150, 6), (200, 10)]:
        pat = rand_text(rng, m)
        for e in range(0, k + 3):
            text = np.concatenate([mutate(rng, pat, e), rand_text(rng, 50)])
            st = sp.Store.from_host(ctx, text)
            p = sp.Pattern(ctx, sp.ALG_MYERS_PREFIX, pat, k)
            assert p.window_size == m + k     # myers_prefix_matcher_restorable.hpp:151-153
            got = sp.search(ctx, st, p)
            we, wr = orc.myers_prefix(text, pat, k)
            assert got["end"].tolist() == we.tolist(), (m, k, e)
            assert got["errors"].tolist() == wr.tolist(), (m, k, e)
            p.close(); st.close()


def test_prefix_matcher_anchored_batch(ctx):
    rng = np.random.default_rng(29)
    n, m, k = 100_000, 60, 3
    text = rand_text(rng, n)
    pats = [text[500:500 + m].copy(), text[9000:9000 + m - 7].copy()]
    anchors = []
    for i in range(40):
        pat = pats[i % 2]
        mp = mutate(rng, pat, int(rng.integers(0, k + 2)))
        s = int(rng.integers(0, n - 2 * m))
        text[s:s + len(mp)] = mp
        anchors.append(s + int(rng.integers(-1, 2)))
    anchors += [0, n - 10, n - 1]
    st = sp.Store.from_host(ctx, text, 96)
    p = sp.Pattern(ctx, sp.ALG_MYERS_PREFIX, pats, k)
    got = sp.search_anchored(ctx, st, p, anchors)
    want = []
    for pid, pat in enumerate(pats):
        for ai, a in enumerate(anchors):
            we, wr = orc.myers_prefix(text[a:], pat, k)
            want += [(pid, ai, int(e) + a, int(r)) for e, r in zip(we, wr)]
    want.sort()
    assert [(int(x["pattern_id"]), int(x["seq_id"]), int(x["end"]), int(x["errors"])) for x in got] == want
    p.close(); st.close()


def test_prefix_matcher_stream_capture_restore(ctx):
    rng = np.random.default_rng(31)
    m, k = 40, 3
    pat = rand_text(rng, m)
    text = np.concatenate([mutate(rng, pat, 2), rand_text(rng, 100)])
    we, wr = orc.myers_prefix(text, pat, k)
    p = sp.Pattern(ctx, sp.ALG_MYERS_PREFIX, pat, k)
    for chunk in (1, 7, 13, 64):
        s = sp.Stream(ctx, p)
        ends, errs = [], []
        for off in range(0, len(text), chunk):
            if off == 2 * chunk:             # capture / restore into a fresh stream mid-way
                state = s.capture()
                s.close()
                s = sp.Stream(ctx, p)
                s.restore(state)
            h = s.feed(text[off:off + chunk])
            ends += [int(e) + off for e in h["end"]]
            errs += [int(e) for e in h["errors"]]
        assert ends == we.tolist() and errs == wr.tolist(), chunk
        s.close()
    p.close()
