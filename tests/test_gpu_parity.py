"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle on the same
seeded inputs -- bit-exact hit lists (end, errors)."""
import numpy as np
import pytest

import libspm_b200 as sp
import oracle as orc
from tests.util import (GOLD_EXACT_BEGIN, GOLD_HAYSTACK, GOLD_MYERS_K1_END, GOLD_NEEDLE, dna4, mutate,
                        rand_text)

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = sp.Context(0)
    yield c
    c.close()


def _diff(got, want, what):
    got = np.asarray(got)
    want = np.asarray(want)
    if got.shape == want.shape and np.array_equal(got, want):
        return
    gs, ws = set(got.tolist()), set(want.tolist())
    miss = sorted(ws - gs)[:10]
    extra = sorted(gs - ws)[:10]
    raise AssertionError(f"{what}: got {len(got)} want {len(want)}; missing {miss} extra {extra}")


def check_exact(ctx, store, text, pat, algs=(sp.ALG_HORSPOOL, sp.ALG_SHIFTOR), tag=""):
    want = orc.exact_bruteforce(text, pat) if len(text) < 3_000_000 else orc.search_mt(2, text, pat)[1]
    for alg in algs:
        p = sp.Pattern(ctx, alg, pat)
        assert p.window_size == len(pat)
        got = sp.search(ctx, store, p)
        _diff(got["end"], want, f"exact alg={alg} m={len(pat)} rl={store.row_len} {tag}")
        assert not got["errors"].any()
        p.close()
    return want


def check_myers(ctx, store, text, pat, k, tag=""):
    if len(text) <= 20000:
        we, wr = orc.sellers(text, pat, k)
    else:
        _, we, wr = orc.search_mt(3, text, pat, k)
    p = sp.Pattern(ctx, sp.ALG_MYERS, pat, k)
    assert p.window_size == len(pat) + k
    got = sp.search(ctx, store, p)
    _diff(got["end"], we, f"myers ends m={len(pat)} k={k} rl={store.row_len} {tag}")
    _diff(got["errors"], wr, f"myers errors m={len(pat)} k={k} rl={store.row_len} {tag}")
    p.close()
    return we


# ----------------------------------------------------------------------------------------
@pytest.mark.parametrize("n,rl", [(1, 0), (31, 0), (44, 0), (1000, 32), (1024, 32), (32768, 0), (100001, 64),
                                  (100001, 0), (1_000_003, 0), (3_000_000, 1056)])
def test_store_roundtrip(ctx, n, rl):
    rng = np.random.default_rng(n)
    text = rand_text(rng, n)
    st = sp.Store.from_host(ctx, text, rl)
    assert len(st) == n
    back = st.read()
    assert np.array_equal(back, text)
    if n > 100:
        assert np.array_equal(st.read(37, 50), text[37:87])
    st.close()


def test_reference_golden_vectors(ctx):
    st = sp.Store.from_host(ctx, GOLD_HAYSTACK)
    m = len(GOLD_NEEDLE)
    for alg in (sp.ALG_HORSPOOL, sp.ALG_SHIFTOR):
        p = sp.Pattern(ctx, alg, GOLD_NEEDLE)
        assert p.window_size == 5
        got = sp.search(ctx, st, p)
        assert [int(e) - m for e in got["end"]] == GOLD_EXACT_BEGIN
    p = sp.Pattern(ctx, sp.ALG_MYERS, GOLD_NEEDLE, 1)
    assert p.window_size == 6
    got = sp.search(ctx, st, p)
    assert [int(e) for e in got["end"]] == GOLD_MYERS_K1_END
    assert [int(e) for e in got["errors"]] == [1, 0, 1] * 3


def test_reference_golden_restorable_chunks(ctx):
    # myers_matcher_restorable_test.cpp:55-74 -- capture / restore around 13-symbol chunks
    p = sp.Pattern(ctx, sp.ALG_MYERS, GOLD_NEEDLE, 1)
    s = sp.Stream(ctx, p)
    state = s.capture()
    got = []
    for off in range(0, len(GOLD_HAYSTACK), 13):
        s.restore(state)
        h = s.feed(GOLD_HAYSTACK[off:off + 13])
        got += [int(e) + off for e in h["end"]]
        state = s.capture()
    assert got == GOLD_MYERS_K1_END
    # restorable Shift-Or, 7-symbol chunks
    p2 = sp.Pattern(ctx, sp.ALG_SHIFTOR, GOLD_NEEDLE)
    s2 = sp.Stream(ctx, p2)
    got = []
    for off in range(0, len(GOLD_HAYSTACK), 7):
        got += [int(e) + off - 5 for e in s2.feed(GOLD_HAYSTACK[off:off + 7])["end"]]
    assert got == GOLD_EXACT_BEGIN


@pytest.mark.parametrize("rl", [32, 96, 1024, 0])
@pytest.mark.parametrize("sigma", [2, 4])
def test_exact_random(ctx, rl, sigma):
    rng = np.random.default_rng(100 + rl + sigma)
    n = 200_000
    text = rand_text(rng, n, sigma)
    R = rl if rl else 1024
    pats = []
    for m in (1, 2, 5, 7, 8, 9, 16, 31, 32, 33, 64, 65, 100, 150, 257):
        start = int(rng.integers(0, n - m))
        pat = text[start:start + m].copy()
        # plant copies straddling row ends, super-tile ends and the text end
        for pos in (R - m // 2, 2 * R - 1, 32 * R - m // 2 - 1, 32 * R - 1, 64 * R - 3, n - m, 0):
            pos = int(pos)
            if 0 <= pos and pos + m <= n:
                text[pos:pos + m] = pat
        pats.append(pat)
    st = sp.Store.from_host(ctx, text, rl)
    total = 0
    for pat in pats:
        total += len(check_exact(ctx, st, text, pat))
    assert total >= len(pats)   # later plants may overwrite earlier ones, but most survive
    st.close()


def test_exact_overlapping_and_dense(ctx):
    text = np.concatenate([dna4("TTGCA"), np.tile(dna4("ACGT"), 16), dna4("GGAT")])
    pat = np.tile(dna4("ACGT"), 8)
    st = sp.Store.from_host(ctx, text)
    assert len(check_exact(ctx, st, text, pat)) == 9
    # every position is a candidate: poly-A text, poly-A needle; also forces the hit buffer to grow
    text = np.zeros(150_000, dtype=np.uint8)
    st = sp.Store.from_host(ctx, text, 64)
    for m in (1, 3, 8, 20, 70):
        pat = np.zeros(m, dtype=np.uint8)
        p = sp.Pattern(ctx, sp.ALG_SHIFTOR, pat)
        h = sp.Hits(ctx, 16)
        got = sp.search(ctx, st, p, h)
        assert len(got) == len(text) - m + 1
        assert np.array_equal(got["end"], np.arange(m, len(text) + 1, dtype=np.uint64))


def test_edge_cases(ctx):
    text = rand_text(np.random.default_rng(1), 500)
    st = sp.Store.from_host(ctx, text)
    # needle longer than the haystack, empty needle
    assert len(sp.search(ctx, st, sp.Pattern(ctx, sp.ALG_HORSPOOL, rand_text(np.random.default_rng(2), 600)))) == 0
    p0 = sp.Pattern(ctx, sp.ALG_SHIFTOR, np.zeros(0, np.uint8))
    assert p0.window_size == 0 and len(sp.search(ctx, st, p0)) == 0
    pm0 = sp.Pattern(ctx, sp.ALG_MYERS, np.zeros(0, np.uint8), 2)
    assert len(sp.search(ctx, st, pm0)) == 0
    # whole text is the needle
    check_exact(ctx, st, text, text.copy())
    # hit at 0 and at the end
    check_exact(ctx, st, text, text[:40].copy())
    check_exact(ctx, st, text, text[-40:].copy())
    check_myers(ctx, st, text, text[:40].copy(), 3)
    check_myers(ctx, st, text, text[-40:].copy(), 3)
    # empty haystack
    st0 = sp.Store.from_host(ctx, np.zeros(0, np.uint8))
    assert len(sp.search(ctx, st0, sp.Pattern(ctx, sp.ALG_HORSPOOL, text[:5]))) == 0
    assert len(sp.search(ctx, st0, sp.Pattern(ctx, sp.ALG_MYERS, text[:5], 1))) == 0


@pytest.mark.parametrize("rl", [32, 160, 1024, 0])
def test_myers_random(ctx, rl):
    rng = np.random.default_rng(500 + rl)
    n = 150_000
    R = rl if rl else 1024
    for (m, k) in [(5, 1), (8, 0), (20, 3), (31, 2), (32, 4), (33, 1), (64, 2), (65, 2), (96, 5), (100, 4),
                   (128, 6), (129, 3), (150, 2), (160, 20), (200, 9), (224, 3), (256, 8), (30, 29)]:
        text = rand_text(rng, n)
        pat = rand_text(rng, m)
        for pos in (0, R - m // 2, 2 * R - 2, 32 * R - m // 3, 32 * R - 1, 64 * R + 5, n - m - 1, n - m // 2 - 1,
                    40_000, 77_777, 120_000):
            mp = mutate(rng, pat, int(rng.integers(0, k + 3)))
            pos = max(0, min(int(pos), n - len(mp)))
            text[pos:pos + len(mp)] = mp
        st = sp.Store.from_host(ctx, text, rl)
        we = check_myers(ctx, st, text, pat, k)
        assert len(we) >= 1
        st.close()


def test_myers_low_complexity(ctx):
    rng = np.random.default_rng(9)
    text = rand_text(rng, 60_000, 2)
    text[10_000:20_000] = 0
    st = sp.Store.from_host(ctx, text, 96)
    for m, k in [(12, 3), (40, 10), (70, 5)]:
        pat = rand_text(rng, m, 2)
        check_myers(ctx, st, text, pat, k)
        check_myers(ctx, st, text, np.zeros(m, np.uint8), k)


def test_pattern_sets(ctx):
    rng = np.random.default_rng(11)
    n = 120_000
    text = rand_text(rng, n)
    pats = []
    for i in range(37):
        m = int(rng.integers(97, 129))
        s = int(rng.integers(0, n - m))
        pats.append(mutate(rng, text[s:s + m].copy(), int(rng.integers(0, 5)))[:128])
    pats = [p for p in pats if 97 <= len(p) <= 128]
    st = sp.Store.from_host(ctx, text, 512)
    ps = sp.Pattern(ctx, sp.ALG_MYERS, pats, 3)
    assert ps.count == len(pats)
    got = sp.search(ctx, st, ps)
    for i, p in enumerate(pats):
        _, we, wr = orc.search_mt(3, text, p, 3)
        sel = got[got["pattern_id"] == i]
        _diff(sel["end"], we, f"set myers pattern {i}")
        _diff(sel["errors"], wr, f"set myers errors {i}")
    es = sp.Pattern(ctx, sp.ALG_HORSPOOL, [text[5:25].copy(), text[1000:1064].copy(), text[n - 9:].copy()])
    got = sp.search(ctx, st, es)
    for i, p in enumerate([text[5:25], text[1000:1064], text[n - 9:]]):
        _diff(got[got["pattern_id"] == i]["end"], orc.exact_bruteforce(text, p), f"set exact {i}")


def test_journaled_store(ctx):
    rng = np.random.default_rng(21)
    n = 300_000
    ref = rand_text(rng, n)
    # ~1% events: 80% SNP, 10% insertion U[1,10], 10% deletion U[1,10]
    pos = np.sort(rng.choice(n - 20, size=3000, replace=False))
    ev, pool = [], []
    last_end = 0
    for p in pos:
        p = int(p)
        if p < last_end:
            continue
        u = rng.random()
        if u < 0.8:
            alt = (int(ref[p]) + 1 + int(rng.integers(0, 3))) % 4
            ev.append((p, 1, 1, len(pool))); pool.append(alt); last_end = p + 1
        elif u < 0.9:
            L = int(rng.integers(1, 11))
            ev.append((p, 0, L, len(pool))); pool += list(rng.integers(0, 4, size=L)); last_end = p
            last_end = p + 1  # keep events on distinct reference positions
        else:
            L = int(rng.integers(1, 11))
            ev.append((p, L, 0, 0)); last_end = p + L
    ev = np.array(ev, dtype=sp.EVENT_DTYPE)
    pool = np.array(pool, dtype=np.uint8)
    derived = orc.journal_apply(ref, ev.astype(orc.EVENT_DTYPE), pool)
    base = sp.Store.from_host(ctx, ref)
    for rl in (0, 96):
        js = base.journaled(ev, pool, rl)
        assert len(js) == len(derived)
        assert np.array_equal(js.read(), derived)
        # needle cut from the derived string across an insertion
        ins = ev[ev["ins_len"] > 3][0]
        dpos = int(ins["ref_pos"]) + int((ev["ins_len"].astype(np.int64) - ev["del_len"])[ev["ref_pos"] < ins["ref_pos"]].sum())
        pat = derived[dpos - 30:dpos + 34].copy()
        want = check_exact(ctx, js, derived, pat, tag="journaled")
        assert len(want) >= 1
        check_myers(ctx, js, derived, pat, 2, tag="journaled")
        js.close()
    # empty journal = identity
    js = base.journaled(np.zeros(0, dtype=sp.EVENT_DTYPE), np.zeros(0, np.uint8))
    assert np.array_equal(js.read(), ref)


def test_stream_chunks_equal_unchunked(ctx):
    rng = np.random.default_rng(31)
    text = rand_text(rng, 30_000)
    pat = text[5000:5100].copy()
    text[20_000:20_100] = pat
    text[20_050] ^= 3
    we, wr = orc.myers(text, pat, 4)
    p = sp.Pattern(ctx, sp.ALG_MYERS, pat, 4)
    for chunk in (1000, 4999, 103, 104):
        s = sp.Stream(ctx, p)
        ge, gr = [], []
        for off in range(0, len(text), chunk):
            h = s.feed(text[off:off + chunk])
            ge += [int(e) + off for e in h["end"]]
            gr += [int(e) for e in h["errors"]]
        _diff(ge, we, f"stream chunk={chunk}")
        _diff(gr, wr, f"stream chunk={chunk} errors")
        s.close()


def test_search_host_call_shape(ctx):
    rng = np.random.default_rng(41)
    text = rand_text(rng, 400_000)
    pat = text[1234:1234 + 32].copy()
    p = sp.Pattern(ctx, sp.ALG_HORSPOOL, pat)
    got = sp.search_host(ctx, text, p)
    _diff(got["end"], orc.exact_bruteforce(text, pat), "search_host")


def test_int_peak_probe(ctx):
    r = ctx.measure_int_peak()
    assert r["lop3_gops"] > 1000 and r["iadd3_gops"] > 1000


def test_journal_validation_errors(ctx):
    ref = rand_text(np.random.default_rng(3), 10_000)
    base = sp.Store.from_host(ctx, ref)
    pool = np.zeros(4, np.uint8)
    bad_sets = [
        [(500, 1, 1, 0), (400, 1, 1, 1)],          # unsorted
        [(500, 10, 0, 0), (505, 1, 1, 0)],         # overlapping
        [(9_995, 10, 0, 0)],                       # deletion runs off the reference
        [(100, 0, 3, 2)],                          # insertion outside the pool
    ]
    for evs in bad_sets:
        with pytest.raises(sp.SpmbError) as ei:
            base.journaled(np.array(evs, dtype=sp.EVENT_DTYPE), pool)
        assert ei.value.code == 1
    # adjacent events (deletion directly followed by an insertion) are legal
    ev = np.array([(100, 5, 0, 0), (105, 0, 4, 0), (9_999, 1, 1, 0)], dtype=sp.EVENT_DTYPE)
    js = base.journaled(ev, pool)
    want = orc.journal_apply(ref, ev.astype(orc.EVENT_DTYPE), pool)
    assert np.array_equal(js.read(), want)


@pytest.mark.parametrize("n,rl,m", [(170_000_000, 0, 64), (170_000_000, 1504, 20), (165_000_001, 4096, 7), (200_000_000, 8192, 33)])
def test_exact_large_store(ctx, n, rl, m):
    """stores of 165-200 Mbp (grid-stride exact kernel over many CTA tiles), unsharded and
    sharded, against the multi-threaded oracle; needles planted across thread-row, row and super-tile seams."""
    import torch
    from libspm_b200 import synth
    dev = torch.device("cuda", 0)
    t = synth.text_device(n, n % 1000 + m, dev)
    host = t.cpu().numpy()
    rng = np.random.default_rng(m)
    pat = rand_text(rng, m)
    st = sp.Store.from_device(ctx, t.data_ptr(), n, rl)
    ctx.sync()
    RL = st.row_len
    st.close()
    spots = [0, n - m, RL - m // 2, 32 * RL - 3, 32 * RL - m + 1, 5 * 32 * RL + RL - 1, 1024 * 37 - 5, 256 * 32 * 9 - m // 2]
    spots += [int(x) for x in rng.integers(0, n - m, size=200)]
    for s_ in spots:
        if 0 <= s_ <= n - m:
            host[s_:s_ + m] = pat
    t.copy_(torch.from_numpy(host))
    st = sp.Store.from_device(ctx, t.data_ptr(), n, rl)
    ctx.sync()
    del t
    _, want, _ = orc.search_mt(2, host, pat, 0)
    p = sp.Pattern(ctx, sp.ALG_SHIFTOR, pat)
    got = sp.search(ctx, st, p)
    _diff(got["end"], want, f"large store n={n} rl={RL} m={m}")
    parts = []
    for sh in range(3):
        ctx.set_shard(sh, 3)
        parts.append(sp.search(ctx, st, p)["end"])
    ctx.set_shard(0, 1)
    _diff(np.sort(np.concatenate(parts)), want, f"large store sharded n={n} rl={RL} m={m}")
    p.close(); st.close()


def test_pipelined_launch_finish_two_hit_lists(ctx):
    """spmb_search_launch / spmb_search_finish interleaved over two hit lists (launch i+1 is queued before
    the hits of i are collected): every list equals the one-call result -- short lists (counters + hits in
    one copy), lists longer than the speculative copy, and a list that outgrows its buffer."""
    rng = np.random.default_rng(404)
    text = rand_text(rng, 3_000_000)
    store = sp.Store.from_host(ctx, text)
    pats = []
    for i in range(6):
        p = text[100_000 * (i + 1):100_000 * (i + 1) + 48].copy()
        for c in range(3 + 700 * (i % 3)):                      # 3, 703, 1403 planted copies
            pos = int(rng.integers(0, len(text) - 64))
            text[pos:pos + 48] = p
        pats.append(p)
    store.close()
    store = sp.Store.from_host(ctx, text)
    pm = [sp.Pattern(ctx, sp.ALG_SHIFTOR if i % 2 else sp.ALG_MYERS, p, 0 if i % 2 else 2) for i, p in enumerate(pats)]
    want = [sp.search(ctx, store, p).copy() for p in pm]
    ha, hb = sp.Hits(ctx, 16), sp.Hits(ctx, 1 << 12)            # ha starts too small: regrowth inside finish
    hl = (ha, hb)
    sp.search_launch(ctx, store, pm[0], hl[0])
    for i in range(len(pm)):
        if i + 1 < len(pm):
            sp.search_launch(ctx, store, pm[i + 1], hl[(i + 1) & 1])
        got = sp.search_finish(ctx, hl[i & 1])
        assert len(got) == len(want[i]) and len(got) >= 1
        for f in ("end", "errors", "pattern_id", "seq_id"):
            assert np.array_equal(got[f], want[i][f]), (i, f)
    assert max(len(w) for w in want) > 1024 and min(len(w) for w in want) < 16
    ha.close(); hb.close(); store.close()
