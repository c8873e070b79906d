"""GPU parity tests for the round-2 kernel work: the whole full-size C4 hit list against the committed oracle
fixture, the exact kernel's filter window at any needle offset (low-complexity text, row seams, text start),
dense journals through the warp-cooperative event batch, hit-dense Myers through the warp-aggregated append,
and the cut-off switch on 4-bit alphabets."""
import hashlib
import os

import numpy as np
import pytest

import libspm_b200 as sp
import oracle as orc
from libspm_b200 import synth
from tests.util import mutate, rand_text

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


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


# ---------------------------------------------------------------------------------------- C4 at full size
def test_c4_full_size_hit_list_equals_the_oracle_fixture(ctx):
    """BASELINE configs[3] as specified: 2504 haplotypes x 50 Mbp, 1 M variants, Myers m=150 k=2.  The oracle's
    per-haplotype pass over all 2504 haplotypes (125 Gbp of CPU Myers, tests/golden/make_c4_golden.py) is committed
    as tests/golden/c4_hits.npz; the CUDA path's WHOLE hit list must equal it."""
    g = np.load(os.path.join(ROOT, "tests", "golden", "c4_hits.npz"))
    d = synth.c4()
    h = hashlib.sha256()
    for key in ("ref", "variants", "pool", "coverage", "pattern"):
        h.update(np.ascontiguousarray(d[key]).view(np.uint8).tobytes())
    assert h.hexdigest() == str(g["input_sha256"]), "the generator no longer produces the inputs the fixture was made from"
    store = sp.Store.from_host(ctx, d["ref"])
    jst = sp.Jst(ctx, store, d["variants"], d["pool"], d["coverage"], d["n_hap"])
    p = sp.Pattern(ctx, sp.ALG_MYERS, d["pattern"], d["k"])
    got = jst.search(p)
    assert len(got) == len(g["hap"]), (len(got), len(g["hap"]))
    assert np.array_equal(got["seq_id"], g["hap"]) and np.array_equal(got["end"], g["end"])
    assert np.array_equal(got["errors"], g["err"].astype(np.uint32))
    # the same list from 4 shards of the window array (what ranks do at N > 1)
    parts = []
    for s in range(4):
        ctx.set_shard(s, 4)
        parts.append(jst.search(p))
    ctx.set_shard(0, 1)
    merged = np.concatenate(parts)
    merged = merged[np.lexsort((merged["end"], merged["seq_id"]))]
    assert np.array_equal(merged["seq_id"], g["hap"]) and np.array_equal(merged["end"], g["end"])
    jst.close(); p.close(); store.close()


# ---------------------------------------------------------------------------------------- exact: filter window
def low_complexity_text(rng, n, frac=0.05):
    """iid text with runs of poly-A, period-2 (AC)n and period-3 repeats covering ~frac of it."""
    t = rand_text(rng, n)
    pos = 0
    while pos < n - 400:
        pos += int(rng.integers(200, int(2 * 300 / frac)))
        ln = int(rng.integers(20, 300))
        kind = int(rng.integers(0, 3))
        unit = [np.array([0]), np.array([0, 1]), np.array([3, 3, 2])][kind]
        if pos + ln < n:
            t[pos:pos + ln] = np.tile(unit, ln // len(unit) + 1)[:ln]
    return t


@pytest.mark.parametrize("rl", [32, 64, 1024, 0])
def test_exact_needle_that_starts_in_a_repeat(ctx, rl):
    """the needle's first symbols lie in a poly-A / (AC)n run: with the filter on the first eight symbols every
    alignment of every such run reaches verification; the filter window now sits on the least repetitive eight
    symbols, anywhere in the needle.  Hit list = oracle, for windows at offsets far beyond a row (rl = 32)."""
    rng = np.random.default_rng(1000 + rl)
    n = 400_000
    text = low_complexity_text(rng, n, 0.08)
    store = sp.Store.from_host(ctx, text, rl)
    for head, unit in ((12, [0]), (40, [0, 1]), (150, [0]), (9, [3, 3, 2])):
        tail = rand_text(rng, int(rng.integers(8, 60)))
        pat = np.concatenate([np.tile(np.array(unit, dtype=np.uint8), head)[:head], tail]).astype(np.uint8)
        # plant it a few times, also right at the start and the end of the text
        for s in (0, 1001, 2031, 3032, 4033, 5000, 77_777, n - len(pat)):
            text[s:s + len(pat)] = pat
        store.close()
        store = sp.Store.from_host(ctx, text, rl)
        want = orc.exact_bruteforce(text, pat)
        assert len(want) >= 8
        for alg in (sp.ALG_HORSPOOL, sp.ALG_SHIFTOR):
            p = sp.Pattern(ctx, alg, pat)
            got = sp.search(ctx, store, p)
            _same(got["end"], want, f"alg={alg} head={head} unit={unit} m={len(pat)} rl={rl}")
            p.close()
    # a needle that is low-complexity throughout: every window ties, hits are dense
    pat = np.zeros(20, dtype=np.uint8)
    want = orc.exact_bruteforce(text, pat)
    p = sp.Pattern(ctx, sp.ALG_SHIFTOR, pat)
    _same(sp.search(ctx, store, p)["end"], want, f"poly-A needle rl={rl}")
    p.close(); store.close()


def test_exact_window_offset_with_shards_and_wide_alphabet(ctx):
    rng = np.random.default_rng(77)
    n = 300_000
    text = low_complexity_text(rng, n)
    pat = np.concatenate([np.zeros(30, np.uint8), rand_text(rng, 25)])
    for s in (0, 10, 4096 - 30, 150_000, n - len(pat)):
        text[s:s + len(pat)] = pat
    want = orc.exact_bruteforce(text, pat)
    store = sp.Store.from_host(ctx, text, 96)
    p = sp.Pattern(ctx, sp.ALG_HORSPOOL, pat)
    parts = []
    for s in range(5):
        ctx.set_shard(s, 5)
        parts.append(sp.search(ctx, store, p)["end"])
    ctx.set_shard(0, 1)
    _same(np.sort(np.concatenate(parts)), want, "shards partition the hit list")
    p.close(); store.close()
    # dna5: the high plane pair takes part in the verification with the same offsets
    t5 = text.copy()
    t5[rng.integers(0, n, size=n // 50)] = 4
    pat5 = pat.copy(); pat5[40] = 4
    for s in (0, 5, 100_000):
        t5[s:s + len(pat5)] = pat5
    want = orc.exact_bruteforce(t5, pat5)
    st = sp.Store.from_host_wide(ctx, t5, 3, 64)
    p = sp.Pattern(ctx, sp.ALG_SHIFTOR, pat5, bits=3)
    _same(sp.search(ctx, st, p)["end"], want, "dna5 with a filter window offset")
    p.close(); st.close()


def test_jst_exact_search_with_window_offset(ctx):
    """windows mode: an alignment must begin inside its own row whatever the filter window offset is."""
    rng = np.random.default_rng(5)
    n_ref, n_hap = 30_000, 40
    ref = low_complexity_text(rng, n_ref, 0.1)
    var, pool, kinds, cov = synth.jst(ref, 900, n_hap, 5, p_snp=0.7, p_ins=0.15, af_max=0.6)
    store = sp.Store.from_host(ctx, ref, 64)
    jst = sp.Jst(ctx, store, var, pool, cov, n_hap)
    hap = orc.jst_haplotype(ref, var.astype(orc.EVENT_DTYPE), pool, cov, 3)
    for s, m in ((5000, 40), (12_000, 25), (20_000, 90)):
        pat = hap[s:s + m].copy()
        pat[:10] = 0                                        # a repeat at the front pushes the window back
        p = sp.Pattern(ctx, sp.ALG_HORSPOOL, pat)
        got = jst.search(p)
        want = orc.jst_search(1, ref, var.astype(orc.EVENT_DTYPE), pool, cov, n_hap, pat, 0)
        assert len(got) == len(want) and np.array_equal(got["seq_id"], want["hap"]) and np.array_equal(got["end"], want["end"])
        p.close()
    jst.close(); store.close()


# ---------------------------------------------------------------------------------------- journal: dense events
@pytest.mark.parametrize("density,max_indel", [(0.01, 10), (0.05, 4), (0.2, 1), (0.3, 1)])
@pytest.mark.parametrize("rl", [0, 96])
def test_journal_materialisation_dense_events(ctx, density, max_indel, rl):
    """the warp-cooperative event batch holds 32 events; at 0.2-0.3 events per symbol a 1024-symbol segment sees
    hundreds, so the batch slides and lanes finish their walk in global memory.  Derived text = oracle."""
    n = 700_000
    ref = synth.text_host(n, 900 + max_indel)
    ev, pool, kinds = synth.journal(ref, 901, density=density, max_indel=max_indel)
    assert len(ev) > 0.6 * density * n
    derived = orc.journal_apply(ref, ev.astype(orc.EVENT_DTYPE), pool)
    base = sp.Store.from_host(ctx, ref)
    js = base.journaled(ev, pool, rl)
    assert len(js) == len(derived)
    got = js.read()
    if not np.array_equal(got, derived):
        bad = np.flatnonzero(got != derived)
        raise AssertionError(f"{len(bad)} symbols differ, first at {bad[:5]}")
    pat = derived[n // 2:n // 2 + 48].copy()
    p = sp.Pattern(ctx, sp.ALG_SHIFTOR, pat)
    _same(sp.search(ctx, js, p)["end"], orc.exact_bruteforce(derived, pat), "search over the dense journal")
    p.close(); js.close(); base.close()


def test_journal_events_clustered_at_one_place(ctx):
    """thousands of no-op and 1-symbol events packed into a few hundred reference positions, the rest of the
    reference untouched: the batch must slide many times for the segments right behind the cluster."""
    rng = np.random.default_rng(3)
    n = 200_000
    ref = rand_text(rng, n)
    pos = np.sort(np.concatenate([np.arange(50_000, 50_000 + 3000, 1), np.array([10, 120_000, 199_000])]))
    ev = np.zeros(len(pos), dtype=sp.EVENT_DTYPE)
    ev["ref_pos"] = pos
    kinds = rng.integers(0, 3, size=len(pos))
    ev["del_len"] = np.where(kinds == 0, 1, 0)
    ev["ins_len"] = np.where(kinds == 1, 1, np.where(kinds == 0, 1, 0))
    il = ev["ins_len"].astype(np.int64)
    ev["ins_off"] = np.cumsum(il) - il
    pool = rng.integers(0, 4, size=int(il.sum()), dtype=np.uint8)
    derived = orc.journal_apply(ref, ev.astype(orc.EVENT_DTYPE), pool)
    base = sp.Store.from_host(ctx, ref)
    for rl in (0, 32, 2048):
        js = base.journaled(ev, pool, rl)
        assert np.array_equal(js.read(), derived), rl
        js.close()
    base.close()


# ---------------------------------------------------------------------------------------- Myers: dense hits
@pytest.mark.parametrize("m,k", [(20, 2), (100, 4), (150, 2)])
def test_myers_hit_dense_text(ctx, m, k):
    """poly-A text, poly-A needle: EVERY column is a hit in every lane (the warp-aggregated append takes one
    atomic per warp and column); a second text mixes dense stretches with random ones (divergent appends)."""
    rng = np.random.default_rng(m)
    n = 300_000
    for mix in (False, True):
        text = np.zeros(n, dtype=np.uint8)
        if mix:
            for s in range(0, n, 20_000):
                text[s:s + 9000] = rand_text(rng, 9000)
        pat = np.zeros(m, dtype=np.uint8)
        pat[m // 2] = 1
        store = sp.Store.from_host(ctx, text)
        _, we, wr = orc.search_mt(3, text, pat, k)
        assert len(we) > n // 4
        p = sp.Pattern(ctx, sp.ALG_MYERS, pat, k)
        got = sp.search(ctx, store, p)
        _same(got["end"], we, f"dense ends m={m} k={k} mix={mix}")
        _same(got["errors"], wr, f"dense errors m={m} k={k} mix={mix}")
        ctx.set_myers_cutoff(False)
        got2 = sp.search(ctx, store, p)
        ctx.set_myers_cutoff(True)
        assert np.array_equal(got2["end"], we) and np.array_equal(got2["errors"], wr)
        p.close(); store.close()


def test_myers_pair_lookup_fast_path_long_rows(ctx):
    """the one-word fast path now reads Eq for two symbols at a time: long rows of random text (where a warp stays
    in that path) with planted near-matches, cut-off result = full recurrence = oracle."""
    rng = np.random.default_rng(8)
    n = 3_000_000
    text = rand_text(rng, n)
    for m, k in ((33, 1), (64, 3), (100, 4), (150, 2), (250, 6)):
        pat = rand_text(rng, m)
        for s in rng.integers(0, n - 2 * m, size=60):
            mp = mutate(rng, pat, int(rng.integers(0, k + 3)))
            text[int(s):int(s) + len(mp)] = mp
        store = sp.Store.from_host(ctx, text)
        _, we, wr = orc.search_mt(3, text, pat, k)
        p = sp.Pattern(ctx, sp.ALG_MYERS, pat, k)
        got = sp.search(ctx, store, p)
        _same(got["end"], we, f"ends m={m} k={k}")
        _same(got["errors"], wr, f"errors m={m} k={k}")
        p.close(); store.close()


# ---------------------------------------------------------------------------------------- 4-bit alphabets
@pytest.mark.parametrize("m,k", [(40, 3), (100, 4), (200, 12)])
def test_wide_alphabet_cutoff_switch_selects_the_full_recurrence(ctx, m, k):
    """spmb_context_set_myers_cutoff(0) was ignored for dna5 / dna15 stores (the cut-off kernel ran
    either way).  Both settings must give the oracle's list; the launch counts show they are different kernels
    only indirectly, so the check here is parity of both."""
    rng = np.random.default_rng(m * 7 + k)
    n = 150_000
    text = rand_text(rng, n, 15)
    pat = rand_text(rng, m, 15)
    for s in rng.integers(0, n - 2 * m, size=30):
        mp = mutate(rng, pat, int(rng.integers(0, k + 2)))
        text[int(s):int(s) + len(mp)] = mp
    _, we, wr = orc.search_mt(3, text, pat, k)
    st = sp.Store.from_host_wide(ctx, text, 4)
    p = sp.Pattern(ctx, sp.ALG_MYERS, pat, k, bits=4)
    for cut in (True, False):
        ctx.set_myers_cutoff(cut)
        got = sp.search(ctx, st, p)
        _same(got["end"], we, f"wide m={m} k={k} cutoff={cut}")
        _same(got["errors"], wr, f"wide errors m={m} k={k} cutoff={cut}")
    ctx.set_myers_cutoff(True)
    p.close(); st.close()


# ---------------------------------------------------------------------------------------- seed filter shares
@pytest.mark.parametrize("n,rl,shards", [(40_000, 32, 1), (700_000, 0, 3), (5_000_000, 0, 8), (5_000_000, 8192, 1)])
def test_seed_filter_equal_shares_cover_every_position(ctx, n, rl, shards):
    """every warp of the seed filter takes a contiguous share of chunk-columns that may start and end anywhere
    (mid super-tile, mid row group): filtered Myers over every shard = the direct kernel = the oracle."""
    rng = np.random.default_rng(n % 977 + shards)
    text = rand_text(rng, n)
    pats = []
    for i in range(24):
        s = int(rng.integers(0, n - 200))
        pats.append(mutate(rng, text[s:s + 100].copy(), int(rng.integers(0, 4)))[:128])
        if len(pats[-1]) < 97:
            pats[-1] = np.concatenate([pats[-1], text[s:s + 97 - len(pats[-1])]])
    store = sp.Store.from_host(ctx, text, rl)
    pf = sp.Pattern(ctx, sp.ALG_MYERS_FILTERED, pats, 3)
    assert pf.filter_info["filtered"]
    parts = []
    for s in range(shards):
        ctx.set_shard(s, shards)
        parts.append(sp.search(ctx, store, pf))
    ctx.set_shard(0, 1)
    got = np.concatenate(parts)
    got = got[np.lexsort((got["end"], got["pattern_id"]))]
    for i, q in enumerate(pats):
        _, we, wr = orc.search_mt(3, text, q, 3)
        sel = got[got["pattern_id"] == i]
        _same(sel["end"], we, f"pattern {i} ends")
        _same(sel["errors"], wr, f"pattern {i} errors")
    pf.close(); store.close()
