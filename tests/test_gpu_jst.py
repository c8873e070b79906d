"""JST search (host flattening into branch-context windows + one reference scan) against the
oracle that materialises every haplotype and searches it."""
import numpy as np
import pytest

import libspm_b200 as sp
import oracle as orc
from libspm_b200 import synth
from tests.util import mutate, rand_text

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = sp.Context(0)
    yield c
    c.close()


def make_jst(n_ref, n_var, n_hap, seed, **kw):
    rng = np.random.default_rng(seed)
    ref = rand_text(rng, n_ref)
    var, pool, kinds, cov = synth.jst(ref, n_var, n_hap, seed, **kw)
    return ref, var, pool, cov


def compare(got, want, tag):
    g = np.stack([got["seq_id"].astype(np.int64), got["end"].astype(np.int64), got["errors"].astype(np.int64)], 1)
    w = np.stack([want["hap"].astype(np.int64), want["end"].astype(np.int64), want["err"].astype(np.int64)], 1)
    if g.shape == w.shape and np.array_equal(g, w):
        return
    gs, ws = set(map(tuple, g.tolist())), set(map(tuple, w.tolist()))
    raise AssertionError(f"{tag}: got {len(g)} want {len(w)}; missing {sorted(ws - gs)[:8]} extra {sorted(gs - ws)[:8]} "
                         f"dup={len(g) - len(gs)}")


@pytest.mark.parametrize("n_hap,n_var,seed", [(70, 600, 1), (64, 300, 2), (5, 900, 3), (130, 1500, 4)])
def test_jst_vs_per_haplotype_oracle(ctx, n_hap, n_var, seed):
    n_ref = 20_000
    ref, var, pool, cov = make_jst(n_ref, n_var, n_hap, seed, p_snp=0.6, p_ins=0.2, max_indel=8, af_max=0.7)
    rng = np.random.default_rng(seed + 100)
    store = sp.Store.from_host(ctx, ref)
    jst = sp.Jst(ctx, store, var, pool, cov, n_hap)
    # haplotype materialisation sanity (oracle side): lengths differ by the carried indels
    h0 = orc.jst_haplotype(ref, var.astype(orc.EVENT_DTYPE), pool, cov, 0)
    assert abs(len(h0) - n_ref) < 10 * n_var
    cases = []
    for (alg, m, k) in [(sp.ALG_SHIFTOR, 12, 0), (sp.ALG_HORSPOOL, 33, 0), (sp.ALG_MYERS, 24, 2), (sp.ALG_MYERS, 40, 3),
                        (sp.ALG_MYERS, 70, 5)]:
        # needles cut from haplotypes (so that they cross variants) and from the reference
        for src_h in (0, n_hap - 1, None):
            seq = ref if src_h is None else orc.jst_haplotype(ref, var.astype(orc.EVENT_DTYPE), pool, cov, src_h)
            s = int(rng.integers(0, len(seq) - m))
            pat = seq[s:s + m].copy()
            if alg == sp.ALG_MYERS:
                pat = mutate(rng, pat, 1)
            cases.append((alg, pat, k))
    for alg, pat, k in cases:
        p = sp.Pattern(ctx, alg, pat, k)
        got = jst.search(p)
        want = orc.jst_search(alg, ref, var.astype(orc.EVENT_DTYPE), pool, cov, n_hap, pat, k)
        compare(got, want, f"alg={alg} m={len(pat)} k={k} haps={n_hap}")
        assert len(want) >= 1
    st = jst.last_stats()
    assert st["windows"] > 0 and st["buckets"] >= 1
    jst.close()


def test_jst_edges(ctx):
    # variants at the very start / end, adjacent indels, a haplotype without variants, no variants at all
    rng = np.random.default_rng(9)
    n_ref, n_hap = 3_000, 10
    ref = rand_text(rng, n_ref)
    pool = rand_text(rng, 64)
    var = np.array([(0, 0, 3, 0), (0, 1, 1, 3), (5, 2, 0, 0), (7, 0, 4, 4), (40, 1, 1, 8), (41, 1, 1, 9),
                    (1500, 6, 2, 10), (1506, 0, 5, 12), (2990, 3, 0, 0), (2999, 1, 1, 20)], dtype=sp.EVENT_DTYPE)
    cov = np.zeros((len(var), 1), dtype=np.uint64)
    sets = {0: [0, 1, 2, 3], 1: [0], 2: [1, 4, 5], 3: [2, 3, 6, 7], 4: [8, 9], 5: [9], 6: [4], 7: [5, 6], 8: [0, 1, 2, 3, 4, 5, 6, 7, 8, 9]}
    for h, vs in sets.items():          # haplotype 9 carries nothing
        for v in vs:
            cov[v, 0] |= np.uint64(1 << h)
    pool[3] = (ref[0] + 1) % 4
    pool[8] = (ref[40] + 1) % 4
    pool[9] = (ref[41] + 2) % 4
    pool[20] = (ref[2999] + 1) % 4
    store = sp.Store.from_host(ctx, ref)
    jst = sp.Jst(ctx, store, var, pool, cov, n_hap)
    ev = var.astype(orc.EVENT_DTYPE)
    for h in (0, 3, 8):
        seq = orc.jst_haplotype(ref, ev, pool, cov, h)
        for (alg, m, k, s) in [(sp.ALG_SHIFTOR, 9, 0, 0), (sp.ALG_MYERS, 16, 2, 0), (sp.ALG_MYERS, 16, 2, len(seq) - 16),
                               (sp.ALG_HORSPOOL, 30, 0, 1490), (sp.ALG_MYERS, 35, 4, 1480), (sp.ALG_SHIFTOR, 5, 0, len(seq) - 5)]:
            pat = seq[s:s + m].copy()
            got = jst.search(sp.Pattern(ctx, alg, pat, k))
            want = orc.jst_search(alg, ref, ev, pool, cov, n_hap, pat, k)
            compare(got, want, f"edges h={h} alg={alg} m={m} k={k} s={s}")
    jst.close()
    # no variants: every reference hit belongs to every haplotype
    j0 = sp.Jst(ctx, store, np.zeros(0, sp.EVENT_DTYPE), np.zeros(0, np.uint8), np.zeros((0, 1), np.uint64), 3)
    pat = ref[100:120].copy()
    got = j0.search(sp.Pattern(ctx, sp.ALG_HORSPOOL, pat))
    assert len(got) == 3 * len(orc.exact_bruteforce(ref, pat))
    assert sorted(set(got["seq_id"].tolist())) == [0, 1, 2]


@pytest.mark.parametrize("n_hap,n_var,seed", [(70, 600, 11), (130, 1500, 12)])
def test_jst_filtered_myers_equals_oracle(ctx, n_hap, n_var, seed):
    """the pigeonhole-filtered Myers path over JST branch windows (windows-mode filter + verification)."""
    n_ref = 20_000
    ref, var, pool, cov = make_jst(n_ref, n_var, n_hap, seed, p_snp=0.6, p_ins=0.2, max_indel=8, af_max=0.7)
    rng = np.random.default_rng(seed + 100)
    store = sp.Store.from_host(ctx, ref)
    jst = sp.Jst(ctx, store, var, pool, cov, n_hap)
    ev = var.astype(orc.EVENT_DTYPE)
    for (m, k) in [(40, 2), (70, 4), (150, 2)]:
        for src_h in (0, n_hap - 1, None):
            seq = ref if src_h is None else orc.jst_haplotype(ref, ev, pool, cov, src_h)
            s = int(rng.integers(0, len(seq) - m))
            pat = mutate(rng, seq[s:s + m].copy(), 1)
            p = sp.Pattern(ctx, sp.ALG_MYERS_FILTERED, [pat], k)
            assert p.filter_info["filtered"]
            got = jst.search(p)
            want = orc.jst_search(sp.ALG_MYERS, ref, ev, pool, cov, n_hap, pat, k)
            compare(got, want, f"filtered m={len(pat)} k={k} haps={n_hap}")
            assert len(want) >= 1
    jst.close()


def test_jst_small_buckets_beside_large_ones(ctx):
    # enough variants that the common window lengths fill the device (main stream) while the few long windows
    # make small buckets (second stream, jst_enqueue): same hits as the serial order and as the oracle
    import os
    n_ref, n_var, n_hap = 8_000_000, 700_000, 16
    ref, var, pool, cov = make_jst(n_ref, n_var, n_hap, 77, p_snp=0.85, p_ins=0.08, max_indel=60, af_max=0.6)
    store = sp.Store.from_host(ctx, ref)
    jst = sp.Jst(ctx, store, var, pool, cov, n_hap)
    rng = np.random.default_rng(78)
    hap = orc.jst_haplotype(ref, var.astype(orc.EVENT_DTYPE), pool, cov, 3)
    for alg, m, k in [(sp.ALG_MYERS, 40, 2), (sp.ALG_SHIFTOR, 20, 0), (sp.ALG_MYERS, 90, 3)]:
        s = int(rng.integers(0, len(hap) - m))
        pat = hap[s:s + m].copy()
        p = sp.Pattern(ctx, alg, pat, k)
        got = jst.search(p)
        os.environ["SPMB_JST_SERIAL"] = "1"
        try:
            serial = jst.search(p)
        finally:
            del os.environ["SPMB_JST_SERIAL"]
        for f in ("seq_id", "end", "errors"):
            assert np.array_equal(got[f], serial[f]), f
        want = orc.jst_search(alg, ref, var.astype(orc.EVENT_DTYPE), pool, cov, n_hap, pat, k)
        compare(got, want, f"alg={alg} m={m} k={k}")
        assert len(want) >= 1
    st = jst.last_stats()
    assert st["buckets"] >= 2 and st["windows"] > 2 * 16 * 32 * 148   # at least one bucket past 4 CTAs per SM
    jst.close()


@pytest.mark.gpu
def test_jst_create_rejects_overlapping_variants_of_one_haplotype(ctx):
    """a haplotype that carries a deletion [100, 110) and an indel at 105 is not a valid journal; the
    flattening would skip the second variant while the coordinate shifts still count it.  The oracle rejects the
    input (journal_apply: unsorted / overlapping), so does spmb_jst_create -- per haplotype, so the same two
    variants on DIFFERENT haplotypes are fine."""
    rng = np.random.default_rng(11)
    ref = rand_text(rng, 4000)
    var = np.zeros(3, dtype=sp.EVENT_DTYPE)
    var["ref_pos"] = [100, 105, 300]
    var["del_len"] = [10, 0, 1]
    var["ins_len"] = [0, 3, 1]
    var["ins_off"] = [0, 0, 3]
    pool = np.array([1, 2, 3, 0], dtype=np.uint8)
    store = sp.Store.from_host(ctx, ref, 64)
    ok_cov = np.array([[0b01], [0b10], [0b11]], dtype=np.uint64)
    j = sp.Jst(ctx, store, var, pool, ok_cov, 2)
    j.close()
    bad_cov = np.array([[0b01], [0b01], [0b11]], dtype=np.uint64)
    with pytest.raises(sp.SpmbError) as ei:
        sp.Jst(ctx, store, var, pool, bad_cov, 2)
    assert ei.value.code == 1
    with pytest.raises(ValueError):
        orc.journal_apply(ref, var[:2].astype(orc.EVENT_DTYPE), pool)
    # tie order: a replacement before an insertion at the same position is an overlap, the other order is not
    var2 = np.zeros(2, dtype=sp.EVENT_DTYPE)
    var2["ref_pos"] = [50, 50]
    var2["del_len"] = [1, 0]
    var2["ins_len"] = [1, 2]
    var2["ins_off"] = [0, 1]
    both = np.array([[1], [1]], dtype=np.uint64)
    with pytest.raises(sp.SpmbError):
        sp.Jst(ctx, store, var2, pool, both, 1)
    var2["del_len"] = [0, 1]
    var2["ins_len"] = [2, 1]
    var2["ins_off"] = [1, 0]
    sp.Jst(ctx, store, var2, pool, both, 1).close()
    store.close()


def test_jst_launch_finish_in_rotation_equals_the_one_call_search(ctx):
    """spmb_jst_search_launch / _finish: three searches queued on three hit lists before the first is collected, with
    different needles and algorithms (window sizes: the flattened tree is rebuilt between them), sharded and not;
    every list equals the one spmb_jst_search returns and the oracle's."""
    n_ref, n_hap, n_var, seed = 30_000, 70, 900, 31
    ref, var, pool, cov = make_jst(n_ref, n_var, n_hap, seed, p_snp=0.6, p_ins=0.2, max_indel=8, af_max=0.7)
    ev = var.astype(orc.EVENT_DTYPE)
    rng = np.random.default_rng(seed)
    store = sp.Store.from_host(ctx, ref)
    jst = sp.Jst(ctx, store, var, pool, cov, n_hap)
    pats = []
    for alg, m, k, src in [(sp.ALG_MYERS, 40, 3, 0), (sp.ALG_MYERS, 40, 3, n_hap - 1), (sp.ALG_SHIFTOR, 14, 0, 5),
                           (sp.ALG_MYERS, 70, 4, 9), (sp.ALG_MYERS, 40, 3, 20)]:
        seq = orc.jst_haplotype(ref, ev, pool, cov, src)
        s = int(rng.integers(0, len(seq) - m))
        pat = seq[s:s + m].copy()
        if alg == sp.ALG_MYERS:
            pat[m // 2] = (pat[m // 2] + 1) % 4     # one substitution: the length (and with it the window size) stays
        pats.append((alg, pat, k, sp.Pattern(ctx, alg, pat, k)))
    want = [jst.search(p) for (_, _, _, p) in pats]
    for (alg, pat, k, _), w in zip(pats, want):
        compare(w, orc.jst_search(alg, ref, ev, pool, cov, n_hap, pat, k), "one-call search")
        assert len(w) >= 1
    hl = [sp.Hits(ctx, 1 << 10) for _ in range(3)]
    # same window size back to back: three in flight
    same_w = [0, 1, 4, 0, 1, 4, 0]
    got = [None] * len(same_w)
    for i in range(2):
        jst.search_launch(pats[same_w[i]][3], hl[i % 3])
    for i in range(len(same_w)):
        if i + 2 < len(same_w):
            jst.search_launch(pats[same_w[i + 2]][3], hl[(i + 2) % 3])
        got[i] = jst.search_finish(hl[i % 3])
    for i, pi in enumerate(same_w):
        assert np.array_equal(got[i], want[pi]), (i, pi, len(got[i]), len(want[pi]))
    # another window size while a search is in flight is refused (it would rebuild the window stores in use)
    jst.search_launch(pats[0][3], hl[0])
    assert sp.lib().spmb_jst_search_launch(ctx._h, jst._h, pats[3][3]._h, hl[1]._h) == 1
    assert "window size" in (sp.lib().spmb_last_error(ctx._h) or b"").decode()
    assert np.array_equal(jst.search_finish(hl[0]), want[0])
    # window sizes change between searches: one in flight at a time (a launch may rebuild the window stores)
    for pi in (2, 3, 0, 2):
        jst.search_launch(pats[pi][3], hl[0])
        g = jst.search_finish(hl[0])
        assert np.array_equal(g, want[pi]), (pi, len(g), len(want[pi]))
    # shards partition the list
    parts = []
    for sh in range(3):
        ctx.set_shard(sh, 3)
        jst.search_launch(pats[0][3], hl[1])
        parts.append(jst.search_finish(hl[1]))
    ctx.set_shard(0, 1)
    merged = np.concatenate(parts)
    merged = merged[np.lexsort((merged["end"], merged["seq_id"], merged["pattern_id"]))]
    assert np.array_equal(merged, want[0])
    # a finish without a launch is refused
    assert sp.lib().spmb_jst_search_finish(ctx._h, hl[2]._h) == 1
    for h in hl:
        h.close()
    for (_, _, _, p) in pats:
        p.close()
    jst.close(); store.close()


def test_jst_window_groups_dealt_into_stripes(ctx, monkeypatch):
    """inside a length bucket the 32-window groups are dealt into stripes so that the shards of a bucket hold short and
    long windows alike (jst_prepare); with 2 stripes a small tree takes that path too.  Hit lists: unchanged, sharded or not."""
    n_ref, n_hap, n_var, seed = 40_000, 90, 2500, 41
    ref, var, pool, cov = make_jst(n_ref, n_var, n_hap, seed, p_snp=0.6, p_ins=0.2, max_indel=8, af_max=0.7)
    ev = var.astype(orc.EVENT_DTYPE)
    rng = np.random.default_rng(seed)
    store = sp.Store.from_host(ctx, ref)
    seq = orc.jst_haplotype(ref, ev, pool, cov, 7)
    cases = []
    for alg, m, k in [(sp.ALG_MYERS, 45, 3), (sp.ALG_SHIFTOR, 13, 0)]:
        s0 = int(rng.integers(0, len(seq) - m))
        pat = seq[s0:s0 + m].copy()
        cases.append((alg, pat, k, orc.jst_search(alg, ref, ev, pool, cov, n_hap, pat, k)))
    for stripes in ("2", "3", "64"):
        monkeypatch.setenv("SPMB_JST_STRIPES", stripes)
        jst = sp.Jst(ctx, store, var, pool, cov, n_hap)
        for alg, pat, k, want in cases:
            p = sp.Pattern(ctx, alg, pat, k)
            got = jst.search(p)
            assert jst.last_stats()["windows"] > 64 * 3
            compare(got, want, f"stripes={stripes} alg={alg}")
            parts = []
            for sh in range(5):
                ctx.set_shard(sh, 5)
                parts.append(jst.search(p))
            ctx.set_shard(0, 1)
            merged = np.concatenate(parts)
            merged = merged[np.lexsort((merged["end"], merged["seq_id"], merged["pattern_id"]))]
            assert np.array_equal(merged, got), f"stripes={stripes}: shards partition the list"
            p.close()
        jst.close()
    store.close()
