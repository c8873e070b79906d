"""Sharding on the device: the shards of one store partition its hit list (emulated on one GPU by
searching the shards one after the other, like ranks would on their own GPUs)."""
import numpy as np
import pytest

import libspm_b200 as sp
import oracle as orc
from libspm_b200 import sharded, synth
from tests.util import mutate, rand_text

pytestmark = pytest.mark.gpu


@pytest.fixture()
def ctx():
    c = sp.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("n_shards", [2, 3, 8])
def test_shards_partition_the_hit_list(ctx, n_shards):
    rng = np.random.default_rng(n_shards)
    n = 400_000
    text = rand_text(rng, n)
    pat = text[1000:1100].copy()
    for pos in rng.integers(0, n - 200, size=40):
        mp_ = mutate(rng, pat, int(rng.integers(0, 4)))
        text[int(pos):int(pos) + len(mp_)] = mp_
    store = sp.Store.from_host(ctx, text, 256)
    for alg, k in ((sp.ALG_SHIFTOR, 0), (sp.ALG_MYERS, 3)):
        p = sp.Pattern(ctx, alg, pat, k)
        ctx.set_shard(0, 1)
        full = sp.search(ctx, store, p)
        parts = []
        for s in range(n_shards):
            ctx.set_shard(s, n_shards)
            parts.append(sp.search(ctx, store, p))
        ctx.set_shard(0, 1)
        merged = sharded.sort_hits(np.concatenate(parts))
        assert len(merged) == len(full) and np.array_equal(merged["end"], full["end"])
        assert np.array_equal(merged["errors"], full["errors"])
        assert sum(len(x) > 0 for x in parts) >= 2
        if alg == sp.ALG_MYERS:
            _, we, wr = orc.search_mt(3, text, pat, k)
            assert np.array_equal(full["end"], we) and np.array_equal(full["errors"], wr)


def test_sliced_text_ranks_match_unsharded(ctx):
    rng = np.random.default_rng(77)
    n, world = 250_000, 4
    text = rand_text(rng, n)
    pat = text[62_450:62_550].copy()      # straddles the seam between rank 0 and rank 1
    p = sp.Pattern(ctx, sp.ALG_MYERS, pat, 4)
    parts = []
    for r in range(world):
        start, stop, lo, hi = sharded.slice_for_rank(n, r, world, p.window_size)
        st = sp.Store.from_host(ctx, text[start:stop])
        parts.append(sharded.globalize(sp.search(ctx, st, p), start, lo))
        st.close()
    merged = sharded.sort_hits(np.concatenate(parts))
    we, wr = orc.myers(text, pat, 4)
    assert np.array_equal(merged["end"], we) and np.array_equal(merged["errors"], wr)


def test_jst_shards(ctx):
    rng = np.random.default_rng(5)
    n_ref, n_hap = 30_000, 40
    ref = rand_text(rng, n_ref)
    var, pool, kinds, cov = synth.jst(ref, 900, n_hap, 5, p_snp=0.7, p_ins=0.15, af_max=0.6)
    store = sp.Store.from_host(ctx, ref, 64)
    jst = sp.Jst(ctx, store, var, pool, cov, n_hap)
    hap = orc.jst_haplotype(ref, var.astype(orc.EVENT_DTYPE), pool, cov, 3)
    pat = hap[5000:5040].copy()
    p = sp.Pattern(ctx, sp.ALG_MYERS, pat, 2)
    parts = []
    for s in range(3):
        ctx.set_shard(s, 3)
        parts.append(jst.search(p))
    ctx.set_shard(0, 1)
    merged = sharded.sort_hits(np.concatenate(parts))
    want = orc.jst_search(3, ref, var.astype(orc.EVENT_DTYPE), pool, cov, n_hap, pat, 2)
    assert len(merged) == len(want)
    assert np.array_equal(merged["seq_id"], want["hap"]) and np.array_equal(merged["end"], want["end"])
    assert np.array_equal(merged["errors"], want["err"])
