"""The single-process multi-GPU driver of the C ABI (spmb_multi_*; north_star item 4): one host
haystack cut into shards with left context, one context per shard, per-device hit lists merged on the host.
With one GPU the shards share the device (the driver accepts a device more than once), so the cutting, the
journal split and the merge are exercised on any box; with several GPUs the same tests use them all."""
import numpy as np
import pytest
import torch

import libspm_b200 as sp
import oracle as orc
from libspm_b200 import synth
from tests.util import mutate, rand_text

pytestmark = pytest.mark.gpu


def device_lists():
    n = torch.cuda.device_count() if torch.cuda.is_available() else 1
    out = [[0], [0, 0, 0]]
    if n >= 2:
        out.append(list(range(n)))
    return out


@pytest.mark.parametrize("devs", device_lists())
def test_multi_plain_exact_and_myers(devs):
    rng = np.random.default_rng(len(devs))
    n = 900_000
    text = rand_text(rng, n)
    pat = text[12_345:12_345 + 64].copy()
    m = sp.Multi(devs)
    try:
        store = m.store(text)
        assert len(store) == n and store.shards == min(len(devs), n >> 16)
        # copies across every cut and at both ends
        cuts = [n // len(devs) * i // 64 * 64 for i in range(1, len(devs))]
        for c in cuts:
            text[c - 40:c + 24] = pat
        text[:64] = pat
        text[n - 64:] = pat
        store.close()
        store = m.store(text)
        want = orc.exact_bruteforce(text, pat)
        for alg in (sp.ALG_HORSPOOL, sp.ALG_SHIFTOR):
            p = m.pattern(alg, pat)
            got = m.search(store, p)
            assert np.array_equal(got["end"], want), (alg, len(got), len(want))
            p.close()
        _, we, wr = orc.search_mt(3, text, pat, 3)
        p = m.pattern(sp.ALG_MYERS, pat, 3)
        got = m.search(store, p)
        assert np.array_equal(got["end"], we) and np.array_equal(got["errors"], wr)
        # the reference's call shape over all devices
        got = m.search_host(text, p)
        assert np.array_equal(got["end"], we) and np.array_equal(got["errors"], wr)
        st = m.last_stats()
        assert st[0] > 0 and st[2] > 0
        p.close(); store.close()
    finally:
        m.close()


@pytest.mark.parametrize("devs", device_lists())
def test_multi_pattern_set_merge_order(devs):
    rng = np.random.default_rng(7)
    n = 700_000
    text = rand_text(rng, n)
    pats = []
    for i in range(5):
        s0 = int(rng.integers(0, n - 200))
        q = mutate(rng, text[s0:s0 + 100].copy(), 2)[:128]
        if len(q) < 97:
            q = np.concatenate([q, text[s0:s0 + 97 - len(q)]])
        pats.append(q)
        for s1 in rng.integers(0, n - 200, size=6):
            text[int(s1):int(s1) + len(q)] = q
    m = sp.Multi(devs)
    try:
        store = m.store(text, overlap=1024)
        p = m.pattern(sp.ALG_MYERS, pats, 3)
        got = m.search(store, p)
        assert np.all(np.diff(got["pattern_id"].astype(np.int64)) >= 0)
        for i, q in enumerate(pats):
            _, we, wr = orc.search_mt(3, text, q, 3)
            sel = got[got["pattern_id"] == i]
            assert np.array_equal(sel["end"], we) and np.array_equal(sel["errors"], wr), i
        p.close()
        # a window longer than the shards' left context is refused, not silently wrong
        if store.shards > 1:
            big = m.pattern(sp.ALG_MYERS, rand_text(rng, 250), 12)
            tiny = m.store(text, overlap=64)
            with pytest.raises(sp.SpmbError) as ei:
                m.search(tiny, big)
            assert ei.value.code == 4
            tiny.close(); big.close()
        store.close()
    finally:
        m.close()


@pytest.mark.parametrize("devs", device_lists())
@pytest.mark.parametrize("density,max_indel", [(0.01, 10), (0.08, 3)])
def test_multi_journaled_store(devs, density, max_indel):
    """events are split at the shard cuts (slices of the caller's table, coordinates of the whole base); hits come
    back in the coordinates of the whole derived sequence."""
    n = 1_200_000
    ref = synth.text_host(n, 4242)
    ev, pool, kinds = synth.journal(ref, 4243, density=density, max_indel=max_indel)
    derived = orc.journal_apply(ref, ev.astype(orc.EVENT_DTYPE), pool)
    m = sp.Multi(devs)
    try:
        base = m.store(ref)
        js = base.journaled(ev, pool)
        assert len(js) == len(derived)
        rng = np.random.default_rng(1)
        # needles cut from the derived sequence: around every cut (in derived coordinates, roughly) and at random
        places = [int(len(derived) * i / len(devs)) for i in range(1, len(devs))] + [100, len(derived) - 200, 600_000]
        for c in places:
            for m_len, k in ((64, 0), (100, 3)):
                s0 = max(0, min(len(derived) - m_len, c - m_len // 2))
                pat = derived[s0:s0 + m_len].copy()
                if k:
                    _, we, wr = orc.search_mt(3, derived, pat, k)
                    p = m.pattern(sp.ALG_MYERS, pat, k)
                    got = m.search(js, p)
                    assert np.array_equal(got["end"], we) and np.array_equal(got["errors"], wr), (c, m_len)
                else:
                    want = orc.exact_bruteforce(derived, pat)
                    p = m.pattern(sp.ALG_SHIFTOR, pat)
                    got = m.search(js, p)
                    assert np.array_equal(got["end"], want), (c, m_len, got["end"][:5], want[:5])
                    assert len(want) >= 1
                p.close()
        js.close(); base.close()
    finally:
        m.close()


def test_multi_journal_event_across_a_cut():
    """a deletion that covers a shard cut moves the cut behind it; a deletion longer than the overlap is refused."""
    rng = np.random.default_rng(3)
    n = 600_000
    ref = rand_text(rng, n)
    m = sp.Multi([0, 0, 0])
    try:
        base = m.store(ref, overlap=2048)
        cut1 = n // 3 // 64 * 64
        ev = np.zeros(3, dtype=sp.EVENT_DTYPE)
        ev["ref_pos"] = [1000, cut1 - 30, 500_000]
        ev["del_len"] = [5, 100, 0]
        ev["ins_len"] = [0, 2, 7]
        ev["ins_off"] = [0, 0, 2]
        pool = rng.integers(0, 4, size=9, dtype=np.uint8)
        derived = orc.journal_apply(ref, ev.astype(orc.EVENT_DTYPE), pool)
        js = base.journaled(ev, pool)
        assert len(js) == len(derived)
        d_cut = cut1 - 30 - 5                       # derived position of the big event
        for s0 in (d_cut - 50, d_cut - 1, d_cut + 2, d_cut + 40):
            pat = derived[s0:s0 + 80].copy()
            p = m.pattern(sp.ALG_SHIFTOR, pat)
            got = m.search(js, p)
            assert np.array_equal(got["end"], orc.exact_bruteforce(derived, pat)), s0
            p.close()
        js.close()
        ev["del_len"][1] = 5000                     # more than the overlap: not representable in the shards
        with pytest.raises(sp.SpmbError) as ei:
            base.journaled(ev, pool)
        assert ei.value.code == 4
        base.close()
    finally:
        m.close()
