"""The host-haystack upload pipeline (libspm_b200/csrc/upload.cuh + hostpack.cpp): the host packing loop
against numpy on the CPU; on the GPU, stores built through the chunk queue (raw chunks over the copy
engine + host-packed chunks) must read back symbol for symbol and give the oracle's hit lists."""
import os

import numpy as np
import pytest

import libspm_b200 as sp
import oracle as orc
from tests.util import mutate, rand_text


def _pack_numpy(a):
    n = a.size
    b = np.zeros((n + 31) // 32 * 32, dtype=np.uint64)
    b[:n] = a & 3
    b = b.reshape(-1, 32)
    x = np.zeros(len(b), dtype=np.uint64)
    y = np.zeros(len(b), dtype=np.uint64)
    for j in range(16):
        x |= b[:, j] << np.uint64(30 - 2 * j)
        y |= b[:, 16 + j] << np.uint64(30 - 2 * j)
    return x | (y << np.uint64(32))


@pytest.mark.parametrize("n", [0, 1, 31, 32, 33, 95, 127, 128, 129, 160, 1000, 4097, 100_003])
def test_host_pack_chunks_matches_numpy(n):
    rng = np.random.default_rng(n)
    a = rng.integers(0, 256, size=n, dtype=np.uint8)       # values are masked to 2 bits
    assert np.array_equal(sp.host_pack_chunks(a), _pack_numpy(a))


def test_host_pack_chunks_unaligned_source():
    rng = np.random.default_rng(5)
    a = rng.integers(0, 4, size=5000, dtype=np.uint8)
    for off in (1, 3, 17):
        assert np.array_equal(sp.host_pack_chunks(a[off:]), _pack_numpy(a[off:]))


# ------------------------------------------------------------------------------------------
@pytest.fixture
def small_chunks(monkeypatch):
    """force the chunked path on test-sized inputs (the knobs are read when a context first uploads)"""
    monkeypatch.setenv("SPMB_UPLOAD_MIN", "1")
    monkeypatch.setenv("SPMB_UPLOAD_CHUNK", "65536")


@pytest.mark.gpu
@pytest.mark.parametrize("threads", [0, 1, 4])
@pytest.mark.parametrize("n,rl", [(1, 0), (65536, 0), (65537, 64), (1_000_003, 0), (3_000_017, 1056)])
def test_chunked_upload_round_trip(small_chunks, monkeypatch, threads, n, rl):
    monkeypatch.setenv("SPMB_HOST_THREADS", str(threads))
    ctx = sp.Context(0)
    try:
        rng = np.random.default_rng(n + threads)
        text = rand_text(rng, n)
        store = sp.Store.from_host(ctx, text, row_len=rl)
        st = ctx.upload_last_stats()
        assert st["chunk_symbols"] == 65536 and st["host_threads"] == threads
        assert st["raw_chunks"] + st["host_packed_chunks"] == (n + 65535) // 65536
        if threads == 0:
            assert st["host_packed_chunks"] == 0
        assert np.array_equal(store.read(), text)
        store.close()
    finally:
        ctx.close()


@pytest.mark.gpu
def test_chunked_upload_search_parity(small_chunks, monkeypatch):
    monkeypatch.setenv("SPMB_HOST_THREADS", "3")
    ctx = sp.Context(0)
    try:
        rng = np.random.default_rng(77)
        text = rand_text(rng, 2_500_000)
        pat = text[1_234_567:1_234_567 + 64].copy()
        for pos in (100, 65536 - 30, 131072 - 64, 2_400_000):     # copies across chunk seams
            text[pos:pos + 64] = pat
        pm = text[700_000:700_100].copy()
        mm = mutate(rng, pm, 3)
        text[65536 * 7 - 50:65536 * 7 - 50 + len(mm)] = mm
        store = sp.Store.from_host(ctx, text)
        want = orc.search_mt(2, text, pat)[1]
        got = sp.search(ctx, store, sp.Pattern(ctx, sp.ALG_SHIFTOR, pat))
        assert np.array_equal(got["end"], want) and len(want) >= 5
        _, we, wr = orc.search_mt(3, text, pm, 4)
        got = sp.search(ctx, store, sp.Pattern(ctx, sp.ALG_MYERS, pm, 4))
        assert np.array_equal(got["end"], we) and np.array_equal(got["errors"], wr)
        # several uploads through one pool, and the same result from the single-copy path
        again = sp.Store.from_host(ctx, text[:1_000_000])
        assert np.array_equal(again.read(), text[:1_000_000])
    finally:
        ctx.close()
    monkeypatch.setenv("SPMB_UPLOAD_MIN", str(1 << 40))
    ctx2 = sp.Context(0)
    try:
        s2 = sp.Store.from_host(ctx2, text)
        assert ctx2.upload_last_stats()["chunk_symbols"] == 0      # pool never created
        got2 = sp.search(ctx2, s2, sp.Pattern(ctx2, sp.ALG_SHIFTOR, pat))
        assert np.array_equal(got2["end"], want)
    finally:
        ctx2.close()


@pytest.mark.gpu
def test_chunked_upload_default_geometry_large():
    """default knobs (4 Mi-symbol chunks, all host threads) on a 100 Mbp haystack"""
    for k in ("SPMB_UPLOAD_MIN", "SPMB_UPLOAD_CHUNK", "SPMB_HOST_THREADS"):
        os.environ.pop(k, None)
    ctx = sp.Context(0)
    try:
        rng = np.random.default_rng(9)
        text = rand_text(rng, 100_000_001)
        store = sp.Store.from_host(ctx, text)
        st = ctx.upload_last_stats()
        assert st["raw_chunks"] + st["host_packed_chunks"] == 24
        for lo in (0, 4 * (1 << 20) - 1000, 50_000_000, 100_000_001 - 70_000):
            assert np.array_equal(store.read(lo, 70_000), text[lo:lo + 70_000])
        pat = text[99_999_000:99_999_064].copy()
        got = sp.search(ctx, store, sp.Pattern(ctx, sp.ALG_HORSPOOL, pat))
        assert np.array_equal(got["end"], orc.search_mt(1, text, pat)[1])
    finally:
        ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("threads", [0, 3])
def test_event_table_through_upload_queue(small_chunks, monkeypatch, threads):
    """journal events over the chunk queue (raw 24-byte events + chunks rewritten as 8-byte records by
    host threads, expanded on the device): the derived sequence equals the oracle's journal_apply,
    including chunks whose records do not fit the short form."""
    from libspm_b200 import synth
    monkeypatch.setenv("SPMB_HOST_THREADS", str(threads))
    ctx = sp.Context(0)
    try:
        n = 2_000_000
        ref = synth.text_host(n, 31)
        ev, pool, kinds = synth.journal(ref, 32)
        ev = ev.copy()
        # records that do not fit 8 bytes: a 300-symbol deletion, a 260-symbol insertion stored at the END of
        # the pool (offset delta >= 2^16), and a pool offset that runs backwards
        big = np.flatnonzero(kinds == 2)[len(ev) // 40]
        nxt = int(ev["ref_pos"][big + 1])
        ev["del_len"][big] = 300
        keep = (ev["ref_pos"] < ev["ref_pos"][big]) | (ev["ref_pos"] >= ev["ref_pos"][big] + 300) | (np.arange(len(ev)) == big)
        ev = ev[keep]
        kinds = kinds[keep]
        ins = np.flatnonzero(kinds == 1)[len(ev) // 30]
        tail = np.random.default_rng(3).integers(0, 4, size=260, dtype=np.uint8)
        ev["ins_len"][ins] = 260
        ev["ins_off"][ins] = len(pool)
        pool = np.concatenate([pool, tail])
        back = np.flatnonzero(kinds == 0)[len(ev) // 2]
        ev["ins_off"][back] = ev["ins_off"][np.flatnonzero(kinds == 0)[0]]
        assert nxt > 0 and len(ev) * 24 > 65536
        derived = orc.journal_apply(ref, ev.astype(orc.EVENT_DTYPE), pool)
        base = sp.Store.from_host(ctx, ref)
        js = base.journaled(ev, pool)
        st = ctx.upload_last_stats()
        n_chunks = (len(ev) + st["events_per_chunk"] - 1) // st["events_per_chunk"]
        assert st["events_per_chunk"] == 1024
        assert st["event_raw_chunks"] + st["event_host_packed_chunks"] == n_chunks
        if threads == 0:
            assert st["event_host_packed_chunks"] == 0
        assert len(js) == len(derived)
        assert np.array_equal(js.read(), derived)
        pat = derived[len(derived) // 2:len(derived) // 2 + 64].copy()
        got = sp.search(ctx, js, sp.Pattern(ctx, sp.ALG_SHIFTOR, pat))
        assert np.array_equal(got["end"], orc.search_mt(2, derived, pat)[1])
        # an unsorted table is still rejected after the trip through the queue
        bad = ev.copy()
        bad["ref_pos"][100] = bad["ref_pos"][5000]
        with pytest.raises(sp.SpmbError):
            base.journaled(bad, pool)
    finally:
        ctx.close()


# ---------------- the host loop of the event-table upload (no device needed) ----------------
def _events(rng, n, *, big_gap_at=None):
    ev = np.zeros(n, dtype=sp.EVENT_DTYPE)
    kind = rng.integers(0, 10, n)
    gaps = rng.integers(20, 200, n).astype(np.uint64)
    if big_gap_at is not None:
        gaps[big_gap_at] = 1 << 33
    dl = np.where(kind < 8, 1, np.where(kind == 8, 0, rng.integers(1, 11, n))).astype(np.uint32)
    il = np.where(kind < 8, 1, np.where(kind == 8, rng.integers(1, 11, n), 0)).astype(np.uint32)
    ev["ref_pos"] = np.cumsum(gaps + np.concatenate([[0], dl[:-1]]).astype(np.uint64))
    ev["del_len"], ev["ins_len"] = dl, il
    off = np.concatenate([[0], np.cumsum(il)[:-1]]).astype(np.uint64)
    ev["ins_off"] = np.where(il > 0, off, 0)
    return ev, int(il.sum())


def _pack_events_numpy(ev):
    n = len(ev)
    n_blk = (n + 1023) // 1024
    hdr = np.zeros((n_blk, 2), dtype=np.uint64)
    rec = np.zeros(n, dtype=np.uint64)
    for b in range(n_blk):
        blk = ev[b * 1024:(b + 1) * 1024]
        ins = np.nonzero(blk["ins_len"])[0]
        base_pos, base_off = blk["ref_pos"][0], (blk["ins_off"][ins[0]] if len(ins) else 0)
        hdr[b] = (base_pos, base_off)
        dof = np.where(blk["ins_len"] > 0, blk["ins_off"] - np.uint64(base_off), 0).astype(np.uint64)
        rec[b * 1024:(b + 1) * 1024] = ((blk["ref_pos"] - base_pos) | (dof << np.uint64(32)) |
                                        (blk["del_len"].astype(np.uint64) << np.uint64(48)) |
                                        (blk["ins_len"].astype(np.uint64) << np.uint64(56)))
    return hdr, rec


@pytest.mark.parametrize("n", [1, 3, 4, 5, 1023, 1024, 1025, 4099, 70001])
def test_host_pack_events_matches_numpy(n):
    rng = np.random.default_rng(n)
    ev, pool_len = _events(rng, n)
    got = sp.host_pack_events(ev, pool_len)
    assert got is not None
    hdr, rec = _pack_events_numpy(ev)
    assert np.array_equal(got[0], hdr) and np.array_equal(got[1], rec)


def test_host_pack_events_rejects_what_does_not_fit():
    rng = np.random.default_rng(5)
    ev, pool_len = _events(rng, 5000)
    assert sp.host_pack_events(ev, pool_len) is not None
    for field, idx, val in (("del_len", 4001, 256), ("ins_len", 77, 300), ("del_len", 3, 1 << 20)):
        bad = ev.copy()
        bad[field][idx] = val
        assert sp.host_pack_events(bad, pool_len + 1000) is None, (field, idx)
    bad, pl = _events(rng, 5000, big_gap_at=2500)          # position delta >= 2^32 inside a block
    assert sp.host_pack_events(bad, pl) is None
    bad = ev.copy()                                          # an event that inserts nothing but points past the pool
    i = int(np.nonzero(bad["ins_len"] == 0)[0][10])
    bad["ins_off"][i] = pool_len + 1
    assert sp.host_pack_events(bad, pool_len) is None
    bad["ins_off"][i] = (1 << 63) + 5
    assert sp.host_pack_events(bad, pool_len) is None
    bad = ev.copy()                                          # pool offsets 2^16 apart inside one block
    j = np.nonzero(bad["ins_len"])[0]
    bad["ins_off"][j[50]:] += 1 << 16
    assert sp.host_pack_events(bad, pool_len + (1 << 17)) is None
