"""spmb_search_filtered_noalloc (include/spmb.h): the filtered Myers search with its scratch (candidate list, counters,
duplicate-suppression table) in the library's file-scope device array instead of buffers of a spmb_hits object.
No counterpart in the reference; the checker is the existing entry point, spmb_search, on the same store and set:
same (pattern_id, end, errors) records, bit-exact, on calls in a row that share the array.

Collected last on purpose: it must pass alone in a fresh process and after the
other GPU tests."""
import ctypes as C

import numpy as np
import pytest

import libspm_b200 as sp
from tests.util import mutate, rand_text

pytestmark = pytest.mark.gpu

MAX_CANDIDATES = 131072      # SPMB_NOALLOC_MAX_CANDIDATES
MAX_HITS = 1 << 24           # SPMB_NOALLOC_MAX_HITS


@pytest.fixture(scope="module")
def ctx():
    c = sp.Context(0)
    yield c
    c.close()


class DeviceOut:
    """the caller's outputs: device memory the TEST allocates (torch), poisoned so stale contents would show."""

    def __init__(self, capacity):
        import torch
        self.torch = torch
        self.capacity = capacity
        self.hits = torch.full((max(capacity, 1) * sp.HIT_DTYPE.itemsize,), 0xAB, dtype=torch.uint8, device="cuda:0")
        self.count = torch.full((1,), -1, dtype=torch.int64, device="cuda:0")
        torch.cuda.synchronize()

    def call(self, ctx, store, pattern, capacity=None):
        n = C.c_uint64(0xDEAD)
        rc = sp.lib().spmb_search_filtered_noalloc(ctx._h, store._h, pattern._h, C.c_void_p(self.hits.data_ptr()),
                                                   self.capacity if capacity is None else capacity,
                                                   C.c_void_p(self.count.data_ptr()), C.byref(n))
        return rc, int(n.value), (sp.lib().spmb_last_error(ctx._h) or b"").decode()

    def records(self, n):
        assert int(self.count.cpu()[0]) == n, "device counter and host count agree"
        raw = self.hits[: n * sp.HIT_DTYPE.itemsize].cpu().numpy()
        return np.frombuffer(raw.tobytes(), dtype=sp.HIT_DTYPE, count=n)


def normalise(h):
    """what spmb_search's ordering does with the raw records: sort by (pattern_id, end), smallest error count of a key."""
    assert np.all(h["seq_id"] == 0) and np.all(h["reserved"] == 0)
    order = np.lexsort((h["errors"], h["end"], h["pattern_id"]))
    h = h[order]
    if len(h) == 0:
        return h
    first = np.ones(len(h), dtype=bool)
    first[1:] = (h["pattern_id"][1:] != h["pattern_id"][:-1]) | (h["end"][1:] != h["end"][:-1])
    return h[first]


def same(got, want, what):
    assert len(want) > 0, what
    for f in ("pattern_id", "end", "errors"):
        assert np.array_equal(got[f], want[f]), f"{what}: field {f} differs ({len(got)} records, want {len(want)})"


def workload(seed, n, n_pat, m, k, copies):
    rng = np.random.default_rng(seed)
    text = rand_text(rng, n)
    pats = []
    for _ in range(n_pat):
        s = int(rng.integers(0, n - m))
        pats.append(text[s:s + m].copy())
    for pat in pats:   # mutated copies: hits with 0..k errors, some beyond k
        for s in rng.integers(0, n - 2 * m, size=copies):
            mp = mutate(rng, pat, int(rng.integers(0, k + 2)))
            text[int(s):int(s) + len(mp)] = mp
    return text, pats


def test_noalloc_equals_spmb_search_on_calls_in_a_row(ctx):
    text, pats = workload(11, 6_000_000, 48, 100, 3, 12)
    store = sp.Store.from_host(ctx, text)
    p = sp.Pattern(ctx, sp.ALG_MYERS_FILTERED, pats, 3)
    assert p.filter_info["step"] != 0, "the seed filter is in use"
    out = DeviceOut(1 << 16)
    # a store nobody searched yet has no lane-chunk copy: refused, nothing allocated behind the caller's back
    rc, _, msg = out.call(ctx, store, p)
    assert rc == 1 and "lane-chunk" in msg
    want = sp.search(ctx, store, p)          # the existing entry point (also derives the lane-chunk copy: set-up)
    assert len(want) > 48 * 6
    for call in range(2):                    # two calls in a row share the scratch array
        rc, n, msg = out.call(ctx, store, p)
        assert rc == 0, msg
        assert ctx.last_launch_count == 3    # filter, verification, minimum write-back
        same(normalise(out.records(n)), want, f"call {call}")
    # another set and another store right behind it: nothing of the last call leaks through the array
    text2, pats2 = workload(12, 3_000_000, 20, 64, 2, 20)
    store2 = sp.Store.from_host(ctx, text2, row_len=4096)
    p2 = sp.Pattern(ctx, sp.ALG_MYERS_FILTERED, pats2, 2)
    want2 = sp.search(ctx, store2, p2)
    rc, n, msg = out.call(ctx, store2, p2)
    assert rc == 0, msg
    same(normalise(out.records(n)), want2, "second workload")
    rc, n, msg = out.call(ctx, store, p)
    assert rc == 0, msg
    same(normalise(out.records(n)), want, "first workload again")
    # shards partition the list like they do for spmb_search
    parts = []
    for sh in range(3):
        ctx.set_shard(sh, 3)
        rc, n, msg = out.call(ctx, store, p)
        assert rc == 0, msg
        parts.append(out.records(n).copy())
    ctx.set_shard(0, 1)
    same(normalise(np.concatenate(parts)), want, "three shards")
    p2.close(); store2.close(); p.close(); store.close()


def test_noalloc_limits_fail_loudly_and_leave_the_array_usable(ctx):
    text, pats = workload(13, 2_000_000, 16, 80, 2, 10)
    store = sp.Store.from_host(ctx, text)
    p = sp.Pattern(ctx, sp.ALG_MYERS_FILTERED, pats, 2)
    want = sp.search(ctx, store, p)
    out = DeviceOut(1 << 15)
    # sizes beyond the stated maximum: refused before anything runs
    rc, _, msg = out.call(ctx, store, p, capacity=MAX_HITS + 1)
    assert rc == 4 and "SPMB_NOALLOC_MAX_HITS" in msg
    # other algorithms / sets without a usable filter
    pm = sp.Pattern(ctx, sp.ALG_MYERS, pats, 2)
    rc, _, msg = out.call(ctx, store, pm)
    assert rc == 4 and "SPMB_ALG_MYERS_FILTERED" in msg
    pm.close()
    # a hit buffer that is too small: the count says how many records there are
    rc, n, msg = out.call(ctx, store, p, capacity=8)
    assert rc == 3 and n >= len(want) and "caller's buffer" in msg
    # more filter survivors than the array holds: poly-A text, poly-A needles
    flat = np.zeros(1_000_000, dtype=np.uint8)
    fstore = sp.Store.from_host(ctx, flat)
    fp = sp.Pattern(ctx, sp.ALG_MYERS_FILTERED, [np.zeros(60, dtype=np.uint8)], 1)
    assert fp.filter_info["step"] != 0
    sp.search(ctx, fstore, fp)
    rc, _, msg = out.call(ctx, fstore, fp)
    assert rc == 3 and "filter candidates" in msg and str(MAX_CANDIDATES) in msg
    fp.close(); fstore.close()
    # the array is as good as new afterwards
    rc, n, msg = out.call(ctx, store, p)
    assert rc == 0, msg
    same(normalise(out.records(n)), want, "after the refused calls")
    p.close(); store.close()
