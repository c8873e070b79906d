"""N>1 host logic on CPU: world_size-2 gloo processes shard a text (sliced, with window_size-1 of
left context), "search" their slice with the oracle (no GPU here), convert to global coordinates
and gather the hit lists -- the result must equal the unsharded search.  On the GPU box the same
driver code runs with the CUDA path (tests/test_gpu_sharded.py)."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import oracle as orc
    from libspm_b200 import HIT_DTYPE, sharded
    rng = np.random.default_rng(123)
    n = 300_000
    text = rng.integers(0, 4, size=n, dtype=np.uint8)
    pat = text[149_990:150_090].copy()          # straddles the shard seam at n/2
    text[10:110] = pat
    text[n - 100:] = pat
    out = {}
    for alg, k in ((2, 0), (3, 4)):
        ws = len(pat) + k
        start, stop, own_lo, own_hi = sharded.slice_for_rank(n, rank, world, ws)
        if alg == 3:
            e, r = orc.myers(text[start:stop], pat, k)
        else:
            e, r = orc.shiftor(text[start:stop], pat), None
        loc = np.zeros(len(e), dtype=HIT_DTYPE)
        loc["end"] = e
        if r is not None:
            loc["errors"] = r
        glob = sharded.globalize(loc, start, own_lo)
        allh = sharded.gather_hits(glob, dst=0)
        if rank == 0:
            if alg == 3:
                we, wr = orc.myers(text, pat, k)
            else:
                we, wr = orc.shiftor(text, pat), None
            ok = np.array_equal(allh["end"], we) and (wr is None or np.array_equal(allh["errors"], wr))
            out[alg] = (bool(ok), int(len(we)))
        else:
            assert allh is None
    lo, hi = sharded.pattern_slice(10, rank, world)
    out["pslice"] = (lo, hi)
    if rank == 0:
        q.put(out)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharding_and_gather():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = q.get(timeout=180)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert out[2][0] and out[2][1] >= 3
    assert out[3][0] and out[3][1] >= 3
    assert out["pslice"] == (0, 5)


def test_slice_math_covers_everything():
    from libspm_b200 import sharded
    for n in (0, 1, 99, 1000, 12345):
        for world in (1, 2, 3, 8):
            owned = []
            for r in range(world):
                start, stop, lo, hi = sharded.slice_for_rank(n, r, world, 17)
                assert start <= lo <= hi == stop <= n
                assert lo - start <= 16 and (start == 0 or lo - start == 16)
                owned.append((lo, hi))
            assert owned[0][0] == 0 and owned[-1][1] == n
            for a, b in zip(owned, owned[1:]):
                assert a[1] == b[0]


def _shm_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    os.environ["LOCAL_WORLD_SIZE"] = str(world)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import bench
    from libspm_b200 import HIT_DTYPE
    assert bench.HitGatherShm.usable(world)
    g = bench.make_gather(torch, dist, torch.device("cpu"), world, rank)
    assert isinstance(g, bench.HitGatherShm)
    passes = 4
    rng = np.random.default_rng(7)
    lists = [[None] * passes for _ in range(world)]   # every rank draws the same lists, writes only its own
    sizes = [[int(rng.integers(0, 90)) for _ in range(passes)] for _ in range(world)]
    ok = True
    for rnd in range(3):                              # three gathers in a row over the two generations of slots
        for r in range(world):
            for i in range(passes):
                h = np.zeros(sizes[r][i] if rnd != 1 else sizes[r][i] // 2, dtype=HIT_DTYPE)
                h["end"] = rng.integers(0, 1 << 40, size=len(h))
                h["pattern_id"] = r * 100 + i + rnd
                lists[r][i] = h
        if rnd == 0:
            g.prepare(max(max(s) for s in sizes), passes)
        g.barrier()                                   # what bench.py does in front of its timed region
        for i in range(passes):
            g.add(i, lists[rank][i])
        total = g.finish()
        ok &= total == sum(len(lists[r][i]) for r in range(world) for i in range(passes))
        for i in range(passes):
            want = np.concatenate([lists[r][i] for r in range(world)])
            ok &= bool(np.array_equal(g.merged(HIT_DTYPE, i), want))
        dist.barrier()
    q.put((rank, bool(ok), g.gathers))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_shared_memory_gather():
    """bench.py's final gather of the ranks of one host (HitGatherShm): every rank ends up with every rank's lists."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_shm_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted(q.get(timeout=180) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got == [(0, True, 4), (1, True, 4)]     # prepare's warm-up gather + three
