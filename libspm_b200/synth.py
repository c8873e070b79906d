"""Synthetic workloads of BASELINE.json's configs, produced by the ONE generator of this
repository: libspmsynth.so (libspm_b200/csrc/synth.cpp, include/spmb_synth.h) -- xoshiro256** seeded via
splitmix64, counter-based per (seed, stream, block), so the oracle side and the GPU harness, any rank and any
machine get the same bytes.  Harness helpers for tests, tools and bench.py; no search logic here and nothing
from oracle/ is imported."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import EVENT_DTYPE

SEEDS = {"C1": 0x5EED0001, "C2": 0x5EED0002, "C3": 0x5EED0003, "C4": 0x5EED0004, "C5": 0x5EED0005}
STREAM_USER = 16

_SO = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", "libspmsynth.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            raise ImportError(f"{_SO} not found: build it with `make`")
        L = C.CDLL(_SO)
        u64, u32, vp, dbl = C.c_uint64, C.c_uint32, C.c_void_p, C.c_double
        L.spms_u64.restype = u64
        L.spms_u64.argtypes = [u64, u64, u64]
        L.spms_uniform.restype = None
        L.spms_uniform.argtypes = [u64, u64, u64, u64, u64, u64, vp]
        L.spms_text.restype = None
        L.spms_text.argtypes = [u64, u64, u64, vp]
        L.spms_mutate.restype = u32
        L.spms_mutate.argtypes = [u64, u64, vp, u32, u32, vp]
        L.spms_journal.restype = u64
        L.spms_journal.argtypes = [u64, vp, u64, dbl, dbl, dbl, u32, vp, u64, vp, u64, C.POINTER(u64), vp]
        L.spms_jst.restype = u64
        L.spms_jst.argtypes = [u64, vp, u64, u64, u32, dbl, dbl, u32, dbl, vp, vp, u64, C.POINTER(u64), vp, vp]
        L.spms_set_threads.restype = None
        L.spms_set_threads.argtypes = [C.c_int]
        L.spms_set_threads(len(os.sched_getaffinity(0)))     # launchers export OMP_NUM_THREADS=1
        _lib = L
    return _lib


# ------------------------------------------------------------------------------------------ primitives
def text_into(out: np.ndarray, seed: int, begin: int = 0):
    """fill `out` (uint8, contiguous) with text positions [begin, begin + len(out)) of stream `seed`."""
    assert out.dtype == np.uint8 and out.flags.c_contiguous
    lib().spms_text(seed, begin, out.size, out.ctypes.data)
    return out


def text_host(n: int, seed: int, begin: int = 0) -> np.ndarray:
    return text_into(np.empty(n, dtype=np.uint8), seed, begin)


def text_device(n: int, seed: int, device):
    """the same text in device memory (torch is plumbing here: the bytes come from the host generator)."""
    import torch
    out = torch.empty(n, dtype=torch.uint8, device=device)
    step = 1 << 28
    for o in range(0, n, step):
        e = min(n, o + step)
        out[o:e] = torch.from_numpy(text_host(e - o, seed, o)).to(device)
    return out


def uniform(seed: int, stream: int, n: int, lo: int, hi: int, index0: int = 0) -> np.ndarray:
    """n stateless draws in [lo, hi)."""
    out = np.empty(n, dtype=np.uint64)
    lib().spms_uniform(seed, stream, index0, n, lo, hi, out.ctypes.data)
    return out.astype(np.int64)


def mutate_seeded(seed: int, stream: int, pat: np.ndarray, e: int) -> np.ndarray:
    """e random edits (sub / ins / del equiprobable) drawn from (seed, stream)."""
    p = np.ascontiguousarray(pat, dtype=np.uint8)
    out = np.empty(p.size + e + 1, dtype=np.uint8)
    m = lib().spms_mutate(seed, stream, p.ctypes.data, p.size, e, out.ctypes.data)
    return out[:m].copy()


def mutate(rng, pat: np.ndarray, e: int) -> np.ndarray:
    """ad-hoc edits for tests and tools that carry their own numpy rng: the stream id is drawn from it."""
    return mutate_seeded(int(rng.integers(0, 1 << 62)), STREAM_USER, pat, e)


def journal(ref: np.ndarray, seed: int, density: float = 0.01, p_snp: float = 0.8, p_ins: float = 0.1,
            max_indel: int = 10):
    """sorted, non-overlapping events at ~density per reference symbol: 80% SNP (alt != ref), 10% insertion
    U[1,max], 10% deletion U[1,max].  Returns (events[EVENT_DTYPE], pool, kinds)."""
    ref = np.ascontiguousarray(ref, dtype=np.uint8)
    L = lib()
    pl = C.c_uint64(0)
    n_ev = L.spms_journal(seed, ref.ctypes.data, ref.size, density, p_snp, p_ins, max_indel, None, 0, None, 0, C.byref(pl), None)
    ev = np.zeros(n_ev, dtype=EVENT_DTYPE)
    pool = np.zeros(max(pl.value, 1), dtype=np.uint8)
    kinds = np.zeros(n_ev, dtype=np.uint8)
    n2 = L.spms_journal(seed, ref.ctypes.data, ref.size, density, p_snp, p_ins, max_indel, ev.ctypes.data, n_ev,
                        pool.ctypes.data, pool.size, C.byref(pl), kinds.ctypes.data)
    assert n2 == n_ev
    return ev, pool[:pl.value], kinds


def jst(ref: np.ndarray, n_var: int, n_hap: int, seed: int, p_snp: float = 0.9, p_ins: float = 0.05,
        max_indel: int = 10, af_max: float = 0.5):
    """C4-style journaled string tree: variants at sorted, non-overlapping positions (90% SNP, 5% insertion,
    5% deletion U[1,max]); allele frequency ~ truncated 1/x on [1/n_hap, af_max]; membership Bernoulli(AF) per
    haplotype (no LD).  Returns (variants[EVENT_DTYPE], pool, kinds, coverage[n_var, ceil(n_hap/64)])."""
    ref = np.ascontiguousarray(ref, dtype=np.uint8)
    L = lib()
    wpr = (n_hap + 63) // 64
    var = np.zeros(n_var, dtype=EVENT_DTYPE)
    kinds = np.zeros(n_var, dtype=np.uint8)
    pool = np.zeros(n_var * max_indel + 1, dtype=np.uint8)
    cov = np.zeros((n_var, wpr), dtype=np.uint64)
    pl = C.c_uint64(0)
    n = L.spms_jst(seed, ref.ctypes.data, ref.size, n_var, n_hap, p_snp, p_ins, max_indel, af_max, var.ctypes.data,
                   pool.ctypes.data, pool.size, C.byref(pl), kinds.ctypes.data, cov.ctypes.data)
    return var[:n].copy(), pool[:pl.value].copy(), kinds[:n].copy(), cov[:n].copy()


def apply_local(ref: np.ndarray, ev: np.ndarray, pool: np.ndarray, lo: int, hi: int, carried=None) -> np.ndarray:
    """the derived symbols of reference window [lo, hi) under the events that start inside it (optionally only
    those with carried[i] true): a few hundred symbols, plain Python -- used to cut needles."""
    i0 = int(np.searchsorted(ev["ref_pos"], lo))
    out = []
    p = lo
    i = i0
    while i < len(ev) and int(ev["ref_pos"][i]) < hi:
        if carried is None or carried(i):
            q = int(ev["ref_pos"][i])
            if q >= p:
                out.append(ref[p:q])
                o, il = int(ev["ins_off"][i]), int(ev["ins_len"][i])
                out.append(pool[o:o + il])
                p = q + int(ev["del_len"][i])
        i += 1
    if p < hi:
        out.append(ref[p:hi])
    return np.concatenate(out).astype(np.uint8) if out else np.zeros(0, np.uint8)


# ------------------------------------------------------------------------------------------ configs
def c1(n: int = 1_000_000):
    """configs[0]: Horspool m=32 over a 1 Mbp text; pattern = text[333333..+32), 99 more copies planted at seeded
    non-overlapping offsets, and one 64-bp run of period 4 (ACGT x 16) searched with ACGT x 8 (9 overlapping hits)."""
    seed = SEEDS["C1"]
    text = text_host(n, seed)
    p0 = min(333_333, max(0, n // 3))
    pat = text[p0:p0 + 32].copy()
    slots = 101
    sl = n // slots
    offs = uniform(seed, STREAM_USER, slots, 0, max(1, sl - 96))
    planted = []
    for j in range(slots - 1):
        s = j * sl + int(offs[j])
        if len(planted) == 99 or abs(s - p0) < 64:
            continue
        text[s:s + 32] = pat
        planted.append(s)
    run_at = (slots - 1) * sl + int(offs[slots - 1])
    text[run_at:run_at + 64] = np.tile(np.array([0, 1, 2, 3], dtype=np.uint8), 16)
    return {"text": text, "pattern": pat, "pattern_overlap": np.tile(np.array([0, 1, 2, 3], dtype=np.uint8), 8),
            "planted": sorted(planted + [p0]), "run_at": run_at, "seed": seed}


def c2(n: int = 50_000_000, seed: int | None = None, m: int = 64, n_plant: int = 100, ref: np.ndarray | None = None):
    """configs[1]: reference + 1% journal (80/10/10 SNP/ins/del); needle of m symbols cut from the DERIVED
    sequence; n_plant more copies planted into the derived sequence, at least 10 % of them straddling an edit
    (the reference and, for SNPs / insertions, the event's inserted symbols are rewritten so that the journal
    produces the needle there).  `ref` may be a preallocated (pinned) buffer of n bytes to fill."""
    seed = SEEDS["C2"] if seed is None else seed
    ref = text_host(n, seed) if ref is None else text_into(ref, seed)
    ev, pool, kinds = journal(ref, seed)
    pos = ev["ref_pos"].astype(np.int64)
    ne = len(ev)
    # the needle: derived symbols around a seeded event
    e0 = int(uniform(seed, STREAM_USER + 1, 1, ne // 3, ne // 3 + max(1, ne // 8))[0]) if ne else 0
    lo = int(pos[e0]) - 20 if ne else n // 3
    pat = apply_local(ref, ev, pool, lo, lo + m + 64)[:m].copy()
    keep_lo, keep_hi = lo - 2 * m, lo + 3 * m + 64
    # planted copies: one per slot, around a seeded event of the slot whose neighbours leave room
    picks = uniform(seed, STREAM_USER + 2, n_plant, 0, 1 << 40)
    offs = uniform(seed, STREAM_USER + 3, n_plant, 8, m - 24)
    planted, straddling = 0, 0
    room = m + 16
    for j in range(n_plant):
        if ne < 3 * n_plant:
            break
        a, b = j * ne // n_plant, (j + 1) * ne // n_plant
        i = a + int(picks[j]) % max(1, b - a)
        tries = 0
        while tries < 200:
            i = a + (i - a + 1) % max(1, b - a)
            tries += 1
            if i < 1 or i + 1 >= ne:
                continue
            q, dl, il = int(pos[i]), int(ev["del_len"][i]), int(ev["ins_len"][i])
            if q - (int(pos[i - 1]) + int(ev["del_len"][i - 1])) < room or int(pos[i + 1]) - (q + dl) < room:
                continue
            if keep_lo < q < keep_hi:
                continue
            break
        else:
            continue
        want_straddle = (j % 4 == 0)            # 25 % of the copies cross their event
        off = int(offs[j])
        if want_straddle:
            k = int(kinds[i])
            p = q - off
            if k == 0:                           # SNP: alt = needle symbol, reference symbol differs
                ref[p:p + m] = pat
                ref[q] = (int(pat[off]) + 1) & 3
                pool[int(ev["ins_off"][i])] = pat[off]
            elif k == 1:                         # insertion: the inserted symbols are part of the needle
                ref[p:q] = pat[:off]
                pool[int(ev["ins_off"][i]):int(ev["ins_off"][i]) + il] = pat[off:off + il]
                ref[q:q + m - off - il] = pat[off + il:]
            else:                                # deletion: the needle continues behind the deleted symbols
                ref[p:q] = pat[:off]
                ref[q + dl:q + dl + m - off] = pat[off:]
            straddling += 1
        else:                                    # clean copy right behind the event
            p = q + dl + 4
            ref[p:p + m] = pat
        planted += 1
    return {"ref": ref, "events": ev, "pool": pool, "kinds": kinds, "pattern": pat, "planted": planted,
            "straddling": straddling, "seed": seed}


def c3(n: int = 250_000_000, seed: int | None = None, m: int = 100, k: int = 4, n_plant: int = 1000, text: np.ndarray | None = None):
    """configs[2]: pattern of m symbols cut from the text; n_plant copies, each with e in U{0..6} random edits
    (0..4 hit at k=4, 5..6 are decoys), one per slot."""
    seed = SEEDS["C3"] if seed is None else seed
    text = text_host(n, seed) if text is None else text_into(text, seed)
    s0 = int(uniform(seed, STREAM_USER, 1, 0, n - m)[0])
    pat = text[s0:s0 + m].copy()
    sl = n // n_plant
    if sl >= 4 * m:
        offs = uniform(seed, STREAM_USER + 1, n_plant, 0, sl - 3 * m)
        edits = uniform(seed, STREAM_USER + 2, n_plant, 0, 7)
        for j in range(n_plant):
            s = j * sl + int(offs[j])
            if abs(s - s0) < 3 * m:
                continue
            mp = mutate_seeded(seed, STREAM_USER + 1000 + j, pat, int(edits[j]))
            text[s:s + len(mp)] = mp
    return {"text": text, "pattern": pat, "k": k, "source": s0, "seed": seed}


def c4(scale: float = 1.0, n_hap: int = 2504, m: int = 150, k: int = 2):
    """configs[3]: 50 Mbp reference, 1 M variants (90/5/5), 2504 haplotypes, needle of 150 symbols cut from a
    seeded haplotype around a seeded variant it carries, k=2."""
    seed = SEEDS["C4"]
    n_ref, n_var = int(50_000_000 * scale), int(1_000_000 * scale)
    ref = text_host(n_ref, seed)
    var, pool, kinds, cov = jst(ref, n_var, n_hap, seed)
    nv = len(var)
    hap = int(uniform(seed, STREAM_USER, 1, 0, n_hap)[0])
    v0 = int(uniform(seed, STREAM_USER + 1, 1, nv // 4, nv // 2)[0])
    while not (int(cov[v0, hap >> 6]) >> (hap & 63)) & 1:
        v0 += 1
    lo = int(var["ref_pos"][v0]) - m // 2
    seg = apply_local(ref, var, pool, lo, lo + 2 * m + 64, carried=lambda i: (int(cov[i, hap >> 6]) >> (hap & 63)) & 1)
    pat = seg[:m].copy()
    return {"ref": ref, "variants": var, "pool": pool, "kinds": kinds, "coverage": cov, "n_hap": n_hap,
            "pattern": pat, "k": k, "source_haplotype": hap, "seed": seed}


def c5_patterns(text_at, n: int, n_pat: int = 10_000, m: int = 100, seed: int | None = None):
    """configs[4] needles: n_pat needles of m symbols cut at seeded sorted offsets and mutated with e in U{0..5}
    edits (kept at 97..128 symbols: one Myers word count per set).  text_at(offsets) -> [n_pat, m] ranks."""
    seed = SEEDS["C5"] if seed is None else seed
    offs = np.sort(uniform(seed, STREAM_USER, n_pat, 0, n - 2 * m))
    cut = text_at(offs, m)
    edits = uniform(seed, STREAM_USER + 1, n_pat, 0, 6)
    pats = []
    for i in range(n_pat):
        q = mutate_seeded(seed, STREAM_USER + 1000 + i, cut[i], int(edits[i]))
        if len(q) < 97:
            q = np.concatenate([q, cut[i][:97 - len(q)]])
        pats.append(q[:128])
    return offs, pats
