"""Synthetic workloads of BASELINE.json's configs: deterministic, seeded, iid
uniform ACGT ranks.  Host-side helpers for tests and bench.py; no search logic here."""
from __future__ import annotations

import numpy as np

from . import EVENT_DTYPE

SEEDS = {"C1": 0x5EED0001, "C2": 0x5EED0002, "C3": 0x5EED0003, "C4": 0x5EED0004, "C5": 0x5EED0005}


def text_host(n: int, seed: int) -> np.ndarray:
    return np.random.default_rng(seed).integers(0, 4, size=n, dtype=np.uint8)


def text_device(n: int, seed: int, device):
    """iid ranks generated on the GPU (torch is plumbing here: device memory + RNG)."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    out = torch.empty(n, dtype=torch.uint8, device=device)
    step = 1 << 28
    for o in range(0, n, step):
        e = min(n, o + step)
        out[o:e] = torch.randint(0, 4, (e - o,), dtype=torch.uint8, device=device, generator=g)
    return out


def mutate(rng, pat: np.ndarray, e: int) -> np.ndarray:
    p = list(pat)
    for _ in range(e):
        if not p:
            break
        op = int(rng.integers(0, 3))
        i = int(rng.integers(0, len(p)))
        if op == 0:
            p[i] = (p[i] + 1 + int(rng.integers(0, 3))) % 4
        elif op == 1:
            p.insert(i, int(rng.integers(0, 4)))
        else:
            del p[i]
    return np.array(p, dtype=np.uint8)


def journal(n_ref: int, seed: int, density: float = 0.01, p_snp: float = 0.8, p_ins: float = 0.1,
            max_indel: int = 10):
    """sorted, non-overlapping events at ~density per reference symbol:
    80% SNP (alt != ref is enforced by the caller via `snp_alts`), 10% insertion U[1,max], 10%
    deletion U[1,max].  Returns (events[EVENT_DTYPE], pool, kinds); SNP alt symbols in the pool are
    offsets 1..3 to be added to the reference symbol (see finish_snps)."""
    rng = np.random.default_rng(seed)
    mean_gap = 1.0 / density
    n_ev = int(n_ref * density * 1.02) + 16
    lo = max_indel + 2
    gaps = rng.integers(lo, int(2 * mean_gap) - lo + 1, size=n_ev, dtype=np.int64)
    pos = np.cumsum(gaps)
    pos = pos[pos < n_ref - 2 * max_indel - 2]
    n_ev = len(pos)
    u = rng.random(n_ev)
    kinds = np.where(u < p_snp, 0, np.where(u < p_snp + p_ins, 1, 2)).astype(np.uint8)
    lens = rng.integers(1, max_indel + 1, size=n_ev, dtype=np.uint32)
    ev = np.zeros(n_ev, dtype=EVENT_DTYPE)
    ev["ref_pos"] = pos
    ev["del_len"] = np.where(kinds == 0, 1, np.where(kinds == 2, lens, 0))
    ev["ins_len"] = np.where(kinds == 0, 1, np.where(kinds == 1, lens, 0))
    ins_len = ev["ins_len"].astype(np.int64)
    off = np.cumsum(ins_len) - ins_len
    ev["ins_off"] = off
    pool = rng.integers(0, 4, size=int(ins_len.sum()), dtype=np.uint8)
    return ev, pool, kinds


def finish_snps(ev, pool, kinds, ref_at_snp: np.ndarray):
    """make every SNP's alt differ from the reference symbol: alt = (ref + 1 + pool%3) % 4."""
    snp = kinds == 0
    o = ev["ins_off"][snp].astype(np.int64)
    pool[o] = (ref_at_snp.astype(np.int64) + 1 + (pool[o] % 3)) % 4
    return pool


def jst(n_ref: int, n_var: int, n_hap: int, seed: int, p_snp: float = 0.9, p_ins: float = 0.05,
        max_indel: int = 10, af_max: float = 0.5, chunk: int = 8192):
    """C4-style journaled string tree: n_var variants at sorted, non-overlapping
    seeded positions (90% SNP, 5% insertion U[1,max], 5% deletion U[1,max]); allele frequency ~
    truncated 1/x on [1/n_hap, af_max]; membership Bernoulli(AF) per haplotype (no LD).
    Returns (variants[EVENT_DTYPE], pool, kinds, coverage[n_var, ceil(n_hap/64)] uint64).
    SNP alts are offsets to be fixed with finish_snps()."""
    rng = np.random.default_rng(seed)
    mean_gap = n_ref / (n_var + 1)
    lo = max_indel + 2
    hi = max(lo + 1, int(2 * mean_gap) - lo)
    gaps = rng.integers(lo, hi + 1, size=n_var, dtype=np.int64)
    pos = np.cumsum(gaps)
    pos = pos[pos < n_ref - 2 * max_indel - 2]
    n_var = len(pos)
    u = rng.random(n_var)
    kinds = np.where(u < p_snp, 0, np.where(u < p_snp + p_ins, 1, 2)).astype(np.uint8)
    lens = rng.integers(1, max_indel + 1, size=n_var, dtype=np.uint32)
    var = np.zeros(n_var, dtype=EVENT_DTYPE)
    var["ref_pos"] = pos
    var["del_len"] = np.where(kinds == 0, 1, np.where(kinds == 2, lens, 0))
    var["ins_len"] = np.where(kinds == 0, 1, np.where(kinds == 1, lens, 0))
    ins_len = var["ins_len"].astype(np.int64)
    var["ins_off"] = np.cumsum(ins_len) - ins_len
    pool = rng.integers(0, 4, size=int(ins_len.sum()), dtype=np.uint8)
    af = np.exp(rng.uniform(np.log(1.0 / n_hap), np.log(af_max), size=n_var))
    wpr = (n_hap + 63) // 64
    cov = np.zeros((n_var, wpr), dtype=np.uint64)
    thr = np.minimum((af * 65536.0).astype(np.int64), 65535).astype(np.uint16)
    for o in range(0, n_var, chunk):
        e = min(n_var, o + chunk)
        bits = rng.integers(0, 65536, size=(e - o, wpr * 64), dtype=np.uint16) < thr[o:e, None]
        bits[:, n_hap:] = False
        # every variant is carried by at least one haplotype
        none = ~bits.any(axis=1)
        if none.any():
            bits[np.nonzero(none)[0], rng.integers(0, n_hap, size=int(none.sum()))] = True
        cov[o:e] = np.packbits(bits, axis=1, bitorder="little").view(np.uint64)
    return var, pool, kinds, cov
