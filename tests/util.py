"""shared helpers for the tests: the reference's golden fixture and seeded inputs."""
import numpy as np

_RANK = {"A": 0, "C": 1, "G": 2, "T": 3}


def dna4(s: str) -> np.ndarray:
    return np.array([_RANK[c] for c in s], dtype=np.uint8)


# the one fixture every reference matcher test shares
# (/root/reference/test/api/libspm/matcher/horspool_matcher_test.cpp:19-31)
GOLD_HAYSTACK = dna4("ACGTGACTAGCACGTGACTAGCACGTGACTAGCACGTGACTAGC")
GOLD_NEEDLE = dna4("GCACG")
GOLD_EXACT_BEGIN = [9, 20, 31]                       # horspool/shiftor tests :26
GOLD_MYERS_K1_END = [13, 14, 15, 24, 25, 26, 35, 36, 37]   # myers_matcher_test.cpp:27


def rand_text(rng, n, sigma=4):
    return rng.integers(0, sigma, size=n, dtype=np.uint8)


def mutate(rng, pat, e):
    """apply e random edits (sub/ins/del equiprobable)."""
    p = list(pat)
    for _ in range(e):
        if not p:
            break
        op = rng.integers(0, 3)
        i = int(rng.integers(0, len(p)))
        if op == 0:
            p[i] = (p[i] + 1 + int(rng.integers(0, 3))) % 4
        elif op == 1:
            p.insert(i, int(rng.integers(0, 4)))
        else:
            del p[i]
    return np.array(p, dtype=np.uint8)
