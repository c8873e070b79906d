"""The oracle on alphabets wider than dna4 (spm::dna5 / spm::dna15, libspm/libspm/seqan/alphabet.hpp:119-120):
the restated searchers against the definitional ones (all-occurrence compare, Sellers DP)."""
import numpy as np
import pytest

import oracle as orc


@pytest.mark.parametrize("sigma", [5, 15, 16])
def test_exact_oracles_agree_on_wide_alphabets(sigma):
    rng = np.random.default_rng(sigma)
    text = rng.integers(0, sigma, 30000, dtype=np.uint8)
    for m in (1, 3, 9, 32, 33, 80):
        pat = text[777:777 + m].copy()
        text[20000:20000 + m] = pat
        want = orc.exact_bruteforce(text, pat)
        assert len(want) >= 2
        assert np.array_equal(orc.horspool(text, pat), want)
        assert np.array_equal(orc.shiftor(text, pat), want)


@pytest.mark.parametrize("sigma", [5, 15])
@pytest.mark.parametrize("m,k", [(10, 2), (64, 4), (65, 3), (130, 6)])
def test_myers_oracle_agrees_with_sellers_on_wide_alphabets(sigma, m, k):
    rng = np.random.default_rng(100 * sigma + m)
    text = rng.integers(0, sigma, 12000, dtype=np.uint8)
    pat = rng.integers(0, sigma, m, dtype=np.uint8)
    for s in (100, 5000, 9000):
        q = pat.copy()
        q[m // 2] = (q[m // 2] + 1) % sigma
        text[s:s + m] = q
    we, wr = orc.sellers(text, pat, k)
    ge, gr = orc.myers(text, pat, k)
    assert len(we) >= 3
    assert np.array_equal(ge, we) and np.array_equal(gr, wr)


def test_symbols_with_equal_low_bits_are_distinct():
    text = np.array([1, 5, 9, 13, 1, 5], dtype=np.uint8)
    assert orc.shiftor(text, np.array([5], dtype=np.uint8)).tolist() == [2, 6]
    assert orc.horspool(text, np.array([1, 5], dtype=np.uint8)).tolist() == [2, 6]
