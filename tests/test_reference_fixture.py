"""Known-answer test from the reference's OWN data (test/data/datasources.cmake:72-102): the
10 kb simulated reference + mason VCF must reproduce the 100 shipped haplotypes.  Pins the
journal / JST semantics of a14 (oracle on CPU; device store and JST search on the GPU).
Fixtures: tests/golden/*.npz, made by tests/golden/make_jst_fixture.py."""
import hashlib
import os

import numpy as np
import pytest

import oracle as orc

HERE = os.path.dirname(os.path.abspath(__file__))
TAGS = ("SNPs", "SNP_INDELs")


def load(tag):
    z = np.load(os.path.join(HERE, "golden", f"sim_ref_10Kb_{tag}.npz"))
    return z, int(z["n_hap"])


def carried(var, cov, h):
    return var[(cov[:, h // 64] >> np.uint64(h % 64)) & np.uint64(1) == 1]


@pytest.mark.parametrize("tag", TAGS)
def test_oracle_reproduces_shipped_haplotypes(tag):
    z, n_hap = load(tag)
    assert n_hap == 100
    ev = z["var"].astype(orc.EVENT_DTYPE)
    for h in range(n_hap):
        seq = orc.jst_haplotype(z["ref"], ev, z["pool"], z["cov"], h)
        assert len(seq) == z["hap_len"][h]
        assert hashlib.sha256(seq.tobytes()).hexdigest() == z["hap_sha"][h], (tag, h)
    assert np.array_equal(orc.jst_haplotype(z["ref"], ev, z["pool"], z["cov"], 37), z["hap37"])
    if tag == "SNP_INDELs":
        assert z["hap_len"].min() == 9401 and z["hap_len"].max() == 10930


@pytest.mark.gpu
@pytest.mark.parametrize("tag", TAGS)
def test_device_journal_reproduces_shipped_haplotypes(tag):
    import libspm_b200 as sp
    z, n_hap = load(tag)
    ctx = sp.Context(0)
    base = sp.Store.from_host(ctx, z["ref"])
    for h in range(n_hap):
        js = base.journaled(carried(z["var"], z["cov"], h), z["pool"])
        seq = js.read()
        assert hashlib.sha256(seq.tobytes()).hexdigest() == z["hap_sha"][h], (tag, h)
        js.close()


@pytest.mark.gpu
@pytest.mark.parametrize("tag", TAGS)
def test_device_jst_search_on_reference_fixture(tag):
    import libspm_b200 as sp
    z, n_hap = load(tag)
    ctx = sp.Context(0)
    base = sp.Store.from_host(ctx, z["ref"])
    jst = sp.Jst(ctx, base, z["var"], z["pool"], z["cov"], n_hap)
    ev = z["var"].astype(orc.EVENT_DTYPE)
    rng = np.random.default_rng(1)
    for name in ("hap0", "hap37", "hap99"):
        hap = z[name]
        for (alg, m, k) in ((sp.ALG_HORSPOOL, 32, 0), (sp.ALG_SHIFTOR, 64, 0), (sp.ALG_MYERS, 100, 4), (sp.ALG_MYERS, 150, 2)):
            s = int(rng.integers(0, len(hap) - m))
            pat = hap[s:s + m].copy()
            got = jst.search(sp.Pattern(ctx, alg, pat, k))
            want = orc.jst_search(alg, z["ref"], ev, z["pool"], z["cov"], n_hap, pat, k)
            assert len(want) >= 1
            assert len(got) == len(want), (tag, name, alg, m, len(got), len(want))
            assert np.array_equal(got["seq_id"], want["hap"]) and np.array_equal(got["end"], want["end"])
            assert np.array_equal(got["errors"], want["err"])
