"""FASTA / VCF ingest (spmb_fasta_read / spmb_vcf_read, host-side, no device needed) against the
reference's own test data: the arrays parsed here must equal the committed fixtures made from the same
files by tests/golden/make_jst_fixture.py, which the oracle turns into the 100 shipped haplotypes
(tests/test_reference_fixture.py).  /root/reference exists only in the build container; the synthetic
cases below run everywhere."""
import gzip
import os

import numpy as np
import pytest

import libspm_b200 as sp
import oracle as orc

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DATA = "/root/reference/test/data/"


def to_str(r):
    return "".join("ACGT"[x] for x in r)


def test_fasta_plain_and_gzip(tmp_path):
    recs = [("chr1", "ACGTNNACGT" * 13), ("two words here", "acgu"), ("empty", ""), ("last", "T" * 201)]
    text = ""
    for name, seq in recs:
        text += f">{name}\n"
        for o in range(0, len(seq), 60):
            text += seq[o:o + 60] + "\r\n"
    p1 = tmp_path / "a.fa"
    p1.write_text(text)
    p2 = tmp_path / "a.fa.gz"
    with gzip.open(p2, "wt") as f:
        f.write(text)
    for p in (p1, p2):
        got, unknown = sp.read_fasta(str(p))
        assert [g[0] for g in got] == ["chr1", "two", "empty", "last"]
        assert to_str(got[0][1]) == ("ACGTAAACGT" * 13)          # N -> A like seqan3::dna4
        assert to_str(got[1][1]) == "ACGT" and len(got[2][1]) == 0 and to_str(got[3][1]) == "T" * 201
        assert unknown == 26
    with pytest.raises(sp.SpmbError):
        sp.read_fasta(str(tmp_path / "missing.fa"))


def test_vcf_events_and_coverage(tmp_path):
    vcf = "\n".join([
        "##fileformat=VCFv4.1",
        "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\tFORMAT\ts0\ts1\ts2",
        "1\t3\t.\tG\tA,T\t.\tPASS\t.\tGT\t1|0\t2|2\t0|1",          # SNPs at 0-based 2
        "1\t5\t.\tC\tCGGA\t.\tPASS\t.\tGT:DP\t0|1:7\t0|0:3\t.|.:0",  # insertion of GGA after position 4
        "1\t5\t.\tCTT\tC\t.\tPASS\t.\tGT\t1|0\t0|0\t0/0",            # deletion of 2 at 0-based 5
        "1\t9\t.\tA\t<DEL>\t.\tPASS\t.\tGT\t1|1\t0|0\t0|0",          # symbolic: skipped
        "2\t1\t.\tA\tC\t.\tPASS\t.\tGT\t1|1\t1|1\t1|1",              # other contig
    ]) + "\n"
    p = tmp_path / "a.vcf"
    p.write_text(vcf)
    r = sp.read_vcf(str(p), None, 12)
    assert r["n_haplotypes"] == 6 and r["samples"] == ["s0", "s1", "s2"]
    assert r["stats"] == {"records": 4, "skipped": 1, "unknown_symbols": 0, "dropped_calls": 0}
    v = r["variants"]
    assert [(int(x["ref_pos"]), int(x["del_len"]), int(x["ins_len"])) for x in v] == [(2, 1, 1), (2, 1, 1), (5, 0, 3), (5, 2, 0)]
    assert to_str(r["pool"]) == "ATGGA"
    cov = [[(int(c[0]) >> h) & 1 for h in range(6)] for c in r["coverage"]]
    assert cov == [[1, 0, 0, 0, 0, 1], [0, 0, 1, 1, 0, 0], [0, 1, 0, 0, 0, 0], [1, 0, 0, 0, 0, 0]]
    # the parsed arrays drive the oracle: haplotype 0 = ref with SNP G->A and the 2-symbol deletion
    ref = np.array([0, 1, 2, 3, 1, 1, 3, 3, 0, 0, 2, 2], dtype=np.uint8)        # ACGTCCTTAAGG
    h0 = orc.jst_haplotype(ref, v.astype(orc.EVENT_DTYPE), r["pool"], r["coverage"], 0)
    assert to_str(h0) == "ACATC" + "TAAGG"
    h1 = orc.jst_haplotype(ref, v.astype(orc.EVENT_DTYPE), r["pool"], r["coverage"], 1)
    assert to_str(h1) == "ACGTC" + "GGA" + "CTTAAGG"
    assert sp.read_vcf(str(p), "2")["variants"]["ref_pos"].tolist() == [0]


@pytest.mark.skipif(not os.path.isdir(REF_DATA), reason="the reference's test data lives only in the build container")
@pytest.mark.parametrize("tag", ["SNPs", "SNP_INDELs"])
def test_ingest_of_reference_data_equals_committed_fixture(tag):
    z = np.load(os.path.join(HERE, "golden", f"sim_ref_10Kb_{tag}.npz"))
    recs, unknown = sp.read_fasta(REF_DATA + "sim_ref_10Kb.fasta.gz")
    assert unknown == 0 and len(recs) == 1 and np.array_equal(recs[0][1], z["ref"])
    r = sp.read_vcf(REF_DATA + f"sim_ref_10Kb_{tag}.vcf", None, len(recs[0][1]))
    assert r["n_haplotypes"] == int(z["n_hap"]) == 100 and r["stats"]["skipped"] == 0
    assert np.array_equal(r["variants"], z["var"]) and np.array_equal(r["pool"], z["pool"])
    assert np.array_equal(r["coverage"], z["cov"])
    # and the shipped haplotype FASTA is what reference + carried variants gives (known answer)
    haps, _ = sp.read_fasta(REF_DATA + f"sim_ref_10Kb_{tag}_haplotypes.fasta.gz")
    assert len(haps) == 100
    ev = r["variants"].astype(orc.EVENT_DTYPE)
    for h in (0, 37, 99):
        assert np.array_equal(orc.jst_haplotype(recs[0][1], ev, r["pool"], r["coverage"], h), haps[h][1])


def test_fasta_wide_alphabets(tmp_path):
    """spmb_fasta_read_alphabet: seqan3's rank orders for dna5 (ACGNT) and dna15 (ABCDGHKMNRSTVWY)
    (libspm/libspm/seqan/alphabet.hpp:119-120); characters outside the alphabet become N."""
    p = tmp_path / "w.fa"
    p.write_text(">r\nACGTNacgtnRYKMSWBDHVU-x\n")
    got5, unk5 = sp.read_fasta(str(p), bits=3)
    assert "".join("ACGNT"[x] for x in got5[0][1]) == "ACGTNACGTN" + "N" * 10 + "T" + "NN"
    assert unk5 == 12
    got15, unk15 = sp.read_fasta(str(p), bits=4)
    assert "".join("ABCDGHKMNRSTVWY"[x] for x in got15[0][1]) == "ACGTNACGTNRYKMSWBDHVT" + "NN"
    assert unk15 == 2
    got4, unk4 = sp.read_fasta(str(p))
    assert to_str(got4[0][1]) == "ACGTAACGTA" + "A" * 10 + "T" + "AA" and unk4 == 14


def test_vcf_overlapping_calls_first_wins(tmp_path):
    """one haplotype calls a 5-symbol deletion AND a SNP inside it (and an insertion behind a replacement at the
    same position): the later call is cleared for that haplotype and counted, so every haplotype's variants form a
    valid journal -- the input spmb_jst_create demands (it rejects overlapping variants, as the oracle does)."""
    vcf = "\n".join([
        "##fileformat=VCFv4.1",
        "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\tFORMAT\ts0\ts1",
        "1\t4\t.\tACGTAC\tA\t.\tPASS\t.\tGT\t1|0\t0|0",     # deletion of 0-based [4, 9) on haplotype 0
        "1\t7\t.\tT\tG\t.\tPASS\t.\tGT\t1|1\t0|1",          # SNP at 6: inside the deletion for haplotype 0
        "1\t12\t.\tC\tT\t.\tPASS\t.\tGT\t0|0\t1|0",         # SNP at 11 on haplotype 2 ...
        "1\t12\t.\tC\tCA\t.\tPASS\t.\tGT\t0|0\t1|0",        # ... and an insertion at 12 on the same haplotype: fine
    ]) + "\n"
    p = tmp_path / "o.vcf"
    p.write_text(vcf)
    r = sp.read_vcf(str(p), None, 40)
    v = r["variants"]
    assert list(v["ref_pos"]) == [4, 6, 11, 12]
    cov = r["coverage"][:, 0]
    assert list(cov) == [0b0001, 0b1010, 0b0100, 0b0100]      # haplotype 0 lost the SNP at 6
    assert r["stats"]["dropped_calls"] == 1
    # every haplotype's list is now a valid journal for the oracle
    import oracle as orc
    ref = np.arange(40, dtype=np.uint8) % 4
    for h in range(4):
        hap = orc.jst_haplotype(ref, v.astype(orc.EVENT_DTYPE), r["pool"], r["coverage"], h)
        assert len(hap) == 40 + sum(int(v["ins_len"][i]) - int(v["del_len"][i]) for i in range(4) if (int(cov[i]) >> h) & 1)
