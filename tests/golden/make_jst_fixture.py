"""Generates tests/golden/sim_ref_10Kb_{SNPs,SNP_INDELs}.npz from the reference's OWN test data
(run in the build container, where /root/reference exists; the .npz files are committed because
the GPU box has no /root/reference):

    /root/reference/test/data/sim_ref_10Kb.fasta.gz                       10 kb simulated reference
    /root/reference/test/data/sim_ref_10Kb_{SNPs,SNP_INDELs}.vcf          mason variants, 50 samples x 2
    /root/reference/test/data/sim_ref_10Kb_{...}_haplotypes.fasta.gz      the 100 resulting haplotypes
    (declared in test/data/datasources.cmake:72-102; unused by every reference test)

Each VCF record becomes one journal-event-like variant per ALT allele (common prefix stripped),
with a coverage bit per haplotype; the shipped haplotype FASTA is the known answer for
"reference + carried variants -> haplotype" (known-answer fixture for the store).
Stored: reference ranks, variants, insertion pool, coverage matrix, and length + sha256 of every
shipped haplotype (as rank bytes), plus three full haplotypes for needle cutting."""
import gzip
import hashlib
import os

import numpy as np

D = "/root/reference/test/data/"
HERE = os.path.dirname(os.path.abspath(__file__))
RANK = {"A": 0, "C": 1, "G": 2, "T": 3}
EVENT_DTYPE = np.dtype([("ref_pos", "<u8"), ("del_len", "<u4"), ("ins_len", "<u4"), ("ins_off", "<u8")])


def fasta(path):
    recs, name, seq = [], None, []
    for line in gzip.open(path, "rt"):
        line = line.strip()
        if line.startswith(">"):
            if name is not None:
                recs.append((name, "".join(seq)))
            name, seq = line[1:], []
        else:
            seq.append(line)
    recs.append((name, "".join(seq)))
    return recs


def ranks(s):
    return np.array([RANK[c] for c in s], dtype=np.uint8)


ref = fasta(D + "sim_ref_10Kb.fasta.gz")[0][1]
for tag in ("SNPs", "SNP_INDELs"):
    haps = fasta(D + f"sim_ref_10Kb_{tag}_haplotypes.fasta.gz")
    events, covs = [], []
    n_hap = None
    for line in open(D + f"sim_ref_10Kb_{tag}.vcf"):
        if line.startswith("##"):
            continue
        f = line.rstrip("\n").split("\t")
        if line.startswith("#"):
            n_hap = 2 * (len(f) - 9)
            continue
        pos, r, alts = int(f[1]) - 1, f[3], f[4].split(",")
        gts = [g.split("|") for g in f[9:]]
        for ai, a in enumerate(alts, start=1):
            c = 0
            while c < len(r) and c < len(a) and r[c] == a[c]:
                c += 1
            events.append((pos + c, len(r) - c, a[c:]))
            covs.append([int(g[h]) == ai for g in gts for h in (0, 1)])
    order = sorted(range(len(events)), key=lambda i: (events[i][0], events[i][1] != 0))
    var = np.zeros(len(events), dtype=EVENT_DTYPE)
    pool = []
    wpr = (n_hap + 63) // 64
    bits = np.zeros((len(events), wpr * 64), dtype=bool)
    for j, i in enumerate(order):
        p, d, ins = events[i]
        var[j] = (p, d, len(ins), len(pool))
        pool += [RANK[c] for c in ins]
        bits[j, :n_hap] = covs[i]
    cov = np.packbits(bits, axis=1, bitorder="little").view(np.uint64)
    hap_len = np.array([len(s) for _, s in haps], dtype=np.int64)
    hap_sha = np.array([hashlib.sha256(ranks(s).tobytes()).hexdigest() for _, s in haps])
    np.savez_compressed(os.path.join(HERE, f"sim_ref_10Kb_{tag}.npz"), ref=ranks(ref), var=var,
                        pool=np.array(pool, dtype=np.uint8), cov=cov, n_hap=np.int64(n_hap), hap_len=hap_len,
                        hap_sha=hap_sha, hap0=ranks(haps[0][1]), hap37=ranks(haps[37][1]), hap99=ranks(haps[99][1]))
    print(tag, len(var), "variants", n_hap, "haplotypes", hap_len.min(), hap_len.max())
