"""Full-size config C4 (BASELINE.json configs[3]: 2504 haplotypes x 50 Mbp reference, 1 M variants, Myers
m=150 k=2) through the CPU oracle ONCE: every haplotype is materialised and searched (about 125 Gbp of
three-block Myers).  The resulting hit list (haplotype, exclusive end in that haplotype's coordinates, errors)
is committed as tests/golden/c4_hits.npz so that a -m gpu test and bench.py can compare the WHOLE C4 hit list of
the CUDA path without repeating the CPU pass.

    python tests/golden/make_c4_golden.py            # writes tests/golden/c4_hits.npz (needs only the CPU)

The inputs come from libspm_b200.synth.c4() (the shared xoshiro256** generator, seed 0x5EED0004); their digest is
stored next to the hits, so a generator change that would invalidate the fixture is detected instead of
producing a silent mismatch."""
import hashlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import oracle as orc  # noqa: E402
from libspm_b200 import synth  # noqa: E402


def input_digest(d) -> str:
    h = hashlib.sha256()
    for key in ("ref", "variants", "pool", "coverage", "pattern"):
        h.update(np.ascontiguousarray(d[key]).view(np.uint8).tobytes())
    return h.hexdigest()


def main():
    scale = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
    out = os.path.join(ROOT, "tests", "golden", "c4_hits.npz" if scale == 1.0 else f"c4_hits_scale{scale:g}.npz")
    t0 = time.time()
    d = synth.c4(scale)
    print(f"generated in {time.time() - t0:.1f} s: {len(d['ref'])} bp, {len(d['variants'])} variants", flush=True)
    threads = orc.use_all_cores()
    t0 = time.time()
    want = orc.jst_search(3, d["ref"], d["variants"].astype(orc.EVENT_DTYPE), d["pool"], d["coverage"], d["n_hap"],
                          d["pattern"], d["k"], threads=threads)
    dt = time.time() - t0
    eff = float(len(d["ref"])) * d["n_hap"]
    print(f"oracle: {len(want)} hits in {dt:.1f} s on {threads} threads = {eff / dt / 1e9:.2f} effective Gbp/s", flush=True)
    np.savez_compressed(out, hap=want["hap"].astype(np.uint32), end=want["end"].astype(np.uint64),
                        err=want["err"].astype(np.uint8), input_sha256=np.array(input_digest(d)),
                        oracle_seconds=np.array(dt), oracle_threads=np.array(threads), scale=np.array(scale),
                        hits_sha256=np.array(hashlib.sha256(np.ascontiguousarray(want).view(np.uint8).tobytes()).hexdigest()))
    print("wrote", out)


if __name__ == "__main__":
    main()
