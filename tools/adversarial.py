"""Adversarial and realistic inputs for the measured set: every figure that bench.py reports is
on iid uniform ACGT; these are the cases where the kernels' fast paths do not apply.

    python tools/adversarial.py [--out profiles/r02_adversarial.json] [--n 800000000]

  exact_low_complexity   text with 5 % poly-A / (AC)n / (TTG)n runs, needle whose first 8 symbols lie in such a run:
                         filter window at needle offset 0 (SPMB_EXACT_WINDOW=first, round 1) against the least
                         repetitive window (default) -- child processes, the switch is read once per process
  myers_high_k           m=100, k=8 (k/m = 0.08) on iid text: cut-off against full recurrence
  myers_tandem           m=100, k=4 on a text with 20 % tandem repeats (period 37..61) and the needle cut from one
  myers_hit_dense        >= 10^6 hits: a 3 Mbp poly-A stretch, needle A^100 with one C, k=4 (warp-aggregated append)
Each case checks its hit list against the CPU oracle on a slice."""
import argparse
import json
import os
import subprocess
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np


def low_complexity(n, seed, frac=0.05):
    from libspm_b200 import synth
    t = synth.text_host(n, seed)
    rng = np.random.default_rng(seed)
    mean_gap = int(150 / frac)
    n_runs = n // mean_gap
    starts = np.sort(rng.integers(0, n - 400, size=n_runs))
    lens = rng.integers(20, 280, size=n_runs)
    kinds = rng.integers(0, 3, size=n_runs)
    units = [np.array([0], np.uint8), np.array([0, 1], np.uint8), np.array([3, 3, 2], np.uint8)]
    tiles = [np.tile(u, 300 // len(u) + 1) for u in units]
    for s, ln, kd in zip(starts.tolist(), lens.tolist(), kinds.tolist()):
        t[s:s + ln] = tiles[kd][:ln]
    return t, starts, lens, kinds


def time_kernel(sp, torch, ctx, store, p, hits, reps=8):
    sp.search(ctx, store, p, hits)
    sp.search(ctx, store, p, hits)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ks, tot = [], []
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e0.record()
        sp.search_launch(ctx, store, p, hits)
        e1.record()
        h = sp.search_finish(ctx, hits)
        torch.cuda.synchronize()
        tot.append((time.perf_counter() - t0) * 1e3)
        ks.append(e0.elapsed_time(e1))
    return float(np.median(ks)), float(np.median(tot)), h


def exact_case(args):
    import torch
    import libspm_b200 as sp
    import oracle as orc
    n = args.n
    text, starts, lens, kinds = low_complexity(n, 11)
    rng = np.random.default_rng(5)
    res = {}
    with torch.cuda.stream(torch.cuda.Stream()):
        ctx = sp.Context(0, stream=torch.cuda.current_stream().cuda_stream)
        hits = sp.Hits(ctx, 1 << 20)
        store = sp.Store.from_host(ctx, text)
        for name, head in (("polyA_head", np.zeros(12, np.uint8)), ("AC_head", np.tile(np.array([0, 1], np.uint8), 8)),
                           ("random", None)):
            tail = rng.integers(0, 4, size=52, dtype=np.uint8)
            pat = tail if head is None else np.concatenate([head, tail])[:64]
            p = sp.Pattern(ctx, sp.ALG_SHIFTOR, pat)
            k_ms, t_ms, h = time_kernel(sp, torch, ctx, store, p, hits)
            sl = text[:50_000_000]
            want = orc.search_mt(2, sl, pat, 0)[1]
            got = h["end"][h["end"] <= len(sl)]
            res[name] = {"kernel_ms": k_ms, "step_ms": t_ms, "gbp_per_s_kernel": n / k_ms / 1e6, "hits": int(len(h)),
                         "parity_on_50Mbp_slice": bool(np.array_equal(got, want))}
            p.close()
        store.close()
    res["window"] = os.environ.get("SPMB_EXACT_WINDOW", "best")
    res["n"] = n
    print(json.dumps(res))


def myers_cases(args):
    import torch
    import libspm_b200 as sp
    import oracle as orc
    from libspm_b200 import synth
    n = 250_000_000
    out = {}
    with torch.cuda.stream(torch.cuda.Stream()):
        ctx = sp.Context(0, stream=torch.cuda.current_stream().cuda_stream)
        hits = sp.Hits(ctx, 1 << 22)

        def both(store, pat, k):
            p = sp.Pattern(ctx, sp.ALG_MYERS, pat, k)
            ck, ct, h = time_kernel(sp, torch, ctx, store, p, hits)
            ctx.set_myers_cutoff(False)
            fk, ft, h2 = time_kernel(sp, torch, ctx, store, p, hits, reps=3)
            ctx.set_myers_cutoff(True)
            p.close()
            return {"cutoff_kernel_ms": ck, "cutoff_step_ms": ct, "full_kernel_ms": fk, "hits": int(len(h)),
                    "cutoff_equals_full": bool(np.array_equal(h["end"], h2["end"]) and np.array_equal(h["errors"], h2["errors"]))}, h

        # k/m = 0.08 on iid text
        cfg = synth.c3(n, seed=77, m=100, k=8)
        store = sp.Store.from_host(ctx, cfg["text"])
        r, h = both(store, cfg["pattern"], 8)
        sl = cfg["text"][:30_000_000]
        _, we, wr = orc.search_mt(3, sl, cfg["pattern"], 8)
        g = h[h["end"] <= len(sl)]
        r["parity_on_30Mbp_slice"] = bool(np.array_equal(g["end"], we) and np.array_equal(g["errors"], wr))
        out["myers_high_k"] = dict(r, workload="m=100 k=8 (k/m = 0.08), 250 Mbp iid + planted copies")
        # the same text at k=4 for reference
        r4, _ = both(store, cfg["pattern"], 4)
        out["myers_k4_same_text"] = r4
        store.close()
        # tandem repeats
        text = synth.text_host(n, 78)
        rng = np.random.default_rng(78)
        pos = 0
        n_rep = 0
        while pos < n - 5000:
            pos += int(rng.integers(2000, 14000))
            period = int(rng.integers(37, 62))
            ln = int(rng.integers(400, 3000))
            unit = rng.integers(0, 4, size=period, dtype=np.uint8)
            rep = np.tile(unit, ln // period + 1)[:ln].copy()
            flips = rng.integers(0, ln, size=ln // 50)
            rep[flips] = (rep[flips] + 1) & 3                       # 2 % divergence between copies
            text[pos:pos + ln] = rep
            if n_rep == 1000:
                pat = np.tile(unit, 4)[:100].copy()
            n_rep += 1
            pos += ln
        store = sp.Store.from_host(ctx, text)
        r, h = both(store, pat, 4)
        sl = text[:30_000_000]
        _, we, wr = orc.search_mt(3, sl, pat, 4)
        g = h[h["end"] <= len(sl)]
        r["parity_on_30Mbp_slice"] = bool(np.array_equal(g["end"], we) and np.array_equal(g["errors"], wr))
        out["myers_tandem"] = dict(r, workload=f"m=100 k=4, 250 Mbp with {n_rep} tandem repeats (period 37..61, ~20 % of the text), "
                                               "needle = four copies of one repeat unit")
        store.close()
        # hit-dense
        text = synth.text_host(n, 79)
        text[100_000_000:103_000_000] = 0
        pat = np.zeros(100, np.uint8)
        pat[50] = 1
        store = sp.Store.from_host(ctx, text)
        r, h = both(store, pat, 4)
        sl = text[99_000_000:104_000_000]
        _, we, wr = orc.search_mt(3, sl, pat, 4)
        g = h[(h["end"] > 99_000_000 + 200) & (h["end"] <= 104_000_000)]
        keep = we > 200
        r["parity_on_dense_stretch"] = bool(np.array_equal(g["end"] - 99_000_000, we[keep]) and np.array_equal(g["errors"], wr[keep]))
        out["myers_hit_dense"] = dict(r, workload="m=100 k=4, 250 Mbp with a 3 Mbp poly-A stretch, needle A^50 C A^49")
        store.close()
    print(json.dumps(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="profiles/r02_adversarial.json")
    ap.add_argument("--n", type=int, default=800_000_000)
    ap.add_argument("--case", default="all")
    args = ap.parse_args()
    if args.case == "exact":
        return exact_case(args)
    if args.case == "myers":
        return myers_cases(args)
    res = {}
    for win in ("first", "best"):
        env = dict(os.environ)
        if win == "first":
            env["SPMB_EXACT_WINDOW"] = "first"
        else:
            env.pop("SPMB_EXACT_WINDOW", None)
        r = subprocess.run([sys.executable, __file__, "--case", "exact", "--n", str(args.n)], env=env, capture_output=True, text=True)
        if r.returncode:
            res[f"exact_low_complexity_{win}_window"] = {"error": r.stderr[-2000:]}
        else:
            res[f"exact_low_complexity_{win}_window"] = json.loads(r.stdout.strip().splitlines()[-1])
    r = subprocess.run([sys.executable, __file__, "--case", "myers"], capture_output=True, text=True)
    if r.returncode:
        res["myers"] = {"error": r.stderr[-2000:]}
    else:
        res.update(json.loads(r.stdout.strip().splitlines()[-1]))
    os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
    json.dump(res, open(args.out, "w"), indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
