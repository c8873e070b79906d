"""Runs BASELINE.json's configs C3, C4 and C5 at their specified sizes on one GPU and checks the
hit lists against the CPU oracle (C1/C2/C3 are also checked inside bench.py).  Writes one JSON
document (default profiles/r01_configs.json).   python tools/run_configs.py [c4] [c5] [--out path]
C4 oracle: a fixed-seed sample of 32 haplotypes (full per-haplotype oracle = 125 Gbp of CPU Myers).
C5 oracle: all 10k patterns x a 3 Mbp slice, and 16 patterns x the full 3 Gbp."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import libspm_b200 as sp
import oracle as orc
from libspm_b200 import synth

args = [a for a in sys.argv[1:] if not a.startswith("--")]
out_path = "profiles/r01_configs.json"
if "--out" in sys.argv:
    out_path = sys.argv[sys.argv.index("--out") + 1]
which = set(args) or {"c4", "c5"}
scale = 1.0
if "--scale" in sys.argv:
    scale = float(sys.argv[sys.argv.index("--scale") + 1])
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev)
res = {}


def timed(fn, reps=3):
    best = 1e30
    out = None
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        out = fn()
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
    return best, out


with torch.cuda.stream(stream):
    ctx = sp.Context(0, stream=stream.cuda_stream)
    hits = sp.Hits(ctx, 1 << 20)
    if "c4" in which:
        n_ref, n_var, n_hap, m, k = int(50_000_000 * scale), int(1_000_000 * scale), 2504, 150, 2
        seed = synth.SEEDS["C4"]
        t0 = time.perf_counter()
        ref = synth.text_host(n_ref, seed)
        var, pool, kinds, cov = synth.jst(n_ref, n_var, n_hap, seed)
        pool = synth.finish_snps(var, pool, kinds, ref[var["ref_pos"][kinds == 0].astype(np.int64)])
        gen_s = time.perf_counter() - t0
        ev = var.astype(orc.EVENT_DTYPE)
        rng = np.random.default_rng(seed)
        src_h = int(rng.integers(0, n_hap))
        hap = orc.jst_haplotype(ref, ev, pool, cov, src_h)
        s0 = int(rng.integers(0, len(hap) - m))
        pat = hap[s0:s0 + m].copy()
        store = sp.Store.from_host(ctx, ref)
        t0 = time.perf_counter()
        jst = sp.Jst(ctx, store, var, pool, cov, n_hap)
        create_s = time.perf_counter() - t0
        p = sp.Pattern(ctx, sp.ALG_MYERS, pat, k)
        t0 = time.perf_counter()
        got = jst.search(p, hits)
        first_s = time.perf_counter() - t0
        st_first = jst.last_stats()
        warm_s, got2 = timed(lambda: jst.search(p, hits))
        st_warm = jst.last_stats()
        sample = np.sort(np.random.default_rng(seed + 1).choice(n_hap, size=32, replace=False)).astype(np.uint32)
        sample = np.unique(np.append(sample, np.uint32(src_h)))
        t0 = time.perf_counter()
        want = orc.jst_search(3, ref, ev, pool, cov, n_hap, pat, k, haps=sample)
        oracle_s = time.perf_counter() - t0
        sel = got[np.isin(got["seq_id"], sample)]
        ok = (len(sel) == len(want) and np.array_equal(sel["seq_id"], want["hap"]) and
              np.array_equal(sel["end"], want["end"]) and np.array_equal(sel["errors"], want["err"]))
        hap_len_sum = float(n_ref) * n_hap  # effective bp (ins/del shifts average out)
        streamed = n_ref + st_first["window_symbols"]
        res["C4"] = {
            "workload": f"JST: {n_hap} haplotypes, {n_ref} bp reference, {len(var)} variants, Myers m={m} k={k}",
            "hits": int(len(got)), "haplotypes_with_hits": int(len(np.unique(got["seq_id"]))),
            "parity_sample_haplotypes": int(len(sample)), "parity": bool(ok), "oracle_hits_in_sample": int(len(want)),
            "windows": int(st_first["windows"]), "window_symbols": int(st_first["window_symbols"]),
            "buckets": int(st_first["buckets"]), "raw_device_hits": int(st_first["raw_hits"]),
            "first_search_s": first_s, "flatten_ms": st_first["flatten_ms"], "materialise_ms": st_first["materialise_ms"],
            "kernels_ms": st_warm["kernels_ms"], "expand_ms": st_warm["expand_ms"], "warm_search_s": warm_s,
            "streamed_bp": int(streamed), "effective_bp": hap_len_sum,
            "streamed_gbp_per_s_kernels": streamed / (st_warm["kernels_ms"] * 1e-3) / 1e9,
            "effective_gbp_per_s_warm": hap_len_sum / warm_s / 1e9,
            "int32_instr_per_bp": 10 * ((m + 31) // 32) + 4,
            "oracle_sample_s": oracle_s, "oracle_threads": orc.max_threads(),
            "cpu_effective_gbp_per_s": float(len(sample)) * n_ref / oracle_s / 1e9,
            "generate_s": gen_s, "jst_create_s": create_s,
        }
        print(json.dumps(res["C4"]), flush=True)
        jst.close(); store.close()
        del ref, var, cov, hap
    if "c5" in which:
        n, n_pat, m, k = int(3_000_000_000 * scale), int(10_000 * min(1.0, scale * 10)), 100, 3
        seed = synth.SEEDS["C5"]
        text = synth.text_device(n, seed, dev)
        rng = np.random.default_rng(seed)
        offs = np.sort(rng.integers(0, n - 2 * m, size=n_pat))
        pats = []
        offs_t = torch.from_numpy(offs).to(dev)
        idx = (offs_t[:, None] + torch.arange(m, device=dev)[None, :]).reshape(-1)
        cut = text[idx].cpu().numpy().reshape(n_pat, m)
        for i in range(n_pat):
            q = synth.mutate(rng, cut[i], int(rng.integers(0, 6)))
            if len(q) < 97:
                q = np.concatenate([q, cut[i][:97 - len(q)]])
            pats.append(q[:128])
        store = sp.Store.from_device(ctx, text.data_ptr(), n)
        ctx.sync()
        # (i) all patterns x a 3 Mbp slice starting at a pattern-rich offset
        sl_n = min(n, 3_000_000)
        sl_start = int(offs[n_pat // 2]) // 1024 * 1024
        sl_start = min(sl_start, n - sl_n)
        slice_host = text[sl_start:sl_start + sl_n].cpu().numpy()
        st_slice = sp.Store.from_host(ctx, slice_host)
        pset = sp.Pattern(ctx, sp.ALG_MYERS, pats, k)
        t_slice, g = timed(lambda: sp.search(ctx, st_slice, pset, hits), reps=2)
        t0 = time.perf_counter()
        bad = 0
        n_or = 0
        for i, q in enumerate(pats):
            _, we, wr = orc.search_mt(3, slice_host, q, k, threads=orc.max_threads())
            sel = g[g["pattern_id"] == i]
            n_or += len(we)
            if not (np.array_equal(sel["end"], we) and np.array_equal(sel["errors"], wr)):
                bad += 1
        oracle_slice_s = time.perf_counter() - t0
        # (ii) 16 patterns x the full text
        sub = [pats[i] for i in range(0, n_pat, max(1, n_pat // 16))][:16]
        pset16 = sp.Pattern(ctx, sp.ALG_MYERS, sub, k)
        t16, g16 = timed(lambda: sp.search(ctx, store, pset16, hits), reps=2)
        host = text.cpu().numpy()
        t0 = time.perf_counter()
        bad16 = 0
        full_oracle = []
        for i, q in enumerate(sub):
            _, we, wr = orc.search_mt(3, host, q, k, threads=orc.max_threads())
            full_oracle.append((we, wr))
            sel = g16[g16["pattern_id"] == i]
            if not (np.array_equal(sel["end"], we) and np.array_equal(sel["errors"], wr)):
                bad16 += 1
        oracle16_s = time.perf_counter() - t0
        del host
        # (iii) the whole config in one call through the pigeonhole seed filter (SPMB_ALG_MYERS_FILTERED):
        # all patterns x the full text; checked against (i) on the slice, (ii) on the 16 patterns, and
        # against the direct kernel's hits for those 16
        psetf = sp.Pattern(ctx, sp.ALG_MYERS_FILTERED, pats, k)
        finfo = psetf.filter_info
        gs = sp.search(ctx, st_slice, psetf, hits)
        bad_slice_f = int(not (np.array_equal(gs["pattern_id"], g["pattern_id"]) and np.array_equal(gs["end"], g["end"])
                               and np.array_equal(gs["errors"], g["errors"])))
        sp.search(ctx, store, psetf, hits)     # builds the lane-chunk copy of the store once
        tf, gf = timed(lambda: sp.search(ctx, store, psetf, hits), reps=3)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        kms = []
        for _ in range(3):
            e0.record(stream)
            sp.search_launch(ctx, store, psetf, hits)
            e1.record(stream)
            sp.search_finish(ctx, hits)
            kms.append(e0.elapsed_time(e1))
        step = max(1, n_pat // 16)
        bad_f = 0
        for j, (we, wr) in enumerate(full_oracle):
            sel = gf[gf["pattern_id"] == j * step]
            if not (np.array_equal(sel["end"], we) and np.array_equal(sel["errors"], wr)):
                bad_f += 1
        res["C5_filtered"] = {
            "workload": f"the whole config in one call: {n_pat} patterns x {n} bp, k={k}, SPMB_ALG_MYERS_FILTERED "
                        "(pigeonhole seed filter + windowed Myers verification)",
            "filter": finfo, "hits": int(len(gf)), "patterns_with_hits": int(len(np.unique(gf["pattern_id"]))),
            "search_s": tf, "kernels_ms": float(np.min(kms)),
            "searched_gbp_per_s": n / tf / 1e9, "cell_rate_T_per_s": n_pat * n / tf / 1e12,
            "text_gbytes_per_s_kernels": n * 0.25 / (float(np.min(kms)) * 1e-3) / 1e9,
            "equals_direct_kernel_on_slice_all_patterns": bool(bad_slice_f == 0),
            "patterns_mismatching_full_text_oracle": bad_f, "parity": bool(bad_f == 0 and bad_slice_f == 0),
            "speedup_vs_direct_kernel_extrapolated": (n_pat * n / (len(sub) * n / t16)) / tf,
        }
        print(json.dumps(res["C5_filtered"]), flush=True)
        ops = 10 * 4 + 4
        res["C5"] = {
            "workload": f"batched multi-pattern: {n_pat} patterns of ~{m} bp, k={k}, over {n} bp (one GPU; 8-GPU run = 8 text shards)",
            "slice_check": {"patterns": n_pat, "slice_bp": sl_n, "gpu_s": t_slice, "hits": int(len(g)), "oracle_hits": int(n_or),
                            "patterns_mismatching": bad, "oracle_s": oracle_slice_s,
                            "cell_rate_T_per_s": n_pat * sl_n / t_slice / 1e12,
                            "int32_Tinstr_per_s": n_pat * sl_n * ops / t_slice / 1e12},
            "full_text_check": {"patterns": len(sub), "text_bp": n, "gpu_s": t16, "hits": int(len(g16)),
                                "patterns_mismatching": bad16, "oracle_s": oracle16_s,
                                "cell_rate_T_per_s": len(sub) * n / t16 / 1e12,
                                "int32_Tinstr_per_s": len(sub) * n * ops / t16 / 1e12,
                                "cpu_cell_rate_T_per_s": len(sub) * n / oracle16_s / 1e12},
            "parity": bool(bad == 0 and bad16 == 0),
            "extrapolated_full_c5_one_gpu_s": n_pat * n / (len(sub) * n / t16),
        }
        print(json.dumps(res["C5"]), flush=True)
os.makedirs(os.path.dirname(out_path) or ".", exist_ok=True)
prev = {}
if os.path.exists(out_path):
    try:
        prev = json.load(open(out_path))
    except Exception:
        prev = {}
prev.update(res)
json.dump(prev, open(out_path, "w"), indent=1)
