#!/usr/bin/env python
"""bench.py -- searched Gbp/s of the online pattern-matching hot path on B200.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU algorithms (oracle port)

A "step" is one search pass (kernel launch -> sorted hit list on the host) over one resident
sequence store.  Headline workload = BASELINE.json configs[1] at steady-state scale ("C2x"):
Shift-Or exact search, m=64, over a journaled sequence (reference + 1% SNP/indel journal),
N = 3.2 Gbp per GPU so that the packed text (0.25 B/bp = 800 MB) exceeds the 126 MB L2.
The same line also carries the Myers k-error measurement (configs[2], "C3": m=100, k=4,
250 Mbp) with its INT-pipe roofline, and the as-specified small configs C1/C2.
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ALG_NAMES = {1: "horspool", 2: "shiftor", 3: "myers"}
BYTES_PER_BP = 0.25                 # 2-bit packed text read once


def myers_ops_per_bp(m):            # INT32 instructions per (bp x pattern), SURVEY.md 8(d)
    return 10 * ((m + 31) // 32) + 4


# ------------------------------------------------------------------------------------------
def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)", d
    return 6650.0, "fallback (B200_PROFILING.md)", {}


def ncu_traffic(key):
    """per-launch DRAM bytes of the dominant kernel from the committed ncu --set full capture."""
    p = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get(key)
        except Exception:
            return None
    return None


class ClockSampler:
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.lines = []
        self.proc = None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "50", "-i", str(self.index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        return self

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.06)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons, pw = [], [], set(), []
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(pw)}


# ------------------------------------------------------------------------------------------
def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def build_c2x(sp, synth, torch, ctx, dev, n, seed, m=64):
    """reference (device RNG) + 1% journal (host RNG) -> base store + journaled store + pattern."""
    ref = synth.text_device(n, seed, dev)
    rng = np.random.default_rng(seed ^ 0xABCDEF)
    pat = rng.integers(0, 4, size=m, dtype=np.uint8)
    # plant copies of the needle at seeded reference offsets (about half survive the journal)
    n_plant = 100
    starts = np.sort(rng.integers(0, n - m, size=n_plant)).astype(np.int64)
    idx = (torch.from_numpy(starts).to(dev)[:, None] + torch.arange(m, device=dev)[None, :]).reshape(-1)
    ref[idx] = torch.from_numpy(np.tile(pat, n_plant)).to(dev)
    ev, pool, kinds = synth.journal(n, seed)
    snp_pos = torch.from_numpy(ev["ref_pos"][kinds == 0].astype(np.int64)).to(dev)
    pool = synth.finish_snps(ev, pool, kinds, ref[snp_pos].cpu().numpy())
    return ref, ev, pool, pat


def time_search(sp, torch, ctx, stream, store, pattern, hits, steps, warmup, flush=None):
    """returns (total_ms for `steps` launch+finish passes, mean kernel ms, n_hits, launches/step).
    The passes are pipelined the way the launch/finish pair of the C ABI is meant to be used: passes i+1 and
    i+2 are queued (on their own hit lists) before pass i's hits are collected, so the host's share of a pass
    (D2H of the counters + hit list, ordering) runs under the next kernels.  Every pass still delivers its
    sorted hit list to the host."""
    DEPTH = int(os.environ.get("SPMB_BENCH_DEPTH", "3"))   # passes in flight: two queued behind the one being collected
    hl = (hits,) + tuple(sp.Hits(ctx, 1 << 16) for _ in range(DEPTH - 1))
    for w in range(max(warmup, 2 * DEPTH)):     # every hit list twice: buffers sized, long-list ordering armed
        if flush is not None:
            flush.add_(1)
        sp.search_launch(ctx, store, pattern, hl[w % DEPTH])
        sp.search_finish(ctx, hl[w % DEPTH])
    ks = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    ke = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    flush_ms = 0.0
    e0.record(stream)
    n_hits = 0

    def launch(i):
        if flush is not None:
            flush.add_(1)       # evict L2 between iterations (buffer > L2)
        ks[i].record(stream)
        sp.search_launch(ctx, store, pattern, hl[i % DEPTH])
        ke[i].record(stream)

    for i in range(min(DEPTH - 1, steps)):
        launch(i)
    for i in range(steps):
        if i + DEPTH - 1 < steps:
            launch(i + DEPTH - 1)
        n_hits = len(sp.search_finish(ctx, hl[i % DEPTH]))
    e1.record(stream)
    torch.cuda.synchronize()
    kern = [ks[i].elapsed_time(ke[i]) for i in range(steps)]
    total = e0.elapsed_time(e1)
    if flush is not None:       # take the flush writes out of the step time
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record(stream)
        for _ in range(steps):
            flush.add_(1)
        f1.record(stream)
        torch.cuda.synchronize()
        flush_ms = f0.elapsed_time(f1)
    return max(total - flush_ms, 1e-6), float(np.mean(kern)), n_hits, ctx.last_launch_count


def run_ours(args):
    import torch
    import libspm_b200 as sp
    from libspm_b200 import synth

    rank, world, local = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- libspm_b200 has no CPU path to measure")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    stream = torch.cuda.Stream(device=dev)
    hbm_peak, peak_src, _ = measured_peaks()
    out = {}
    if world > 1 and "SPMB_HOST_THREADS" not in os.environ:
        # the ranks of one box share its host cores: split the upload's packing threads between them
        os.environ["SPMB_HOST_THREADS"] = str(max(0, len(os.sched_getaffinity(0)) // world - 1))
    with torch.cuda.stream(stream):
        ctx = sp.Context(local, stream=stream.cuda_stream)
        hits = sp.Hits(ctx, 1 << 16)
        n = args.n
        m = 64
        # ---------------- headline: C2x exact search over a journaled sequence ----------------
        t0 = time.perf_counter()
        ref, ev, pool, pat = build_c2x(sp, synth, torch, ctx, dev, n, synth.SEEDS["C2"] + rank, m)
        host_ref = torch.empty(n, dtype=torch.uint8, pin_memory=True)
        host_ref.copy_(ref)
        torch.cuda.synchronize()
        base = sp.Store.from_device(ctx, ref.data_ptr(), n)
        ctx.sync()
        del ref
        torch.cuda.empty_cache()
        j0, j1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        j0.record(stream)
        js = base.journaled(ev, pool)
        j1.record(stream)
        torch.cuda.synchronize()
        journal_ms = j0.elapsed_time(j1)
        n_der = len(js)
        pattern = sp.Pattern(ctx, sp.ALG_SHIFTOR, pat)
        setup_s = time.perf_counter() - t0

        if world > 1:
            dist.barrier()
        # clocks are sampled from the warm-up to the end of the e2e legs (the timed region of the
        # headline alone is a few ms -- shorter than nvidia-smi's sampling period)
        clk = ClockSampler(local)
        clk.__enter__()
        total_ms, kern_ms, n_hits, launches = time_search(sp, torch, ctx, stream, js, pattern, hits,
                                                          args.steps, args.warmup)
        if world > 1:
            t = torch.tensor([total_ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            total_ms = float(t.item())
        value = world * n_der * args.steps / (total_ms * 1e-3) / 1e9
        achieved = n_der * BYTES_PER_BP / (kern_ms * 1e-3) / 1e9
        out.update({
            "metric": "searched_gbp_per_s", "value": value, "unit": "Gbp/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32",
            "data": "synthetic",
            "config": {"workload": "C2x: Shift-Or exact search m=64 over a journaled sequence "
                                   f"({n} bp reference + 1% SNP/indel journal, {len(ev)} events) per GPU; "
                                   "BASELINE configs[1] generator at steady-state scale",
                       "searched_bp_per_step_per_gpu": n_der, "pattern_len": m, "hits": n_hits,
                       "l2": f"inputs larger than L2 ({n_der * BYTES_PER_BP / 1e6:.0f} MB packed text vs 126 MB)",
                       "row_len": js.row_len, "journal_materialise_ms": journal_ms,
                       "step": "kernel launch + hit count/list D2H + host sort; passes i+1 and i+2 are queued before pass i's hits are "
                               "collected (spmb_search_launch / spmb_search_finish on two hit lists)"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                         "frac": achieved / hbm_peak, "traffic": ncu_traffic("exact_c2x"), "peak_source": peak_src,
                         "kernel": "exact_kernel<plain>", "kernel_ms": kern_ms,
                         "algorithmic_bytes_per_launch": n_der * BYTES_PER_BP},
            "gpu_launches": launches * args.steps,
        })

        # ---------------- e2e: host buffers -> hits through the C ABI ----------------
        e2e_steps = max(1, min(args.steps, args.e2e_steps))
        hp = host_ref.data_ptr()
        ev_bytes = ev.nbytes + pool.nbytes
        # the step's host inputs live in pinned memory: reference ranks, journal events, insertion pool
        ev_pin = torch.empty(ev.nbytes, dtype=torch.uint8, pin_memory=True)
        ev_pin.numpy()[:] = ev.view(np.uint8)
        ev = ev_pin.numpy().view(sp.EVENT_DTYPE)
        pool_pin = torch.empty(max(pool.nbytes, 1), dtype=torch.uint8, pin_memory=True)
        pool_pin.numpy()[:pool.nbytes] = pool
        pool = pool_pin.numpy()[:pool.nbytes]

        phase = np.zeros(3)                                  # upload+pack | journal | search (each call syncs)

        def e2e_once():
            ta = time.perf_counter()
            b2 = sp.Store(ctx, _store_from_ptr(sp, ctx, hp, n))
            tb = time.perf_counter()
            j2 = b2.journaled(ev, pool)
            tc = time.perf_counter()
            h = sp.search(ctx, j2, pattern, hits)
            td = time.perf_counter()
            phase[:] += (tb - ta, tc - tb, td - tc)
            j2.close(); b2.close()
            return len(h)

        e2e_once()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        phase[:] = 0
        text_h2d = 0
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            nh = e2e_once()
            us = ctx.upload_last_stats()
            # bytes that crossed PCIe for the haystack: raw chunks at 1 B/symbol, host-packed ones at 0.25
            text_h2d += (min(n, us["raw_chunks"] * us["chunk_symbols"]) + us["host_packed_chunks"] * us["chunk_symbols"] // 4
                         if us["chunk_symbols"] else n)
        torch.cuda.synchronize()
        e2e_s = time.perf_counter() - t0
        text_h2d //= e2e_steps
        us = ctx.upload_last_stats()
        if us["events_per_chunk"]:                           # events: 24 B raw, 8 B (+16 per chunk) host-packed
            epc = us["events_per_chunk"]
            ev_h2d = (len(ev) - min(len(ev), us["event_host_packed_chunks"] * epc)) * 24 \
                + us["event_host_packed_chunks"] * (16 + 8 * epc) + pool.nbytes
        else:
            ev_h2d = ev_bytes
        if world > 1:
            t = torch.tensor([e2e_s], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e2e_s = float(t.item())
        out["e2e"] = {"value": world * n_der * e2e_steps / e2e_s / 1e9, "unit": "Gbp/s",
                      "h2d_bytes_per_step": int(text_h2d + ev_h2d + m), "d2h_bytes_per_step": int(8 + nh * 24),
                      "steps": e2e_steps, "ms_per_step": e2e_s / e2e_steps * 1e3,
                      "host_input_bytes_per_step": int(n + ev_bytes + m),
                      "upload": ctx.upload_last_stats(),
                      "phase_ms": {"store_create(upload+pack)": phase[0] / e2e_steps * 1e3,
                                   "store_create_journaled": phase[1] / e2e_steps * 1e3,
                                   "search": phase[2] / e2e_steps * 1e3},
                      "path": "spmb_store_create(host ranks: raw chunks over the copy engine + chunks 2-bit packed by "
                              "host threads) + spmb_store_create_journaled + spmb_search"}
        # resident-store e2e: the store is the persistent object; per step only the needle goes up
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            p2 = sp.Pattern(ctx, sp.ALG_SHIFTOR, pat)
            nh = len(sp.search(ctx, js, p2, hits))
            p2.close()
        e2e_res_s = time.perf_counter() - t0
        out["e2e_resident"] = {"value": world * n_der * args.steps / e2e_res_s / 1e9, "unit": "Gbp/s",
                               "h2d_bytes_per_step": m, "d2h_bytes_per_step": int(8 + nh * 24),
                               "path": "spmb_pattern_create(host needle) + spmb_search on the resident store"}

        clk.__exit__()
        out["clocks"] = clk.summary()
        out["clocks"]["window"] = "warm-up + timed steps + e2e legs of the headline workload"

        # ---------------- CPU baseline + cross-check (rank 0, N=1 only) ----------------
        if rank == 0 and world == 1 and not args.no_cpu:
            import oracle as orc
            s_n = int(min(n_der, args.cpu_sample))
            sample = js.read(0, s_n)
            threads = orc.use_all_cores()
            best = 1e30
            for _ in range(3):
                t0 = time.perf_counter()
                cnt, ce, _ = orc.search_mt(2, sample, pat, 0, threads=threads)
                best = min(best, time.perf_counter() - t0)
            g = sp.search(ctx, js, pattern, hits)
            ge = g["end"][g["end"] <= s_n]
            out["cpu_baseline"] = {"value": s_n / best / 1e9, "unit": "Gbp/s", "cores": threads, "kind": "port",
                                   "sample": f"first {s_n} bp of the derived sequence, oracle Shift-Or (32-bit blocks), "
                                             f"{threads} threads, best of 3; journal already applied",
                                   "hits_match_gpu": bool(np.array_equal(ge, ce))}
            del sample
        js.close(); base.close(); del host_ref

        # ---------------- Myers k-error: C3 (m=100, k=4, 250 Mbp) ----------------
        if not args.no_myers:
            out["myers"] = bench_myers(sp, synth, torch, ctx, dev, stream, hits, args, rank, world, hbm_peak)
        if not args.no_c5:
            out["c5"] = bench_c5(sp, synth, torch, ctx, dev, stream, hits, args, rank, world)
        if rank == 0 and world == 1 and not args.no_small:
            out["small_configs"] = bench_small(sp, synth, torch, ctx, dev, stream, hits, args)
        out["setup_s"] = setup_s
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        emit(out)


def _store_from_ptr(sp, ctx, ptr, n):
    import ctypes as C
    h = C.c_void_p()
    ctx._check(sp.lib().spmb_store_create(ctx._h, C.c_void_p(ptr), n, 0, C.byref(h)), "spmb_store_create")
    return h


def bench_myers(sp, synth, torch, ctx, dev, stream, hits, args, rank, world, hbm_peak):
    n, m, k = args.n_myers, 100, 4
    seed = synth.SEEDS["C3"] + rank
    text = synth.text_device(n, seed, dev)
    rng = np.random.default_rng(seed)
    s0 = int(rng.integers(0, n - m))
    pat = text[s0:s0 + m].cpu().numpy().copy()
    # 1000 planted copies, each with e in U{0..6} random edits (0..4 hit, 5..6 are decoys)
    starts = np.sort(rng.integers(0, n - 2 * m, size=1000)).astype(np.int64)
    for st in starts:
        mp = synth.mutate(rng, pat, int(rng.integers(0, 7)))
        text[int(st):int(st) + len(mp)] = torch.from_numpy(mp).to(dev)
    store = sp.Store.from_device(ctx, text.data_ptr(), n)
    ctx.sync()
    pattern = sp.Pattern(ctx, sp.ALG_MYERS, pat, k)
    flush = torch.zeros(64 << 20, dtype=torch.int32, device=dev)   # 256 MB > L2
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    with ClockSampler(torch.cuda.current_device()) as clk:
        total_ms, kern_ms, n_hits, launches = time_search(sp, torch, ctx, stream, store, pattern, hits, args.steps,
                                                          args.warmup, flush=flush)
    # the same search with the full recurrence (all ceil(m/32) words every column): the kernel the INT-pipe
    # roofline is stated on; the default above is its Ukkonen cut-off form (myers_cut.cuh)
    ctx.set_myers_cutoff(False)
    full_total_ms, full_kern_ms, full_hits, _ = time_search(sp, torch, ctx, stream, store, pattern, hits, args.steps,
                                                            args.warmup, flush=flush)
    ctx.set_myers_cutoff(True)
    if world > 1:
        t = torch.tensor([total_ms, full_total_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms, full_total_ms = float(t[0].item()), float(t[1].item())
    peak = ctx.measure_int_peak()
    ops = myers_ops_per_bp(m)
    ops_cut = 11            # one active word: 10 for the column update + the symbol pop (state of a warp on random text)
    achieved = n * ops / (full_kern_ms * 1e-3) / 1e9
    achieved_cut = n * ops_cut / (kern_ms * 1e-3) / 1e9
    res = {"workload": f"C3: Myers m={m} k={k} over {n} bp per GPU, 1000 planted mutated copies",
           "value": world * n * args.steps / (total_ms * 1e-3) / 1e9, "unit": "Gbp/s", "ms_per_step": total_ms / args.steps,
           "kernel_ms": kern_ms, "kernel_gbp_per_s": n / kern_ms / 1e6, "kernel": "myers_cut_kernel<4,plain>",
           "hits": n_hits, "row_len": store.row_len, "l2": "256 MB flush write between iterations",
           "roofline": {"bound": "int_alu", "achieved": achieved_cut, "peak": peak["lop3_gops"], "unit": "Ginstr/s",
                        "frac": achieved_cut / peak["lop3_gops"], "traffic": ncu_traffic("myers_cut_c3"),
                        "peak_source": "measured live: spmb_measure_int_peak (LOP3 chains, all SMs)",
                        "int_peak_iadd3": peak["iadd3_gops"], "algorithmic_int32_instr_per_bp": ops_cut,
                        "note": "cut-off kernel: a warp updates only the words that can hold a cell <= k; 11 = one word "
                                "(10) + symbol pop, the state of >= 99.9 % of the 8-column chunks on random DNA at k=4",
                        "kernel": "myers_cut_kernel<4,plain>"},
           "full_recurrence": {"value": world * n * args.steps / (full_total_ms * 1e-3) / 1e9, "unit": "Gbp/s",
                               "kernel_ms": full_kern_ms, "kernel_gbp_per_s": n / full_kern_ms / 1e6,
                               "hits_equal_cutoff": bool(full_hits == n_hits),
                               "roofline": {"bound": "int_alu", "achieved": achieved, "peak": peak["lop3_gops"],
                                            "unit": "Ginstr/s", "frac": achieved / peak["lop3_gops"],
                                            "traffic": ncu_traffic("myers_c3"),
                                            "algorithmic_int32_instr_per_bp": ops, "kernel": "myers_kernel<4,plain>"}},
           "speedup_cutoff_vs_full_kernel": full_kern_ms / kern_ms,
           "clocks": clk.summary()}
    if rank == 0 and world == 1 and not args.no_cpu:
        import oracle as orc
        host = text.cpu().numpy()
        threads = orc.use_all_cores()
        best = 1e30
        for _ in range(2):
            t0 = time.perf_counter()
            cnt, ce, cr = orc.search_mt(3, host, pat, k, threads=threads)
            best = min(best, time.perf_counter() - t0)
        g = sp.search(ctx, store, pattern, hits)
        res["cpu_baseline"] = {"value": n / best / 1e9, "unit": "Gbp/s", "cores": threads, "kind": "port",
                               "sample": f"whole {n} bp text, oracle Myers (64-bit blocks, Ukkonen cut-off), {threads} threads",
                               "hits_match_gpu": bool(np.array_equal(g["end"], ce) and np.array_equal(g["errors"], cr))}
    store.close()
    del text, flush
    torch.cuda.empty_cache()
    return res


def bench_c5(sp, synth, torch, ctx, dev, stream, hits, args, rank, world):
    """configs[4]: 10k patterns of ~100 bp, k=3, over a 3 Gbp text, in ONE call through the pigeonhole seed
    filter (SPMB_ALG_MYERS_FILTERED); with N ranks the SAME text is sharded (strong scaling), each rank
    searching shard rank/N, no collective.  Cross-checked against the direct Myers kernel (16 patterns,
    this rank's shard) and, at N=1, the CPU oracle on a 3 Mbp slice."""
    n, n_pat, m, k = args.n_c5, 10_000, 100, 3
    seed = synth.SEEDS["C5"]
    text = synth.text_device(n, seed, dev)
    rng = np.random.default_rng(seed)
    offs = np.sort(rng.integers(0, n - 2 * m, size=n_pat))
    idx = (torch.from_numpy(offs).to(dev)[:, None] + torch.arange(m, device=dev)[None, :]).reshape(-1)
    cut = text[idx].cpu().numpy().reshape(n_pat, m)
    pats = []
    for i in range(n_pat):
        q = synth.mutate(rng, cut[i], int(rng.integers(0, 6)))
        if len(q) < 97:
            q = np.concatenate([q, cut[i][:97 - len(q)]])
        pats.append(q[:128])
    store = sp.Store.from_device(ctx, text.data_ptr(), n)
    ctx.sync()
    pset = sp.Pattern(ctx, sp.ALG_MYERS_FILTERED, pats, k)
    ctx.set_shard(rank, world)
    sp.search(ctx, store, pset, hits)          # builds the lane-chunk copy of the store
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    total_ms, kern_ms, n_hits, launches = time_search(sp, torch, ctx, stream, store, pset, hits, args.steps, args.warmup)
    if world > 1:
        t = torch.tensor([total_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
        c = torch.tensor([n_hits], device=dev, dtype=torch.int64)
        dist.all_reduce(c)
        n_hits = int(c.item())
    got = sp.search(ctx, store, pset, hits)
    step = n_pat // 16
    sub = [pats[i] for i in range(0, n_pat, step)][:16]
    direct = sp.Pattern(ctx, sp.ALG_MYERS, sub, k)
    d = sp.search(ctx, store, direct, hits)
    sel = got[np.isin(got["pattern_id"], np.arange(0, n_pat, step)[:16])]
    same = bool(len(sel) == len(d) and np.array_equal(sel["pattern_id"] // step, d["pattern_id"]) and
                np.array_equal(sel["end"], d["end"]) and np.array_equal(sel["errors"], d["errors"]))
    ctx.set_shard(0, 1)
    res = {"workload": f"C5: {n_pat} patterns of ~{m} bp, k={k}, over {n} bp in one call (seed filter + verification); "
                       f"the text is sharded over {world} GPU(s)",
           "value": n * args.steps / (total_ms * 1e-3) / 1e9, "unit": "Gbp/s (x 10k patterns)", "ms_per_step": total_ms / args.steps,
           "kernels_ms": kern_ms, "cell_rate_T_per_s": n * n_pat * args.steps / (total_ms * 1e-3) / 1e12,
           "hits": n_hits, "filter": pset.filter_info, "scaling": "strong",
           "equals_direct_myers_kernel_16_patterns": same,
           "text_gbytes_per_s_kernels": n / world * 0.25 / (kern_ms * 1e-3) / 1e9}
    if rank == 0 and world == 1 and not args.no_cpu:
        import oracle as orc
        sl_n = 3_000_000
        sl0 = int(offs[n_pat // 2]) // 1024 * 1024
        sl0 = min(sl0, n - sl_n)
        host = text[sl0:sl0 + sl_n].cpu().numpy()
        st2 = sp.Store.from_host(ctx, host)
        g2 = sp.search(ctx, st2, pset, hits)
        ids = list(range(n_pat // 2 - 150, n_pat // 2 + 150))
        threads = orc.use_all_cores()
        t0 = time.perf_counter()
        bad = 0
        for i in ids:
            _, we, wr = orc.search_mt(3, host, pats[i], k, threads=threads)
            s2 = g2[g2["pattern_id"] == i]
            bad += int(not (np.array_equal(s2["end"], we) and np.array_equal(s2["errors"], wr)))
        cpu_s = time.perf_counter() - t0
        res["cpu_baseline"] = {"value": sl_n * len(ids) / cpu_s / 1e12, "unit": "T cells/s", "cores": threads, "kind": "port",
                               "sample": f"{len(ids)} patterns x a {sl_n} bp slice, oracle Myers, {threads} threads",
                               "patterns_mismatching_gpu": bad}
        st2.close()
    store.close()
    del text
    torch.cuda.empty_cache()
    return res


def bench_small(sp, synth, torch, ctx, dev, stream, hits, args):
    """C1 and C2 exactly as BASELINE.json specifies them (they fit in L2: flushed between iterations)."""
    import oracle as orc
    res = {}
    flush = torch.zeros(64 << 20, dtype=torch.int32, device=dev)
    # C1: Horspool m=32 over 1 Mbp
    n = 1_000_000
    text = synth.text_host(n, synth.SEEDS["C1"])
    pat = text[333_333:333_333 + 32].copy()
    rng = np.random.default_rng(synth.SEEDS["C1"])
    for st in rng.integers(0, n - 32, size=99):
        text[int(st):int(st) + 32] = pat
    store = sp.Store.from_host(ctx, text)
    p = sp.Pattern(ctx, sp.ALG_HORSPOOL, pat)
    tot, kern, nh, _ = time_search(sp, torch, ctx, stream, store, p, hits, args.steps, args.warmup, flush=flush)
    want = orc.horspool(text, pat)
    got = sp.search(ctx, store, p, hits)
    t0 = time.perf_counter(); orc.horspool(text, pat); cpu_s = time.perf_counter() - t0
    res["C1"] = {"workload": "Horspool m=32, 1 Mbp", "kernel_us": kern * 1e3, "step_us": tot / args.steps * 1e3,
                 "gbp_per_s_kernel": n / (kern * 1e-3) / 1e9, "hits": nh, "parity": bool(np.array_equal(got["end"], want)),
                 "cpu_1thread_gbp_per_s": n / cpu_s / 1e9}
    store.close()
    # C2: Shift-Or m=64 over 50 Mbp + 1% journal
    n = 50_000_000
    ref = synth.text_host(n, synth.SEEDS["C2"])
    ev, pool, kinds = synth.journal(n, synth.SEEDS["C2"])
    pool = synth.finish_snps(ev, pool, kinds, ref[ev["ref_pos"][kinds == 0].astype(np.int64)])
    base = sp.Store.from_host(ctx, ref)
    js = base.journaled(ev, pool)
    nd = len(js)
    pat = js.read(nd // 3, 64)
    p = sp.Pattern(ctx, sp.ALG_SHIFTOR, pat)
    tot, kern, nh, _ = time_search(sp, torch, ctx, stream, js, p, hits, args.steps, args.warmup, flush=flush)
    derived = orc.journal_apply(ref, ev.astype(orc.EVENT_DTYPE), pool)
    same_text = bool(np.array_equal(js.read(), derived))
    cnt, ce, _ = orc.search_mt(2, derived, pat, 0)
    got = sp.search(ctx, js, p, hits)
    res["C2"] = {"workload": "Shift-Or m=64, 50 Mbp + 1% journal", "kernel_us": kern * 1e3,
                 "step_us": tot / args.steps * 1e3, "gbp_per_s_kernel": nd / (kern * 1e-3) / 1e9, "hits": nh,
                 "journal_events": int(len(ev)), "derived_text_matches_oracle": same_text,
                 "parity": bool(np.array_equal(got["end"], ce))}
    js.close(); base.close()
    return res


# ------------------------------------------------------------------------------------------
def run_reference(args):
    """The reference's CPU algorithms for the path (the oracle port: SeqAn2 is not on this machine,
    SURVEY.md 8(c)), all host threads, bounded sample of the same workload per step."""
    rank, world, _ = dist_env()
    if rank != 0:
        return
    import oracle as orc
    from libspm_b200 import synth
    m = 64
    s_n = int(min(args.n, args.cpu_sample))
    seed = synth.SEEDS["C2"]
    ref = synth.text_host(s_n, seed)
    ev, pool, kinds = synth.journal(s_n, seed)
    pool = synth.finish_snps(ev, pool, kinds, ref[ev["ref_pos"][kinds == 0].astype(np.int64)])
    rng = np.random.default_rng(seed ^ 0xABCDEF)
    pat = rng.integers(0, 4, size=m, dtype=np.uint8)
    threads = orc.use_all_cores()
    times = []
    hits = 0
    evo = ev.astype(orc.EVENT_DTYPE)
    times, times_search = [], []
    hits = 0
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        derived = orc.journal_apply(ref, evo, pool)          # host buffers in -> journaled haystack
        t1 = time.perf_counter()
        hits, _, _ = orc.search_mt(2, derived, pat, 0, threads=threads, store=False)
        t2 = time.perf_counter()
        if i >= args.warmup:
            times.append(t2 - t0)
            times_search.append(t2 - t1)
    tot = sum(times)
    v = len(derived) * args.steps / tot / 1e9
    v_search = len(derived) * args.steps / sum(times_search) / 1e9
    sample = (f"{s_n} bp reference + 1% journal (bounded sample of the {args.n} bp workload); each step = host buffers "
              f"in, hit count out: oracle journal apply + oracle Shift-Or (32-bit blocks), both on {threads} threads")
    emit({
        "impl": "reference", "metric": "searched_gbp_per_s", "value": v, "unit": "Gbp/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": tot / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": "C2x: Shift-Or exact search m=64 over a journaled sequence (CPU, bounded sample)",
                   "searched_bp_per_step": int(len(derived)), "pattern_len": m, "hits": int(hits),
                   "search_only_gbp_per_s": v_search},
        "cpu_baseline": {"value": v, "unit": "Gbp/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "Gbp/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


def _claim_stdout():
    """libraries (NCCL prints its version) must not pollute the one-JSON-line contract: route fd 1
    to stderr for the whole run and hand back a private handle on the real stdout."""
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(os.dup(2), "w", buffering=1)
    return os.fdopen(real, "w", buffering=1)


_REAL_STDOUT = None


def emit(obj):
    _REAL_STDOUT.write(json.dumps(obj) + "\n")
    _REAL_STDOUT.flush()


def main():
    global _REAL_STDOUT
    _REAL_STDOUT = _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", type=int, default=3_200_000_000, help="reference symbols per GPU (C2x)")
    ap.add_argument("--n-myers", type=int, default=250_000_000)
    ap.add_argument("--cpu-sample", type=int, default=1_000_000_000)
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-myers", action="store_true")
    ap.add_argument("--no-small", action="store_true")
    ap.add_argument("--no-c5", action="store_true")
    ap.add_argument("--n-c5", type=int, default=3_000_000_000)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
