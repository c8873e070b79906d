"""does the nvidia-smi clock sampler disturb the timed region?  The headline's timed region (20 passes, ~2.8 ms) run
many times with the sampler of bench.py on (every 50 ms, every 200 ms) and off; prints median / max / outliers."""
import os, sys, json, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
import libspm_b200 as sp
from libspm_b200 import synth

n = 800_000_000
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev)
res = {}
with torch.cuda.stream(stream):
    ctx = sp.Context(0, stream=stream.cuda_stream)
    hits = sp.Hits(ctx, 1 << 16)
    text = synth.text_host(n, 7)
    pat = text[12345:12345 + 64].copy()
    store = sp.Store.from_host(ctx, text)
    p = sp.Pattern(ctx, sp.ALG_SHIFTOR, pat, 0)
    sp.search(ctx, store, p, hits)
    for mode in ("off", "50", "200", "off", "50"):
        smp = None
        if mode != "off":
            bench.ClockSampler.PERIOD_MS = mode
            smp = bench.ClockSampler(0)
            smp.__enter__()
            time.sleep(0.3)
        tot = []
        t_end = time.perf_counter() + 6.0
        while time.perf_counter() < t_end:
            t, k, nh, _ = bench.time_search(sp, torch, ctx, stream, store, p, hits, 20, 3)
            tot.append(t)
        if smp:
            smp.__exit__()
        tot = np.array(tot)
        med = float(np.median(tot))
        r = {"regions": len(tot), "median_ms": round(med, 4), "p99_ms": round(float(np.percentile(tot, 99)), 4), "max_ms": round(float(tot.max()), 4),
             "over_1.1x_median": int((tot > 1.1 * med).sum()), "over_1.3x_median": int((tot > 1.3 * med).sum()),
             "samples": (len(smp.lines) if smp else 0)}
        res.setdefault(mode, []).append(r)
        print(mode, r, flush=True)
json.dump(res, open("gpurun_out/sampler_jitter.json", "w"), indent=1)
