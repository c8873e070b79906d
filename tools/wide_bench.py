"""4-bit alphabets (spm::dna5 / spm::dna15) on resident stores: exact search (HBM-bound, 0.5 B per bp: two plane
pairs) and Myers (cut-off kernel), CUDA-event timed; hit lists checked against the CPU oracle on a slice.
    python tools/wide_bench.py [n_exact] [n_myers]   -> one JSON line"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import libspm_b200 as sp
import oracle as orc

n_exact = int(sys.argv[1]) if len(sys.argv) > 1 else 1_600_000_000
n_myers = int(sys.argv[2]) if len(sys.argv) > 2 else 250_000_000
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev)
peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"] \
    if os.path.exists(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else 6650.0
out = {"hbm_peak_gbs": peak, "cases": []}


def timed(ctx, st, p, hits, reps=5):
    for _ in range(2):
        h = sp.search(ctx, st, p, hits)
    ms = []
    for _ in range(reps):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(stream); sp.search_launch(ctx, st, p, hits); e1.record(stream)
        h = sp.search_finish(ctx, hits)
        ms.append(e0.elapsed_time(e1))
    return float(np.mean(ms)), h


with torch.cuda.stream(stream):
    ctx = sp.Context(0, stream=stream.cuda_stream)
    hits = sp.Hits(ctx)
    for sigma, bits, name in ((5, 3, "dna5"), (15, 4, "dna15")):
        g = torch.Generator(device=dev); g.manual_seed(sigma)
        t = torch.randint(0, sigma, (n_exact,), dtype=torch.uint8, device=dev, generator=g)
        pat = t[n_exact // 3:n_exact // 3 + 64].cpu().numpy().copy()
        for o in (12345, n_exact // 2, n_exact - 1000):
            t[o:o + 64] = torch.from_numpy(pat).to(dev)
        st = sp.Store.from_device_wide(ctx, t.data_ptr(), n_exact, bits); ctx.sync()
        p = sp.Pattern(ctx, sp.ALG_SHIFTOR, pat, bits=bits)
        ms, h = timed(ctx, st, p, hits)
        sl = t[:20_000_000].cpu().numpy()
        want = orc.search_mt(2, sl, pat)[1]
        ok = bool(np.array_equal(h["end"][h["end"] <= len(sl)], want))
        out["cases"].append({"alphabet": name, "search": "Shift-Or m=64", "n": n_exact, "kernel_ms": ms, "hits": len(h),
                             "tbp_per_s": n_exact / ms / 1e9, "algorithmic_bytes_per_bp": 0.5,
                             "achieved_gbs": n_exact * 0.5 / ms / 1e6, "hbm_frac": n_exact * 0.5 / ms / 1e6 / peak,
                             "parity_first_20Mbp": ok})
        st.close(); p.close()
        tm = t[:n_myers].contiguous()
        pat = tm[n_myers // 3:n_myers // 3 + 100].cpu().numpy().copy()
        st = sp.Store.from_device_wide(ctx, tm.data_ptr(), n_myers, bits); ctx.sync()
        p = sp.Pattern(ctx, sp.ALG_MYERS, pat, 4, bits=bits)
        ms, h = timed(ctx, st, p, hits)
        sl = tm[:20_000_000].cpu().numpy()
        _, we, wr = orc.search_mt(3, sl, pat, 4)
        keep = h["end"] <= len(sl)
        ok = bool(np.array_equal(h["end"][keep], we) and np.array_equal(h["errors"][keep], wr))
        out["cases"].append({"alphabet": name, "search": "Myers m=100 k=4 (cut-off kernel)", "n": n_myers, "kernel_ms": ms,
                             "hits": len(h), "gbp_per_s": n_myers / ms / 1e6, "parity_first_20Mbp": ok})
        st.close(); p.close(); del t, tm
        torch.cuda.empty_cache()
print(json.dumps(out))
