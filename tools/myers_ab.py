"""Myers kernel timing on resident stores (CUDA events on the context's stream), for A/B runs:
    python tools/myers_ab.py            # cut-off kernel (default)
    SPMB_MYERS_FULL=1 python tools/myers_ab.py   # full recurrence
Cases: C3 (250 Mbp, m=100, k=4), m=150 k=2 (C4's needle) and k=8 on random text, and the worst case for the
cut-off (poly-A text and needle: every word active)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import libspm_b200 as sp
from libspm_b200 import synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 250_000_000
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev)
out = {"full_recurrence": bool(os.environ.get("SPMB_MYERS_FULL")), "n": n, "cases": []}
with torch.cuda.stream(stream):
    ctx = sp.Context(0, stream=stream.cuda_stream)
    hits = sp.Hits(ctx)
    t = synth.text_device(n, 2, dev)
    st = sp.Store.from_device(ctx, t.data_ptr(), n); ctx.sync()
    for (m, k) in ((100, 4), (150, 2), (100, 8), (64, 3), (250, 10)):
        pat = st.read(n // 3, m)
        p = sp.Pattern(ctx, sp.ALG_MYERS, pat, k)
        for _ in range(2):
            h = sp.search(ctx, st, p, hits)
        ms = []
        for _ in range(5):
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            e0.record(stream); sp.search_launch(ctx, st, p, hits); e1.record(stream)
            h = sp.search_finish(ctx, hits)
            ms.append(e0.elapsed_time(e1))
        out["cases"].append({"text": "random", "m": m, "k": k, "hits": len(h), "ms": min(ms), "gbp_per_s": n / min(ms) / 1e6})
    st.close(); del t
    n2 = n // 8
    t = torch.zeros(n2, dtype=torch.uint8, device=dev)
    st = sp.Store.from_device(ctx, t.data_ptr(), n2); ctx.sync()
    p = sp.Pattern(ctx, sp.ALG_MYERS, np.ones(100, dtype=np.uint8), 4)   # no hit, but rows stay shallow? no: all mismatch
    for name, needle in (("polyA/needle C", np.ones(100, dtype=np.uint8)),):
        p = sp.Pattern(ctx, sp.ALG_MYERS, needle, 4)
        h = sp.search(ctx, st, p, hits)
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(stream); sp.search_launch(ctx, st, p, hits); e1.record(stream)
        h = sp.search_finish(ctx, hits)
        out["cases"].append({"text": name, "m": 100, "k": 4, "hits": len(h), "n": n2, "ms": e0.elapsed_time(e1),
                             "gbp_per_s": n2 / e0.elapsed_time(e1) / 1e6})
    # worst case: periodic text, needle from it with 5 substitutions spread out -> deep active region, few hits
    per = torch.arange(n2, device=dev, dtype=torch.int64)
    t = ((per * 2654435761 >> 7) % 4).to(torch.uint8)
    blk = t[:1000].clone()
    t = blk.repeat(n2 // 1000 + 1)[:n2].contiguous()
    st2 = sp.Store.from_device(ctx, t.data_ptr(), n2); ctx.sync()
    needle = blk[100:200].cpu().numpy().copy()
    for i in (10, 30, 50, 70, 90):
        needle[i] = (needle[i] + 1) % 4
    p = sp.Pattern(ctx, sp.ALG_MYERS, needle, 4)
    h = sp.search(ctx, st2, p, hits)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(stream); sp.search_launch(ctx, st2, p, hits); e1.record(stream)
    h = sp.search_finish(ctx, hits)
    out["cases"].append({"text": "period-1000, needle = 5 substitutions off", "m": 100, "k": 4, "hits": len(h), "n": n2,
                         "ms": e0.elapsed_time(e1), "gbp_per_s": n2 / e0.elapsed_time(e1) / 1e6})
print(json.dumps(out))
