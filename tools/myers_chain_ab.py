"""lane-chained Myers scan (one warp per row, words across the lanes) against the one-lane kernels at the same needle:
    SPMB_MYERS_CHAIN=1 python tools/myers_chain_ab.py   vs   python tools/myers_chain_ab.py
prints kernel ms for m in (100, 256) over 250 Mbp and checks the hit lists against each other via a digest."""
import hashlib, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import libspm_b200 as sp
from libspm_b200 import synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 250_000_000
dev = torch.device("cuda", 0)
ctx = sp.Context(0)
hits = sp.Hits(ctx)
t = synth.text_device(n, 3, dev)
st = sp.Store.from_device(ctx, t.data_ptr(), n); ctx.sync(); del t
res = {"chain": bool(os.environ.get("SPMB_MYERS_CHAIN")), "n": n}
for m, k in ((100, 4), (256, 10)):
    pat = st.read(n // 3, m)
    pat[m // 2] ^= 1
    p = sp.Pattern(ctx, sp.ALG_MYERS, pat, k)
    for full in (False, True):
        ctx.set_myers_cutoff(not full)
        h = sp.search(ctx, st, p, hits)
        ctx.sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            sp.search_launch(ctx, st, p, hits); sp.search_finish(ctx, hits)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 3
        dig = hashlib.sha256(np.ascontiguousarray(h[["end", "errors"]]).tobytes()).hexdigest()[:16]
        res[f"m{m}_k{k}_{'full' if full else 'cut'}"] = {"ms_per_search": ms, "hits": int(len(h)), "digest": dig}
    ctx.set_myers_cutoff(True)
    p.close()
print(json.dumps(res))
