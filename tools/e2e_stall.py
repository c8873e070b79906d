"""distribution of the search call's host time in the e2e sequence (store_create, journaled, search, close)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import libspm_b200 as sp
from libspm_b200 import synth
n = 3_200_000_000
dev = torch.device("cuda", 0)
ctx = sp.Context(0)
hits = sp.Hits(ctx)
host = torch.empty(n, dtype=torch.uint8, pin_memory=True)
cfg = synth.c2(n, seed=synth.SEEDS["C2"], m=64, ref=host.numpy())
ev, pool, pat = cfg["events"], cfg["pool"], cfg["pattern"]
evp = torch.empty(ev.nbytes, dtype=torch.uint8, pin_memory=True); evp.numpy()[:] = ev.view(np.uint8)
evh = evp.numpy().view(sp.EVENT_DTYPE)
p = sp.Pattern(ctx, sp.ALG_SHIFTOR, pat)
for it in range(int(sys.argv[1]) if len(sys.argv) > 1 else 24):
    t0 = time.perf_counter()
    b = sp.Store.from_host(ctx, host.numpy()); t1 = time.perf_counter()
    j = b.journaled(evh, pool); t2 = time.perf_counter()
    sp.search_launch(ctx, j, p, hits); t3 = time.perf_counter()
    ctx.sync(); t4 = time.perf_counter()
    h = sp.search_finish(ctx, hits); t5 = time.perf_counter()
    j.close(); b.close(); t6 = time.perf_counter()
    print(f"{it:2d} create {1e3*(t1-t0):6.2f} journaled {1e3*(t2-t1):6.2f} launch {1e3*(t3-t2):7.3f} sync {1e3*(t4-t3):7.3f} finish {1e3*(t5-t4):7.3f} close {1e3*(t6-t5):6.2f} hits {len(h)}", flush=True)
