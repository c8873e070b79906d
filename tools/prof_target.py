"""small fixed target for ncu: a few launches of the exact and Myers kernels on resident stores.
    python tools/prof_target.py [exact|myers|both] [n_exact] [n_myers]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import libspm_b200 as sp
from libspm_b200 import synth

what = sys.argv[1] if len(sys.argv) > 1 else "both"
n_exact = int(sys.argv[2]) if len(sys.argv) > 2 else 3_200_000_000
n_myers = int(sys.argv[3]) if len(sys.argv) > 3 else 250_000_000
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev)
with torch.cuda.stream(stream):
    ctx = sp.Context(0, stream=stream.cuda_stream)
    hits = sp.Hits(ctx)
    if what in ("exact", "both"):
        t = synth.text_device(n_exact, 1, dev)
        st = sp.Store.from_device(ctx, t.data_ptr(), n_exact); ctx.sync(); del t
        pat = st.read(n_exact // 3, 64)
        p = sp.Pattern(ctx, sp.ALG_SHIFTOR, pat)
        for _ in range(6):
            h = sp.search(ctx, st, p, hits)
        print("exact hits", len(h), "row_len", st.row_len)
        st.close()
    if what in ("myers", "both"):
        t = synth.text_device(n_myers, 2, dev)
        st = sp.Store.from_device(ctx, t.data_ptr(), n_myers); ctx.sync(); del t
        pat = st.read(n_myers // 3, 100)
        p = sp.Pattern(ctx, sp.ALG_MYERS, pat, 4)
        for _ in range(6):
            h = sp.search(ctx, st, p, hits)
        print("myers (cut-off) hits", len(h), "row_len", st.row_len)
        ctx.set_myers_cutoff(False)      # the full recurrence: myers_kernel<4,plain>
        for _ in range(6):
            h = sp.search(ctx, st, p, hits)
        ctx.set_myers_cutoff(True)
        print("myers (full recurrence) hits", len(h))
        st.close()
    if what == "filter":
        # C5 shape: 10k patterns of 100 bp, k=3 through the pigeonhole seed filter
        n_f = n_exact
        t = synth.text_device(n_f, 5, dev)
        rng = np.random.default_rng(5)
        offs = np.sort(rng.integers(0, n_f - 200, size=10_000))
        idx = (torch.from_numpy(offs).to(dev)[:, None] + torch.arange(100, device=dev)[None, :]).reshape(-1)
        cut = t[idx].cpu().numpy().reshape(-1, 100)
        pats = [synth.mutate(rng, c, int(rng.integers(0, 4))) for c in cut]
        pats = [np.concatenate([q, c[:max(0, 97 - len(q))]]) for q, c in zip(pats, cut)]
        st = sp.Store.from_device(ctx, t.data_ptr(), n_f); ctx.sync(); del t
        p = sp.Pattern(ctx, sp.ALG_MYERS_FILTERED, pats, 3)
        import time
        for _ in range(4):
            h = sp.search(ctx, st, p, hits)
        print("filtered hits", len(h), p.filter_info)
        for _ in range(3):
            torch.cuda.synchronize(); t0 = time.perf_counter()
            sp.search_launch(ctx, st, p, hits); t1 = time.perf_counter()
            ctx.sync(); t2 = time.perf_counter()
            h = sp.search_finish(ctx, hits); t3 = time.perf_counter()
            print(f"launch {1e3*(t1-t0):.3f} ms  kernels {1e3*(t2-t1):.3f} ms  finish {1e3*(t3-t2):.3f} ms  total {1e3*(t3-t0):.3f} ms")
        # pipelined: two passes queued behind the one being collected; host time of every call
        hl = [hits, sp.Hits(ctx, 1 << 16), sp.Hits(ctx, 1 << 16)]
        for hh in hl + hl:
            sp.search(ctx, st, p, hh)
        steps = 12
        torch.cuda.synchronize(); T0 = time.perf_counter()
        log = []
        for i in range(2):
            a0 = time.perf_counter(); sp.search_launch(ctx, st, p, hl[i % 3]); log.append(("L", i, a0 - T0, time.perf_counter() - T0))
        for i in range(steps):
            if i + 2 < steps:
                a0 = time.perf_counter(); sp.search_launch(ctx, st, p, hl[(i + 2) % 3]); log.append(("L", i + 2, a0 - T0, time.perf_counter() - T0))
            a0 = time.perf_counter(); sp.search_finish(ctx, hl[i % 3]); log.append(("F", i, a0 - T0, time.perf_counter() - T0))
        torch.cuda.synchronize(); T1 = time.perf_counter()
        print(f"pipelined depth 3: {1e3*(T1-T0)/steps:.3f} ms per step")
        for kind, i, a, b in log:
            print(f"  {kind}{i:2d}  start {1e3*a:7.3f}  end {1e3*b:7.3f}  ({1e3*(b-a):.3f} ms)")
        st.close()
    if what == "e2e":
        # per-call times of the headline e2e path: host ranks + journal -> hits
        import time
        n = n_exact
        ref = synth.text_device(n, 2, dev)
        ev, pool, kinds = synth.journal(ref.cpu().numpy(), 2)
        host = torch.empty(n, dtype=torch.uint8, pin_memory=True); host.copy_(ref); del ref
        evp = torch.empty(ev.nbytes, dtype=torch.uint8, pin_memory=True); evp.numpy()[:] = ev.view(np.uint8)
        evh = evp.numpy().view(sp.EVENT_DTYPE)
        pp = torch.empty(pool.nbytes, dtype=torch.uint8, pin_memory=True); pp.numpy()[:] = pool
        pat = np.random.default_rng(1).integers(0, 4, 64, dtype=np.uint8)
        p = sp.Pattern(ctx, sp.ALG_SHIFTOR, pat)
        for it in range(4):
            torch.cuda.synchronize(); t0 = time.perf_counter()
            b = sp.Store.from_host(ctx, host.numpy()); t1 = time.perf_counter()
            j = b.journaled(evh, pp.numpy()); t2 = time.perf_counter()
            h = sp.search(ctx, j, p, hits); t3 = time.perf_counter()
            j.close(); b.close(); torch.cuda.synchronize(); t4 = time.perf_counter()
            print(f"store_create {1e3*(t1-t0):.2f} ms ({n/(t1-t0)/1e9:.1f} GB/s)  journaled {1e3*(t2-t1):.2f} ms ({(ev.nbytes+pool.nbytes)/1e6:.0f} MB up)  search {1e3*(t3-t2):.3f} ms  close {1e3*(t4-t3):.2f} ms  total {1e3*(t4-t0):.2f} ms")
