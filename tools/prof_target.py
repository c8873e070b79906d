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
        print("myers hits", len(h), "row_len", st.row_len)
        st.close()
