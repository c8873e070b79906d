"""per-call host times of a short submit/collect stream (SPMB_UPLOAD_TRACE=1 adds the movers' per-job lines)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import libspm_b200 as sp
from libspm_b200 import synth

chunk = int(sys.argv[1]) if len(sys.argv) > 1 else 64 << 20
n_chunks = int(sys.argv[2]) if len(sys.argv) > 2 else 10
n = chunk * n_chunks
host = torch.empty(n, dtype=torch.uint8, pin_memory=True)
synth.text_into(host.numpy(), 0x5EED0002)
ctx = sp.Context(0)
pat = host.numpy()[n // 3:n // 3 + 64].copy()
p = sp.Pattern(ctx, sp.ALG_SHIFTOR, pat)
base = host.numpy().ctypes.data
for rep in range(2):
    st = sp.Stream(ctx, p)
    log = []
    ctx.sync()
    T0 = time.perf_counter()
    inflight = 0
    for i in range(n_chunks):
        if inflight == 2:
            a = time.perf_counter(); st.collect(); log.append(("C", a - T0, time.perf_counter() - T0)); inflight -= 1
        a = time.perf_counter(); st.submit(ptr=base + i * chunk, n=chunk); log.append(("S", a - T0, time.perf_counter() - T0)); inflight += 1
    while inflight:
        a = time.perf_counter(); st.collect(); log.append(("C", a - T0, time.perf_counter() - T0)); inflight -= 1
    T1 = time.perf_counter()
    st.close()
    print(f"rep {rep}: {n / (T1 - T0) / 1e9:.1f} Gbp/s, {1e3 * (T1 - T0) / n_chunks:.3f} ms per chunk", flush=True)
    if rep == 1:
        for kind, a, b in log:
            print(f"  {kind} start {1e3 * a:8.3f} end {1e3 * b:8.3f} ({1e3 * (b - a):.3f} ms)")
