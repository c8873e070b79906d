"""stream leg alone (bench.py's `stream` key) at a few chunk sizes; SPMB_UPLOAD_TRACE=1 for per-job timings.
    python tools/stream_ab.py [n] [chunk ...]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import libspm_b200 as sp
from libspm_b200 import synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 3_200_000_000
chunks = [int(x) for x in sys.argv[2:]] or [64 << 20]
host = torch.empty(n, dtype=torch.uint8, pin_memory=True)
synth.text_into(host.numpy(), 0x5EED0002)
ctx = sp.Context(0)
hits = sp.Hits(ctx)
pat = host.numpy()[n // 3:n // 3 + 64].copy()
out = {}
for c in chunks:
    r = sp.stream_leg(ctx, host.numpy(), pat, hits, chunk=c)
    out[str(c)] = r
    print(c, json.dumps({k: r[k] for k in ("value", "one_shot_gbp_per_s", "ratio_to_one_shot", "sync_feed_gbp_per_s", "hits_equal_one_shot")}), flush=True)
# Myers stream (lane chunks built as well)
r = sp.stream_leg(ctx, host.numpy()[:1_600_000_000], host.numpy()[1000:1100].copy(), hits, chunk=64 << 20, k=4, alg=sp.ALG_MYERS)
out["myers_64Mi"] = r
print("myers", json.dumps({k: r[k] for k in ("value", "one_shot_gbp_per_s", "ratio_to_one_shot", "hits_equal_one_shot")}), flush=True)
if len(sys.argv) > 1:
    json.dump(out, open("gpurun_out/stream_ab.json", "w"), indent=1)
