"""C5 strong scaling measured on ONE GPU: rank r of N searches shard r/N of the same 3 Gbp store, so
ctx.set_shard(0, N) on one device does exactly the per-rank work of an N-GPU run (no collective is involved).
Prints ms/step (pipelined launch/finish as in bench.py) and kernel ms for N = 1, 2, 4, 8.
    python tools/c5_shard.py [--n 3000000000] [--patterns 10000] [--out profiles/r02_c5_shards.json]"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import bench
import libspm_b200 as sp
from libspm_b200 import synth

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=3_000_000_000)
ap.add_argument("--patterns", type=int, default=10_000)
ap.add_argument("--steps", type=int, default=20)
ap.add_argument("--out", default="profiles/r02_c5_shards.json")
ap.add_argument("--only", type=int, default=0, help="one shard count only (launch lists)")
args = ap.parse_args()
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev)
res = {"n": args.n, "patterns": args.patterns, "shards": {}}
with torch.cuda.stream(stream):
    ctx = sp.Context(0, stream=stream.cuda_stream)
    hits = sp.Hits(ctx, 1 << 16)
    text = synth.text_host(args.n, synth.SEEDS["C5"])
    offs, pats = synth.c5_patterns(lambda o, m: np.stack([text[int(x):int(x) + m] for x in o]), args.n, args.patterns, 100)
    store = sp.Store.from_host(ctx, text)
    del text
    pset = sp.Pattern(ctx, sp.ALG_MYERS_FILTERED, pats, 3)
    sp.search(ctx, store, pset, hits)
    for N in ((args.only,) if args.only else (1, 2, 4, 8)):
        worst = None
        for r in sorted({0, N // 2, N - 1}):
            ctx.set_shard(r, N)
            tot, kern, nh, launches = bench.time_search(sp, torch, ctx, stream, store, pset, hits, args.steps, 3)
            cur = {"rank": r, "ms_per_step": tot / args.steps, "kernels_ms": kern, "hits": nh, "launches": launches}
            if worst is None or cur["ms_per_step"] > worst["ms_per_step"]:
                worst = cur
        res["shards"][str(N)] = worst
        print(N, worst, flush=True)
    ctx.set_shard(0, 1)
    one = res["shards"].get("1", {}).get("ms_per_step", 0.0)
    for N in (() if args.only else (2, 4, 8)):
        res["shards"][str(N)]["efficiency"] = one / (N * res["shards"][str(N)]["ms_per_step"])
os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
json.dump(res, open(args.out, "w"), indent=1)
print(json.dumps(res))
