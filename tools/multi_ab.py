"""the e2e_multi leg of bench.py alone: ONE 3.2 Gbp host haystack (+ journal) through the single-process multi-GPU
driver at 1, 2, 4, 8 devices (as many as the box shows).
    python tools/multi_ab.py [n] [--threads T]   (T = SPMB_MULTI_HOST_THREADS: packing threads per device)"""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
import libspm_b200 as sp
from libspm_b200 import synth

ap = argparse.ArgumentParser()
ap.add_argument("n", type=int, nargs="?", default=3_200_000_000)
ap.add_argument("--threads", type=int, default=-1)
ap.add_argument("--out", default="gpurun_out/multi_ab.json")
a = ap.parse_args()
if a.threads >= 0:
    os.environ["SPMB_MULTI_HOST_THREADS"] = str(a.threads)
host = torch.empty(a.n, dtype=torch.uint8, pin_memory=True)
cfg = synth.c2(a.n, seed=synth.SEEDS["C2"], m=64, ref=host.numpy())
ev, pool, pat = cfg["events"], cfg["pool"], cfg["pattern"]
# like bench.py's e2e legs: all of the step's host inputs in pinned memory (reference ranks, journal events, insertion pool)
ev_pin = torch.empty(ev.nbytes, dtype=torch.uint8, pin_memory=True)
ev_pin.numpy()[:] = ev.view(np.uint8)
ev = ev_pin.numpy().view(sp.EVENT_DTYPE)
pool_pin = torch.empty(max(pool.nbytes, 1), dtype=torch.uint8, pin_memory=True)
pool_pin.numpy()[:pool.nbytes] = pool
pool = pool_pin.numpy()[:pool.nbytes]
# derived length: reference + sum(ins - del)
n_der = int(a.n + (ev["ins_len"].astype(np.int64) - ev["del_len"].astype(np.int64)).sum())
args = argparse.Namespace(steps=5, e2e_steps=5)
res = {}
nd = torch.cuda.device_count()
for n_dev in [d for d in (1, 2, 4, 8) if d <= nd]:
    r = bench.bench_multi(sp, torch, args, n_dev, host.numpy(), ev, pool, pat, n_der)
    res[str(n_dev)] = r
    print(n_dev, json.dumps({"e2e_gbp_s": round(r["value"], 1), "ms": round(r["ms_per_step"], 2), "phase_ms": {k: round(v, 2) for k, v in r["phase_ms"].items()},
                             "plain_gbp_s": round(r["plain"]["value"], 1), "plain_ms": round(r["plain"]["ms_per_step"], 2), "hits": r["hits"]}), flush=True)
json.dump(res, open(a.out, "w"), indent=1)
