#!/usr/bin/env python
"""bench.py -- searched Gbp/s of the online pattern-matching hot path on B200.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU algorithms (oracle port)

A "step" is one search pass (kernel launch -> sorted hit list on the host) over one resident sequence store; at N > 1 the
hit lists of all passes are gathered to every rank behind the last pass, inside the timed region (HitGatherShm).  Headline workload = BASELINE.json configs[1] at
steady-state scale ("C2x"): Shift-Or exact search, m=64, over a journaled sequence (reference + 1% SNP/indel
journal), N = 3.2 Gbp per GPU so that the packed text (0.25 B/bp = 800 MB) exceeds the 126 MB L2.
The same line also carries: the end-to-end legs (host buffers in, hits out: one GPU per rank, ONE host haystack
over all GPUs through the single-process multi-GPU driver, chunked streaming), Myers k-error (configs[2], "C3"),
the JST config (configs[3], "C4"), the batched multi-pattern config (configs[4], "C5") and the as-specified
small configs C1/C2.  All inputs come from the one xoshiro256** generator (libspm_b200/synth.py).
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import gc
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BYTES_PER_BP = 0.25                 # 2-bit packed text read once
M_EXACT = 64


def myers_ops_per_bp(m):            # INT32 instructions per (bp x pattern), SURVEY.md 8(d)
    return 10 * ((m + 31) // 32) + 4


def headline_config(n):
    """identical in both arms (--impl ours / reference): the workload, not the run."""
    return {"workload": f"C2x: Shift-Or exact search m={M_EXACT} over a journaled sequence ({n} bp reference + 1% SNP/indel "
                        "journal; needle cut from the derived sequence, 100 more copies planted, 25 of them across an edit) "
                        "per GPU; BASELINE configs[1] generator (xoshiro256**, seed 0x5EED0002 + rank) at steady-state scale",
            "pattern_len": M_EXACT, "reference_bp_per_gpu": n, "journal_density": 0.01,
            "l2": f"inputs larger than L2 ({n * BYTES_PER_BP / 1e6:.0f} MB packed text vs 126 MB)"}


# ------------------------------------------------------------------------------------------
def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)", d
    return 6650.0, "fallback (B200_PROFILING.md)", {}


def ncu_capture(key):
    """figures of the committed ncu --set full captures (profiles/roofline_traffic.json): constants READ FROM
    THE PROFILE, not measured in this run -- every use is labelled so."""
    p = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get(key)
        except Exception:
            return None
    return None


class ClockSampler:
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    PERIOD_MS = "50"

    def __init__(self, index):
        self.index = index
        self.lines = []
        self.proc = None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", str(self.PERIOD_MS), "-i", str(self.index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        return self

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.06)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons, pw = [], [], set(), []
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(pw)}


# ------------------------------------------------------------------------------------------
def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


class HitGather:
    """the final gather of the ranks' small hit lists (north_star item 4: "only a final gather of the
    small hit lists"), INSIDE the timed region: every pass's sorted hit list is copied into a pinned buffer on its
    rank as it arrives; when the last pass has been collected the lists of all passes travel in one piece -- pinned
    -> device -> ONE all_gather -> pinned -- and every rank holds the merged lists before the region's closing event
    is recorded.  No collective touches the search itself.  Buffers are sized during warm-up (largest per-pass list
    over the ranks, with slack).  (A gather per pass was measured first: five torch calls and a collective per
    0.14 ms pass made the step host-bound, 0.31 ms at N = 8.)"""
    def __init__(self, torch, dist, dev, world):
        self.torch, self.dist, self.dev, self.world = torch, dist, dev, world
        self.stream = torch.cuda.Stream(device=dev)
        self.cap = 0
        self.gathers = 0
        self.overflow = False

    def prepare(self, n_first, passes):
        """warm-up: agree on the per-pass capacity, allocate, run the collective once"""
        torch, dist = self.torch, self.dist
        with torch.cuda.stream(self.stream):
            c = torch.tensor([int(n_first)], dtype=torch.int64, device=self.dev)
            dist.all_reduce(c, op=dist.ReduceOp.MAX)
            self.cap = int(c.item()) * 5 // 4 + 64
        self.passes = passes
        self.rec_bytes = self.cap * 24
        nbytes = passes * (8 + self.rec_bytes)
        self.pin_in = torch.zeros(nbytes, dtype=torch.uint8, pin_memory=True)
        self.dev_in = torch.zeros(nbytes, dtype=torch.uint8, device=self.dev)
        self.dev_out = torch.zeros(nbytes * self.world, dtype=torch.uint8, device=self.dev)
        self.pin_out = torch.zeros(nbytes * self.world, dtype=torch.uint8, pin_memory=True)
        self.nbytes = nbytes
        self.finish()

    def add(self, i, hits):
        n = len(hits)
        if n > self.cap:
            self.overflow = True
            n = self.cap
        buf = self.pin_in.numpy()
        o = i * (8 + self.rec_bytes)
        buf[o:o + 8].view(np.uint64)[0] = n
        if n:
            buf[o + 8:o + 8 + n * 24] = hits[:n].view(np.uint8).reshape(-1)

    def finish(self):
        """all ranks: one all_gather of every pass's list; returns the merged record count over ranks and passes"""
        torch = self.torch
        with torch.cuda.stream(self.stream):
            self.dev_in.copy_(self.pin_in, non_blocking=True)
            self.dist.all_gather_into_tensor(self.dev_out, self.dev_in)
            self.pin_out.copy_(self.dev_out, non_blocking=True)
        self.stream.synchronize()
        counts = self.pin_out.numpy().reshape(self.world * self.passes, 8 + self.rec_bytes)[:, :8].copy().view(np.uint64)
        self.gathers += 1
        return int(counts.sum())

    def barrier(self):
        self.dist.barrier()

    def merged(self, hit_dtype, i=0):
        """the merged list of pass i (rank-major), for checks outside the timed region"""
        raw = self.pin_out.numpy()
        parts = []
        for r in range(self.world):
            o = r * self.nbytes + i * (8 + self.rec_bytes)
            n = int(raw[o:o + 8].view(np.uint64)[0])
            parts.append(raw[o + 8:o + 8 + n * 24].view(hit_dtype))
        return np.concatenate(parts)


class HitGatherShm:
    """the same final gather for ranks that share ONE host (torchrun, one node: LOCAL_WORLD_SIZE == WORLD_SIZE).
    spmb_search_finish has already delivered every pass's sorted list to HOST memory, so the gather is a host-side
    exchange: every rank writes its lists straight into its slot of a POSIX shared-memory segment (/dev/shm) as the
    passes arrive, publishes a generation flag behind its last pass and waits for the flags of the others -- from then
    on every rank holds (has mapped) the lists of all ranks and passes.  No device bounce: the NCCL form (HitGather)
    moved host -> device -> all_gather -> host and cost 0.17 ms per timed region for 101-hit lists and 0.4-1.2 ms for the
    C5 lists (profiles/r02_gather_n2.txt).  Two generations of slots alternate, so a rank that is one gather ahead never
    overwrites what a slower rank still reads.  x86 only (stores stay in program order: data, then flag); other
    hosts and multi-node jobs take the NCCL form."""
    FLAG = 64

    def __init__(self, torch, dist, dev, world, rank):
        self.torch, self.dist, self.dev, self.world, self.rank = torch, dist, dev, world, rank
        self.gen = 0
        self.gathers = 0
        self.overflow = False
        self.mm = None
        self._fb = None      # the NCCL form, when /dev/shm has no room for the segment

    @staticmethod
    def usable(world):
        import platform
        return (platform.machine() in ("x86_64", "AMD64") and os.path.isdir("/dev/shm")
                and int(os.environ.get("LOCAL_WORLD_SIZE", "0")) == world
                and os.environ.get("SPMB_BENCH_GATHER", "shm") == "shm")

    def prepare(self, n_first, passes):
        """warm-up (untimed): agree on the per-pass capacity, create and map the segment, run one gather"""
        torch, dist = self.torch, self.dist
        c = torch.tensor([int(n_first)], dtype=torch.int64, device=self.dev)
        dist.all_reduce(c, op=dist.ReduceOp.MAX)
        self.cap = int(c.item()) * 5 // 4 + 64
        self.passes = passes
        self.rec_bytes = self.cap * 24
        self.region = (self.FLAG + 8 * passes + passes * self.rec_bytes + 63) // 64 * 64   # flag | counts | records
        total = 2 * self.world * self.region
        # /dev/shm may be small (containers: 64 MB): writing past its free space is a SIGBUS, so every rank looks, the
        # ranks agree (MIN), and without room the NCCL form takes over
        try:
            st = os.statvfs("/dev/shm")
            room = st.f_bavail * st.f_frsize >= 2 * total + (8 << 20)
        except OSError:
            room = False
        if os.environ.get("SPMB_BENCH_SHM_NO_ROOM"):    # test switch
            room = False
        ok = torch.tensor([1 if room else 0], dtype=torch.int64, device=self.dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok.item()) == 0:
            self._fb = HitGather(torch, dist, self.dev, self.world)
            self._fb.prepare(n_first, passes)
            return
        HitGatherShm.serial = getattr(HitGatherShm, "serial", 0) + 1
        path = "/dev/shm/spmb_gather_%s_%s_%d" % (os.environ.get("MASTER_PORT", "0"),
                                                  os.environ.get("TORCHELASTIC_RUN_ID", "0"), HitGatherShm.serial)
        if self.rank == 0:
            self.mm = np.memmap(path, dtype=np.uint8, mode="w+", shape=(total,))
            self.mm[:] = 0
        dist.barrier()
        if self.rank != 0:
            self.mm = np.memmap(path, dtype=np.uint8, mode="r+", shape=(total,))
        dist.barrier()
        if self.rank == 0:
            os.unlink(path)                       # the segment lives as long as it is mapped
        self.grid = self.mm.reshape(2 * self.world, self.region)
        for g in (0, 1):                          # this rank's slots: page-table entries made now, not in the timed region
            self.grid[self._row(g, self.rank), self.FLAG:] = 0
        self.flags = self.grid[:, :8].view(np.uint64).reshape(-1)
        self.counts = self.grid[:, self.FLAG:self.FLAG + 8 * passes].view(np.uint64)          # [2 * world, passes]
        self.gen = 0
        self.finish()

    def _row(self, gen, rank):
        return (gen & 1) * self.world + rank

    def add(self, i, hits):
        if self._fb:
            return self._fb.add(i, hits)
        n = len(hits)
        if n > self.cap:
            self.overflow = True
            n = self.cap
        row = self._row(self.gen + 1, self.rank)
        self.counts[row, i] = n
        if n:
            o = self.FLAG + 8 * self.passes + i * self.rec_bytes
            self.grid[row, o:o + n * 24] = hits[:n].view(np.uint8).reshape(-1)

    def finish(self):
        """publish this rank's lists, wait for every rank's: returns the merged record count over ranks and passes"""
        if self._fb:
            return self._fb.finish()
        self.gen += 1
        g = self.gen
        base = (g & 1) * self.world
        self.flags[base + self.rank] = g
        fl = self.flags[base:base + self.world]
        t_end = time.perf_counter() + 60.0
        while not (fl == g).all():
            if time.perf_counter() > t_end:
                raise RuntimeError("HitGatherShm: a rank did not publish its hit lists within 60 s")
        self.gathers += 1
        return int(self.counts[base:base + self.world].sum())

    def barrier(self):
        """the ranks leave within microseconds of each other (a flag exchange, no lists)"""
        if self._fb:
            return self._fb.barrier()
        self.counts[self._row(self.gen + 1, self.rank), :] = 0
        self.finish()
        self.gathers -= 1

    def merged(self, hit_dtype, i=0):
        """the merged list of pass i of the last gather (rank-major), for checks outside the timed region"""
        if self._fb:
            return self._fb.merged(hit_dtype, i)
        base = (self.gen & 1) * self.world
        parts = []
        o = self.FLAG + 8 * self.passes + i * self.rec_bytes
        for r in range(self.world):
            n = int(self.counts[base + r, i])
            parts.append(np.array(self.grid[base + r, o:o + n * 24]).view(hit_dtype))
        return np.concatenate(parts)


def gather_note(gather):
    if gather is None:
        return False
    if isinstance(gather, HitGatherShm) and gather._fb is None:
        return ("all passes' lists gathered inside the timed region: ranks of one host write their (host-resident) lists into a "
                "shared-memory segment and wait for each other's flags behind the last pass; no device bounce")
    return "all passes' lists in one final all_gather (NCCL) inside the timed region"


def make_gather(torch, dist, dev, world, rank):
    if world <= 1:
        return None
    if HitGatherShm.usable(world):
        return HitGatherShm(torch, dist, dev, world, rank)
    return HitGather(torch, dist, dev, world)


def time_search(sp, torch, ctx, stream, store, pattern, hits, steps, warmup, flush=None, gather=None, launch_fn=None,
                finish_fn=None):
    """returns (total_ms for `steps` launch+finish passes (+ the final gather), mean kernel ms, n_hits, launches/step).
    The passes are pipelined the way the launch/finish pair of the C ABI is meant to be used: passes i+1 and
    i+2 are queued (on their own hit lists) before pass i's hits are collected, so the host's share of a pass
    (D2H of the counters + hit list, ordering) runs under the next kernels.  Every pass still delivers its
    sorted hit list to the host; with `gather` (N > 1) the lists of all passes travel to every rank in one final
    all_gather before the closing event (HitGather)."""
    DEPTH = int(os.environ.get("SPMB_BENCH_DEPTH", "3"))   # passes in flight: two queued behind the one being collected
    if launch_fn is None:       # a plain store; the JST leg passes spmb_jst_search_launch / _finish
        launch_fn = lambda h_: sp.search_launch(ctx, store, pattern, h_)
        finish_fn = lambda h_: sp.search_finish(ctx, h_, copy=False)   # the library's host list itself (spmb_hits_data)
    hl = (hits,) + tuple(sp.Hits(ctx, 1 << 16) for _ in range(DEPTH - 1))
    for w in range(max(warmup, 2 * DEPTH)):     # every hit list twice: buffers sized, long-list ordering armed
        if flush is not None:
            flush.add_(1)
        launch_fn(hl[w % DEPTH])
        h = finish_fn(hl[w % DEPTH])
        if gather is not None and w == 0:
            gather.prepare(len(h), steps)     # warm: capacity agreed, buffers, communicator
    ks = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    ke = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    em = torch.cuda.Event(enable_timing=True)
    for ev_ in ks + ke + [e0, e1, em]:
        ev_.record(stream)      # torch creates the CUDA event at its first record: here, not inside the timed region
    torch.cuda.synchronize()
    # no cyclic garbage collection inside the few milliseconds of the timed region (one generation-2 pass of the
    # interpreter on one rank is a stall of the whole job: the other ranks wait for it in the final gather)
    gc.collect()
    gc_was_on = gc.isenabled()
    gc.disable()
    if gather is not None:
        gather.barrier()        # the ranks enter the timed region together (the warm-up above leaves them skewed)
    flush_ms = 0.0
    e0.record(stream)
    n_hits = 0
    merged_hits = None

    trace = [] if os.environ.get("SPMB_BENCH_TRACE") else None   # host timeline of the passes (tools/r02_trace.sh)
    t_origin = time.perf_counter()

    def launch(i):
        if flush is not None:
            flush.add_(1)       # evict L2 between iterations (buffer > L2)
        ks[i].record(stream)
        t_a = time.perf_counter()
        launch_fn(hl[i % DEPTH])
        if trace is not None:
            trace.append(("launch", i, (t_a - t_origin) * 1e3, (time.perf_counter() - t_origin) * 1e3))
        ke[i].record(stream)

    for i in range(min(DEPTH - 1, steps)):
        launch(i)
    for i in range(steps):
        if i + DEPTH - 1 < steps:
            launch(i + DEPTH - 1)
        t_a = time.perf_counter()
        h = finish_fn(hl[i % DEPTH])
        n_hits = len(h)
        t_b = time.perf_counter()
        if gather is not None:
            gather.add(i, h)
        if trace is not None:
            trace.append(("finish", i, (t_a - t_origin) * 1e3, (t_b - t_origin) * 1e3, (time.perf_counter() - t_origin) * 1e3))
    em.record(stream)                                          # last pass collected: what follows is the gather alone
    if gather is not None:                                     # the final gather: all passes' lists, inside the timed region
        merged_hits = gather.finish() // max(1, steps)
    e1.record(stream)
    torch.cuda.synchronize()
    if gc_was_on:
        gc.enable()
    kern = [ks[i].elapsed_time(ke[i]) for i in range(steps)]
    total = e0.elapsed_time(e1)
    time_search.gather_ms = em.elapsed_time(e1) if gather is not None else None
    if trace is not None:
        gpu = [(round(e0.elapsed_time(ks[i]), 3), round(e0.elapsed_time(ke[i]), 3)) for i in range(steps)]
        print(json.dumps({"trace_host_ms": [[t[0], t[1]] + [round(x, 3) for x in t[2:]] for t in trace], "gpu_launch_ms": gpu,
                          "total_ms": total}), file=sys.stderr, flush=True)
    if flush is not None:       # take the flush writes out of the step time
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record(stream)
        for _ in range(steps):
            flush.add_(1)
        f1.record(stream)
        torch.cuda.synchronize()
        flush_ms = f0.elapsed_time(f1)
    for h in hl[1:]:
        h.close()
    time_search.merged_hits = merged_hits
    return max(total - flush_ms, 1e-6), float(np.mean(kern)), n_hits, ctx.last_launch_count


def allmax(torch, dev, world, *vals):
    if world == 1:
        return vals if len(vals) > 1 else vals[0]
    import torch.distributed as dist
    t = torch.tensor(list(vals), device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    out = [float(x) for x in t.tolist()]
    return out if len(out) > 1 else out[0]


def per_rank(torch, dev, world, **vals):
    """{name: [value of rank 0, rank 1, ...]} -- the spread behind a max-over-ranks figure (N > 1 only)"""
    if world == 1:
        return None
    import torch.distributed as dist
    names = sorted(vals)
    t = torch.tensor([float(vals[k] if vals[k] is not None else -1.0) for k in names], device=dev, dtype=torch.float64)
    out = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(out, t)
    return {k: [round(float(o[i]), 4) for o in out] for i, k in enumerate(names)}


def host_barrier(world, group):
    """ranks that only wait use a gloo (CPU) barrier: an NCCL barrier would park a spinning kernel on their GPUs
    while rank 0 drives all of them through the single-process multi-GPU driver."""
    if world > 1:
        import torch.distributed as dist
        dist.barrier(group=group)


def run_ours(args):
    import torch
    import libspm_b200 as sp
    from libspm_b200 import synth

    rank, world, local = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- libspm_b200 has no CPU path to measure")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    gloo = None
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
        gloo = dist.new_group(backend="gloo")
    stream = torch.cuda.Stream(device=dev)
    hbm_peak, peak_src, _ = measured_peaks()
    out = {}
    cpus = len(os.sched_getaffinity(0))
    if world > 1 and "SPMB_HOST_THREADS" not in os.environ:
        # the ranks of one box share its host cores: split the upload's packing threads between them
        os.environ["SPMB_HOST_THREADS"] = str(max(0, cpus // world - 1))
    synth.lib().spms_set_threads(max(1, cpus // world))
    with torch.cuda.stream(stream):
        ctx = sp.Context(local, stream=stream.cuda_stream)
        hits = sp.Hits(ctx, 1 << 16)
        n = args.n
        m = M_EXACT
        # ---------------- headline: C2x exact search over a journaled sequence ----------------
        t0 = time.perf_counter()
        host_ref = torch.empty(n, dtype=torch.uint8, pin_memory=True)
        cfg = synth.c2(n, seed=synth.SEEDS["C2"] + rank, m=m, ref=host_ref.numpy())
        ev, pool, pat = cfg["events"], cfg["pool"], cfg["pattern"]
        gen_s = time.perf_counter() - t0
        hp = host_ref.data_ptr()
        base = sp.Store(ctx, _store_from_ptr(sp, ctx, hp, n))
        j0, j1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        js = base.journaled(ev, pool)          # first call sizes the context's scratch
        js.close()
        torch.cuda.synchronize()
        j0.record(stream)
        js = base.journaled(ev, pool)
        j1.record(stream)
        torch.cuda.synchronize()
        journal_ms = j0.elapsed_time(j1)
        n_der = len(js)
        pattern = sp.Pattern(ctx, sp.ALG_SHIFTOR, pat)
        setup_s = time.perf_counter() - t0

        gather = make_gather(torch, dist, dev, world, rank)
        if world > 1:
            dist.barrier()
        # clocks are sampled from the warm-up to the end of the e2e legs (the timed region of the
        # headline alone is a few ms -- shorter than nvidia-smi's sampling period)
        clk = ClockSampler(local)
        clk.__enter__()
        total_ms, kern_ms, n_hits, launches = time_search(sp, torch, ctx, stream, js, pattern, hits,
                                                          args.steps, args.warmup, gather=gather)
        spread = per_rank(torch, dev, world, total_ms=total_ms, kernel_ms=kern_ms, final_gather_ms=getattr(time_search, "gather_ms", None))
        total_ms = allmax(torch, dev, world, total_ms)
        value = world * n_der * args.steps / (total_ms * 1e-3) / 1e9
        achieved = n_der * BYTES_PER_BP / (kern_ms * 1e-3) / 1e9
        cap = ncu_capture("exact_c2x") or {}
        out.update({
            "metric": "searched_gbp_per_s", "value": value, "unit": "Gbp/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32",
            "data": "synthetic",
            "config": headline_config(n),
            "detail": {"searched_bp_per_step_per_gpu": n_der, "hits": n_hits, "planted": cfg["planted"],
                       "planted_across_an_edit": cfg["straddling"], "journal_events": int(len(ev)),
                       "row_len": js.row_len, "journal_materialise_ms": journal_ms,
                       "journal_materialise_note": "spmb_store_create_journaled on a resident base, events from pinned host "
                                                   "memory: event upload + prefix sums + journal_rec_kernel + journal_derive_kernel + pack_chunks_kernel",
                       "hit_lists_gathered": gather_note(gather),
                       "merged_hits_per_pass_rank0": getattr(time_search, "merged_hits", None),
                       "final_gather_ms_this_rank": getattr(time_search, "gather_ms", None),
                       "per_rank_ms": spread,
                       "step": "kernel launch + hit count/list D2H + host sort; passes i+1 and i+2 are queued before pass i's "
                               "hits are collected (spmb_search_launch / spmb_search_finish on three hit lists)"
                               + ("; the hit lists of all passes are gathered to every rank inside the timed region" if gather else "")},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                         "frac": achieved / hbm_peak,
                         "traffic": cap.get("dram_bytes") if isinstance(cap, dict) else cap,
                         "traffic_source": "constant read from the committed ncu --set full capture (profiles/roofline_traffic.json), "
                                           "not measured in this run",
                         "peak_source": peak_src,
                         "kernel": "exact_kernel<plain>", "kernel_ms": kern_ms,
                         "algorithmic_bytes_per_launch": n_der * BYTES_PER_BP},
            "gpu_launches": launches * args.steps,
        })

        # ---------------- e2e: host buffers -> hits through the C ABI ----------------
        e2e_steps = max(1, min(args.steps, args.e2e_steps))
        ev_bytes = ev.nbytes + pool.nbytes
        # the step's host inputs live in pinned memory: reference ranks, journal events, insertion pool
        ev_pin = torch.empty(ev.nbytes, dtype=torch.uint8, pin_memory=True)
        ev_pin.numpy()[:] = ev.view(np.uint8)
        ev = ev_pin.numpy().view(sp.EVENT_DTYPE)
        pool_pin = torch.empty(max(pool.nbytes, 1), dtype=torch.uint8, pin_memory=True)
        pool_pin.numpy()[:pool.nbytes] = pool
        pool = pool_pin.numpy()[:pool.nbytes]

        phase = np.zeros(3)                                  # upload+pack | journal | search (each call syncs)

        def e2e_once(ptr=hp):
            ta = time.perf_counter()
            b2 = sp.Store(ctx, _store_from_ptr(sp, ctx, ptr, n))
            tb = time.perf_counter()
            j2 = b2.journaled(ev, pool)
            tc = time.perf_counter()
            h = sp.search(ctx, j2, pattern, hits)
            td = time.perf_counter()
            phase[:] += (tb - ta, tc - tb, td - tc)
            j2.close(); b2.close()
            return len(h)

        e2e_once()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        phase[:] = 0
        text_h2d = 0
        step_ms = []
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            ts = time.perf_counter()
            nh = e2e_once()
            step_ms.append((time.perf_counter() - ts) * 1e3)
            us = ctx.upload_last_stats()
            # bytes that crossed PCIe for the haystack: raw chunks at 1 B/symbol, host-packed ones at 0.25
            text_h2d += (min(n, us["raw_chunks"] * us["chunk_symbols"]) + us["host_packed_chunks"] * us["chunk_symbols"] // 4
                         if us["chunk_symbols"] else n)
        torch.cuda.synchronize()
        e2e_s = time.perf_counter() - t0
        text_h2d //= e2e_steps
        us = ctx.upload_last_stats()
        if us["events_per_chunk"]:                           # events: 24 B raw, 8 B (+16 per chunk) host-packed
            epc = us["events_per_chunk"]
            ev_h2d = (len(ev) - min(len(ev), us["event_host_packed_chunks"] * epc)) * 24 \
                + us["event_host_packed_chunks"] * (16 + 8 * epc) + pool.nbytes
        else:
            ev_h2d = ev_bytes
        e2e_s = allmax(torch, dev, world, e2e_s)
        out["e2e"] = {"value": world * n_der * e2e_steps / e2e_s / 1e9, "unit": "Gbp/s",
                      "h2d_bytes_per_step": int(text_h2d + ev_h2d + m), "d2h_bytes_per_step": int(8 + nh * 24),
                      "steps": e2e_steps, "ms_per_step": e2e_s / e2e_steps * 1e3,
                      "ms_per_step_median_this_rank": float(np.median(step_ms)), "ms_per_step_max_this_rank": float(max(step_ms)),
                      "host_input_bytes_per_step": int(n + ev_bytes + m),
                      "upload": ctx.upload_last_stats(),
                      "phase_ms": {"store_create(upload+pack)": phase[0] / e2e_steps * 1e3,
                                   "store_create_journaled": phase[1] / e2e_steps * 1e3,
                                   "search": phase[2] / e2e_steps * 1e3},
                      "path": "spmb_store_create(host ranks: raw chunks over the copy engine + chunks 2-bit packed by "
                              "host threads) + spmb_store_create_journaled + spmb_search; inputs in pinned host memory"}
        # the same step with the haystack in PAGEABLE memory (what a libspm caller's std::vector<spm::dna4> is)
        if rank == 0 and world == 1 and not args.no_pageable:
            pageable = np.empty(n, dtype=np.uint8)
            pageable[:] = host_ref.numpy()
            e2e_once(pageable.ctypes.data)
            t0 = time.perf_counter()
            for _ in range(2):
                e2e_once(pageable.ctypes.data)
            torch.cuda.synchronize()
            dt = (time.perf_counter() - t0) / 2
            out["e2e_pageable"] = {"value": n_der / dt / 1e9, "unit": "Gbp/s", "ms_per_step": dt * 1e3, "steps": 2,
                                   "path": "the e2e step with the reference ranks in pageable host memory (numpy array); "
                                           "events and pool pinned as before"}
            del pageable
        # resident-store e2e: the store is the persistent object; per step only the needle goes up
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            p2 = sp.Pattern(ctx, sp.ALG_SHIFTOR, pat)
            nh = len(sp.search(ctx, js, p2, hits))
            p2.close()
        e2e_res_s = time.perf_counter() - t0
        out["e2e_resident"] = {"value": world * n_der * args.steps / e2e_res_s / 1e9, "unit": "Gbp/s",
                               "h2d_bytes_per_step": m, "d2h_bytes_per_step": int(8 + nh * 24),
                               "path": "spmb_pattern_create(host needle) + spmb_search on the resident store"}

        clk.__exit__()
        out["clocks"] = clk.summary()
        out["clocks"]["window"] = "warm-up + timed steps + e2e legs of the headline workload"

        # ---------------- CPU baseline + cross-check (rank 0, N=1 only) ----------------
        want_hits = None
        if rank == 0 and world == 1 and not args.no_cpu:
            import oracle as orc
            s_n = int(min(n_der, args.cpu_sample))
            sample = js.read(0, s_n)
            threads = orc.use_all_cores()
            best = 1e30
            for _ in range(3):
                t0 = time.perf_counter()
                cnt, ce, _ = orc.search_mt(2, sample, pat, 0, threads=threads)
                best = min(best, time.perf_counter() - t0)
            s1 = sample[:min(s_n, 200_000_000)]
            t0 = time.perf_counter()
            orc.search_mt(2, s1, pat, 0, threads=1, store=False)
            one = time.perf_counter() - t0
            g = sp.search(ctx, js, pattern, hits)
            ge = g["end"][g["end"] <= s_n]
            out["cpu_baseline"] = {"value": s_n / best / 1e9, "unit": "Gbp/s", "cores": threads, "kind": "port",
                                   "sample": f"first {s_n} bp of the derived sequence, oracle Shift-Or (32-bit blocks), "
                                             f"{threads} threads, best of 3; journal already applied",
                                   "one_thread_gbp_per_s": len(s1) / one / 1e9,
                                   "one_thread_sample": f"first {len(s1)} bp, 1 thread (the reference is single-threaded)",
                                   "hits_match_gpu": bool(np.array_equal(ge, ce)), "hits_in_sample": int(len(ce))}
            del sample, s1

        # ---------------- streaming (f1) and the single-process multi-GPU driver: rank 0 only ----------------
        plain_host = host_ref.numpy()
        if rank == 0 and not args.no_stream and hasattr(sp, "stream_leg"):
            out["stream"] = sp.stream_leg(ctx, plain_host, pat, hits, chunk=args.stream_chunk)
        js.close(); base.close()
        if not args.no_multi and hasattr(sp, "Multi"):
            host_barrier(world, gloo)                        # every rank has released its big stores
            if rank == 0:
                out["e2e_multi"] = bench_multi(sp, torch, args, world, plain_host, ev, pool, pat, n_der)
            host_barrier(world, gloo)
        del host_ref, plain_host

        # ---------------- Myers k-error: C3 (m=100, k=4, 250 Mbp) ----------------
        if not args.no_myers:
            out["myers"] = bench_myers(sp, synth, torch, ctx, dev, stream, hits, args, rank, world, hbm_peak)
        if not args.no_c4:
            out["c4"] = bench_c4(sp, synth, torch, ctx, dev, stream, hits, args, rank, world)
        if not args.no_c5:
            out["c5"] = bench_c5(sp, synth, torch, ctx, dev, stream, hits, args, rank, world)
        if rank == 0 and world == 1 and not args.no_small:
            out["small_configs"] = bench_small(sp, synth, torch, ctx, dev, stream, hits, args)
        out["setup_s"] = setup_s
        out["generate_s"] = gen_s
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        emit(out)


def _store_from_ptr(sp, ctx, ptr, n):
    import ctypes as C
    h = C.c_void_p()
    ctx._check(sp.lib().spmb_store_create(ctx._h, C.c_void_p(ptr), n, 0, C.byref(h)), "spmb_store_create")
    return h


def bench_multi(sp, torch, args, n_dev, host, ev, pool, pat, n_der):
    """ONE host haystack (+ its journal) -> hits through the single-process multi-GPU driver of the C ABI
    (spmb_multi_*): the haystack is cut into n_dev contiguous shards with window_size-1 symbols of left context,
    every shard crosses its own PCIe link, the journal's events are split at the cuts, every device searches its
    shard, the hit lists come back per device and are merged on the host -- the final gather, no NCCL."""
    n_dev = max(1, min(n_dev, torch.cuda.device_count()))
    mg = sp.Multi(list(range(n_dev)))
    try:
        pset = mg.pattern(sp.ALG_SHIFTOR, pat)
        res = None

        def once():
            base = mg.store(host)
            js = base.journaled(ev, pool)
            h = mg.search(js, pset)
            st = mg.last_stats()
            js.close(); base.close()
            return h, st

        once()
        steps = max(1, min(args.steps, args.e2e_steps))
        acc = np.zeros(4)
        t0 = time.perf_counter()
        for _ in range(steps):
            h, st = once()
            acc += st
        dt = (time.perf_counter() - t0) / steps
        # the plain (unjournaled) call shape too: spmb_multi_search_host
        mg.search_host(host, pset)
        t0 = time.perf_counter()
        for _ in range(steps):
            hp = mg.search_host(host, pset)
        dt_plain = (time.perf_counter() - t0) / steps
        res = {"value": n_der / dt / 1e9, "unit": "Gbp/s", "n_devices": n_dev, "ms_per_step": dt * 1e3, "steps": steps,
               "hits": int(len(h)), "scaling": "strong (one host haystack of %d bp whatever the device count)" % len(host),
               "phase_ms": {"upload": acc[0] / steps, "journal": acc[1] / steps, "search": acc[2] / steps, "merge": acc[3] / steps},
               "plain": {"value": len(host) / dt_plain / 1e9, "unit": "Gbp/s", "ms_per_step": dt_plain * 1e3, "hits": int(len(hp)),
                         "path": "spmb_multi_search_host: host ranks -> shards -> hits, no journal"},
               "path": "spmb_multi_store_create + spmb_multi_store_create_journaled + spmb_multi_search (one process, "
                       "one context and upload pipeline per device, host merge of the per-device hit lists inside the step)"}
        pset.close()
        return res
    finally:
        mg.close()


def bench_myers(sp, synth, torch, ctx, dev, stream, hits, args, rank, world, hbm_peak):
    n, m, k = args.n_myers, 100, 4
    cfg = synth.c3(n, seed=synth.SEEDS["C3"] + rank, m=m, k=k)
    pat = cfg["pattern"]
    store = sp.Store.from_host(ctx, cfg["text"])
    ctx.sync()
    pattern = sp.Pattern(ctx, sp.ALG_MYERS, pat, k)
    flush = torch.zeros(64 << 20, dtype=torch.int32, device=dev)   # 256 MB > L2
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    with ClockSampler(torch.cuda.current_device()) as clk:
        total_ms, kern_ms, n_hits, launches = time_search(sp, torch, ctx, stream, store, pattern, hits, args.steps,
                                                          args.warmup, flush=flush)
    # the same search with the full recurrence (all ceil(m/32) words every column): the kernel the INT-pipe
    # roofline is stated on; the default above is its Ukkonen cut-off form (myers_cut.cuh)
    ctx.set_myers_cutoff(False)
    full_total_ms, full_kern_ms, full_hits, _ = time_search(sp, torch, ctx, stream, store, pattern, hits, args.steps,
                                                            args.warmup, flush=flush)
    ctx.set_myers_cutoff(True)
    total_ms, full_total_ms = allmax(torch, dev, world, total_ms, full_total_ms)
    peak = ctx.measure_int_peak()
    ops = myers_ops_per_bp(m)
    achieved = n * ops / (full_kern_ms * 1e-3) / 1e9
    sm = ctx.device_info()
    issue_ceiling = 128.0 * sm["sm_count"] * sm["sm_clock_khz"] * 1e3 / 1e9   # Ginstr/s: 4 schedulers x 32 lanes per clock
    cut_cap = ncu_capture("myers_cut_c3_counters") or {}
    res = {"workload": f"C3: Myers m={m} k={k} over {n} bp per GPU, 1000 planted copies with 0..6 random edits",
           "value": world * n * args.steps / (total_ms * 1e-3) / 1e9, "unit": "Gbp/s", "ms_per_step": total_ms / args.steps,
           "kernel_ms": kern_ms, "kernel_gbp_per_s": n / kern_ms / 1e6, "kernel": "myers_cut_kernel<4,plain> (Ukkonen cut-off, the default)",
           "hits": n_hits, "row_len": store.row_len, "l2": "256 MB flush write between iterations",
           # THE INT roofline figure: the full recurrence on SURVEY 8(d)'s 10*ceil(m/32)+4 instructions per bp
           "roofline": {"bound": "int_alu", "achieved": achieved, "peak": peak["lop3_gops"], "unit": "Ginstr/s",
                        "frac": achieved / peak["lop3_gops"], "traffic": None,
                        "peak_source": "measured live: spmb_measure_int_peak (LOP3 chains, all SMs) = the ALU pipe, 64 lanes/SM/clk",
                        "issue_slot_ceiling": issue_ceiling, "int_peak_iadd3": peak["iadd3_gops"],
                        "algorithmic_int32_instr_per_bp": ops, "kernel": "myers_kernel<4,plain> (full recurrence)",
                        "kernel_ms": full_kern_ms},
           "full_recurrence": {"value": world * n * args.steps / (full_total_ms * 1e-3) / 1e9, "unit": "Gbp/s",
                               "kernel_ms": full_kern_ms, "kernel_gbp_per_s": n / full_kern_ms / 1e6,
                               "hits_equal_cutoff": bool(full_hits == n_hits)},
           # the cut-off kernel skips words, so an algorithmic-count fraction would exceed 1: what it executes instead
           "cutoff_kernel": {"kernel_ms": kern_ms, "speedup_vs_full_kernel": full_kern_ms / kern_ms,
                             "equivalent_fraction_on_full_count": n * ops / (kern_ms * 1e-3) / 1e9 / peak["lop3_gops"],
                             "equivalent_fraction_note": "44 instr/bp the kernel did NOT execute, over the ALU peak: > 1 means words skipped, not a roofline fraction",
                             "ncu": cut_cap,
                             "ncu_source": "constants of the committed ncu --set full capture of this kernel on this workload "
                                           "(profiles/roofline_traffic.json), not measured in this run",
                             "alu_pipe_ceiling_ginstr": peak["lop3_gops"], "issue_slot_ceiling_ginstr": issue_ceiling},
           "clocks": clk.summary()}
    if isinstance(cut_cap, dict) and cut_cap.get("executed_lane_instr_per_bp"):
        res["cutoff_kernel"]["executed_ginstr_per_s"] = n * cut_cap["executed_lane_instr_per_bp"] / (kern_ms * 1e-3) / 1e9
        res["cutoff_kernel"]["issue_slot_fraction"] = res["cutoff_kernel"]["executed_ginstr_per_s"] / issue_ceiling
    if rank == 0 and world == 1 and not args.no_cpu:
        import oracle as orc
        host = cfg["text"]
        threads = orc.use_all_cores()
        best = 1e30
        for _ in range(2):
            t0 = time.perf_counter()
            cnt, ce, cr = orc.search_mt(3, host, pat, k, threads=threads)
            best = min(best, time.perf_counter() - t0)
        s1 = host[:min(n, 50_000_000)]
        t0 = time.perf_counter()
        orc.search_mt(3, s1, pat, k, threads=1, store=False)
        one = time.perf_counter() - t0
        g = sp.search(ctx, store, pattern, hits)
        res["cpu_baseline"] = {"value": n / best / 1e9, "unit": "Gbp/s", "cores": threads, "kind": "port",
                               "sample": f"whole {n} bp text, oracle Myers (64-bit blocks, Ukkonen cut-off), {threads} threads",
                               "one_thread_gbp_per_s": len(s1) / one / 1e9,
                               "one_thread_sample": f"first {len(s1)} bp, 1 thread",
                               "hits_match_gpu": bool(np.array_equal(g["end"], ce) and np.array_equal(g["errors"], cr))}
    store.close()
    del flush
    torch.cuda.empty_cache()
    return res


def bench_c4(sp, synth, torch, ctx, dev, stream, hits, args, rank, world):
    """configs[3]: JST search, 2504 haplotypes over a 50 Mbp reference with 1 M variants, Myers m=150 k=2.
    Every rank builds the tree (host flattening + window materialisation, both reported) and searches shard
    rank/N of the reference scan and of every window bucket (spmb_context_set_shard): strong scaling, no
    collective; the timed step is the warm search (kernels + expansion of the raw hits to haplotypes + sort)."""
    import hashlib
    d = synth.c4(args.c4_scale)
    n_ref, n_hap, m, k = len(d["ref"]), d["n_hap"], len(d["pattern"]), d["k"]
    store = sp.Store.from_host(ctx, d["ref"])
    t0 = time.perf_counter()
    jst = sp.Jst(ctx, store, d["variants"], d["pool"], d["coverage"], n_hap)
    create_s = time.perf_counter() - t0
    p = sp.Pattern(ctx, sp.ALG_MYERS, d["pattern"], k)
    whole = jst.search(p, hits)                 # unsharded, once: flatten + materialise + search
    st_first = jst.last_stats()
    ctx.set_shard(rank, world)
    jst.search(p, hits)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    steps = args.steps
    # one synchronous call at a time (the reference's call shape): latency of a search, its kernels and its host half
    kms, ems = [], []
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(min(steps, 5)):
        got = jst.search(p, hits)
        st = jst.last_stats()
        kms.append(st["kernels_ms"]); ems.append(st["expand_ms"])
    torch.cuda.synchronize()
    sync_ms = (time.perf_counter() - t0) * 1e3 / min(steps, 5)
    sync_kern_ms = float(np.mean(kms))
    # the timed step: searches back to back through spmb_jst_search_launch / _finish on three hit lists, so the host
    # half of a search (D2H, expansion to haplotypes, ordering) runs under the kernels of the next one
    gather = make_gather(torch, dist, dev, world, rank) if world > 1 else None
    total_ms, kern_ms, n_got, _ = time_search(sp, torch, ctx, stream, None, None, hits, steps, args.warmup, gather=gather,
                                               launch_fn=lambda h_: jst.search_launch(p, h_),
                                               finish_fn=lambda h_: jst.search_finish(h_, copy=False))
    c4_gather_ms = getattr(time_search, "gather_ms", None)
    jst.search_launch(p, hits)
    piped = jst.search_finish(hits)
    piped_same = bool(len(piped) == len(got) and np.array_equal(piped, got))
    warm_ms = total_ms / steps
    spread = per_rank(torch, dev, world, total_ms=total_ms, kernels_ms=kern_ms, final_gather_ms=c4_gather_ms)
    warm_ms, kern_ms, sync_ms = allmax(torch, dev, world, warm_ms, kern_ms, sync_ms)
    # the full recurrence on this shard (the kernel SURVEY 8(d)'s 54 instr/bp is stated on)
    ctx.set_myers_cutoff(False)
    jst.search(p, hits)
    full_k = []
    for _ in range(3):
        jst.search(p, hits)
        full_k.append(jst.last_stats()["kernels_ms"])
    ctx.set_myers_cutoff(True)
    ctx.set_shard(0, 1)
    full_kern_ms = allmax(torch, dev, world, float(np.mean(full_k)))
    streamed = n_ref + st_first["window_symbols"]
    effective = float(n_ref) * n_hap
    peak = ctx.measure_int_peak()
    ops = myers_ops_per_bp(m)
    res = {"workload": f"C4: JST search, {n_hap} haplotypes over a {n_ref} bp reference with {len(d['variants'])} variants, "
                       f"Myers m={m} k={k}; window array sharded over {world} GPU(s)",
           "value": effective / (warm_ms * 1e-3) / 1e9, "unit": "effective Gbp/s (sum of haplotype lengths per warm search)",
           "ms_per_step": warm_ms, "scaling": "strong", "steps": steps,
           "kernels_ms": kern_ms, "expand_ms": float(np.mean(ems)),
           "step": "spmb_jst_search_launch / spmb_jst_search_finish on three hit lists in rotation: every search delivers its "
                   "expanded, ordered hit list to the host; the host half of search i runs under the kernels of search i+1"
                   + ("; the hit lists of all searches are gathered to every rank inside the timed region" if gather else ""),
           "synchronous_call_ms": sync_ms, "synchronous_call_kernels_ms": sync_kern_ms,
           "pipelined_hits_equal_synchronous": piped_same,
           "hit_lists_gathered": gather_note(gather), "final_gather_ms_this_rank": c4_gather_ms, "per_rank_ms": spread,
           "streamed_bp": int(streamed), "effective_bp": effective,
           "streamed_gbp_per_s_kernels": streamed / (kern_ms * 1e-3) / 1e9,
           "one_off": {"jst_create_s": create_s, "flatten_ms": st_first["flatten_ms"], "materialise_ms": st_first["materialise_ms"],
                       "windows": int(st_first["windows"]), "window_symbols": int(st_first["window_symbols"]),
                       "buckets": int(st_first["buckets"])},
           "hits": int(len(whole)), "hits_this_shard": int(n_got), "raw_device_hits": int(st_first["raw_hits"]),
           "full_recurrence": {"kernels_ms": full_kern_ms, "int32_instr_per_bp": ops,
                               "achieved_ginstr_per_s": streamed / world * ops / (full_kern_ms * 1e-3) / 1e9,
                               "frac_of_alu_peak": streamed / world * ops / (full_kern_ms * 1e-3) / 1e9 / peak["lop3_gops"],
                               "note": "all launches of the search (reference scan + every window bucket, two streams) "
                                       "over the streamed symbols of this shard"},
           "cutoff_kernel_executed_lane_instr_per_bp": (ncu_capture("myers_cut_c3_counters") or {}).get("executed_lane_instr_per_bp")}
    # parity: the WHOLE hit list against the committed full-size oracle pass (tests/golden/c4_hits.npz)
    gp = os.path.join(ROOT, "tests", "golden", "c4_hits.npz")
    if args.c4_scale == 1.0 and os.path.exists(gp):
        g = np.load(gp)
        h = hashlib.sha256()
        for key in ("ref", "variants", "pool", "coverage", "pattern"):
            h.update(np.ascontiguousarray(d[key]).view(np.uint8).tobytes())
        same_in = h.hexdigest() == str(g["input_sha256"])
        res["parity"] = bool(same_in and len(whole) == len(g["hap"]) and np.array_equal(whole["seq_id"], g["hap"]) and
                             np.array_equal(whole["end"], g["end"]) and np.array_equal(whole["errors"], g["err"].astype(np.uint32)))
        res["parity_against"] = (f"all {len(g['hap'])} hits of the oracle's pass over all {n_hap} haplotypes "
                                 f"(tests/golden/c4_hits.npz, {float(g['oracle_seconds']):.0f} s on {int(g['oracle_threads'])} threads); "
                                 f"inputs identical: {same_in}")
    if rank == 0 and world == 1 and not args.no_cpu:
        import oracle as orc
        sample = np.sort(np.unique(np.append(synth.uniform(d["seed"], synth.STREAM_USER + 9, 32, 0, n_hap).astype(np.uint32),
                                             np.uint32(d["source_haplotype"]))))
        threads = orc.use_all_cores()
        t0 = time.perf_counter()
        want = orc.jst_search(3, d["ref"], d["variants"].astype(orc.EVENT_DTYPE), d["pool"], d["coverage"], n_hap,
                              d["pattern"], k, haps=sample, threads=threads)
        dt = time.perf_counter() - t0
        sel = whole[np.isin(whole["seq_id"], sample)]
        res["cpu_baseline"] = {"value": float(len(sample)) * n_ref / dt / 1e9, "unit": "effective Gbp/s", "cores": threads,
                               "kind": "port", "sample": f"{len(sample)} sampled haplotypes materialised and searched by the "
                                                         f"oracle (Myers, 64-bit blocks, cut-off), {threads} threads",
                               "hits_match_gpu": bool(len(sel) == len(want) and np.array_equal(sel["seq_id"], want["hap"]) and
                                                      np.array_equal(sel["end"], want["end"]) and
                                                      np.array_equal(sel["errors"], want["err"])),
                               "hits_in_sample": int(len(want))}
    jst.close(); p.close(); store.close()
    torch.cuda.empty_cache()
    return res


def bench_c5(sp, synth, torch, ctx, dev, stream, hits, args, rank, world):
    """configs[4]: 10k patterns of ~100 bp, k=3, over a 3 Gbp text, in ONE call through the pigeonhole seed
    filter (SPMB_ALG_MYERS_FILTERED); with N ranks the SAME text is sharded (strong scaling), each rank
    searching shard rank/N, no collective on the search; every pass's hit list is gathered to rank 0.
    Parity at N=1 as SURVEY.md 8(c) specifies: all patterns x a 3 Mbp slice and 16 patterns x the full text,
    both against the CPU oracle; at N > 1 per-pattern equality with the direct Myers kernel on this rank's shard."""
    n, n_pat, m, k = args.n_c5, args.c5_patterns, 100, 3
    seed = synth.SEEDS["C5"]
    text = synth.text_host(n, seed)

    def text_at(offs, mm):
        return np.stack([text[int(o):int(o) + mm] for o in offs])

    offs, pats = synth.c5_patterns(text_at, n, n_pat, m, seed)
    store = sp.Store.from_host(ctx, text)
    ctx.sync()
    pset = sp.Pattern(ctx, sp.ALG_MYERS_FILTERED, pats, k)
    ctx.set_shard(rank, world)
    sp.search(ctx, store, pset, hits)          # builds the lane-chunk copy of the store
    gather = None
    if world > 1:
        import torch.distributed as dist
        gather = make_gather(torch, dist, dev, world, rank)
        dist.barrier()
    total_ms, kern_ms, n_hits, launches = time_search(sp, torch, ctx, stream, store, pset, hits, args.steps, args.warmup,
                                                      gather=gather)
    spread = per_rank(torch, dev, world, total_ms=total_ms, kernels_ms=kern_ms, final_gather_ms=getattr(time_search, "gather_ms", None),
                      hits_per_pass=float(n_hits))
    total_ms = allmax(torch, dev, world, total_ms)
    if world > 1:
        c = torch.tensor([n_hits], device=dev, dtype=torch.int64)
        dist.all_reduce(c)
        n_hits = int(c.item())
    got = sp.search(ctx, store, pset, hits)
    step = max(1, n_pat // 16)
    ids16 = list(range(0, n_pat, step))[:16]
    sub = [pats[i] for i in ids16]
    direct = sp.Pattern(ctx, sp.ALG_MYERS, sub, k)
    dd = sp.search(ctx, store, direct, hits)
    sel = got[np.isin(got["pattern_id"], ids16)]
    same = bool(len(sel) == len(dd) and np.array_equal(sel["pattern_id"] // step, dd["pattern_id"]) and
                np.array_equal(sel["end"], dd["end"]) and np.array_equal(sel["errors"], dd["errors"]))
    ctx.set_shard(0, 1)
    res = {"workload": f"C5: {n_pat} patterns of ~{m} bp, k={k}, over {n} bp in one call (seed filter + verification); "
                       f"the text is sharded over {world} GPU(s)",
           "value": n * args.steps / (total_ms * 1e-3) / 1e9, "unit": "Gbp/s (x 10k patterns)", "ms_per_step": total_ms / args.steps,
           "kernels_ms": kern_ms, "cell_rate_T_per_s": n * n_pat * args.steps / (total_ms * 1e-3) / 1e12,
           "hits": n_hits, "filter": pset.filter_info, "scaling": "strong",
           "hit_lists_gathered": gather_note(gather),
           "merged_hits_per_pass_rank0": getattr(time_search, "merged_hits", None),
           "final_gather_ms_this_rank": getattr(time_search, "gather_ms", None),
           "per_rank_ms": spread,
           "equals_direct_myers_kernel_16_patterns_this_shard": same,
           "text_gbytes_per_s_kernels": n / world * 0.25 / (kern_ms * 1e-3) / 1e9}
    if rank == 0 and world == 1 and not args.no_cpu:
        import oracle as orc
        threads = orc.use_all_cores()
        # (i) ALL patterns x a 3 Mbp slice
        sl_n = min(n, 3_000_000)
        sl0 = min(int(offs[n_pat // 2]) // 1024 * 1024, n - sl_n)
        host = text[sl0:sl0 + sl_n].copy()
        st2 = sp.Store.from_host(ctx, host)
        g2 = sp.search(ctx, st2, pset, hits)
        order = np.argsort(g2["pattern_id"], kind="stable")
        g2 = g2[order]
        lo = np.searchsorted(g2["pattern_id"], np.arange(n_pat), "left")
        hi = np.searchsorted(g2["pattern_id"], np.arange(n_pat), "right")
        t0 = time.perf_counter()
        bad = 0
        slice_hits = 0
        for i in range(n_pat):
            _, we, wr = orc.search_mt(3, host, pats[i], k, threads=threads)
            s2 = g2[lo[i]:hi[i]]
            slice_hits += len(we)
            bad += int(not (np.array_equal(s2["end"], we) and np.array_equal(s2["errors"], wr)))
        cpu_s = time.perf_counter() - t0
        st2.close()
        # (ii) 16 patterns x the FULL text
        t0 = time.perf_counter()
        bad16 = 0
        for j, i in enumerate(ids16):
            _, we, wr = orc.search_mt(3, text, pats[i], k, threads=threads)
            s2 = got[got["pattern_id"] == i]
            bad16 += int(not (np.array_equal(s2["end"], we) and np.array_equal(s2["errors"], wr)))
        cpu16_s = time.perf_counter() - t0
        res["cpu_baseline"] = {"value": (sl_n * n_pat + n * len(ids16)) / (cpu_s + cpu16_s) / 1e12, "unit": "T cells/s",
                               "cores": threads, "kind": "port",
                               "sample": f"all {n_pat} patterns x a {sl_n} bp slice ({cpu_s:.1f} s) + {len(ids16)} patterns x the full "
                                         f"{n} bp text ({cpu16_s:.1f} s), oracle Myers, {threads} threads",
                               "slice_patterns_mismatching_gpu": bad, "slice_hits": slice_hits,
                               "full_text_patterns_mismatching_gpu": bad16}
        res["parity"] = bool(bad == 0 and bad16 == 0 and same)
    store.close()
    del text
    torch.cuda.empty_cache()
    return res


def bench_small(sp, synth, torch, ctx, dev, stream, hits, args):
    """C1 and C2 exactly as SURVEY.md 8(d) specifies them (they fit in L2: flushed between iterations)."""
    import oracle as orc
    res = {}
    flush = torch.zeros(64 << 20, dtype=torch.int32, device=dev)
    # C1: Horspool m=32 over 1 Mbp, 100 occurrences; the period-4 run with 9 overlapping hits of ACGTx8
    d = synth.c1()
    text, pat = d["text"], d["pattern"]
    n = len(text)
    store = sp.Store.from_host(ctx, text)
    p = sp.Pattern(ctx, sp.ALG_HORSPOOL, pat)
    tot, kern, nh, _ = time_search(sp, torch, ctx, stream, store, p, hits, args.steps, args.warmup, flush=flush)
    want = orc.horspool(text, pat)
    got = sp.search(ctx, store, p, hits)
    p2 = sp.Pattern(ctx, sp.ALG_HORSPOOL, d["pattern_overlap"])
    got2 = sp.search(ctx, store, p2, hits)
    want2 = orc.horspool(text, d["pattern_overlap"])
    t0 = time.perf_counter(); orc.horspool(text, pat); cpu_s = time.perf_counter() - t0
    res["C1"] = {"workload": "Horspool m=32, 1 Mbp, needle + 99 planted copies; ACGTx8 over a planted ACGTx16 run",
                 "kernel_us": kern * 1e3, "step_us": tot / args.steps * 1e3,
                 "gbp_per_s_kernel": n / (kern * 1e-3) / 1e9, "hits": nh, "overlapping_hits": int(len(got2)),
                 "parity": bool(np.array_equal(got["end"], want) and np.array_equal(got2["end"], want2) and len(want) >= 100 and len(want2) >= 9),
                 "cpu_1thread_gbp_per_s": n / cpu_s / 1e9}
    store.close(); p.close(); p2.close()
    # C2: Shift-Or m=64 over 50 Mbp + 1% journal, 1 + 100 occurrences, 25 across an edit
    d = synth.c2()
    ref, ev, pool, pat = d["ref"], d["events"], d["pool"], d["pattern"]
    base = sp.Store.from_host(ctx, ref)
    js = base.journaled(ev, pool)
    nd = len(js)
    p = sp.Pattern(ctx, sp.ALG_SHIFTOR, pat)
    tot, kern, nh, _ = time_search(sp, torch, ctx, stream, js, p, hits, args.steps, args.warmup, flush=flush)
    derived = orc.journal_apply(ref, ev.astype(orc.EVENT_DTYPE), pool)
    same_text = bool(np.array_equal(js.read(), derived))
    cnt, ce, _ = orc.search_mt(2, derived, pat, 0)
    got = sp.search(ctx, js, p, hits)
    t0 = time.perf_counter(); orc.search_mt(2, derived, pat, 0, threads=1, store=False); cpu_s = time.perf_counter() - t0
    res["C2"] = {"workload": "Shift-Or m=64, 50 Mbp + 1% journal; needle from the derived sequence + 100 planted copies",
                 "kernel_us": kern * 1e3, "step_us": tot / args.steps * 1e3, "gbp_per_s_kernel": nd / (kern * 1e-3) / 1e9,
                 "hits": nh, "planted": d["planted"], "planted_across_an_edit": d["straddling"],
                 "journal_events": int(len(ev)), "derived_text_matches_oracle": same_text,
                 "parity": bool(np.array_equal(got["end"], ce) and len(ce) >= 1 + d["planted"]),
                 "cpu_1thread_gbp_per_s": nd / cpu_s / 1e9}
    js.close(); base.close()
    return res


# ------------------------------------------------------------------------------------------
def run_reference(args):
    """The reference's CPU algorithms for the path (the oracle port: SeqAn2 is not on this machine,
    SURVEY.md 8(c)), all host threads.  Same workload, same generator, same planted needles as the CUDA arm:
    each step is a bounded sample -- the first `--cpu-sample` bp of the SAME reference stream with the journal
    events and the planted copies that fall into it."""
    rank, world, _ = dist_env()
    if rank != 0:
        return
    import oracle as orc
    from libspm_b200 import synth
    m = M_EXACT
    s_n = int(min(args.n, args.cpu_sample))
    cfg = synth.c2(s_n, seed=synth.SEEDS["C2"], m=m)
    ref, ev, pool, pat = cfg["ref"], cfg["events"], cfg["pool"], cfg["pattern"]
    threads = orc.use_all_cores()
    evo = ev.astype(orc.EVENT_DTYPE)
    times, times_search = [], []
    hits = 0
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        derived = orc.journal_apply(ref, evo, pool)          # host buffers in -> journaled haystack
        t1 = time.perf_counter()
        hits, _, _ = orc.search_mt(2, derived, pat, 0, threads=threads, store=False)
        t2 = time.perf_counter()
        if i >= args.warmup:
            times.append(t2 - t0)
            times_search.append(t2 - t1)
    tot = sum(times)
    v = len(derived) * args.steps / tot / 1e9
    v_search = len(derived) * args.steps / sum(times_search) / 1e9
    t0 = time.perf_counter()
    orc.search_mt(2, derived[:min(len(derived), 200_000_000)], pat, 0, threads=1, store=False)
    one = min(len(derived), 200_000_000) / (time.perf_counter() - t0) / 1e9
    sample = (f"bounded sample: the first {s_n} bp of the {args.n} bp reference stream with its journal ({len(ev)} events) and "
              f"planted needles ({cfg['planted']} copies, {cfg['straddling']} across an edit); each step = host buffers in, hit count "
              f"out: oracle journal apply + oracle Shift-Or (32-bit blocks), both on {threads} threads")
    emit({
        "impl": "reference", "metric": "searched_gbp_per_s", "value": v, "unit": "Gbp/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": tot / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": headline_config(args.n),
        "detail": {"searched_bp_per_step": int(len(derived)), "hits": int(hits), "search_only_gbp_per_s": v_search,
                   "one_thread_search_gbp_per_s": one},
        "cpu_baseline": {"value": v, "unit": "Gbp/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "Gbp/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


def _claim_stdout():
    """libraries (NCCL prints its version) must not pollute the one-JSON-line contract: route fd 1
    to stderr for the whole run and hand back a private handle on the real stdout."""
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(os.dup(2), "w", buffering=1)
    return os.fdopen(real, "w", buffering=1)


_REAL_STDOUT = None


def emit(obj):
    _REAL_STDOUT.write(json.dumps(obj) + "\n")
    _REAL_STDOUT.flush()


def main():
    global _REAL_STDOUT
    _REAL_STDOUT = _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", type=int, default=3_200_000_000, help="reference symbols per GPU (C2x)")
    ap.add_argument("--n-myers", type=int, default=250_000_000)
    ap.add_argument("--cpu-sample", type=int, default=1_000_000_000)
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--stream-chunk", type=int, default=64 << 20)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-myers", action="store_true")
    ap.add_argument("--no-small", action="store_true")
    ap.add_argument("--no-c4", action="store_true")
    ap.add_argument("--no-c5", action="store_true")
    ap.add_argument("--no-multi", action="store_true")
    ap.add_argument("--no-stream", action="store_true")
    ap.add_argument("--no-pageable", action="store_true")
    ap.add_argument("--n-c5", type=int, default=3_000_000_000)
    ap.add_argument("--c5-patterns", type=int, default=10_000)
    ap.add_argument("--c4-scale", type=float, default=1.0)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
