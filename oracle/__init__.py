"""CPU ORACLE loader (TEST INFRASTRUCTURE -- never imported by libspm_b200).

ctypes binding of oracle/liborc.so (built from oracle/spm_oracle.c by oracle/Makefile).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module.  See oracle/spm_oracle.h for the parity status.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liborc.so")


def build(force: bool = False) -> str:
    src = [os.path.join(_HERE, f) for f in ("spm_oracle.c", "spm_oracle.h", "Makefile")]
    stale = (not os.path.exists(_SO)) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in src)
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "-B", "liborc.so"], stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        u8p, u64p, u32p = C.c_void_p, C.c_void_p, C.c_void_p
        L.orc_exact_bruteforce.restype = C.c_uint64
        L.orc_exact_bruteforce.argtypes = [u8p, C.c_uint64, u8p, C.c_uint32, u64p, C.c_uint64]
        L.orc_sellers.restype = C.c_uint64
        L.orc_sellers.argtypes = [u8p, C.c_uint64, u8p, C.c_uint32, C.c_uint32, u64p, u32p, C.c_uint64]
        L.orc_horspool.restype = C.c_uint64
        L.orc_horspool.argtypes = [u8p, C.c_uint64, u8p, C.c_uint32, u64p, C.c_uint64]
        L.orc_shiftor.restype = C.c_uint64
        L.orc_shiftor.argtypes = [u8p, C.c_uint64, u8p, C.c_uint32, u64p, C.c_uint64]
        L.orc_myers.restype = C.c_uint64
        L.orc_myers.argtypes = [u8p, C.c_uint64, u8p, C.c_uint32, C.c_uint32, u64p, u32p, C.c_uint64]
        L.orc_shiftor_init.restype = C.c_int
        L.orc_shiftor_init.argtypes = [C.c_void_p, u8p, C.c_uint32]
        L.orc_shiftor_run.restype = C.c_uint64
        L.orc_shiftor_run.argtypes = [C.c_void_p, u8p, C.c_uint64, C.c_uint64, u64p, C.c_uint64]
        L.orc_myers_init.restype = C.c_int
        L.orc_myers_init.argtypes = [C.c_void_p, u8p, C.c_uint32, C.c_uint32]
        L.orc_myers_run.restype = C.c_uint64
        L.orc_myers_run.argtypes = [C.c_void_p, u8p, C.c_uint64, C.c_uint64, u64p, u32p, C.c_uint64]
        L.orc_search_mt.restype = C.c_uint64
        L.orc_search_mt.argtypes = [C.c_int, u8p, C.c_uint64, u8p, C.c_uint32, C.c_uint32, C.c_int,
                                    u64p, u32p, C.c_uint64]
        L.orc_max_threads.restype = C.c_int
        L.orc_set_threads.argtypes = [C.c_int]
        L.orc_journal_apply.restype = C.c_uint64
        L.orc_journal_apply.argtypes = [u8p, C.c_uint64, C.c_void_p, C.c_uint64, u8p, u8p, C.c_uint64]
        L.orc_jst_search.restype = C.c_uint64
        L.orc_jst_search.argtypes = [C.c_int, u8p, C.c_uint64, C.c_void_p, C.c_uint64, u8p, C.c_void_p, C.c_uint32,
                                     C.c_uint32, C.c_void_p, C.c_uint32, u8p, C.c_uint32, C.c_uint32, C.c_int,
                                     C.c_void_p, C.c_uint64]
        L.orc_jst_haplotype.restype = C.c_uint64
        L.orc_jst_haplotype.argtypes = [u8p, C.c_uint64, C.c_void_p, C.c_uint64, u8p, C.c_void_p, C.c_uint32,
                                        C.c_uint32, u8p, C.c_uint64]
        L.orc_pigeonhole_q.restype = C.c_uint32
        L.orc_pigeonhole_q.argtypes = [C.c_void_p, C.c_uint32, C.c_double]
        L.orc_pigeonhole.restype = C.c_uint64
        L.orc_pigeonhole.argtypes = [u8p, C.c_uint64, u8p, C.c_void_p, C.c_uint32, C.c_double, u64p, u32p, u32p,
                                     C.c_uint64]
        L.orc_myers_prefix.restype = C.c_uint64
        L.orc_myers_prefix.argtypes = [u8p, C.c_uint64, u8p, C.c_uint32, C.c_uint32, u64p, u32p, C.c_uint64]
        _lib = L
    return _lib


EVENT_DTYPE = np.dtype([("ref_pos", "<u8"), ("del_len", "<u4"), ("ins_len", "<u4"), ("ins_off", "<u8")])

_SHIFTOR_BYTES = 4 * 2 + 16 * 32 * 4 + 32 * 4
_MYERS_BYTES = 32768  # >= sizeof(orc_myers_t) (128 blocks: 19.5 KB)


def _u8(a):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    return a, a.ctypes.data


def _run(fn, text, pat, extra=(), with_err=False, cap=None):
    text, tp = _u8(text)
    pat, pp = _u8(pat)
    cap = int(cap if cap is not None else max(1024, min(len(text), 1 << 22)))
    while True:
        ends = np.empty(cap, dtype=np.uint64)
        errs = np.zeros(cap, dtype=np.uint32)
        args = [tp, len(text), pp, len(pat), *extra, ends.ctypes.data]
        if with_err:
            args.append(errs.ctypes.data)
        args.append(cap)
        cnt = int(fn(*args))
        if cnt <= cap:
            return ends[:cnt].copy(), errs[:cnt].copy()
        cap = cnt


def exact_bruteforce(text, pat):
    return _run(lib().orc_exact_bruteforce, text, pat)[0]


def sellers(text, pat, k):
    return _run(lib().orc_sellers, text, pat, (int(k),), with_err=True)


def horspool(text, pat):
    return _run(lib().orc_horspool, text, pat)[0]


def shiftor(text, pat):
    return _run(lib().orc_shiftor, text, pat)[0]


def myers(text, pat, k):
    return _run(lib().orc_myers, text, pat, (int(k),), with_err=True)


def search_mt(alg, text, pat, k=0, threads=None, store=True, cap=None):
    """alg: 1 horspool, 2 shiftor, 3 myers.  Returns (count, ends, errs) (ends sorted)."""
    L = lib()
    text, tp = _u8(text)
    pat, pp = _u8(pat)
    threads = int(threads or L.orc_max_threads())
    if not store:
        cnt = int(L.orc_search_mt(alg, tp, len(text), pp, len(pat), int(k), threads, None, None, 0))
        return cnt, None, None
    cap = int(cap or (1 << 20))
    while True:
        ends = np.empty(cap, dtype=np.uint64)
        errs = np.zeros(cap, dtype=np.uint32)
        cnt = int(L.orc_search_mt(alg, tp, len(text), pp, len(pat), int(k), threads,
                                  ends.ctypes.data, errs.ctypes.data, cap))
        if cnt <= cap:
            return cnt, ends[:cnt].copy(), errs[:cnt].copy()
        cap = cnt


def max_threads():
    return int(lib().orc_max_threads())


def use_all_cores():
    """torchrun exports OMP_NUM_THREADS=1; the CPU baseline is entitled to every core it may run on."""
    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    lib().orc_set_threads(n)
    return n


class ShiftOrState:
    """resumable Shift-Or (capture/restore = copy of the state struct)."""

    def __init__(self, pat):
        self.buf = C.create_string_buffer(_SHIFTOR_BYTES)
        pat, pp = _u8(pat)
        assert lib().orc_shiftor_init(self.buf, pp, len(pat)) == 0

    def capture(self):
        return bytes(self.buf.raw)

    def restore(self, st):
        C.memmove(self.buf, st, len(st))

    def run(self, chunk, base=0):
        chunk, cp = _u8(chunk)
        cap = len(chunk) + 1
        ends = np.empty(cap, dtype=np.uint64)
        cnt = int(lib().orc_shiftor_run(self.buf, cp, len(chunk), int(base), ends.ctypes.data, cap))
        return ends[:cnt].copy()


class MyersState:
    """resumable Myers (capture/restore = copy of the whole pattern state)."""

    def __init__(self, pat, k):
        self.buf = C.create_string_buffer(_MYERS_BYTES)
        pat, pp = _u8(pat)
        assert lib().orc_myers_init(self.buf, pp, len(pat), int(k)) == 0

    def capture(self):
        return bytes(self.buf.raw)

    def restore(self, st):
        C.memmove(self.buf, st, len(st))

    def run(self, chunk, base=0):
        chunk, cp = _u8(chunk)
        cap = len(chunk) + 1
        ends = np.empty(cap, dtype=np.uint64)
        errs = np.empty(cap, dtype=np.uint32)
        cnt = int(lib().orc_myers_run(self.buf, cp, len(chunk), int(base), ends.ctypes.data,
                                      errs.ctypes.data, cap))
        return ends[:cnt].copy(), errs[:cnt].copy()


def journal_apply(ref, events, pool):
    """events: structured array of EVENT_DTYPE sorted by ref_pos."""
    L = lib()
    ref, rp = _u8(ref)
    pool, pp = _u8(pool if len(pool) else np.zeros(1, np.uint8))
    ev = np.ascontiguousarray(events, dtype=EVENT_DTYPE)
    n = int(L.orc_journal_apply(rp, len(ref), ev.ctypes.data, len(ev), pp, None, 0))
    if n == (1 << 64) - 1:
        raise ValueError("journal events unsorted or overlapping")
    out = np.empty(n, dtype=np.uint8)
    L.orc_journal_apply(rp, len(ref), ev.ctypes.data, len(ev), pp, out.ctypes.data, n)
    return out


JST_HIT_DTYPE = np.dtype([("end", "<u8"), ("hap", "<u4"), ("err", "<u4")])


def jst_haplotype(ref, variants, pool, cov, hap):
    """cov: uint64 array [n_var, words_per_row]."""
    L = lib()
    ref, rp = _u8(ref)
    pool, pp = _u8(pool if len(pool) else np.zeros(1, np.uint8))
    var = np.ascontiguousarray(variants, dtype=EVENT_DTYPE)
    cov = np.ascontiguousarray(cov, dtype=np.uint64)
    wpr = cov.shape[1]
    n = int(L.orc_jst_haplotype(rp, len(ref), var.ctypes.data, len(var), pp, cov.ctypes.data, wpr, int(hap), None, 0))
    out = np.empty(n, dtype=np.uint8)
    L.orc_jst_haplotype(rp, len(ref), var.ctypes.data, len(var), pp, cov.ctypes.data, wpr, int(hap), out.ctypes.data, n)
    return out


def jst_search(alg, ref, variants, pool, cov, n_hap, pat, k=0, haps=None, threads=None):
    """materialise every (listed) haplotype and search it; returns JST_HIT_DTYPE sorted by (hap, end)."""
    L = lib()
    ref, rp = _u8(ref)
    pool, pp = _u8(pool if len(pool) else np.zeros(1, np.uint8))
    pat, tp = _u8(pat)
    var = np.ascontiguousarray(variants, dtype=EVENT_DTYPE)
    cov = np.ascontiguousarray(cov, dtype=np.uint64)
    wpr = cov.shape[1]
    hl = None if haps is None else np.ascontiguousarray(haps, dtype=np.uint32)
    threads = int(threads or L.orc_max_threads())
    cap = 1 << 16
    while True:
        out = np.empty(cap, dtype=JST_HIT_DTYPE)
        cnt = int(L.orc_jst_search(alg, rp, len(ref), var.ctypes.data, len(var), pp, cov.ctypes.data, int(n_hap), wpr,
                                   None if hl is None else hl.ctypes.data, 0 if hl is None else len(hl), tp, len(pat),
                                   int(k), threads, out.ctypes.data, cap))
        if cnt <= cap:
            return out[:cnt].copy()
        cap = cnt


SEED_DTYPE = np.dtype([("begin", "<u8"), ("needle", "<u4"), ("offset", "<u4")])


def pigeonhole_q(lengths, error_rate=0.0):
    ln = np.ascontiguousarray(lengths, dtype=np.uint32)
    return int(lib().orc_pigeonhole_q(ln.ctypes.data, len(ln), float(error_rate)))


def pigeonhole(text, needles, error_rate=0.0):
    """seed hits of spm::pigeonhole_matcher{needles, error_rate}: SEED_DTYPE in haystack order."""
    L = lib()
    text, tp = _u8(text)
    parts = [np.ascontiguousarray(x, dtype=np.uint8) for x in needles]
    ln = np.array([len(x) for x in parts], dtype=np.uint32)
    cat, cp = _u8(np.concatenate(parts) if parts else np.zeros(1, np.uint8))
    cap = 1 << 16
    while True:
        b = np.empty(cap, dtype=np.uint64)
        i = np.empty(cap, dtype=np.uint32)
        o = np.empty(cap, dtype=np.uint32)
        cnt = int(L.orc_pigeonhole(tp, len(text), cp, ln.ctypes.data, len(ln), float(error_rate), b.ctypes.data,
                                   i.ctypes.data, o.ctypes.data, cap))
        if cnt <= cap:
            out = np.empty(cnt, dtype=SEED_DTYPE)
            out["begin"], out["needle"], out["offset"] = b[:cnt], i[:cnt], o[:cnt]
            return out
        cap = cnt


def myers_prefix(text, pat, k):
    """restorable_myers_prefix_matcher: (ends, errors) of the needle against prefixes of text."""
    return _run(lib().orc_myers_prefix, text, pat, (int(k),), with_err=True, cap=len(pat) + int(k) + 2)
