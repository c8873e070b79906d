"""libspm_b200 -- B200-native online pattern matching behind the libspm matcher API.

This Python module is a thin ctypes binding of the C ABI in include/spmb.h (libspmb.so, built
from libspm_b200/csrc by `make`), used by tests/ and bench.py.  The reference is a C++ library,
so the user-facing host side is the C++ facade include/libspm_b200/spm.hpp; nothing here
computes anything on the CPU.  If the CUDA library is missing or no GPU is present the calls
fail loudly -- there is no fallback path.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

__all__ = ["Context", "Store", "Pattern", "Hits", "Stream", "Jst", "Multi", "SpmbError", "HIT_DTYPE", "EVENT_DTYPE",
           "ALG_HORSPOOL", "ALG_SHIFTOR", "ALG_MYERS", "ALG_PIGEONHOLE", "ALG_MYERS_FILTERED",
           "ALG_MYERS_PREFIX", "Pigeonhole", "search_anchored", "read_fasta", "read_vcf", "lib", "lib_path"]

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "lib", "libspmb.so")

ALG_HORSPOOL, ALG_SHIFTOR, ALG_MYERS = 1, 2, 3
ALG_PIGEONHOLE, ALG_MYERS_FILTERED, ALG_MYERS_PREFIX = 4, 5, 6

HIT_DTYPE = np.dtype([("end", "<u8"), ("seq_id", "<u4"), ("pattern_id", "<u4"), ("errors", "<u4"),
                      ("reserved", "<u4")])
EVENT_DTYPE = np.dtype([("ref_pos", "<u8"), ("del_len", "<u4"), ("ins_len", "<u4"), ("ins_off", "<u8")])

# every symbol include/spmb.h declares (tests check the library exports all of them)
EXPORTS = [
    "spmb_version", "spmb_status_string", "spmb_context_create", "spmb_context_destroy",
    "spmb_context_stream", "spmb_last_error", "spmb_context_sync", "spmb_store_create",
    "spmb_store_create_from_device", "spmb_store_create_journaled", "spmb_store_destroy",
    "spmb_store_length", "spmb_store_row_len", "spmb_store_device_bytes", "spmb_store_read",
    "spmb_pattern_create", "spmb_pattern_set_create", "spmb_pattern_destroy",
    "spmb_pattern_window_size", "spmb_pattern_count", "spmb_hits_create", "spmb_hits_destroy",
    "spmb_hits_count", "spmb_hits_data", "spmb_search_launch", "spmb_search_finish", "spmb_search",
    "spmb_search_host", "spmb_last_launch_count", "spmb_stream_create", "spmb_stream_destroy",
    "spmb_stream_feed", "spmb_stream_capture", "spmb_stream_restore", "spmb_measure_int_peak",
    "spmb_device_info", "spmb_jst_create", "spmb_jst_destroy", "spmb_jst_search", "spmb_jst_last_stats",
    "spmb_fasta_read", "spmb_fasta_destroy", "spmb_fasta_count", "spmb_fasta_name", "spmb_fasta_ranks",
    "spmb_fasta_unknown_symbols", "spmb_vcf_read", "spmb_vcf_destroy", "spmb_vcf_variant_count", "spmb_vcf_variants",
    "spmb_vcf_pool", "spmb_vcf_coverage", "spmb_vcf_haplotype_count", "spmb_vcf_sample_count", "spmb_vcf_sample_name",
    "spmb_vcf_stats",
    "spmb_context_set_shard", "spmb_context_set_myers_cutoff", "spmb_pigeonhole_create", "spmb_pattern_filter_info", "spmb_search_anchored",
    "spmb_upload_last_stats", "spmb_host_pack_chunks", "spmb_context_trim",
    "spmb_store_create_wide", "spmb_store_create_wide_from_device", "spmb_store_symbol_bits",
    "spmb_pattern_create_wide", "spmb_pattern_symbol_bits", "spmb_search_host_wide", "spmb_fasta_read_alphabet",
    "spmb_host_pack_events", "spmb_vcf_dropped_calls",
    "spmb_context_set_upload_threads", "spmb_device_count",
    "spmb_stream_submit", "spmb_stream_collect", "spmb_stream_errors", "spmb_stream_offset",
    "spmb_multi_create", "spmb_multi_destroy", "spmb_multi_device_count", "spmb_multi_context", "spmb_multi_last_error",
    "spmb_multi_store_create", "spmb_multi_store_create_journaled", "spmb_multi_store_destroy", "spmb_multi_store_length",
    "spmb_multi_store_shards", "spmb_multi_pattern_create", "spmb_multi_pattern_destroy", "spmb_multi_search",
    "spmb_multi_search_host", "spmb_multi_last_stats",
    "spmb_search_filtered_noalloc", "spmb_jst_search_launch", "spmb_jst_search_finish",
]


class SpmbError(RuntimeError):
    def __init__(self, code, where, detail=""):
        self.code = code
        super().__init__(f"{where}: status {code} ({_status(code)}) {detail}".strip())


_lib = None


def lib_path() -> str:
    return _SO


def lib():
    """load libspmb.so (raises if it has not been built -- there is no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            raise ImportError(f"{_SO} not found: build it with `make` (or __graft_entry__.build()); "
                              "libspm_b200 has no CPU fallback")
        L = C.CDLL(_SO)
        vp, u64, u32, i32 = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int
        pp = C.POINTER(C.c_void_p)
        sig = {
            "spmb_version": (i32, []),
            "spmb_status_string": (C.c_char_p, [i32]),
            "spmb_context_create": (i32, [i32, vp, pp]),
            "spmb_context_destroy": (i32, [vp]),
            "spmb_context_stream": (vp, [vp]),
            "spmb_last_error": (C.c_char_p, [vp]),
            "spmb_context_sync": (i32, [vp]),
            "spmb_context_trim": (i32, [vp]),
            "spmb_context_set_shard": (i32, [vp, u32, u32]),
            "spmb_context_set_myers_cutoff": (i32, [vp, i32]),
            "spmb_store_create": (i32, [vp, vp, u64, u32, pp]),
            "spmb_store_create_from_device": (i32, [vp, vp, u64, u32, pp]),
            "spmb_store_create_wide": (i32, [vp, vp, u64, u32, u32, pp]),
            "spmb_store_create_wide_from_device": (i32, [vp, vp, u64, u32, u32, pp]),
            "spmb_store_symbol_bits": (u32, [vp]),
            "spmb_pattern_create_wide": (i32, [vp, i32, vp, u32, u32, u32, pp]),
            "spmb_pattern_symbol_bits": (u32, [vp]),
            "spmb_search_host_wide": (i32, [vp, vp, u64, u32, vp, vp]),
            "spmb_upload_last_stats": (i32, [vp, C.POINTER(u64 * 8)]),
            "spmb_host_pack_chunks": (None, [vp, u64, vp]),
            "spmb_host_pack_events": (i32, [vp, u32, u64, vp]),
            "spmb_store_create_journaled": (i32, [vp, vp, vp, u64, vp, u64, u32, pp]),
            "spmb_store_destroy": (i32, [vp]),
            "spmb_store_length": (u64, [vp]),
            "spmb_store_row_len": (u32, [vp]),
            "spmb_store_device_bytes": (u64, [vp]),
            "spmb_store_read": (i32, [vp, vp, u64, u64, vp]),
            "spmb_pattern_create": (i32, [vp, i32, vp, u32, u32, pp]),
            "spmb_pattern_set_create": (i32, [vp, i32, vp, vp, u32, u32, pp]),
            "spmb_fasta_read": (i32, [C.c_char_p, pp]),
            "spmb_fasta_read_alphabet": (i32, [C.c_char_p, u32, pp]),
            "spmb_fasta_destroy": (i32, [vp]),
            "spmb_fasta_count": (u64, [vp]),
            "spmb_fasta_name": (C.c_char_p, [vp, u64]),
            "spmb_fasta_ranks": (vp, [vp, u64, C.POINTER(u64)]),
            "spmb_fasta_unknown_symbols": (u64, [vp]),
            "spmb_vcf_read": (i32, [C.c_char_p, C.c_char_p, u64, pp]),
            "spmb_vcf_destroy": (i32, [vp]),
            "spmb_vcf_variant_count": (u64, [vp]),
            "spmb_vcf_variants": (vp, [vp]),
            "spmb_vcf_pool": (vp, [vp, C.POINTER(u64)]),
            "spmb_vcf_coverage": (vp, [vp, C.POINTER(u32)]),
            "spmb_vcf_haplotype_count": (u32, [vp]),
            "spmb_vcf_sample_count": (u32, [vp]),
            "spmb_vcf_sample_name": (C.c_char_p, [vp, u32]),
            "spmb_vcf_stats": (i32, [vp, C.POINTER(u64 * 3)]),
            "spmb_vcf_dropped_calls": (u64, [vp]),
            "spmb_context_set_upload_threads": (i32, [vp, i32]),
            "spmb_device_count": (i32, []),
            "spmb_stream_submit": (i32, [vp, vp, vp, u64]),
            "spmb_stream_collect": (i32, [vp, vp, vp]),
            "spmb_stream_errors": (i32, [vp, u32, C.POINTER(u32)]),
            "spmb_stream_offset": (u64, [vp]),
            "spmb_multi_create": (i32, [vp, u32, pp]),
            "spmb_multi_destroy": (i32, [vp]),
            "spmb_multi_device_count": (u32, [vp]),
            "spmb_multi_context": (vp, [vp, u32]),
            "spmb_multi_last_error": (C.c_char_p, [vp]),
            "spmb_multi_store_create": (i32, [vp, vp, u64, u32, pp]),
            "spmb_multi_store_create_journaled": (i32, [vp, vp, vp, u64, vp, u64, pp]),
            "spmb_multi_store_destroy": (i32, [vp]),
            "spmb_multi_store_length": (u64, [vp]),
            "spmb_multi_store_shards": (u32, [vp]),
            "spmb_multi_pattern_create": (i32, [vp, i32, vp, vp, u32, u32, pp]),
            "spmb_multi_pattern_destroy": (i32, [vp]),
            "spmb_multi_search": (i32, [vp, vp, vp, vp]),
            "spmb_multi_search_host": (i32, [vp, vp, u64, vp, vp]),
            "spmb_multi_last_stats": (i32, [vp, C.POINTER(C.c_double * 4)]),
            "spmb_pigeonhole_create": (i32, [vp, vp, vp, u32, C.c_double, pp]),
            "spmb_pattern_filter_info": (i32, [vp, C.POINTER(u32 * 5)]),
            "spmb_search_anchored": (i32, [vp, vp, vp, vp, u64, vp]),
            "spmb_pattern_destroy": (i32, [vp]),
            "spmb_pattern_window_size": (u32, [vp]),
            "spmb_pattern_count": (u32, [vp]),
            "spmb_hits_create": (i32, [vp, u64, pp]),
            "spmb_hits_destroy": (i32, [vp]),
            "spmb_hits_count": (u64, [vp]),
            "spmb_hits_data": (vp, [vp]),
            "spmb_search_launch": (i32, [vp, vp, vp, vp]),
            "spmb_search_finish": (i32, [vp, vp]),
            "spmb_search": (i32, [vp, vp, vp, vp]),
            "spmb_search_host": (i32, [vp, vp, u64, vp, vp]),
            "spmb_search_filtered_noalloc": (i32, [vp, vp, vp, vp, u64, vp, C.POINTER(u64)]),
            "spmb_last_launch_count": (u32, [vp]),
            "spmb_stream_create": (i32, [vp, vp, pp]),
            "spmb_stream_destroy": (i32, [vp]),
            "spmb_stream_feed": (i32, [vp, vp, vp, u64, vp]),
            "spmb_stream_capture": (i32, [vp, vp, C.POINTER(C.c_size_t)]),
            "spmb_stream_restore": (i32, [vp, vp, C.c_size_t]),
            "spmb_measure_int_peak": (i32, [vp, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
            "spmb_device_info": (i32, [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(u64)]),
            "spmb_jst_create": (i32, [vp, vp, vp, u64, vp, u64, vp, u32, pp]),
            "spmb_jst_destroy": (i32, [vp]),
            "spmb_jst_search": (i32, [vp, vp, vp, vp]),
            "spmb_jst_search_launch": (i32, [vp, vp, vp, vp]),
            "spmb_jst_search_finish": (i32, [vp, vp]),
            "spmb_jst_last_stats": (i32, [vp, C.POINTER(u64 * 4), C.POINTER(C.c_double * 4)]),
        }
        for name, (res, args) in sig.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def _status(code):
    try:
        return lib().spmb_status_string(code).decode()
    except Exception:
        return "?"


def host_pack_events(events, pool_len: int):
    """spmb_host_pack_events: (headers uint64[n_blk, 2], records uint64[n]) or None when an event does not fit."""
    ev = np.ascontiguousarray(events, dtype=EVENT_DTYPE)
    n = ev.size
    n_blk = (n + 1023) // 1024
    out = np.zeros(2 * n_blk + n, dtype=np.uint64)
    ok = lib().spmb_host_pack_events(ev.ctypes.data if n else None, n, pool_len, out.ctypes.data)
    if not ok:
        return None
    return out[:2 * n_blk].reshape(n_blk, 2), out[2 * n_blk:]


def host_pack_chunks(ranks) -> np.ndarray:
    """the upload pipeline's host packing loop: 1-byte ranks -> one uint64 per 32 symbols
    (low half = symbols 0..15 with the first on top, high half = symbols 16..31)."""
    a = np.ascontiguousarray(ranks, dtype=np.uint8)
    out = np.zeros((a.size + 31) // 32, dtype=np.uint64)
    lib().spmb_host_pack_chunks(a.ctypes.data, a.size, out.ctypes.data)
    return out


def _ranks(a):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    return a


class Context:
    """one device + one stream (spmb_context).  `stream` is a raw cudaStream_t (int) or None."""

    def __init__(self, device: int = 0, stream: int | None = None):
        self._h = C.c_void_p()
        rc = lib().spmb_context_create(int(device), C.c_void_p(stream) if stream else None, C.byref(self._h))
        if rc:
            raise SpmbError(rc, "spmb_context_create")
        self.device = device

    def _check(self, rc, where):
        if rc:
            raise SpmbError(rc, where, (lib().spmb_last_error(self._h) or b"").decode())

    def sync(self):
        self._check(lib().spmb_context_sync(self._h), "spmb_context_sync")

    def trim(self):
        """release cached scratch / upload threads / pool memory (spmb_context_trim)."""
        self._check(lib().spmb_context_trim(self._h), "spmb_context_trim")

    def set_myers_cutoff(self, enabled: bool):
        """Ukkonen cut-off for multi-word Myers needles (default on); off = the full recurrence."""
        self._check(lib().spmb_context_set_myers_cutoff(self._h, 1 if enabled else 0), "spmb_context_set_myers_cutoff")

    def set_upload_threads(self, n: int):
        """host threads that pack chunks for this context's upload pipeline (-1 = default)."""
        self._check(lib().spmb_context_set_upload_threads(self._h, n), "spmb_context_set_upload_threads")

    def set_shard(self, shard: int, n_shards: int):
        """later searches cover only this shard (contiguous super-tile range) of every store."""
        self._check(lib().spmb_context_set_shard(self._h, shard, n_shards), "spmb_context_set_shard")

    @property
    def stream(self) -> int:
        return int(lib().spmb_context_stream(self._h) or 0)

    @property
    def last_launch_count(self) -> int:
        return int(lib().spmb_last_launch_count(self._h))

    def upload_last_stats(self):
        """figures of the last chunked host upload (spmb_upload_last_stats)."""
        st = (C.c_uint64 * 8)()
        self._check(lib().spmb_upload_last_stats(self._h, C.byref(st)), "spmb_upload_last_stats")
        return {"raw_chunks": st[0], "host_packed_chunks": st[1], "chunk_symbols": st[2], "host_threads": st[3],
                "event_raw_chunks": st[4], "event_host_packed_chunks": st[5], "events_per_chunk": st[6]}

    def device_info(self):
        sm, khz, mem = C.c_int(), C.c_int(), C.c_uint64()
        self._check(lib().spmb_device_info(self._h, C.byref(sm), C.byref(khz), C.byref(mem)), "spmb_device_info")
        return {"sm_count": sm.value, "sm_clock_khz": khz.value, "mem_bytes": mem.value}

    def measure_int_peak(self):
        a, b = C.c_double(), C.c_double()
        self._check(lib().spmb_measure_int_peak(self._h, C.byref(a), C.byref(b)), "spmb_measure_int_peak")
        return {"lop3_gops": a.value, "iadd3_gops": b.value}

    def close(self):
        if self._h:
            lib().spmb_context_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Store:
    """packed sequence in HBM (spmb_store)."""

    def __init__(self, ctx: Context, handle):
        self.ctx = ctx
        self._h = handle

    @classmethod
    def from_host(cls, ctx: Context, ranks, row_len: int = 0):
        r = _ranks(ranks)
        h = C.c_void_p()
        ctx._check(lib().spmb_store_create(ctx._h, r.ctypes.data, r.size, row_len, C.byref(h)), "spmb_store_create")
        return cls(ctx, h)

    @classmethod
    def from_device(cls, ctx: Context, dev_ptr: int, n: int, row_len: int = 0):
        h = C.c_void_p()
        ctx._check(lib().spmb_store_create_from_device(ctx._h, C.c_void_p(dev_ptr), n, row_len, C.byref(h)),
                   "spmb_store_create_from_device")
        return cls(ctx, h)

    @classmethod
    def from_host_wide(cls, ctx: Context, ranks, bits: int = 4, row_len: int = 0):
        """4-bit alphabets (spm::dna5: bits=3, spm::dna15: bits=4): ranks 0..15, one byte each."""
        r = np.ascontiguousarray(ranks, dtype=np.uint8).reshape(-1)
        h = C.c_void_p()
        ctx._check(lib().spmb_store_create_wide(ctx._h, r.ctypes.data, r.size, bits, row_len, C.byref(h)),
                   "spmb_store_create_wide")
        return cls(ctx, h)

    @classmethod
    def from_device_wide(cls, ctx: Context, dev_ptr: int, n: int, bits: int = 4, row_len: int = 0):
        h = C.c_void_p()
        ctx._check(lib().spmb_store_create_wide_from_device(ctx._h, C.c_void_p(dev_ptr), n, bits, row_len, C.byref(h)),
                   "spmb_store_create_wide_from_device")
        return cls(ctx, h)

    @property
    def symbol_bits(self):
        return int(lib().spmb_store_symbol_bits(self._h))

    def journaled(self, events, pool, row_len: int = 0):
        ev = np.ascontiguousarray(events, dtype=EVENT_DTYPE)
        pl = _ranks(pool)
        h = C.c_void_p()
        self.ctx._check(lib().spmb_store_create_journaled(self.ctx._h, self._h, ev.ctypes.data if ev.size else None,
                                                          ev.size, pl.ctypes.data if pl.size else None, pl.size,
                                                          row_len, C.byref(h)), "spmb_store_create_journaled")
        return Store(self.ctx, h)

    def __len__(self):
        return int(lib().spmb_store_length(self._h))

    @property
    def row_len(self):
        return int(lib().spmb_store_row_len(self._h))

    @property
    def device_bytes(self):
        return int(lib().spmb_store_device_bytes(self._h))

    def read(self, begin=0, n=None):
        n = len(self) - begin if n is None else n
        out = np.empty(n, dtype=np.uint8)
        self.ctx._check(lib().spmb_store_read(self.ctx._h, self._h, begin, n, out.ctypes.data), "spmb_store_read")
        return out

    def close(self):
        if self._h:
            lib().spmb_store_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Pattern:
    """compiled needle(s) (spmb_pattern) == horspool_matcher / shiftor_matcher / myers_matcher ctor."""

    def __init__(self, ctx: Context, alg: int, needle, k: int = 0, bits: int = 2):
        self.ctx = ctx
        self._h = C.c_void_p()
        if bits != 2:   # spm::dna5 / spm::dna15 needles: ranks 0..15 (spmb_pattern_create_wide)
            r = np.ascontiguousarray(needle, dtype=np.uint8).reshape(-1)
            ctx._check(lib().spmb_pattern_create_wide(ctx._h, alg, r.ctypes.data if r.size else None, r.size, k, bits,
                                                      C.byref(self._h)), "spmb_pattern_create_wide")
        elif isinstance(needle, (list, tuple)) and len(needle) and not np.isscalar(needle[0]):
            parts = [_ranks(p) for p in needle]
            lens = np.array([p.size for p in parts], dtype=np.uint32)
            cat = np.concatenate(parts) if parts else np.zeros(0, np.uint8)
            ctx._check(lib().spmb_pattern_set_create(ctx._h, alg, cat.ctypes.data, lens.ctypes.data, len(parts), k,
                                                     C.byref(self._h)), "spmb_pattern_set_create")
        else:
            r = _ranks(needle)
            ctx._check(lib().spmb_pattern_create(ctx._h, alg, r.ctypes.data if r.size else None, r.size, k,
                                                 C.byref(self._h)), "spmb_pattern_create")
        self.alg, self.k = alg, k

    @property
    def window_size(self):
        return int(lib().spmb_pattern_window_size(self._h))

    @property
    def filter_info(self):
        """figures of the pigeonhole seed filter behind ALG_PIGEONHOLE / ALG_MYERS_FILTERED patterns."""
        info = (C.c_uint32 * 5)()
        lib().spmb_pattern_filter_info(self._h, C.byref(info))
        return {"q": info[0], "seeds": info[1], "distinct_qgrams": info[2], "log2_bitmap_bits": info[3],
                "filtered": bool(info[4]), "step": info[4]}

    @property
    def count(self):
        return int(lib().spmb_pattern_count(self._h))

    def close(self):
        if self._h:
            lib().spmb_pattern_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Pigeonhole(Pattern):
    """spm::pigeonhole_matcher{needle(s), error_rate}: q-gram seed search (spmb_pigeonhole_create)."""

    def __init__(self, ctx: Context, needles, error_rate: float = 0.0):
        self.ctx = ctx
        self._h = C.c_void_p()
        if not (isinstance(needles, (list, tuple)) and len(needles) and not np.isscalar(needles[0])):
            needles = [needles]
        parts = [_ranks(p) for p in needles]
        lens = np.array([p.size for p in parts], dtype=np.uint32)
        cat = np.concatenate(parts) if parts else np.zeros(0, np.uint8)
        ctx._check(lib().spmb_pigeonhole_create(ctx._h, cat.ctypes.data if cat.size else None, lens.ctypes.data,
                                                len(parts), float(error_rate), C.byref(self._h)),
                   "spmb_pigeonhole_create")
        self.alg, self.k = ALG_PIGEONHOLE, 0


class Hits:
    def __init__(self, ctx: Context, capacity: int = 0):
        self.ctx = ctx
        self._h = C.c_void_p()
        ctx._check(lib().spmb_hits_create(ctx._h, capacity, C.byref(self._h)), "spmb_hits_create")

    def __len__(self):
        return int(lib().spmb_hits_count(self._h))

    def array(self, copy: bool = True) -> np.ndarray:
        """the hit list; copy=False: a view of the library's host memory (spmb_hits_data), valid until the next
        search on this hit list."""
        n = len(self)
        if n == 0:
            return np.zeros(0, dtype=HIT_DTYPE)
        ptr = lib().spmb_hits_data(self._h)
        buf = (C.c_char * (n * HIT_DTYPE.itemsize)).from_address(ptr)
        a = np.frombuffer(buf, dtype=HIT_DTYPE, count=n)
        return a.copy() if copy else a

    def close(self):
        if self._h:
            lib().spmb_hits_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def search(ctx: Context, store: Store, pattern: Pattern, hits: Hits | None = None) -> np.ndarray:
    """seqan_pattern_base::operator(): all hits, sorted by (pattern_id, seq_id, end)."""
    own = hits is None
    hits = hits or Hits(ctx)
    ctx._check(lib().spmb_search(ctx._h, store._h, pattern._h, hits._h), "spmb_search")
    out = hits.array()
    if own:
        hits.close()
    return out


def search_anchored(ctx: Context, store: Store, pattern: Pattern, anchors, hits: Hits | None = None) -> np.ndarray:
    """restorable_myers_prefix_matcher at many anchors: seq_id = anchor index, end absolute."""
    own = hits is None
    hits = hits or Hits(ctx)
    a = np.ascontiguousarray(anchors, dtype=np.uint64)
    ctx._check(lib().spmb_search_anchored(ctx._h, store._h, pattern._h, a.ctypes.data if a.size else None, a.size,
                                          hits._h), "spmb_search_anchored")
    out = hits.array()
    if own:
        hits.close()
    return out


def search_launch(ctx: Context, store: Store, pattern: Pattern, hits: Hits):
    ctx._check(lib().spmb_search_launch(ctx._h, store._h, pattern._h, hits._h), "spmb_search_launch")


def search_finish(ctx: Context, hits: Hits, copy: bool = True) -> np.ndarray:
    ctx._check(lib().spmb_search_finish(ctx._h, hits._h), "spmb_search_finish")
    return hits.array(copy)


def search_host(ctx: Context, haystack, pattern: Pattern, hits: Hits | None = None, ptr: int | None = None,
                n: int | None = None, bits: int = 2) -> np.ndarray:
    """the reference's call shape: host haystack in, hits out (upload + pack + search)."""
    own = hits is None
    hits = hits or Hits(ctx)
    if ptr is None:
        r = _ranks(haystack) if bits == 2 else np.ascontiguousarray(haystack, dtype=np.uint8).reshape(-1)
        ptr, n = r.ctypes.data, r.size
    if bits != 2:
        ctx._check(lib().spmb_search_host_wide(ctx._h, C.c_void_p(ptr), n, bits, pattern._h, hits._h), "spmb_search_host_wide")
    else:
        ctx._check(lib().spmb_search_host(ctx._h, C.c_void_p(ptr), n, pattern._h, hits._h), "spmb_search_host")
    out = hits.array()
    if own:
        hits.close()
    return out


class Stream:
    """resumable search (restorable_myers_matcher / restorable_shiftor_matcher)."""

    def __init__(self, ctx: Context, pattern: Pattern):
        self.ctx, self.pattern = ctx, pattern
        self._h = C.c_void_p()
        ctx._check(lib().spmb_stream_create(ctx._h, pattern._h, C.byref(self._h)), "spmb_stream_create")
        self._hits = Hits(ctx)
        self._held = []   # arrays of the chunks in flight: the C ABI reads them until collect()

    def feed(self, chunk) -> np.ndarray:
        r = _ranks(chunk)
        self.ctx._check(lib().spmb_stream_feed(self.ctx._h, self._h, r.ctypes.data if r.size else None, r.size,
                                               self._hits._h), "spmb_stream_feed")
        return self._hits.array()

    def submit(self, chunk=None, ptr: int | None = None, n: int | None = None):
        """asynchronous feed: the chunk goes to the upload pipeline, the hits come from collect(); a raw `ptr` must
        stay valid until then (arrays passed as `chunk` are held here)."""
        r = None
        if ptr is None:
            r = _ranks(chunk)
            ptr, n = (r.ctypes.data if r.size else None), r.size
        self.ctx._check(lib().spmb_stream_submit(self.ctx._h, self._h, C.c_void_p(ptr) if ptr else None, n), "spmb_stream_submit")
        self._held.append(r)

    def collect(self) -> np.ndarray:
        self.ctx._check(lib().spmb_stream_collect(self.ctx._h, self._h, self._hits._h), "spmb_stream_collect")
        if self._held:
            self._held.pop(0)
        return self._hits.array()

    def errors(self, pattern_index: int = 0) -> int:
        """the score of a needle at the current column (capture().errors of restorable_myers_matcher)."""
        e = C.c_uint32()
        self.ctx._check(lib().spmb_stream_errors(self._h, pattern_index, C.byref(e)), "spmb_stream_errors")
        return int(e.value)

    @property
    def offset(self) -> int:
        return int(lib().spmb_stream_offset(self._h))

    def capture(self) -> bytes:
        sz = C.c_size_t(0)
        lib().spmb_stream_capture(self._h, None, C.byref(sz))
        buf = C.create_string_buffer(sz.value)
        rc = lib().spmb_stream_capture(self._h, buf, C.byref(sz))
        if rc:
            raise SpmbError(rc, "spmb_stream_capture")
        return bytes(buf.raw[:sz.value])

    def restore(self, state: bytes):
        rc = lib().spmb_stream_restore(self._h, state, len(state))
        if rc:
            raise SpmbError(rc, "spmb_stream_restore")

    def close(self):
        if self._h:
            lib().spmb_stream_destroy(self._h)
            self._h = C.c_void_p()
        self._hits.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Jst:
    """journaled string tree over a resident reference store (spmb_jst)."""

    def __init__(self, ctx: Context, reference: Store, variants, pool, coverage, n_haplotypes: int):
        self.ctx, self.reference = ctx, reference
        var = np.ascontiguousarray(variants, dtype=EVENT_DTYPE)
        pl = _ranks(pool)
        cov = np.ascontiguousarray(coverage, dtype=np.uint64)
        wpr = (n_haplotypes + 63) // 64
        assert cov.shape == (len(var), wpr), (cov.shape, len(var), wpr)
        self._h = C.c_void_p()
        ctx._check(lib().spmb_jst_create(ctx._h, reference._h, var.ctypes.data if var.size else None, var.size,
                                         pl.ctypes.data if pl.size else None, pl.size,
                                         cov.ctypes.data if cov.size else None, n_haplotypes, C.byref(self._h)),
                   "spmb_jst_create")
        self.n_haplotypes = n_haplotypes

    def search(self, pattern: Pattern, hits: Hits | None = None) -> np.ndarray:
        own = hits is None
        hits = hits or Hits(self.ctx)
        self.ctx._check(lib().spmb_jst_search(self.ctx._h, self._h, pattern._h, hits._h), "spmb_jst_search")
        out = hits.array()
        if own:
            hits.close()
        return out

    def search_launch(self, pattern: Pattern, hits: Hits):
        """spmb_jst_search_launch: enqueue the kernels of a search and return (finish it with search_finish)."""
        self.ctx._check(lib().spmb_jst_search_launch(self.ctx._h, self._h, pattern._h, hits._h), "spmb_jst_search_launch")

    def search_finish(self, hits: Hits, copy: bool = True) -> np.ndarray:
        self.ctx._check(lib().spmb_jst_search_finish(self.ctx._h, hits._h), "spmb_jst_search_finish")
        return hits.array(copy)

    def last_stats(self):
        st, ms = (C.c_uint64 * 4)(), (C.c_double * 4)()
        lib().spmb_jst_last_stats(self._h, C.byref(st), C.byref(ms))
        return {"windows": st[0], "window_symbols": st[1], "buckets": st[2], "raw_hits": st[3],
                "flatten_ms": ms[0], "materialise_ms": ms[1], "kernels_ms": ms[2], "expand_ms": ms[3]}

    def close(self):
        if self._h:
            lib().spmb_jst_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _np_from(ptr, n, dtype):
    if not ptr or n == 0:
        return np.zeros(0, dtype=dtype)
    dt = np.dtype(dtype)
    buf = (C.c_char * (n * dt.itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dt, count=n).copy()


def read_fasta(path: str, bits: int = 2):
    """[(name, ranks)] of a FASTA file (plain or gzip) through spmb_fasta_read; also returns the number of
    symbols outside the alphabet (bits=2: dna4, stored as rank 0 like seqan3::dna4; bits=3 / 4: dna5 / dna15
    in seqan3's rank order, stored as N)."""
    h = C.c_void_p()
    rc = lib().spmb_fasta_read_alphabet(os.fsencode(path), bits, C.byref(h))
    if rc:
        raise SpmbError(rc, "spmb_fasta_read", path)
    try:
        recs = []
        for i in range(int(lib().spmb_fasta_count(h))):
            n = C.c_uint64()
            p = lib().spmb_fasta_ranks(h, i, C.byref(n))
            recs.append((lib().spmb_fasta_name(h, i).decode(), _np_from(p, n.value, np.uint8)))
        return recs, int(lib().spmb_fasta_unknown_symbols(h))
    finally:
        lib().spmb_fasta_destroy(h)


def read_vcf(path: str, contig: str | None = None, ref_len: int = 0):
    """VCF -> dict(variants[EVENT_DTYPE], pool, coverage[n_var, wpr], n_haplotypes, samples, stats): the
    arguments of Jst(...) (spmb_vcf_read)."""
    h = C.c_void_p()
    rc = lib().spmb_vcf_read(os.fsencode(path), contig.encode() if contig else None, int(ref_len), C.byref(h))
    if rc:
        raise SpmbError(rc, "spmb_vcf_read", path)
    try:
        L = lib()
        nv = int(L.spmb_vcf_variant_count(h))
        pl, wpr = C.c_uint64(), C.c_uint32()
        pool = _np_from(L.spmb_vcf_pool(h, C.byref(pl)), pl.value, np.uint8)
        covp = L.spmb_vcf_coverage(h, C.byref(wpr))
        cov = _np_from(covp, nv * wpr.value, np.uint64).reshape(nv, wpr.value)
        st = (C.c_uint64 * 3)()
        L.spmb_vcf_stats(h, C.byref(st))
        return {"variants": _np_from(L.spmb_vcf_variants(h), nv, EVENT_DTYPE), "pool": pool, "coverage": cov,
                "n_haplotypes": int(L.spmb_vcf_haplotype_count(h)),
                "samples": [L.spmb_vcf_sample_name(h, i).decode() for i in range(int(L.spmb_vcf_sample_count(h)))],
                "stats": {"records": st[0], "skipped": st[1], "unknown_symbols": st[2],
                          "dropped_calls": int(L.spmb_vcf_dropped_calls(h))}}
    finally:
        lib().spmb_vcf_destroy(h)


def stream_leg(ctx: Context, host: np.ndarray, pat, hits: Hits, chunk: int = 64 << 20, k: int = 0, alg: int = ALG_SHIFTOR):
    """bench.py's streaming leg: a host haystack fed in `chunk`-symbol pieces through spmb_stream_submit /
    spmb_stream_collect (three chunks in flight), beside the one-shot spmb_search_host of the same haystack."""
    import time
    n = host.size
    base = host.ctypes.data
    p = Pattern(ctx, alg, pat, k)
    st = Stream(ctx, p)
    total_hits = 0

    def run():
        nonlocal total_hits
        total_hits = 0
        inflight = 0
        for o in range(0, n, chunk):
            ln = min(chunk, n - o)
            if inflight == 3:
                total_hits += len(st.collect())
                inflight -= 1
            st.submit(ptr=base + o, n=ln)
            inflight += 1
        while inflight:
            total_hits += len(st.collect())
            inflight -= 1

    run()                                   # warm: buffers sized
    st.close()
    st = Stream(ctx, p)
    ctx.sync()
    t0 = time.perf_counter()
    run()
    ctx.sync()
    dt = time.perf_counter() - t0
    st.close()
    one = search_host(ctx, None, p, hits, ptr=base, n=n)
    ctx.sync()
    dt1 = 1e30
    for _ in range(2):      # best of two: the boxes are shared hosts, a one-off stall of the host side shows up now and then
        t0 = time.perf_counter()
        one = search_host(ctx, None, p, hits, ptr=base, n=n)
        ctx.sync()
        dt1 = min(dt1, time.perf_counter() - t0)
    # a sync feed() of the same chunks (what a caller of the reference's restorable matchers does)
    st = Stream(ctx, p)
    t0 = time.perf_counter()
    fh = 0
    for o in range(0, n, chunk):
        ln = min(chunk, n - o)
        ctx._check(lib().spmb_stream_feed(ctx._h, st._h, C.c_void_p(base + o), ln, st._hits._h), "spmb_stream_feed")
        fh += len(st._hits)
    dt2 = time.perf_counter() - t0
    st.close()
    p.close()
    return {"value": n / dt / 1e9, "unit": "Gbp/s", "chunk_symbols": chunk, "chunks": (n + chunk - 1) // chunk,
            "ms_total": dt * 1e3, "hits": int(total_hits), "one_shot_gbp_per_s": n / dt1 / 1e9, "one_shot_ms": dt1 * 1e3,
            "one_shot_hits": int(len(one)), "ratio_to_one_shot": dt1 / dt, "hits_equal_one_shot": bool(total_hits == len(one) == fh),
            "sync_feed_gbp_per_s": n / dt2 / 1e9,
            "path": "spmb_stream_submit / spmb_stream_collect, three chunks in flight: device-resident tail, only the new chunk "
                    "crosses PCIe (both movers of the upload queue); one-shot = spmb_search_host of the same plain haystack"}


class _MultiStore:
    def __init__(self, multi, handle):
        self.multi, self._h = multi, handle

    def journaled(self, events, pool):
        ev = np.ascontiguousarray(events, dtype=EVENT_DTYPE)
        pl = _ranks(pool)
        h = C.c_void_p()
        self.multi._check(lib().spmb_multi_store_create_journaled(self.multi._h, self._h, ev.ctypes.data if ev.size else None,
                                                                  ev.size, pl.ctypes.data if pl.size else None, pl.size,
                                                                  C.byref(h)), "spmb_multi_store_create_journaled")
        s = _MultiStore(self.multi, h)
        s._keep = (ev, pl)
        return s

    def __len__(self):
        return int(lib().spmb_multi_store_length(self._h))

    @property
    def shards(self):
        return int(lib().spmb_multi_store_shards(self._h))

    def close(self):
        if self._h:
            lib().spmb_multi_store_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class _MultiPattern:
    def __init__(self, multi, handle):
        self.multi, self._h = multi, handle

    def close(self):
        if self._h:
            lib().spmb_multi_pattern_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Multi:
    """the single-process multi-GPU driver (spmb_multi_*): one context per device, host haystacks sharded with
    left context, per-device hit lists merged on the host."""

    def __init__(self, devices=None):
        self._h = C.c_void_p()
        if devices is None:
            rc = lib().spmb_multi_create(None, 0, C.byref(self._h))
        else:
            arr = (C.c_int * len(devices))(*devices)
            rc = lib().spmb_multi_create(arr, len(devices), C.byref(self._h))
        if rc:
            raise SpmbError(rc, "spmb_multi_create")
        self.n_devices = int(lib().spmb_multi_device_count(self._h))
        ctx0 = Context.__new__(Context)
        ctx0._h = C.c_void_p(lib().spmb_multi_context(self._h, 0))
        ctx0.device = 0
        ctx0.close = lambda: None          # owned by the driver
        self._ctx0 = ctx0
        self._hits = Hits(ctx0, 1 << 12)

    def _check(self, rc, where):
        if rc:
            raise SpmbError(rc, where, (lib().spmb_multi_last_error(self._h) or b"").decode())

    def store(self, host, overlap: int = 0, ptr: int | None = None, n: int | None = None):
        if ptr is None:
            r = _ranks(host)
            ptr, n = r.ctypes.data, r.size
        h = C.c_void_p()
        self._check(lib().spmb_multi_store_create(self._h, C.c_void_p(ptr), n, overlap, C.byref(h)), "spmb_multi_store_create")
        return _MultiStore(self, h)

    def pattern(self, alg: int, needle, k: int = 0):
        if isinstance(needle, (list, tuple)) and len(needle) and not np.isscalar(needle[0]):
            parts = [_ranks(p) for p in needle]
        else:
            parts = [_ranks(needle)]
        lens = np.array([p.size for p in parts], dtype=np.uint32)
        cat = np.concatenate(parts) if parts else np.zeros(0, np.uint8)
        h = C.c_void_p()
        self._check(lib().spmb_multi_pattern_create(self._h, alg, cat.ctypes.data if cat.size else None, lens.ctypes.data,
                                                    len(parts), k, C.byref(h)), "spmb_multi_pattern_create")
        return _MultiPattern(self, h)

    def search(self, store, pattern) -> np.ndarray:
        self._check(lib().spmb_multi_search(self._h, store._h, pattern._h, self._hits._h), "spmb_multi_search")
        return self._hits.array()

    def search_host(self, host, pattern, ptr: int | None = None, n: int | None = None) -> np.ndarray:
        if ptr is None:
            r = _ranks(host)
            ptr, n = r.ctypes.data, r.size
        self._check(lib().spmb_multi_search_host(self._h, C.c_void_p(ptr), n, pattern._h, self._hits._h), "spmb_multi_search_host")
        return self._hits.array()

    def last_stats(self):
        ms = (C.c_double * 4)()
        lib().spmb_multi_last_stats(self._h, C.byref(ms))
        return np.array(list(ms))

    def close(self):
        if self._h:
            self._hits.close()
            lib().spmb_multi_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
