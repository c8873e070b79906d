"""CPU-side checks of the boundary: the C-ABI library loads, exports every symbol
include/spmb.h declares, and refuses to work without a device (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

import libspm_b200 as sp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "spmb.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(spmb_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    names = _declared()
    assert len(names) >= 30
    L = C.CDLL(sp.lib_path())
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    assert sorted(sp.EXPORTS) == names


def test_version_and_status_strings():
    L = sp.lib()
    assert L.spmb_version() == 100
    assert L.spmb_status_string(0) == b"ok"
    assert b"no CPU path" in L.spmb_status_string(5)


def test_no_device_means_error_not_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(sp.SpmbError) as ei:
        sp.Context(0)
    assert ei.value.code == 5


def test_hit_record_layout_matches_header():
    assert sp.HIT_DTYPE.itemsize == 24
    assert sp.HIT_DTYPE.fields["end"][1] == 0 and sp.HIT_DTYPE.fields["errors"][1] == 16
    assert sp.EVENT_DTYPE.itemsize == 24
