"""The C-ABI library loads without a GPU and exports every function include/bi_b200.h declares.  CPU only."""
from __future__ import annotations

import ctypes as C
import re

import pytest

from bayesian_inference_b200 import _lib


@pytest.fixture(scope="module")
def lib():
    _lib.build()          # nvcc cross-compiles sm_100a without a GPU
    return _lib.load()


def _declared():
    text = _lib.HEADER.read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(bi_[a-z_]+)\s*\(", text)))


def test_header_and_export_list_agree():
    assert _declared() == sorted(_lib.EXPORTS)


def test_every_declared_symbol_is_exported(lib):
    for name in _declared():
        assert hasattr(lib, name), name


def test_abi_version(lib):
    assert lib.bi_abi_version() == 1


def test_argument_errors_do_not_touch_the_gpu(lib):
    h = C.c_void_p()
    assert lib.bi_create(None, C.byref(h)) == -1 and not h
    d = _lib.Desc()
    d.abi_version = 99
    assert lib.bi_create(C.byref(d), C.byref(h)) == -1
    d.abi_version = 1           # still no groups / blocks
    assert lib.bi_create(C.byref(d), C.byref(h)) == -1
    assert lib.bi_query(None, 0) == -1
    assert lib.bi_sync(None) == -1
    lib.bi_destroy(None)
    assert lib.bi_last_error(None) == b"null handle"


def test_trainer_argument_errors_do_not_touch_the_gpu(lib):
    import numpy as np
    t = C.c_void_p()
    X = np.zeros((8, 2))
    Y = np.zeros((1, 8))
    xp, yp = X.ctypes.data, Y.ctypes.data
    assert lib.bi_trainer_create(None, 8, 2, yp, 1, 1, 0, 0, 1e-10, 0, C.byref(t)) == -1 and not t
    assert lib.bi_trainer_create(xp, 8, 2, yp, 1, 7, 0, 0, 1e-10, 0, C.byref(t)) == -1          # unknown kernel kind
    assert lib.bi_trainer_create(xp, 8, 17, yp, 1, 1, 0, 0, 1e-10, 0, C.byref(t)) == -1         # d > 16
    assert lib.bi_trainer_create(xp, 992, 2, yp, 1, 1, 0, 0, 1e-10, 0, C.byref(t)) == -1        # N + 1 > 992
    assert lib.bi_trainer_create(xp, 8, 2, yp, 1, 1, 0, 0, -1.0, 0, C.byref(t)) == -1           # negative jitter
    assert lib.bi_trainer_lml_host(None, xp, None, 1, yp, yp) == -1
    assert lib.bi_trainer_launches(None) == 0
    lib.bi_trainer_destroy(None)
    assert lib.bi_trainer_last_error(None) == b"null trainer"


def test_struct_layout_matches_header():
    # 4 int32 + 12 pointers ; 6 int32 + 7 pointers + 2 int32
    assert C.sizeof(_lib.GroupDesc) == 16 + 12 * 8
    assert C.sizeof(_lib.Desc) == 24 + 7 * 8 + 8


def test_no_cpu_fallback_when_library_missing(monkeypatch, tmp_path):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", tmp_path / "nope.so")
    with pytest.raises(_lib.BackendUnavailable):
        _lib.load()


def test_return_types_follow_the_header(lib):
    """Every entry point the header declares as returning ``int`` is typed ``c_int`` in the loader (an explicit list,
    not a slice of the export list), ``int64_t`` -> ``c_int64``, ``const char*`` -> ``c_char_p``, ``void`` -> None."""
    text = re.sub(r"/\*.*?\*/", "", _lib.HEADER.read_text(), flags=re.S)
    decl = dict((name, ret.strip()) for ret, name in re.findall(r"^((?:const\s+char\s*\*|int64_t|int|void))\s+(bi_[a-z_]+)\s*\(", text, flags=re.M))
    assert sorted(decl) == sorted(_lib.EXPORTS)
    want = {"int": C.c_int, "int64_t": C.c_int64, "void": None}
    for name, ret in decl.items():
        got = getattr(lib, name).restype
        if ret.startswith("const"):
            assert got is C.c_char_p, name
        else:
            assert got is want[ret], (name, ret, got)
    assert set(_lib.INT_STATUS_EXPORTS) <= {n for n, r in decl.items() if r == "int"}
