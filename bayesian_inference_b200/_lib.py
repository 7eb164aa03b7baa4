"""Loader for the C-ABI library ``libbi_b200.so`` (include/bi_b200.h) and its in-tree nvcc build.

The library is the product: there is no Python / numpy / torch fallback.  If it is missing or cannot be loaded every
entry point of this package raises ``BackendUnavailable``.
"""
from __future__ import annotations

import ctypes as C
import os
import shutil
import subprocess
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
CSRC = PKG_DIR / "csrc"
LIB_PATH = Path(os.environ.get("BI_B200_LIB") or PKG_DIR / "libbi_b200.so")   # override: developer A/B builds only
HEADER = PKG_DIR.parent / "include" / "bi_b200.h"

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]

EXPORTS = [
    "bi_abi_version", "bi_create", "bi_destroy", "bi_last_error", "bi_query", "bi_sync", "bi_profile", "bi_profile_read",
    "bi_kernel_vectors", "bi_kernel_vectors_host", "bi_predict_pc", "bi_predict_pc_host",
    "bi_predict_group", "bi_predict_group_host", "bi_predict", "bi_predict_host",
    "bi_log_posterior", "bi_log_posterior_host", "bi_loglikelihood_host",
    "bi_stretch_propose", "bi_stretch_accept", "bi_set_data_sets", "bi_log_posterior_sets", "bi_stretch_randoms_philox",
    "bi_trainer_create", "bi_trainer_lml", "bi_trainer_lml_host", "bi_trainer_factor", "bi_trainer_factor_host",
    "bi_trainer_destroy", "bi_trainer_last_error", "bi_trainer_launches",
    "bi_loglikelihood_noalloc", "bi_stretch_accept_record",
]

# entry points that return an int status (0 ok, <0 bad argument, >0 CUDA error code); named one by one so that adding
# an export cannot silently shift a return type
INT_STATUS_EXPORTS = [
    "bi_kernel_vectors", "bi_kernel_vectors_host", "bi_predict_pc", "bi_predict_pc_host",
    "bi_predict_group", "bi_predict_group_host", "bi_predict", "bi_predict_host",
    "bi_log_posterior", "bi_log_posterior_host", "bi_loglikelihood_host",
    "bi_stretch_propose", "bi_stretch_accept", "bi_set_data_sets", "bi_log_posterior_sets", "bi_stretch_randoms_philox",
    "bi_trainer_create", "bi_trainer_lml", "bi_trainer_lml_host", "bi_trainer_factor", "bi_trainer_factor_host",
    "bi_loglikelihood_noalloc", "bi_stretch_accept_record",
]

MODE = {"dense": 0, "structured": 1, "woodbury": 2}


class BackendUnavailable(RuntimeError):
    pass


class GroupDesc(C.Structure):
    _fields_ = [
        ("n_design", C.c_int32), ("n_pc", C.c_int32), ("n_features", C.c_int32), ("n_rest", C.c_int32),
        ("X", C.c_void_p), ("kind", C.c_void_p), ("ell", C.c_void_p), ("constant", C.c_void_p),
        ("noise", C.c_void_p), ("alpha", C.c_void_p), ("L", C.c_void_p), ("S", C.c_void_p),
        ("S_rest", C.c_void_p), ("ev_rest", C.c_void_p), ("scale", C.c_void_p), ("mean", C.c_void_p),
    ]


class Desc(C.Structure):
    _fields_ = [
        ("abi_version", C.c_int32), ("device", C.c_int32), ("d", C.c_int32), ("n_groups", C.c_int32),
        ("n_blocks", C.c_int32), ("n_features", C.c_int32),
        ("groups", C.POINTER(GroupDesc)), ("blocks", C.c_void_p),
        ("prior_min", C.c_void_p), ("prior_max", C.c_void_p),
        ("y", C.c_void_p), ("y_err", C.c_void_p), ("sigma_exp", C.c_void_p),
        ("max_batch", C.c_int32), ("reserved", C.c_int32),
    ]


def sources() -> list[Path]:
    return sorted(CSRC.glob("*.cu")) + sorted(CSRC.glob("*.cuh")) + [HEADER]


def needs_build() -> bool:
    if not LIB_PATH.exists():
        return True
    t = LIB_PATH.stat().st_mtime
    return any(s.stat().st_mtime > t for s in sources())


def build(force: bool = False, verbose: bool = False) -> Path:
    """nvcc -gencode arch=compute_100a,code=sm_100a ... -shared -o libbi_b200.so csrc/bi_api.cu (cross-compiles on CPU)."""
    if not force and not needs_build():
        return LIB_PATH
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    cmd = [nvcc, *NVCC_FLAGS, "-o", str(LIB_PATH), str(CSRC / "bi_api.cu")]
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise BackendUnavailable(f"nvcc failed:\n{' '.join(cmd)}\n{res.stdout}\n{res.stderr}")
    if verbose:
        print(res.stderr)
    return LIB_PATH


_lib = None


def load() -> C.CDLL:
    """dlopen the library and declare the prototypes of include/bi_b200.h.  Raises BackendUnavailable loudly."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise BackendUnavailable(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). There is no CPU fallback.")
    try:
        lib = C.CDLL(str(LIB_PATH))
    except OSError as exc:
        raise BackendUnavailable(f"cannot load {LIB_PATH}: {exc}") from exc
    h = C.c_void_p
    dp = C.c_void_p
    i64 = C.c_int64
    lib.bi_abi_version.restype = C.c_int
    lib.bi_create.argtypes = [C.POINTER(Desc), C.POINTER(h)]
    lib.bi_create.restype = C.c_int
    lib.bi_destroy.argtypes = [h]
    lib.bi_destroy.restype = None
    lib.bi_last_error.argtypes = [h]
    lib.bi_last_error.restype = C.c_char_p
    lib.bi_query.argtypes = [h, C.c_int]
    lib.bi_query.restype = i64
    lib.bi_sync.argtypes = [h]
    lib.bi_sync.restype = C.c_int
    lib.bi_profile.argtypes = [h, C.c_int]
    lib.bi_profile.restype = C.c_int
    lib.bi_profile_read.argtypes = [h, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int64)]
    lib.bi_profile_read.restype = C.c_int
    lib.bi_kernel_vectors.argtypes = [h, C.c_int, dp, i64, dp, C.c_void_p]
    lib.bi_kernel_vectors_host.argtypes = [h, C.c_int, dp, i64, dp]
    lib.bi_predict_pc.argtypes = [h, dp, i64, dp, dp, C.c_void_p]
    lib.bi_predict_pc_host.argtypes = [h, dp, i64, dp, dp]
    lib.bi_predict_group.argtypes = [h, C.c_int, dp, i64, C.c_double, dp, dp, C.c_void_p]
    lib.bi_predict_group_host.argtypes = [h, C.c_int, dp, i64, C.c_double, dp, dp]
    lib.bi_predict.argtypes = [h, dp, i64, C.c_double, dp, dp, C.c_void_p]
    lib.bi_predict_host.argtypes = [h, dp, i64, C.c_double, dp, dp]
    lib.bi_log_posterior.argtypes = [h, dp, i64, C.c_double, C.c_int, dp, dp, C.c_void_p]
    lib.bi_log_posterior_host.argtypes = [h, dp, i64, C.c_double, C.c_int, dp, dp]
    lib.bi_loglikelihood_host.argtypes = [h, dp, dp, i64, C.c_int32, dp, dp]
    lib.bi_stretch_propose.argtypes = [h, dp, dp, dp, dp, i64, dp, C.c_void_p]
    lib.bi_stretch_accept.argtypes = [h, dp, dp, dp, dp, dp, dp, dp, i64, dp, dp, C.c_void_p]
    lib.bi_stretch_accept_record.argtypes = [h, dp, dp, dp, dp, dp, dp, dp, i64, dp, dp, dp, dp, dp, C.c_void_p]
    lib.bi_set_data_sets.argtypes = [h, C.c_int32, dp, dp]
    lib.bi_log_posterior_sets.argtypes = [h, dp, dp, i64, C.c_double, C.c_int, dp, dp, C.c_void_p]
    i32 = C.c_int32
    lib.bi_trainer_create.argtypes = [dp, i32, i32, dp, i32, i32, i32, i32, C.c_double, i32, C.POINTER(C.c_void_p)]
    lib.bi_trainer_lml.argtypes = [h, dp, dp, i64, dp, dp, C.c_void_p]
    lib.bi_trainer_lml_host.argtypes = [h, dp, dp, i64, dp, dp]
    lib.bi_trainer_factor.argtypes = [h, dp, dp, i64, dp, dp, dp, C.c_void_p]
    lib.bi_trainer_factor_host.argtypes = [h, dp, dp, i64, dp, dp, dp]
    lib.bi_trainer_destroy.argtypes = [h]
    lib.bi_trainer_destroy.restype = None
    lib.bi_trainer_last_error.argtypes = [h]
    lib.bi_trainer_last_error.restype = C.c_char_p
    lib.bi_trainer_launches.argtypes = [h]
    lib.bi_trainer_launches.restype = C.c_int64
    lib.bi_stretch_randoms_philox.argtypes = [h, C.c_uint64, C.c_uint64, i32, i32, C.c_double, dp, dp, dp, dp, dp, C.c_void_p]
    for name in INT_STATUS_EXPORTS:
        getattr(lib, name).restype = C.c_int
    if lib.bi_abi_version() != 1:
        raise BackendUnavailable(f"ABI version mismatch: library {lib.bi_abi_version()}, package 1")
    _lib = lib
    return lib
