"""bi_loglikelihood_noalloc (the allocation-free variant of the explicit log-likelihood, added at the request of the
run's operators; no counterpart in the reference) against the existing entry point bi_loglikelihood_host, through the C ABI.

Collected last on purpose (the file name sorts after every other test file): the round's end runs `pytest -x`.
Tolerance: the explicit log-likelihood tolerance of tests/test_gpu_parity.py (1e-11 relative); the two entry points launch
the same kernel, so the values are in fact expected to be bit-equal."""
from __future__ import annotations

import ctypes as C

import numpy as np
import pytest
from conftest import load_golden, rel_err

from bayesian_inference_b200 import _lib, image

pytestmark = pytest.mark.gpu

TOL = 1e-11
BI_E_UNSUPPORTED = -4


def _problem(seed, n, F, bad_row=None):
    rng = np.random.default_rng(seed)
    A = rng.normal(size=(n, F, F))
    cov = A @ np.swapaxes(A, 1, 2) + F * np.eye(F)
    if bad_row is not None:
        cov[bad_row] = np.eye(F)
        cov[bad_row, F - 2, F - 2] = -1.0
    return rng.normal(size=(n, F)), cov


def _noalloc(lib, dev, dy, cov):
    """One call of the new entry point: the caller (this test) owns inputs and outputs on the device."""
    import torch

    n, F = dy.shape
    device = torch.device("cuda", dev.device)
    d_dy = torch.from_numpy(dy).to(device)
    d_cov = torch.from_numpy(cov).to(device)
    d_out = torch.full((n,), 123.0, dtype=torch.float64, device=device)
    d_info = torch.full((n,), 77, dtype=torch.int32, device=device)
    stream = torch.cuda.current_stream(device)
    rc = lib.bi_loglikelihood_noalloc(dev._h, C.c_void_p(d_dy.data_ptr()), C.c_void_p(d_cov.data_ptr()), C.c_int64(n),
                                      C.c_int32(F), C.c_void_p(d_out.data_ptr()), C.c_void_p(d_info.data_ptr()),
                                      C.c_void_p(stream.cuda_stream))
    assert rc == 0, lib.bi_last_error(dev._h)
    stream.synchronize()
    return d_out.cpu().numpy(), d_info.cpu().numpy()


def _same(got, info, want, winfo):
    assert np.array_equal(info, winfo)
    assert np.array_equal(np.isnan(got), np.isnan(want))
    fin = np.isfinite(want)
    assert np.all(np.isfinite(got[fin])) and rel_err(got[fin], want[fin]) < TOL


def test_noalloc_matches_existing_entry_point_twice_in_a_row():
    from bayesian_inference_b200.device import DeviceEmulator

    lib = _lib.load()
    lib.bi_last_error.restype = C.c_char_p
    z, _ = load_golden("toy")
    with DeviceEmulator(image.EmulatorImage.from_flat(z)) as dev:
        # two calls in a row that share the static scratch: more rows than resident CTAs at 4 x 4 tiles (the CTAs loop and
        # reuse their scratch slice), then a different tile count and a non-positive-definite row on the stale contents
        first = _problem(1, 700, 100)
        second = _problem(2, 9, 33, bad_row=4)
        got1 = _noalloc(lib, dev, *first)
        got2 = _noalloc(lib, dev, *second)
        want1 = dev.loglikelihood(*first, return_info=True)
        want2 = dev.loglikelihood(*second, return_info=True)
        _same(*got1, *want1)
        _same(*got2, *want2)
        assert np.isnan(got2[0][4]) and got2[1][4] == 32 and np.all(np.delete(got2[1], 4) == 0)
        # the stated maximum (8 x 8 tiles), again after other contents, and the first problem once more
        third = _problem(3, 40, 256)
        _same(*_noalloc(lib, dev, *third), *dev.loglikelihood(*third, return_info=True))
        again = _noalloc(lib, dev, *first)
        assert np.array_equal(again[0], got1[0]) and np.array_equal(again[1], got1[1])
        # a second handle on the same device shares the array (calls are stream ordered here)
        with DeviceEmulator(image.EmulatorImage.from_flat(z)) as dev2:
            _same(*_noalloc(lib, dev2, *second), *want2)


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    import torch

    from bayesian_inference_b200.device import DeviceEmulator

    lib = _lib.load()
    lib.bi_last_error.restype = C.c_char_p
    z, _ = load_golden("toy")
    with DeviceEmulator(image.EmulatorImage.from_flat(z)) as dev:
        device = torch.device("cuda", dev.device)
        F = 257
        d_dy = torch.zeros((1, F), dtype=torch.float64, device=device)
        d_cov = torch.eye(F, dtype=torch.float64, device=device)[None].contiguous()
        d_out = torch.full((1,), 5.0, dtype=torch.float64, device=device)
        rc = lib.bi_loglikelihood_noalloc(dev._h, C.c_void_p(d_dy.data_ptr()), C.c_void_p(d_cov.data_ptr()), C.c_int64(1),
                                          C.c_int32(F), C.c_void_p(d_out.data_ptr()), None, None)
        assert rc == BI_E_UNSUPPORTED
        msg = lib.bi_last_error(dev._h).decode()
        assert "BI_NOALLOC_MAX_F" in msg and "256" in msg
        torch.cuda.synchronize(device)
        assert float(d_out[0]) == 5.0            # nothing was launched
        # n = 0 is a no-op, NULL pointers are refused
        assert lib.bi_loglikelihood_noalloc(dev._h, C.c_void_p(d_dy.data_ptr()), C.c_void_p(d_cov.data_ptr()), C.c_int64(0),
                                            C.c_int32(8), C.c_void_p(d_out.data_ptr()), None, None) == 0
        assert lib.bi_loglikelihood_noalloc(dev._h, None, None, C.c_int64(1), C.c_int32(8), None, None, None) == -1
