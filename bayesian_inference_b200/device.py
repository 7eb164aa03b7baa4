"""``DeviceEmulator``: the emulator image resident in one GPU's HBM, driven through the C ABI (include/bi_b200.h).

This is the host-side owner of a ``bi_handle``.  It is what ``log_posterior.initialize_pool_variables`` builds once
(the reference pickles the emulators to every pool worker instead, mcmc.py:77-78) and what ``emulation.predict`` /
``log_posterior.log_posterior`` call.  numpy arrays go through the ``*_host`` entry points (copy in, run, copy out);
torch CUDA tensors go through the device-pointer entry points on torch's current stream and never leave the GPU.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np

from . import _lib
from .image import EmulatorImage


def _f64(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.float64)


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class DeviceError(RuntimeError):
    pass


class DeviceEmulator:
    def __init__(self, image: EmulatorImage, device: int = 0, max_batch: int = 0):
        self._lib = _lib.load()
        self._h = C.c_void_p()
        self.image = image
        self.device = int(device)
        self.d = image.d
        self.n_features = image.n_features
        self.group_names = image.group_names
        self.n_pc = [g.n_pc for g in image.groups]
        self.n_pc_total = sum(self.n_pc)
        keep = []   # host arrays must outlive bi_create

        def hold(a, dtype=np.float64):
            a = np.ascontiguousarray(a, dtype=dtype)
            keep.append(a)
            return a.ctypes.data_as(C.c_void_p)

        garr = (_lib.GroupDesc * len(image.groups))()
        for i, g in enumerate(image.groups):
            gd = garr[i]
            gd.n_design, gd.n_pc, gd.n_features, gd.n_rest = g.n_design, g.n_pc, g.n_features, int(g.S_rest.shape[0])
            gd.X, gd.kind, gd.ell = hold(g.X), hold(g.kind, np.int32), hold(g.ell)
            gd.constant, gd.noise, gd.alpha, gd.L = hold(g.const), hold(g.noise), hold(g.alpha), hold(g.L)
            gd.S, gd.scale, gd.mean = hold(g.S), hold(g.scale), hold(g.mean)
            if g.S_rest.shape[0] > 0:
                gd.S_rest, gd.ev_rest = hold(g.S_rest), hold(g.ev_rest)
        desc = _lib.Desc()
        desc.abi_version = 1
        desc.device = self.device
        desc.d = image.d
        desc.n_groups = len(image.groups)
        desc.n_blocks = int(image.blocks.shape[0])
        desc.n_features = image.n_features
        desc.groups = garr
        desc.blocks = hold(image.blocks, np.int64)
        desc.prior_min, desc.prior_max = hold(image.min), hold(image.max)
        self.has_data = image.y is not None and (image.y_err is not None or image.sigma_exp is not None)
        if image.y is not None:
            desc.y = hold(image.y)
        if image.sigma_exp is not None:
            desc.sigma_exp = hold(image.sigma_exp)
        elif image.y_err is not None:
            desc.y_err = hold(image.y_err)
        desc.max_batch = int(max_batch)
        rc = self._lib.bi_create(C.byref(desc), C.byref(self._h))
        if rc != 0 or not self._h:
            self._h = C.c_void_p()
            raise DeviceError(f"bi_create failed with code {rc} (see stderr); there is no CPU fallback")

    # ---- lifetime -------------------------------------------------------------------------------------------
    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.bi_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _check(self, rc: int, what: str) -> None:
        if rc != 0:
            msg = self._lib.bi_last_error(self._h)
            raise DeviceError(f"{what} failed ({rc}): {msg.decode() if msg else ''}")

    def query(self, what: int) -> int:
        return int(self._lib.bi_query(self._h, what))

    @property
    def launches(self) -> int:
        """Kernels launched by this handle since creation."""
        return self.query(7)

    def profile(self, enable: bool = True) -> None:
        self._check(self._lib.bi_profile(self._h, int(enable)), "bi_profile")

    def profile_read(self, kernel: int) -> tuple[float, int]:
        """(summed device ms, launches) of kernel 0 = gp_predict, 1 = chol_loglik since the last read."""
        ms, n = C.c_double(), C.c_int64()
        self._check(self._lib.bi_profile_read(self._h, kernel, C.byref(ms), C.byref(n)), "bi_profile_read")
        return ms.value, n.value

    def sync(self) -> None:
        self._check(self._lib.bi_sync(self._h), "bi_sync")

    def _theta(self, theta) -> np.ndarray:
        th = np.asarray(theta, dtype=np.float64)
        if th.ndim == 1:
            th = th[None, :]
        if th.ndim != 2 or th.shape[1] != self.d:
            raise ValueError(f"theta has shape {th.shape}, expected (W, {self.d})")
        return np.ascontiguousarray(th)

    # ---- numpy (host) API -----------------------------------------------------------------------------------
    def kernel_vectors(self, group: int, theta) -> np.ndarray:
        th = self._theta(theta)
        g = self.image.groups[group]
        out = np.empty((th.shape[0], g.n_pc, g.n_design))
        self._check(self._lib.bi_kernel_vectors_host(self._h, group, _ptr(th), th.shape[0], _ptr(out)), "bi_kernel_vectors")
        return out

    def predict_pc(self, theta) -> tuple[np.ndarray, np.ndarray]:
        th = self._theta(theta)
        m = np.empty((th.shape[0], self.n_pc_total))
        v = np.empty_like(m)
        self._check(self._lib.bi_predict_pc_host(self._h, _ptr(th), th.shape[0], _ptr(m), _ptr(v)), "bi_predict_pc")
        return m, v

    def predict_group(self, group: int, theta, n_div: Optional[float] = None, want_cov: bool = True):
        th = self._theta(theta)
        W, Fg = th.shape[0], self.image.groups[group].n_features
        nd = float(W if n_div is None else n_div)
        mean = np.empty((W, Fg))
        cov = np.empty((W, Fg, Fg)) if want_cov else None
        if W:
            self._check(self._lib.bi_predict_group_host(self._h, group, _ptr(th), W, nd, _ptr(mean), _ptr(cov)),
                        "bi_predict_group")
        return mean, cov

    def predict(self, theta, n_div: Optional[float] = None, want_cov: bool = True):
        th = self._theta(theta)
        W, F = th.shape[0], self.n_features
        nd = float(W if n_div is None else n_div)
        mean = np.empty((W, F))
        cov = np.empty((W, F, F)) if want_cov else None
        if W:
            self._check(self._lib.bi_predict_host(self._h, _ptr(th), W, nd, _ptr(mean), _ptr(cov)), "bi_predict")
        return mean, cov

    def log_posterior(self, theta, n_div: Optional[float] = 1.0, mode: str = "dense", return_info: bool = False):
        """``n_div=None`` -> rows inside the prior box in this call (a batched reference call)."""
        th = self._theta(theta)
        W = th.shape[0]
        out = np.empty(W)
        info = np.zeros(W, dtype=np.int32)
        if W:
            nd = -1.0 if n_div is None else float(n_div)
            self._check(self._lib.bi_log_posterior_host(self._h, _ptr(th), W, nd, _lib.MODE[mode], _ptr(out), _ptr(info)),
                        "bi_log_posterior")
        return (out, info) if return_info else out

    def loglikelihood(self, dy, cov, return_info: bool = False):
        dy = _f64(dy)
        cov = _f64(cov)
        if dy.ndim == 1:
            dy, cov = dy[None], cov[None]
        n, Fd = dy.shape
        if cov.shape != (n, Fd, Fd):
            raise ValueError(f"cov has shape {cov.shape}, expected {(n, Fd, Fd)}")
        out = np.empty(n)
        info = np.zeros(n, dtype=np.int32)
        self._check(self._lib.bi_loglikelihood_host(self._h, _ptr(dy), _ptr(cov), n, Fd, _ptr(out), _ptr(info)),
                    "bi_loglikelihood")
        return (out, info) if return_info else out

    # ---- torch (device-resident) API --------------------------------------------------------------------------
    def log_posterior_tensor(self, theta, n_div: float = 1.0, mode: str = "dense", out=None, info=None):
        """theta: torch.float64 CUDA tensor (W, d) on this device.  Enqueues on torch's current stream, no sync."""
        import torch

        if not (theta.is_cuda and theta.dtype == torch.float64 and theta.is_contiguous() and theta.dim() == 2
                and theta.shape[1] == self.d and theta.device.index == self.device):
            raise ValueError("theta must be a contiguous float64 CUDA tensor (W, d) on the handle's device")
        W = theta.shape[0]
        if out is None:
            out = torch.empty(W, dtype=torch.float64, device=theta.device)
        stream = torch.cuda.current_stream(theta.device).cuda_stream
        self._check(self._lib.bi_log_posterior(self._h, C.c_void_p(theta.data_ptr()), W, float(n_div), _lib.MODE[mode],
                                               C.c_void_p(out.data_ptr()),
                                               C.c_void_p(info.data_ptr()) if info is not None else None,
                                               C.c_void_p(stream)), "bi_log_posterior")
        return out

    def predict_pc_tensor(self, theta):
        import torch

        W = theta.shape[0]
        m = torch.empty((W, self.n_pc_total), dtype=torch.float64, device=theta.device)
        v = torch.empty_like(m)
        stream = torch.cuda.current_stream(theta.device).cuda_stream
        self._check(self._lib.bi_predict_pc(self._h, C.c_void_p(theta.data_ptr()), W, C.c_void_p(m.data_ptr()),
                                            C.c_void_p(v.data_ptr()), C.c_void_p(stream)), "bi_predict_pc")
        return m, v


    def kernel_vectors_tensor(self, group: int, theta):
        """k (W, P_g, N_g) as a CUDA tensor (sklearn kernel_(X, X_train_) per PC); enqueues on torch's current stream."""
        import torch

        g = self.image.groups[group]
        W = theta.shape[0]
        out = torch.empty((W, g.n_pc, g.n_design), dtype=torch.float64, device=theta.device)
        stream = torch.cuda.current_stream(theta.device).cuda_stream
        self._check(self._lib.bi_kernel_vectors(self._h, group, C.c_void_p(theta.data_ptr()), W, C.c_void_p(out.data_ptr()),
                                                C.c_void_p(stream)), "bi_kernel_vectors")
        return out

    def predict_tensor(self, theta, n_div: Optional[float] = None, want_cov: bool = True, mean=None, cov=None):
        """Merged prediction on the device: mean (W, F), cov (W, F, F) CUDA tensors (emulation.predict without the copy back)."""
        import torch

        W, F = theta.shape[0], self.n_features
        nd = float(W if n_div is None else n_div)
        if mean is None:
            mean = torch.empty((W, F), dtype=torch.float64, device=theta.device)
        if want_cov and cov is None:
            cov = torch.empty((W, F, F), dtype=torch.float64, device=theta.device)
        stream = torch.cuda.current_stream(theta.device).cuda_stream
        self._check(self._lib.bi_predict(self._h, C.c_void_p(theta.data_ptr()), W, nd, C.c_void_p(mean.data_ptr()),
                                         C.c_void_p(cov.data_ptr()) if want_cov else None, C.c_void_p(stream)), "bi_predict")
        return mean, cov

    # ---- several experimental data sets on one image (closure tests) -----------------------------------
    def set_data_sets(self, y, y_err) -> int:
        """Load n pseudo-data sets (n, F) side by side; Sigma_exp = diag(y_err**2) per set.  Returns n."""
        y, y_err = _f64(np.atleast_2d(y)), _f64(np.atleast_2d(y_err))
        if y.shape != y_err.shape or y.shape[1] != self.n_features:
            raise ValueError(f"y / y_err must both be (n_sets, {self.n_features})")
        self._check(self._lib.bi_set_data_sets(self._h, y.shape[0], _ptr(y), _ptr(y_err)), "bi_set_data_sets")
        self.n_sets = y.shape[0]
        return self.n_sets

    def log_posterior_sets_tensor(self, theta, set_id, n_div: float = 1.0, mode: str = "dense", out=None):
        """Row w of theta is evaluated against data set set_id[w] (int32 CUDA tensor); enqueues on torch's current stream."""
        import torch

        W = theta.shape[0]
        if not (theta.is_cuda and theta.dtype == torch.float64 and theta.is_contiguous() and set_id.is_cuda
                and set_id.dtype == torch.int32 and set_id.is_contiguous() and set_id.shape[0] == W):
            raise ValueError("theta: contiguous float64 CUDA (W, d); set_id: contiguous int32 CUDA (W)")
        if out is None:
            out = torch.empty(W, dtype=torch.float64, device=theta.device)
        stream = torch.cuda.current_stream(theta.device).cuda_stream
        self._check(self._lib.bi_log_posterior_sets(self._h, C.c_void_p(theta.data_ptr()), C.c_void_p(set_id.data_ptr()), W,
                                                    float(n_div), _lib.MODE[mode], C.c_void_p(out.data_ptr()), None,
                                                    C.c_void_p(stream)), "bi_log_posterior_sets")
        return out

    def log_posterior_sets(self, theta, set_id, n_div: float = 1.0, mode: str = "dense") -> np.ndarray:
        import torch

        th = torch.from_numpy(self._theta(theta)).cuda(self.device)
        sid = torch.from_numpy(np.ascontiguousarray(set_id, dtype=np.int32)).cuda(self.device)
        if sid.numel() and (int(sid.min()) < 0 or int(sid.max()) >= getattr(self, "n_sets", 0)):
            raise ValueError("set_id out of range")
        return self.log_posterior_sets_tensor(th, sid, n_div=n_div, mode=mode).cpu().numpy()

    # ---- stretch-move half-step on device-resident state ------------------------------------------
    def stretch_propose(self, coords, s_idx, c_idx, zz, q_out) -> None:
        """q_out[k] = c - (c - s) * zz[k]; all arguments CUDA tensors (coords f64 (W,d), indices int32 (ns), zz f64 (ns))."""
        import torch

        stream = torch.cuda.current_stream(coords.device).cuda_stream
        self._check(self._lib.bi_stretch_propose(self._h, C.c_void_p(coords.data_ptr()), C.c_void_p(s_idx.data_ptr()),
                                                 C.c_void_p(c_idx.data_ptr()), C.c_void_p(zz.data_ptr()), int(s_idx.shape[0]),
                                                 C.c_void_p(q_out.data_ptr()), C.c_void_p(stream)), "bi_stretch_propose")

    def stretch_accept(self, coords, logp, s_idx, q, new_lp, factors, logu, n_accepted, nan_flag, record=None) -> None:
        """Accept/reject one half-step in place (coords, logp, n_accepted int64 (W), nan_flag int32 (1)).
        record = (other_idx int32 (ns), chain_out f64 (W, d), lp_out f64 (W)): also store the ensemble after the step
        (second half-step of a step; other_idx = the walkers of the first half): bi_stretch_accept_record."""
        import torch

        stream = torch.cuda.current_stream(coords.device).cuda_stream
        if record is not None:
            other_idx, chain_out, lp_out = record
            self._check(self._lib.bi_stretch_accept_record(
                self._h, C.c_void_p(coords.data_ptr()), C.c_void_p(logp.data_ptr()), C.c_void_p(s_idx.data_ptr()),
                C.c_void_p(q.data_ptr()), C.c_void_p(new_lp.data_ptr()), C.c_void_p(factors.data_ptr()),
                C.c_void_p(logu.data_ptr()), int(s_idx.shape[0]), C.c_void_p(n_accepted.data_ptr()),
                C.c_void_p(nan_flag.data_ptr()), C.c_void_p(other_idx.data_ptr()), C.c_void_p(chain_out.data_ptr()),
                C.c_void_p(lp_out.data_ptr()), C.c_void_p(stream)), "bi_stretch_accept_record")
            return
        self._check(self._lib.bi_stretch_accept(self._h, C.c_void_p(coords.data_ptr()), C.c_void_p(logp.data_ptr()),
                                                C.c_void_p(s_idx.data_ptr()), C.c_void_p(q.data_ptr()),
                                                C.c_void_p(new_lp.data_ptr()), C.c_void_p(factors.data_ptr()),
                                                C.c_void_p(logu.data_ptr()), int(s_idx.shape[0]),
                                                C.c_void_p(n_accepted.data_ptr()), C.c_void_p(nan_flag.data_ptr()),
                                                C.c_void_p(stream)), "bi_stretch_accept")

    def stretch_randoms_philox(self, seed: int, step0: int, nsteps: int, n_half: int, a: float, out: dict) -> None:
        """Fill out['s_idx'|'c_idx'|'zz'|'factors'|'logu'] (CUDA tensors (>= nsteps, 2, n_half)) for steps step0 ...
        step0 + nsteps - 1 with counter-based random numbers generated on the device (bi_stretch_randoms_philox)."""
        import torch

        stream = torch.cuda.current_stream(out["zz"].device).cuda_stream
        self._check(self._lib.bi_stretch_randoms_philox(
            self._h, C.c_uint64(seed & 0xFFFFFFFFFFFFFFFF), C.c_uint64(step0), int(nsteps), int(n_half), float(a),
            C.c_void_p(out["s_idx"].data_ptr()), C.c_void_p(out["c_idx"].data_ptr()), C.c_void_p(out["zz"].data_ptr()),
            C.c_void_p(out["factors"].data_ptr()), C.c_void_p(out["logu"].data_ptr()), C.c_void_p(stream)),
            "bi_stretch_randoms_philox")


def from_analysis(emulation_config, emulation_results, *, min, max, experimental_results=None, sigma_exp=None,
                  device: int = 0, max_batch: int = 0) -> DeviceEmulator:
    """Pack what ``mcmc.run_mcmc`` loads (mcmc.py:48-67) and upload it."""
    from . import image as _image

    img = _image.pack(emulation_config, emulation_results, min=min, max=max, experimental_results=experimental_results,
                      sigma_exp=sigma_exp)
    return DeviceEmulator(img, device=device, max_batch=max_batch)
