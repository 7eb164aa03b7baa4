"""Affine-invariant ensemble sampler (Goodman & Weare stretch move) -- the caller of the hot path.

The reference drives ``log_posterior`` with ``emcee.EnsembleSampler`` (mcmc.py:83-107, emcee 3.1.4, pdm.lock:504).
emcee is a third-party dependency that is not vendored in the reference and not installed in this image, so the
sampler is restated here from the published algorithm and emcee 3.1.4's documented behaviour:

* one ``numpy.random.RandomState`` drives everything, consumed per step in emcee's order: move selection (one uniform
  draw), the red/blue split ``arange(W) % 2`` shuffled, then for each half ``rand(n)`` (stretch factors),
  ``randint(n_other, size=n)`` (partners), the log-probability of the proposals, and ``rand()`` per walker (accept);
* stretch factor ``zz = ((a-1) u + 1)^2 / a`` with ``a = 2``; proposal ``c - (c - s) zz``; accept when
  ``(ndim-1) log zz + logp_new - logp_old > log u``;
* NaN log-probability raises ``ValueError('Probability function returned NaN')``; ``-inf`` is allowed.

What changes with respect to the reference's use of emcee: the log-probability callback is *vectorised* -- called once
per half-step with all W/2 proposals (one C-ABI call, optionally sharded over GPUs by ``ShardedLogProb``) instead of
once per walker through a ``multiprocessing`` pool.  With ``log_posterior.log_posterior_walkers`` as the callback every
row is evaluated exactly as the reference evaluates a single walker, so the chain is the one emcee would produce from
the same RandomState.

Parity status: **unpinned against emcee itself** (not installable here); pinned against an independent scalar
re-implementation in tests/test_ensemble.py, and the GPU-vs-oracle chain identity is tested under this same driver.
"""
from __future__ import annotations

import os

from typing import Callable, Optional

import numpy as np


def walkers_independent(coords: np.ndarray) -> bool:
    """emcee 3.1.4's initial-state test (``emcee.ensemble.walkers_independent``, the version the reference pins,
    pdm.lock:504-505): centre the ensemble, scale every column by its largest magnitude and then to unit norm, and ask for
    a condition number <= 1e8 of that matrix.  (emcee 3.0.x tested ``cond(cov(coords)) > 1e8`` instead, which rejects
    ensembles whose parameters merely have very different ranges, e.g. jet_substructure.yaml:127-137.)"""
    coords = np.asarray(coords, dtype=np.float64)
    if not np.all(np.isfinite(coords)):
        return False
    C = coords - np.mean(coords, axis=0)[None, :]
    C_colmax = np.amax(np.abs(C), axis=0)
    if np.any(C_colmax == 0):
        return False
    C = C / C_colmax
    C_colsum = np.sqrt(np.sum(C ** 2, axis=0))
    C = C / C_colsum
    return bool(np.linalg.cond(C.astype(float)) <= 1e8)


def check_initial_state(coords: np.ndarray) -> None:
    if not walkers_independent(coords):
        raise ValueError("Initial state has a large condition number. "
                         "Make sure that your walkers are linearly independent for the best performance")


class StretchEnsemble:
    """Minimal stand-in for ``emcee.EnsembleSampler`` with the default ``StretchMove(a=2)``, vectorised callback."""

    def __init__(self, nwalkers: int, ndim: int, log_prob_fn: Callable[[np.ndarray], np.ndarray], a: float = 2.0,
                 seed: Optional[int] = None, random_state: Optional[np.random.RandomState] = None):
        if nwalkers % 2:
            raise ValueError("the red/blue stretch move needs an even number of walkers")
        if nwalkers < 2 * ndim:
            raise ValueError("It is unadvisable to use a red-blue move with fewer walkers than twice the number of dimensions.")
        self.nwalkers, self.ndim, self.a = int(nwalkers), int(ndim), float(a)
        self.log_prob_fn = log_prob_fn
        self._random = random_state if random_state is not None else np.random.RandomState(seed)
        self.n_log_prob_calls = 0
        self.n_log_prob_rows = 0
        self.reset()

    # The reference pickles the sampler next to mcmc.h5 (mcmc.py:127-132) and plot code calls ``get_chain`` etc. on the
    # unpickled object: everything that makes up the stored chain pickles, what ties the object to a process (the callback,
    # the device handle, communicator and device buffers) is dropped.
    _NOT_PICKLED = ("log_prob_fn", "dev", "group", "_dist", "_buf", "_send", "_recv")

    def __getstate__(self):
        return {k: (None if k in self._NOT_PICKLED else v) for k, v in self.__dict__.items()}

    # ---- storage (emcee backend subset used by mcmc.run_mcmc, mcmc.py:99-116) ---------------------------------
    def reset(self) -> None:
        self._chain: list[np.ndarray] = []
        self._log_prob: list[np.ndarray] = []
        self._accepted = np.zeros(self.nwalkers)
        self.iteration = 0

    def get_chain(self, flat: bool = False, discard: int = 0, thin: int = 1) -> np.ndarray:
        c = np.asarray(self._chain).reshape(-1, self.nwalkers, self.ndim)[discard::thin]
        return c.reshape(-1, self.ndim) if flat else c

    def get_log_prob(self, flat: bool = False, discard: int = 0, thin: int = 1) -> np.ndarray:
        lp = np.asarray(self._log_prob).reshape(-1, self.nwalkers)[discard::thin]
        return lp.reshape(-1) if flat else lp

    @property
    def flatchain(self) -> np.ndarray:
        return self.get_chain(flat=True)

    @property
    def flatlnprobability(self) -> np.ndarray:
        return self.get_log_prob(flat=True)

    @property
    def acceptance_fraction(self) -> np.ndarray:
        return self._accepted / max(self.iteration, 1)

    def get_autocorr_time(self, c: float = 5.0, tol: float = 50.0) -> np.ndarray:
        """Integrated autocorrelation time per parameter (Sokal window, as emcee.autocorr.integrated_time)."""
        x = self.get_chain()
        n_t, n_w, n_d = x.shape
        tau = np.empty(n_d)
        for d in range(n_d):
            f = np.zeros(n_t)
            for k in range(n_w):
                y = x[:, k, d] - np.mean(x[:, k, d])
                n = 1 << int(np.ceil(np.log2(max(2 * n_t, 2))))
                ft = np.fft.fft(y, n=n)
                acf = np.fft.ifft(ft * np.conjugate(ft))[:n_t].real
                f += acf / acf[0] if acf[0] != 0 else 0.0
            f /= n_w
            taus = 2.0 * np.cumsum(f) - 1.0
            m = np.arange(len(taus)) < c * taus
            window = int(np.argmin(m)) if np.any(~m) else len(taus) - 1
            tau[d] = taus[window]
        if np.any(tol * tau > n_t):
            raise RuntimeError(f"The chain is shorter than {tol} times the integrated autocorrelation time "
                               f"for {int(np.sum(tol * tau > n_t))} parameter(s). Use this estimate with caution: tau = {tau}")
        return tau

    # ---- sampling -----------------------------------------------------------------------------------------------
    def compute_log_prob(self, coords: np.ndarray) -> np.ndarray:
        lp = np.asarray(self.log_prob_fn(np.ascontiguousarray(coords)), dtype=np.float64).reshape(-1)
        self.n_log_prob_calls += 1
        self.n_log_prob_rows += coords.shape[0]
        if lp.shape[0] != coords.shape[0]:
            raise ValueError("incompatible input dimensions")
        if np.any(np.isnan(lp)):
            raise ValueError("Probability function returned NaN")
        return lp

    def sample(self, initial_state: np.ndarray, iterations: int = 1, log_prob0: Optional[np.ndarray] = None):
        coords = np.array(initial_state, dtype=np.float64, copy=True)
        if coords.shape != (self.nwalkers, self.ndim):
            raise ValueError("incompatible input dimensions")
        check_initial_state(coords)
        log_prob = self.compute_log_prob(coords) if log_prob0 is None else np.array(log_prob0, dtype=np.float64, copy=True)
        rng = self._random
        W, nd = self.nwalkers, self.ndim
        all_inds = np.arange(W)
        for _ in range(iterations):
            rng.random_sample()                  # emcee: move = random.choice(moves, p=weights) -> one uniform draw
            inds = all_inds % 2
            rng.shuffle(inds)
            for split in range(2):
                S1 = inds == split
                s, c = coords[S1], coords[~S1]
                ns, nc = s.shape[0], c.shape[0]
                zz = ((self.a - 1.0) * rng.rand(ns) + 1.0) ** 2.0 / self.a
                factors = (nd - 1.0) * np.log(zz)
                rint = rng.randint(nc, size=(ns,))
                q = c[rint] - (c[rint] - s) * zz[:, None]
                new_lp = self.compute_log_prob(q)
                lnpdiff = factors + new_lp - log_prob[S1]
                accept = lnpdiff > np.log(rng.rand(ns))
                idx = all_inds[S1][accept]
                coords[idx] = q[accept]
                log_prob[idx] = new_lp[accept]
                self._accepted[idx] += 1
            self.iteration += 1
            self._chain.append(coords.copy())
            self._log_prob.append(log_prob.copy())
            yield coords, log_prob

    def run_mcmc(self, initial_state, nsteps: int, n_logging_steps: int = 100, logger=None, **kwargs):
        """As LoggingEnsembleSampler.run_mcmc (mcmc.py:194-204): returns the last (coords, log_prob)."""
        result = None
        for n, result in enumerate(self.sample(initial_state, iterations=nsteps, **kwargs), start=1):
            if logger is not None and (n % n_logging_steps == 0 or n == nsteps):
                af = self.acceptance_fraction
                logger.info(f"  step {n}: acceptance fraction: mean {af.mean()}, std {af.std()}, min {af.min()}, max {af.max()}")
        return result


def draw_stretch_randoms(rng: np.random.RandomState, nwalkers: int, ndim: int, nsteps: int, a: float = 2.0) -> dict:
    """Everything ``nsteps`` stretch-move steps draw from ``rng``, in emcee's consumption order.

    None of it depends on the chain state: per step one uniform (move selection), the shuffled red/blue split, and per
    half ``rand(ns)`` (stretch factors), ``randint(nc, size=ns)`` (partners) and ``rand(ns)`` (acceptance).  Returned
    arrays have shape (nsteps, 2, W/2): ``s_idx`` the walkers moved, ``c_idx`` their partners (walker ids), ``zz``,
    ``factors = (ndim-1) log zz`` and ``logu = log(u)`` -- the logs are taken here, on the host, so that the device
    accept test compares exactly the numbers the host sampler compares.
    """
    W, ns = nwalkers, nwalkers // 2
    # the loop only draws (the order of the draws is what must be preserved); everything else is vectorised after it.
    # At 8192 walkers this host work is what bounds an ensemble sharded over 8 GPUs (0.9 ms per half-step before it was
    # trimmed: uint8 split labels shuffled in place, the halves by a stable radix argsort instead of two nonzero scans)
    inds_all = np.empty((nsteps, W), np.uint8)
    u_zz, u_acc = np.empty((nsteps, 2, ns)), np.empty((nsteps, 2, ns))
    rint = np.empty((nsteps, 2, ns), np.int64)
    inds_all[:] = (np.arange(W) % 2).astype(np.uint8)
    for t in range(nsteps):
        rng.random_sample()
        rng.shuffle(inds_all[t])            # the swaps depend on the stream and the length only, not on the element type
        for split in range(2):
            u_zz[t, split] = rng.rand(ns)
            rint[t, split] = rng.randint(W - ns, size=(ns,))
            u_acc[t, split] = rng.rand(ns)
    order = np.argsort(inds_all, axis=1, kind="stable")    # walkers of half 0 ascending, then those of half 1
    n0 = W - ns                                            # half 0 has one walker more when W is odd
    if n0 != ns:
        raise ValueError("the device-side stretch move needs an even number of walkers")
    s_idx = np.empty((nsteps, 2, ns), np.int32)
    s_idx[:, 0], s_idx[:, 1] = order[:, :n0], order[:, n0:]
    # partners: walker ids of the complementary half, picked by rint -- one flat take instead of take_along_axis
    comp = s_idx[:, ::-1].reshape(-1)
    rint += (np.arange(nsteps * 2, dtype=np.int64) * ns).reshape(nsteps, 2, 1)
    c_idx = comp.take(rint)
    zz = ((a - 1.0) * u_zz + 1.0) ** 2.0 / a
    with np.errstate(divide="ignore"):
        logu = np.log(u_acc, out=u_acc)
    return {"s_idx": s_idx, "c_idx": c_idx, "zz": zz, "factors": (ndim - 1.0) * np.log(zz), "logu": logu}


class DeviceStretchEnsemble(StretchEnsemble):
    """The same sampler with the chain state resident on the GPU.

    Per half-step: ``bi_stretch_propose`` -> ``bi_log_posterior`` (K1+K2, K4) -> [all-gather of log-probabilities when
    sharded over ranks] -> ``bi_stretch_accept``, all enqueued on torch's current stream; the host only generates the
    chain-independent random numbers (``draw_stretch_randoms``) one block of steps ahead and reads the chain back one
    block at a time.  Same RandomState consumption and the same rounded arithmetic as ``StretchEnsemble``, hence the
    same chain (tests/test_gpu_chain.py).  A NaN log-probability raises ``ValueError`` like emcee, at the end of the
    block it occurred in.

    ``graph=True``: state, scratch and the per-block random-number buffers are fixed allocations, so a block of ``block``
    steps is a fixed launch sequence (the NCCL all-gather included when sharded) and is captured once and replayed as ONE
    CUDA graph; the host then only uploads the next block's random numbers and replays (the
    collective is the only thing between the kernels).  A shorter last block runs eagerly.
    """

    def __init__(self, nwalkers: int, ndim: int, device_emulator, mode: str = "dense", a: float = 2.0,
                 seed: Optional[int] = None, random_state: Optional[np.random.RandomState] = None, block: int = 64,
                 group=None, graph: bool = False):
        self.dev, self.mode, self.block, self.group = device_emulator, mode, int(block), group
        super().__init__(nwalkers, ndim, log_prob_fn=None, a=a, seed=seed, random_state=random_state)
        if ndim != device_emulator.d:
            raise ValueError(f"ndim {ndim} does not match the emulator's parameter dimension {device_emulator.d}")
        self.graph = bool(graph)
        self.n_graph_replays = 0
        # the ensemble after each step is stored by the accept of its second half-step (bi_stretch_accept_record) instead of
        # two device-to-device copies per step; needs the two halves to be of equal size (developer knob BI_RECORD_FUSE=0)
        self._record_fused = nwalkers % 2 == 0 and os.environ.get("BI_RECORD_FUSE", "1") != "0"
        self._buf = None           # state, scratch and (with graph=True) the captured block live as long as the sampler
        self.world, self.rank = 1, 0
        try:
            import torch.distributed as dist

            if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
                self.world, self.rank, self._dist = dist.get_world_size(group), dist.get_rank(group), dist
        except Exception:
            pass

    def _device_log_prob(self, q, out):
        """out[:n] = log posterior of the n rows of q (CUDA tensors), sharded over ranks when distributed."""
        n = q.shape[0]
        if self.world == 1:
            self.dev.log_posterior_tensor(q, n_div=1.0, mode=self.mode, out=out[:n])
        else:
            lo, hi = ShardedLogProb.bounds(n, self.world, self.rank)
            per = (n + self.world - 1) // self.world
            if hi > lo:
                self.dev.log_posterior_tensor(q[lo:hi], n_div=1.0, mode=self.mode, out=self._send[:hi - lo])
            self._dist.all_gather_into_tensor(self._recv[:per * self.world], self._send[:per], group=self.group)
            out[:n].copy_(self._recv[:n])
        self.n_log_prob_calls += 1
        self.n_log_prob_rows += n

    # ---- random numbers of a block of steps -> the fixed device buffers ------------------------------------------------
    def _draw(self, nb: int):
        """Host side of a block: everything ``nb`` steps draw from emcee's stream (chain independent)."""
        return draw_stretch_randoms(self._random, self.nwalkers, self.ndim, nb, self.a)

    def _upload(self, drawn, nb: int, rd: dict) -> None:
        import torch

        for k, v in drawn.items():
            rd[k][:nb].copy_(torch.from_numpy(np.ascontiguousarray(v)), non_blocking=False)

    def _alloc(self, device) -> dict:
        import torch

        W, nd, ns, B = self.nwalkers, self.ndim, self.nwalkers // 2, self.block
        f64 = dict(dtype=torch.float64, device=device)
        per = (W + self.world - 1) // self.world
        self._send = torch.full((per,), float("nan"), **f64)
        self._recv = torch.empty(per * self.world, **f64)
        rd = {k: torch.empty((B, 2, ns), dtype=torch.int32, device=device) for k in ("s_idx", "c_idx")}
        rd.update({k: torch.empty((B, 2, ns), **f64) for k in ("zz", "factors", "logu")})
        return dict(coords=torch.empty((W, nd), **f64), logp=torch.empty(W, **f64), q=torch.empty((ns, nd), **f64),
                    new_lp=torch.empty(ns, **f64), n_acc=torch.zeros(W, dtype=torch.int64, device=device),
                    nan_flag=torch.zeros(1, dtype=torch.int32, device=device), rd=rd,
                    chain_b=torch.empty((B, W, nd), **f64), lp_b=torch.empty((B, W), **f64), graph=None)

    def _run_block(self, nb: int) -> None:
        bf = self._buf
        coords, logp, q, new_lp, n_acc, nan_flag, rd = (bf[k] for k in ("coords", "logp", "q", "new_lp", "n_acc", "nan_flag", "rd"))
        fused = self._record_fused
        for t in range(nb):
            for split in range(2):
                s_idx = rd["s_idx"][t, split]
                self.dev.stretch_propose(coords, s_idx, rd["c_idx"][t, split], rd["zz"][t, split], q)
                self._device_log_prob(q, new_lp)
                # the accept of the second half also stores the ensemble after the step (no copies between the steps)
                record = (rd["s_idx"][t, 0], bf["chain_b"][t], bf["lp_b"][t]) if (fused and split == 1) else None
                self.dev.stretch_accept(coords, logp, s_idx, q, new_lp, rd["factors"][t, split], rd["logu"][t, split],
                                        n_acc, nan_flag, record=record)
            if not fused:
                bf["chain_b"][t].copy_(coords)
                bf["lp_b"][t].copy_(logp)

    def _capture(self, device) -> None:
        """Capture ``block`` steps as one CUDA graph.  Every kernel of the block (and the communicator) has already run once
        -- module loading and NCCL initialisation are not capturable -- and the capture itself executes nothing: the
        state is untouched by it."""
        import torch

        bf = self._buf
        calls, rows = self.n_log_prob_calls, self.n_log_prob_rows
        c2, l2, a2 = bf["coords"].clone(), bf["logp"].clone(), bf["n_acc"].clone()
        rd = bf["rd"]
        self.dev.stretch_propose(c2, rd["s_idx"][0, 0], rd["c_idx"][0, 0], rd["zz"][0, 0], bf["q"])      # throw-away half-step
        self._device_log_prob(bf["q"], bf["new_lp"])
        self.dev.stretch_accept(c2, l2, rd["s_idx"][0, 0], bf["q"], bf["new_lp"], rd["factors"][0, 0], rd["logu"][0, 0], a2,
                                bf["nan_flag"])
        bf["nan_flag"].zero_()
        torch.cuda.synchronize(device)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._run_block(self.block)
        bf["graph"] = g
        self.n_log_prob_calls, self.n_log_prob_rows = calls, rows

    def sample(self, initial_state: np.ndarray, iterations: int = 1, log_prob0: Optional[np.ndarray] = None):
        import torch

        x0 = np.array(initial_state, dtype=np.float64, copy=True)
        W, nd, ns, B = self.nwalkers, self.ndim, self.nwalkers // 2, self.block
        if x0.shape != (W, nd):
            raise ValueError("incompatible input dimensions")
        check_initial_state(x0)
        device = torch.device("cuda", self.dev.device)
        with torch.cuda.device(device):
            if self._buf is None:
                self._buf = self._alloc(device)
            bf = self._buf
            coords, logp, n_acc, nan_flag, rd = (bf[k] for k in ("coords", "logp", "n_acc", "nan_flag", "rd"))
            coords.copy_(torch.from_numpy(x0))
            n_acc.zero_()
            nan_flag.zero_()
            if log_prob0 is None:
                self._device_log_prob(coords, logp)
            else:
                logp.copy_(torch.from_numpy(np.array(log_prob0, dtype=np.float64)))
            if bool(torch.isnan(logp).any()):
                raise ValueError("Probability function returned NaN")
            done = 0
            drawn = self._draw(min(B, iterations))
            while done < iterations:
                nb = min(B, iterations - done)
                self._upload(drawn, nb, rd)
                if self.graph and nb == B and bf["graph"] is None:
                    self._capture(device)
                if bf["graph"] is not None and nb == B:
                    bf["graph"].replay()
                    self.n_graph_replays += 1
                    self.n_log_prob_calls += 2 * nb
                    self.n_log_prob_rows += 2 * nb * ns
                else:
                    self._run_block(nb)
                if done + nb < iterations:      # the next block's random numbers are drawn while the GPU runs this one
                    drawn = self._draw(min(B, iterations - done - nb))
                chain_h, lp_h = bf["chain_b"][:nb].cpu().numpy(), bf["lp_b"][:nb].cpu().numpy()   # synchronises: the block is done
                if int(nan_flag.item()):
                    raise ValueError("Probability function returned NaN")
                self._accepted += n_acc.cpu().numpy()
                n_acc.zero_()
                self._after_block(nb)
                for t in range(nb):
                    self.iteration += 1
                    self._chain.append(chain_h[t])
                    self._log_prob.append(lp_h[t])
                    yield chain_h[t], lp_h[t]
                done += nb

    def _after_block(self, nb: int) -> None:
        pass


class DevicePhiloxEnsemble(DeviceStretchEnsemble):
    """Device-resident stretch move whose random numbers are generated on the GPU as well (the variant
    "reported separately"): counter-based Philox4x32-10 keyed by ``seed`` (``bi_stretch_randoms_philox``), fixed
    even / odd halves instead of emcee's per-step permutation.  Nothing crosses the PCIe bus inside a block of steps;
    with ``graph=True`` a block is replayed as one CUDA graph.

    The chain is a valid stretch-move chain but NOT emcee's for the same seed; it is reproducible from (seed, step) and
    independent of the block size and of ``graph``.
    """

    def __init__(self, nwalkers: int, ndim: int, device_emulator, mode: str = "dense", a: float = 2.0, seed: int = 0,
                 block: int = 64, group=None, graph: bool = False):
        super().__init__(nwalkers, ndim, device_emulator, mode=mode, a=a, seed=0, block=block, group=group, graph=graph)
        if nwalkers % 2:
            raise ValueError("the fixed even / odd split needs an even number of walkers")
        self.seed, self._step = int(seed), 0

    def reset(self, rewind: bool = False) -> None:
        """Forget the chain; buffers and the captured graph are kept.  The random stream carries on (burn-in and
        production of ``mcmc.run_mcmc`` must not see the same numbers) unless ``rewind`` restarts it at step 0."""
        super().reset()
        if rewind:
            self._step = 0

    def _draw(self, nb: int):
        return None                      # nothing comes from the host

    def _upload(self, drawn, nb: int, rd: dict) -> None:
        self.dev.stretch_randoms_philox(self.seed, self._step, nb, self.nwalkers // 2, self.a, rd)

    def _after_block(self, nb: int) -> None:
        self._step += nb


class BatchedClosureEnsemble:
    """Several independent ensembles advanced in lock-step on one GPU, chain c against experimental data set c
    (SURVEY §8 f3).  The reference runs its closure tests -- one MCMC per validation design point, each against that
    point's pseudo-data and the same emulators -- one after the other (steer_analysis.py:168-183); here every half-step
    is one ``bi_stretch_propose`` / ``bi_log_posterior_sets`` / ``bi_stretch_accept`` over the walkers of all chains.
    Chain c consumes its own ``RandomState`` exactly as a stand-alone run would, so it IS the stand-alone chain.

    ``chains[c]`` is a ``StretchEnsemble`` used as the per-chain store (``get_chain``, ``flatchain``,
    ``acceptance_fraction``, ``get_autocorr_time``, ``reset``).
    """

    def __init__(self, nwalkers: int, ndim: int, device_emulator, random_states, mode: str = "dense", a: float = 2.0,
                 block: int = 64):
        self.dev, self.mode, self.block, self.a = device_emulator, mode, int(block), float(a)
        self.nwalkers, self.ndim, self.n_chains = int(nwalkers), int(ndim), len(random_states)
        if getattr(device_emulator, "n_sets", 0) < self.n_chains:
            raise ValueError("load one experimental data set per chain first (DeviceEmulator.set_data_sets)")
        self.chains = [StretchEnsemble(nwalkers, ndim, None, a=a, random_state=rs) for rs in random_states]
        self.n_log_prob_rows = 0

    def reset(self) -> None:
        for c in self.chains:
            c.reset()

    def run_mcmc(self, initial_states, nsteps: int):
        """initial_states: (n_chains, W, d).  Returns the last coordinates (n_chains, W, d)."""
        import torch

        C_, W, nd, ns = self.n_chains, self.nwalkers, self.ndim, self.nwalkers // 2
        x0 = np.array(initial_states, dtype=np.float64, copy=True)
        if x0.shape != (C_, W, nd):
            raise ValueError("incompatible input dimensions")
        for c in range(C_):
            check_initial_state(x0[c])
        device = torch.device("cuda", self.dev.device)
        with torch.cuda.device(device):
            coords = torch.from_numpy(x0.reshape(C_ * W, nd)).to(device)
            logp = torch.empty(C_ * W, dtype=torch.float64, device=device)
            set_all = torch.arange(C_, dtype=torch.int32, device=device).repeat_interleave(W).contiguous()
            set_q = torch.arange(C_, dtype=torch.int32, device=device).repeat_interleave(ns).contiguous()
            self.dev.log_posterior_sets_tensor(coords, set_all, n_div=1.0, mode=self.mode, out=logp)
            self.n_log_prob_rows += C_ * W
            if bool(torch.isnan(logp).any()):
                raise ValueError("Probability function returned NaN")
            q = torch.empty((C_ * ns, nd), dtype=torch.float64, device=device)
            new_lp = torch.empty(C_ * ns, dtype=torch.float64, device=device)
            n_acc = torch.zeros(C_ * W, dtype=torch.int64, device=device)
            nan_flag = torch.zeros(1, dtype=torch.int32, device=device)
            def draw(nb):
                per = [draw_stretch_randoms(ch._random, W, nd, nb, self.a) for ch in self.chains]
                r = {}
                for k in per[0]:
                    off = (lambda c: c * W) if k in ("s_idx", "c_idx") else (lambda c: 0)
                    r[k] = np.ascontiguousarray(np.concatenate([per[c][k] + off(c) for c in range(C_)], axis=2).astype(per[0][k].dtype))
                return r

            done = 0
            r = draw(min(self.block, nsteps))
            while done < nsteps:
                nb = min(self.block, nsteps - done)
                rd = {k: torch.from_numpy(v).pin_memory().to(device, non_blocking=True) for k, v in r.items()}
                chain_b = torch.empty((nb, C_ * W, nd), dtype=torch.float64, device=device)
                lp_b = torch.empty((nb, C_ * W), dtype=torch.float64, device=device)
                for t in range(nb):
                    for split in range(2):
                        s_idx = rd["s_idx"][t, split]
                        self.dev.stretch_propose(coords, s_idx, rd["c_idx"][t, split], rd["zz"][t, split], q)
                        self.dev.log_posterior_sets_tensor(q, set_q, n_div=1.0, mode=self.mode, out=new_lp)
                        self.dev.stretch_accept(coords, logp, s_idx, q, new_lp, rd["factors"][t, split], rd["logu"][t, split],
                                                n_acc, nan_flag)
                    chain_b[t].copy_(coords)
                    lp_b[t].copy_(logp)
                self.n_log_prob_rows += nb * 2 * C_ * ns
                if done + nb < nsteps:          # drawn while the GPU runs this block
                    r = draw(min(self.block, nsteps - done - nb))
                chain_h = chain_b.cpu().numpy().reshape(nb, C_, W, nd)
                lp_h = lp_b.cpu().numpy().reshape(nb, C_, W)
                if int(nan_flag.item()):
                    raise ValueError("Probability function returned NaN")
                acc_h = n_acc.cpu().numpy().reshape(C_, W)
                n_acc.zero_()
                for c, ch in enumerate(self.chains):
                    ch._accepted += acc_h[c]
                    for t in range(nb):
                        ch.iteration += 1
                        ch._chain.append(chain_h[t, c])
                        ch._log_prob.append(lp_h[t, c])
                done += nb
            return coords.cpu().numpy().reshape(C_, W, nd)


class ShardedLogProb:
    """Walker sharding over ranks: each rank evaluates a contiguous slice of the rows, one all-gather of
    the log-probabilities closes the call.  Every rank runs the identical sampler (same seed), so proposals and
    accept/reject are replicated and the all-gather is the only collective."""

    def __init__(self, local_fn: Callable[[np.ndarray], np.ndarray], group=None):
        import torch.distributed as dist

        self.fn, self.dist, self.group = local_fn, dist, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.cuda = dist.get_backend(group) == "nccl"

    @staticmethod
    def bounds(n: int, world: int, rank: int) -> tuple[int, int]:
        per = (n + world - 1) // world
        return min(n, rank * per), min(n, (rank + 1) * per)

    def __call__(self, X: np.ndarray) -> np.ndarray:
        import torch

        n = X.shape[0]
        lo, hi = self.bounds(n, self.world, self.rank)
        per = (n + self.world - 1) // self.world
        local = np.full(per, np.nan)
        if hi > lo:
            local[:hi - lo] = self.fn(X[lo:hi])
        dev = torch.device("cuda", torch.cuda.current_device()) if self.cuda else torch.device("cpu")
        send = torch.from_numpy(local).to(dev)
        recv = torch.empty(per * self.world, dtype=torch.float64, device=dev)
        self.dist.all_gather_into_tensor(recv, send, group=self.group)
        return recv.cpu().numpy()[:n]   # rank r's rows sit at [r*per, r*per + count): contiguous in row order
