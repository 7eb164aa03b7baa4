"""Accepted-chain identity: the same sampler and seed driven by the CUDA log-posterior and by the numpy oracle must
produce the same chain for >= 1000 steps (BASELINE north_star).  A 1e-13 relative difference in logp can only flip an
accept decision at a near-tie; if that ever happens the test reports the step and the margin instead of hiding it."""
from __future__ import annotations

import numpy as np
import pytest
from conftest import load_golden

from bayesian_inference_b200 import image
from bayesian_inference_b200.ensemble import StretchEnsemble
from oracle import closed_form as cf

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode", ["dense", "structured"])
def test_chain_trace_identical_for_1000_steps(mode):
    from bayesian_inference_b200.device import DeviceEmulator

    z, _ = load_golden("mid")                      # d=4, 2 groups, F=148: the oracle needs ~0.5 ms per row
    W, nsteps, seed = 16, 1000, 2026
    x0 = np.random.RandomState(9).uniform(0.2, 0.8, (W, int(z["d"])))
    with DeviceEmulator(image.EmulatorImage.from_flat(z)) as dev:
        gpu = StretchEnsemble(W, int(z["d"]), lambda X: dev.log_posterior(X, n_div=1.0, mode=mode), seed=seed)
        gpu.run_mcmc(x0, nsteps)
    cpu = StretchEnsemble(W, int(z["d"]), lambda X: cf.log_posterior(z, X, n_div=1.0), seed=seed)
    cpu.run_mcmc(x0, nsteps)
    a, b = gpu.get_chain(), cpu.get_chain()
    same = np.all(a == b, axis=(1, 2))
    if not same.all():
        step = int(np.argmin(same))
        lp_a, lp_b = gpu.get_log_prob()[step], cpu.get_log_prob()[step]
        pytest.fail(f"chains diverge at step {step}: max |dlogp| there {np.nanmax(np.abs(lp_a - lp_b)):.3e}")
    assert np.array_equal(gpu.acceptance_fraction, cpu.acceptance_fraction)
    fin = np.isfinite(cpu.get_log_prob())
    assert np.max(np.abs(gpu.get_log_prob()[fin] - cpu.get_log_prob()[fin]) / np.abs(cpu.get_log_prob()[fin])) < 1e-9
    assert 0.05 < gpu.acceptance_fraction.mean() < 0.95


def test_chain_trace_at_the_c1_shape_vs_oracle():
    """BASELINE configs[0] (jet_substructure.yaml-shaped: d=4, n_design=230, n_pc=30, n_obs=200; 100 walkers, seed 42;
    reference driver mcmc.py:83-107): 200 steps of the same StretchEnsemble driven by the CUDA log-posterior (host-side and
    device-resident sampler) and by the oracle evaluated row by row on a pool of host processes.  The 1,000-step run of the
    same comparison is tools/chain_trace.py -> profiles/r02_chain_trace_c1.json."""
    from bayesian_inference_b200 import synthetic
    from bayesian_inference_b200.device import DeviceEmulator
    from bayesian_inference_b200.ensemble import DeviceStretchEnsemble
    from oracle.pool import OraclePool

    a = synthetic.generate("C1")
    img = image.pack(a.emulation_config, a.emulation_results, min=a.min, max=a.max,
                     experimental_results=a.experimental_results)
    z = img.to_flat()
    W, d, nsteps, seed = 100, 4, 200, 42
    assert (img.n_features, img.groups[0].n_design, img.groups[0].n_pc, img.d) == (200, 230, 30, 4)
    x0 = np.random.RandomState(seed).uniform(a.min, a.max, (W, d))           # mcmc.py:88
    with OraclePool(z) as pool:
        cpu = StretchEnsemble(W, d, pool.log_posterior, seed=seed)
        cpu.run_mcmc(x0, nsteps)
    with DeviceEmulator(img) as dev:
        host = StretchEnsemble(W, d, lambda X: dev.log_posterior(X, n_div=1.0, mode="dense"), seed=seed)
        host.run_mcmc(x0, nsteps)
        devs = DeviceStretchEnsemble(W, d, dev, mode="dense", seed=seed)
        devs.run_mcmc(x0, nsteps)
    for s in (host, devs):
        same = np.all(s.get_chain() == cpu.get_chain(), axis=(1, 2))
        if not same.all():
            step = int(np.argmin(same))
            pytest.fail(f"chains diverge at step {step}: max |dlogp| there "
                        f"{np.nanmax(np.abs(s.get_log_prob()[step] - cpu.get_log_prob()[step])):.3e}")
        assert np.array_equal(s.acceptance_fraction, cpu.acceptance_fraction)
        fin = np.isfinite(cpu.get_log_prob())
        assert np.array_equal(np.isfinite(s.get_log_prob()), fin)
        assert np.max(np.abs(s.get_log_prob()[fin] - cpu.get_log_prob()[fin]) / np.abs(cpu.get_log_prob()[fin])) < 1e-9
    assert 0.05 < cpu.acceptance_fraction.mean() < 0.95


@pytest.mark.parametrize("mode", ["dense", "woodbury"])
def test_device_resident_sampler_matches_host_sampler(mode):
    """DeviceStretchEnsemble (propose / accept kernels, state in HBM) == StretchEnsemble driven by the same CUDA
    log-posterior: identical chain, log-probabilities and acceptance counts for 1000 steps (bit for bit)."""
    from bayesian_inference_b200.device import DeviceEmulator
    from bayesian_inference_b200.ensemble import DeviceStretchEnsemble

    z, _ = load_golden("mid")
    W, nsteps, seed, d = 32, 1000, 77, int(z["d"])
    x0 = np.random.RandomState(3).uniform(0.1, 0.9, (W, d))
    x0[0, 0] = 1.5      # one walker starts outside the prior box: logp = -inf, it must move in or stay, never NaN
    with DeviceEmulator(image.EmulatorImage.from_flat(z)) as dev:
        host = StretchEnsemble(W, d, lambda X: dev.log_posterior(X, n_div=1.0, mode=mode), seed=seed)
        host.run_mcmc(x0, nsteps)
        devs = DeviceStretchEnsemble(W, d, dev, mode=mode, seed=seed, block=96)     # 1000 = 10 x 96 + 40: ragged last block
        last = devs.run_mcmc(x0, nsteps)
    assert np.array_equal(devs.get_chain(), host.get_chain())
    assert np.array_equal(devs.get_log_prob(), host.get_log_prob())
    assert np.array_equal(devs.acceptance_fraction, host.acceptance_fraction)
    assert np.array_equal(last[0], host.get_chain()[-1])
    assert devs.n_log_prob_rows == host.n_log_prob_rows


def test_emcee_stream_sampler_as_cuda_graph_is_the_same_chain():
    """graph=True on the emcee-stream sampler: blocks of steps replayed as one CUDA graph (random numbers uploaded per
    block into fixed buffers) give the host sampler's chain bit for bit, ragged last block included."""
    from bayesian_inference_b200.device import DeviceEmulator
    from bayesian_inference_b200.ensemble import DeviceStretchEnsemble

    z, _ = load_golden("mid")
    W, nsteps, seed, d = 32, 150, 77, int(z["d"])
    x0 = np.random.RandomState(3).uniform(0.1, 0.9, (W, d))
    with DeviceEmulator(image.EmulatorImage.from_flat(z)) as dev:
        host = StretchEnsemble(W, d, lambda X: dev.log_posterior(X, n_div=1.0), seed=seed)
        host.run_mcmc(x0, nsteps)
        g = DeviceStretchEnsemble(W, d, dev, seed=seed, block=32, graph=True)         # 150 = 4 x 32 + 22
        g.run_mcmc(x0, nsteps)
        # other divisors on the same handle between replays must not disturb the captured block (A tiles of n_div = 1 are
        # never rewritten)
        dev.log_posterior(x0, n_div=None)
        dev.log_posterior(x0[:7], n_div=3.0)
        g2 = DeviceStretchEnsemble(W, d, dev, seed=seed, block=32, graph=True)
        g2.run_mcmc(x0, 64)
        dev.log_posterior(x0, n_div=5.0)
        g2.run_mcmc(g2.get_chain()[-1], 64, log_prob0=g2.get_log_prob()[-1])
    assert g.n_graph_replays == 4
    assert np.array_equal(g.get_chain(), host.get_chain()) and np.array_equal(g.get_log_prob(), host.get_log_prob())
    assert np.array_equal(g.acceptance_fraction, host.acceptance_fraction)
    assert g.n_log_prob_rows == host.n_log_prob_rows
    assert g2.n_graph_replays == 4 and np.array_equal(g2.get_chain(), host.get_chain()[:128])


def test_accept_record_equals_accept_followed_by_copies():
    """bi_stretch_accept_record (the accept of a step's second half-step, which also stores the ensemble) against
    bi_stretch_accept followed by copies of the state -- the same coords, logp, counts and the same record, bit for bit."""
    import torch

    from bayesian_inference_b200.device import DeviceEmulator

    z, _ = load_golden("mid")
    d, W = int(z["d"]), 48
    ns = W // 2
    rng = np.random.default_rng(11)
    perm = rng.permutation(W).astype(np.int32)
    first, second = np.sort(perm[:ns]), np.sort(perm[ns:])
    with DeviceEmulator(image.EmulatorImage.from_flat(z)) as dev:
        dv = torch.device("cuda", dev.device)
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dv)       # noqa: E731
        coords0, logp0 = rng.uniform(0.2, 0.8, (W, d)), rng.normal(-50.0, 5.0, W)
        q, new_lp = rng.uniform(0.2, 0.8, (ns, d)), rng.normal(-50.0, 5.0, ns)
        factors, logu = rng.normal(0.0, 1.0, ns), np.log(rng.uniform(size=ns))
        out = []
        for record in (False, True):
            coords, logp = t(coords0), t(logp0)
            n_acc = torch.zeros(W, dtype=torch.int64, device=dv)
            nan_flag = torch.zeros(1, dtype=torch.int32, device=dv)
            chain = torch.full((W, d), float("nan"), dtype=torch.float64, device=dv)
            lp = torch.full((W,), float("nan"), dtype=torch.float64, device=dv)
            rec = (t(first), chain, lp) if record else None
            dev.stretch_accept(coords, logp, t(second), t(q), t(new_lp), t(factors), t(logu), n_acc, nan_flag, record=rec)
            if not record:
                chain.copy_(coords)
                lp.copy_(logp)
            torch.cuda.synchronize(dv)
            out.append([x.cpu().numpy() for x in (coords, logp, n_acc, chain, lp)])
    for a, b in zip(*out):
        assert np.array_equal(a, b)
    assert 0 < out[0][2].sum() < ns and not np.isnan(out[1][3]).any()        # some accepted, some not; every row recorded


def test_device_sampler_raises_on_nan_like_emcee():
    """A covariance that is not positive definite gives NaN; emcee raises ValueError, so does the device sampler."""
    from bayesian_inference_b200.device import DeviceEmulator
    from bayesian_inference_b200.ensemble import DeviceStretchEnsemble

    z, _ = load_golden("toy")
    z = dict(z)
    z["y_err"] = np.zeros_like(z["y_err"])
    z["A"] = None if "A" not in z else z["A"]
    img = image.EmulatorImage.from_flat(z)
    img.y_err = np.zeros_like(img.y_err)
    for g in img.groups:          # no truncated-PC covariance either: C = T diag(v) T^T is rank deficient
        g.S_rest = g.S_rest[:0]
        g.ev_rest = g.ev_rest[:0]
    d = img.d
    x0 = np.random.RandomState(1).uniform(0.2, 0.8, (16, d))
    with DeviceEmulator(img) as dev:
        s = DeviceStretchEnsemble(16, d, dev, seed=5)
        with pytest.raises(ValueError, match="returned NaN"):
            s.run_mcmc(x0, 5)


def test_philox_randoms_match_numpy_restatement():
    """bi_stretch_randoms_philox == oracle/philox.py: indices and zz bit for bit, the logarithms to 2 ulp."""
    import torch

    from bayesian_inference_b200.device import DeviceEmulator
    from oracle import philox as ph

    z, _ = load_golden("mid")
    d, W, nb = int(z["d"]), 48, 7
    with DeviceEmulator(image.EmulatorImage.from_flat(z)) as dev:
        rd = {k: torch.empty((nb, 2, W // 2), dtype=torch.int32, device="cuda") for k in ("s_idx", "c_idx")}
        rd.update({k: torch.empty((nb, 2, W // 2), dtype=torch.float64, device="cuda") for k in ("zz", "factors", "logu")})
        dev.stretch_randoms_philox(0x1234567890ABCDEF, (1 << 32) - 3, nb, W // 2, 2.0, rd)       # the step counter crosses 2^32
        got = {k: v.cpu().numpy() for k, v in rd.items()}
    want = ph.stretch_randoms(0x1234567890ABCDEF, (1 << 32) - 3, nb, W, d, 2.0)
    for k in ("s_idx", "c_idx", "zz"):
        assert np.array_equal(got[k], want[k]), k
    for k in ("factors", "logu"):
        assert np.max(np.abs(got[k] - want[k]) / np.maximum(np.abs(want[k]), 1e-300)) < 5e-16, k


def test_philox_sampler_is_the_host_replay_and_graph_invariant():
    """DevicePhiloxEnsemble == the pre-drawn device sampler fed with the numpy-restated random numbers (same chain bit
    for bit), whatever the block size, and identical when the blocks are replayed as CUDA graphs."""
    import torch

    from bayesian_inference_b200.device import DeviceEmulator
    from bayesian_inference_b200.ensemble import DevicePhiloxEnsemble
    from oracle import philox as ph

    z, _ = load_golden("mid")
    W, nsteps, seed, d = 32, 300, 99, int(z["d"])
    x0 = np.random.RandomState(3).uniform(0.1, 0.9, (W, d))

    class Replay(DevicePhiloxEnsemble):
        def _upload(self, drawn, nb, rd):
            r = ph.stretch_randoms(self.seed, self._step, nb, self.nwalkers, self.ndim, self.a)
            for k, v in r.items():
                rd[k][:nb].copy_(torch.from_numpy(v))

    with DeviceEmulator(image.EmulatorImage.from_flat(z)) as dev:
        a = DevicePhiloxEnsemble(W, d, dev, seed=seed, block=64)
        a.run_mcmc(x0, nsteps)
        b = Replay(W, d, dev, seed=seed, block=50)
        b.run_mcmc(x0, nsteps)
        g = DevicePhiloxEnsemble(W, d, dev, seed=seed, block=64, graph=True)
        g.run_mcmc(x0, 200)
        g.run_mcmc(g.get_chain()[-1], 100, log_prob0=g.get_log_prob()[-1])      # the step counter carries on
    assert np.array_equal(a.get_chain(), b.get_chain()) and np.array_equal(a.get_log_prob(), b.get_log_prob())
    assert np.array_equal(a.acceptance_fraction, b.acceptance_fraction)
    assert g.n_graph_replays == 4 and np.array_equal(g.get_chain(), a.get_chain())
    assert g.n_log_prob_rows == a.n_log_prob_rows              # warm-up and capture are not counted as evaluations
    assert 0.15 < a.acceptance_fraction.mean() < 0.85


def test_philox_sampler_samples_the_same_posterior():
    """Not emcee's stream, but the same target: posterior means of the Philox chain and of the RandomState chain agree
    within a fraction of the posterior width (32 walkers x 800 steps, first 200 dropped)."""
    from bayesian_inference_b200.device import DeviceEmulator
    from bayesian_inference_b200.ensemble import DevicePhiloxEnsemble, DeviceStretchEnsemble

    z, _ = load_golden("mid")
    W, d = 32, int(z["d"])
    x0 = np.random.RandomState(3).uniform(0.2, 0.8, (W, d))
    with DeviceEmulator(image.EmulatorImage.from_flat(z)) as dev:
        p = DevicePhiloxEnsemble(W, d, dev, seed=1, block=100, graph=True)
        p.run_mcmc(x0, 800)
        r = DeviceStretchEnsemble(W, d, dev, seed=1, block=100)
        r.run_mcmc(x0, 800)
    cp, cr = p.get_chain()[200:].reshape(-1, d), r.get_chain()[200:].reshape(-1, d)
    assert np.all(np.abs(cp.mean(0) - cr.mean(0)) < 0.35 * cr.std(0))
    assert np.all(np.abs(cp.std(0) / cr.std(0) - 1.0) < 0.35)
