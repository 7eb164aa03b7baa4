"""The stretch-move ensemble sampler and the walker sharding (CPU: gloo, world_size 2)."""
from __future__ import annotations

import os
import socket
import types

import numpy as np
import pytest
from conftest import load_golden

from bayesian_inference_b200.ensemble import ShardedLogProb, StretchEnsemble
from oracle import closed_form as cf


def _gauss(x):
    return -0.5 * np.sum((x - 0.3) ** 2 / 0.04, axis=1)


def _scalar_reference_chain(logp, x0, nsteps, seed, a=2.0):
    """Independent restatement: per-walker Python loops, the same RandomState consumption order."""
    rng = np.random.RandomState(seed)
    x = x0.copy()
    W, nd = x.shape
    lp = np.array([logp(x[i:i + 1])[0] for i in range(W)])
    out = []
    for _ in range(nsteps):
        rng.random_sample()
        inds = np.arange(W) % 2
        rng.shuffle(inds)
        for split in (0, 1):
            mine = [i for i in range(W) if inds[i] == split]
            other = [i for i in range(W) if inds[i] != split]
            u = rng.rand(len(mine))
            partners = rng.randint(len(other), size=(len(mine),))
            props, fac = [], []
            for k, i in enumerate(mine):
                zz = ((a - 1.0) * u[k] + 1.0) ** 2.0 / a
                cpos = x[other[partners[k]]]
                props.append(cpos - (cpos - x[i]) * zz)
                fac.append((nd - 1.0) * np.log(zz))
            new = [logp(p[None, :])[0] for p in props]
            for k, i in enumerate(mine):          # snapshot semantics: proposals were built before any update
                if fac[k] + new[k] - lp[i] > np.log(rng.rand()):
                    x[i], lp[i] = props[k], new[k]
        out.append(x.copy())
    return np.asarray(out)


def test_vectorised_sampler_equals_scalar_restatement():
    rng = np.random.RandomState(1)
    x0 = rng.uniform(0, 1, (12, 3))
    s = StretchEnsemble(12, 3, _gauss, seed=77)
    s.run_mcmc(x0, 60)
    assert np.array_equal(s.get_chain(), _scalar_reference_chain(_gauss, x0, 60, 77))
    assert s.get_chain().shape == (60, 12, 3) and s.get_log_prob().shape == (60, 12)
    assert 0.1 < s.acceptance_fraction.mean() < 0.9
    assert s.flatchain.shape == (720, 3) and s.flatlnprobability.shape == (720,)


def test_sampler_contract():
    with pytest.raises(ValueError):
        StretchEnsemble(4, 3, _gauss)                      # fewer than 2*ndim walkers
    s = StretchEnsemble(8, 2, lambda x: np.full(len(x), np.nan), seed=0)
    with pytest.raises(ValueError, match="NaN"):
        s.run_mcmc(np.random.RandomState(0).rand(8, 2), 1)
    s = StretchEnsemble(8, 2, _gauss, seed=0)
    with pytest.raises(ValueError, match="condition number"):
        s.run_mcmc(np.ones((8, 2)), 1)
    box = lambda x: np.where(np.all((x > 0) & (x < 1), axis=1), 0.0, -np.inf)   # noqa: E731
    s = StretchEnsemble(20, 2, box, seed=3)
    s.run_mcmc(np.random.RandomState(5).uniform(0.2, 0.8, (20, 2)), 200)
    assert np.all((s.flatchain > 0) & (s.flatchain < 1))   # -inf proposals are never accepted
    tau = StretchEnsemble(20, 2, _gauss, seed=3)
    tau.run_mcmc(np.random.RandomState(5).uniform(0, 1, (20, 2)), 2000)
    t = tau.get_autocorr_time(tol=0)
    assert t.shape == (2,) and np.all(t > 1) and np.all(t < 200)


def test_run_mcmc_driver_sequence(monkeypatch, tmp_path):
    """mcmc.run_mcmc with the oracle standing in for the GPU: burn-in, re-seed at the best distinct points, production."""
    from bayesian_inference_b200 import log_posterior, mcmc

    z, _ = load_golden("toy")
    monkeypatch.setattr(log_posterior, "initialize_pool_variables", lambda *a, **k: None)
    monkeypatch.setattr(log_posterior, "log_posterior_walkers", lambda X: cf.log_posterior(z, X, n_div=1.0))
    cfg = types.SimpleNamespace(
        analysis_config={"parameterization": {"p": {"names": ["a", "b", "c"], "min": list(z["min"]), "max": list(z["max"])}}},
        parameterization="p", analysis_name="toy", config_file="", output_dir=str(tmp_path), mcmc_output_dir=str(tmp_path),
        n_walkers=8, n_burn_steps=6, n_sampling_steps=5, n_logging_steps=2, sampler_outputfile=str(tmp_path / "s.pkl"))
    out = mcmc.run_mcmc(cfg, emulation_config=object(), emulation_results={}, experimental_results={}, seed=5, sampler="host")
    assert out["chain"].shape == (5, 8, 3) and out["log_prob"].shape == (5, 8) and out["acceptance_fraction"].shape == (8,)
    assert out["autocorrelation_time"] is None            # chain far too short: recorded as None, like the reference
    assert out["n_log_prob_rows"] == (8 + 3 * 8) + (8 + 3 * 8) + (8 + 5 * 8)   # two burn-in halves + production, W rows per step
    again = mcmc.run_mcmc(cfg, emulation_config=object(), emulation_results={}, experimental_results={}, seed=5, write=False, sampler="host")
    assert np.array_equal(out["chain"], again["chain"])
    assert (tmp_path / "mcmc.npz").exists() or (tmp_path / "mcmc.h5").exists()
    import pickle

    with open(tmp_path / "s.pkl", "rb") as f:              # mcmc.py:127-132: the pickled object answers like the sampler
        back = pickle.load(f)
    assert np.array_equal(back.get_chain(), out["chain"]) and np.array_equal(back.get_log_prob(), out["log_prob"])
    assert np.array_equal(back.acceptance_fraction, out["acceptance_fraction"])
    assert back.get_chain(flat=True, discard=1, thin=2).shape == (2 * 8, 3) and back.log_prob_fn is None
    with pytest.raises(ValueError, match="sigma_exp"):
        mcmc.run_mcmc_closures(cfg, [{}], emulation_config=object(), emulation_results={}, sigma_exp=np.eye(3))


def test_write_results_only_falls_back_on_missing_modules(monkeypatch, tmp_path):
    """I/O errors of the HDF5 writers propagate: only ImportError selects the next writer."""
    import builtins
    import sys

    from bayesian_inference_b200 import mcmc

    cfg = types.SimpleNamespace(mcmc_output_dir=str(tmp_path))
    fake = types.ModuleType("bayesian_inference")
    fake.data_IO = types.SimpleNamespace(write_dict_to_h5=lambda *a, **k: (_ for _ in ()).throw(OSError("disk full")))
    monkeypatch.setitem(sys.modules, "bayesian_inference", fake)
    monkeypatch.setitem(sys.modules, "bayesian_inference.data_IO", fake.data_IO)
    with pytest.raises(OSError, match="disk full"):
        mcmc._write_results({"chain": np.zeros((2, 4, 3))}, cfg)
    assert not (tmp_path / "mcmc.npz").exists()
    monkeypatch.delitem(sys.modules, "bayesian_inference")
    monkeypatch.delitem(sys.modules, "bayesian_inference.data_IO")
    real_import = builtins.__import__

    def no_hdf5(name, *a, **k):
        if name.split(".")[0] in ("bayesian_inference", "h5py", "silx"):
            raise ImportError(name)
        return real_import(name, *a, **k)

    monkeypatch.setattr(builtins, "__import__", no_hdf5)
    path = mcmc._write_results({"chain": np.zeros((2, 4, 3)), "autocorrelation_time": None,
                                "experimental_pseudodata": {"y": np.ones(5), "y_err": np.ones(5)}}, cfg)
    assert path.endswith("mcmc.npz")
    with np.load(path) as z:
        assert sorted(z.files) == ["chain", "experimental_pseudodata/y", "experimental_pseudodata/y_err"]


def test_mcmc_h5_layout_round_trip(tmp_path):
    """mcmc.h5 with the reference's keys (mcmc.py:111-125) written by the h5py restatement of data_IO.write_dict_to_h5."""
    pytest.importorskip("h5py")
    from bayesian_inference_b200 import mcmc

    d = {"chain": np.random.RandomState(0).rand(5, 8, 3), "acceptance_fraction": np.linspace(0, 1, 8),
         "log_prob": np.random.RandomState(1).rand(5, 8), "autocorrelation_time": None, "design_point": np.arange(3.0),
         "experimental_pseudodata": {"y": np.ones(4), "y_err": np.full(4, 0.1)}}
    mcmc.write_dict_to_h5(d, str(tmp_path), "mcmc.h5")
    mcmc.write_dict_to_h5({"chain": d["chain"] * 2}, str(tmp_path), "mcmc.h5")          # update_mode='modify': replaced
    back = mcmc.read_dict_from_h5(str(tmp_path), "mcmc.h5")
    assert sorted(back) == ["acceptance_fraction", "chain", "design_point", "experimental_pseudodata", "log_prob"]
    assert np.array_equal(back["chain"], d["chain"] * 2) and np.array_equal(back["experimental_pseudodata"]["y_err"], d["experimental_pseudodata"]["y_err"])


# ---- world_size-2 gloo: sharded log-probability and replicated sampler ------------------------------------------------
def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        calls = []

        def local(x):
            calls.append(x.shape[0])
            return _gauss(x)

        fn = ShardedLogProb(local)
        X = np.random.RandomState(0).rand(7, 3)           # ragged: 4 + 3
        full = fn(X)
        s = StretchEnsemble(10, 3, fn, seed=11)
        s.run_mcmc(np.random.RandomState(2).rand(10, 3), 25)
        q.put((rank, full, s.get_chain(), calls[0]))
    finally:
        dist.destroy_process_group()


def test_sharded_logprob_gloo_world2():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    res = sorted([q.get(timeout=120) for _ in procs], key=lambda t: t[0])
    [p.join(60) for p in procs]
    X = np.random.RandomState(0).rand(7, 3)
    single = StretchEnsemble(10, 3, _gauss, seed=11)
    single.run_mcmc(np.random.RandomState(2).rand(10, 3), 25)
    for rank, full, chain, first_rows in res:
        assert np.array_equal(full, _gauss(X))                       # gathered in row order, ragged tail handled
        assert np.array_equal(chain, single.get_chain())             # every rank holds the same chain as 1 process
        assert first_rows == (4 if rank == 0 else 3)                 # each rank evaluated only its slice
    assert ShardedLogProb.bounds(7, 2, 0) == (0, 4) and ShardedLogProb.bounds(7, 2, 1) == (4, 7)
    assert ShardedLogProb.bounds(2, 4, 3) == (2, 2)                  # more ranks than rows: empty slice


def test_predrawn_randoms_reproduce_the_sampler_stream():
    """draw_stretch_randoms consumes the RandomState exactly as StretchEnsemble.sample does: replaying its arrays with
    numpy gives the identical chain (this is what the device-resident sampler uploads)."""
    from bayesian_inference_b200.ensemble import StretchEnsemble, draw_stretch_randoms

    W, nd, steps = 20, 3, 60
    f = lambda X: -0.5 * np.sum(X ** 2, axis=1)
    x0 = np.random.RandomState(7).normal(size=(W, nd))
    s = StretchEnsemble(W, nd, f, random_state=np.random.RandomState(11))
    s.run_mcmc(x0, steps)
    rng = np.random.RandomState(11)
    r = draw_stretch_randoms(rng, W, nd, 25)
    r2 = draw_stretch_randoms(rng, W, nd, steps - 25)        # block boundaries do not disturb the stream
    r = {k: np.concatenate([r[k], r2[k]]) for k in r}
    coords, lp, chain = x0.copy(), f(x0), []
    for t in range(steps):
        for sp in range(2):
            si, ci = r["s_idx"][t, sp], r["c_idx"][t, sp]
            assert len(set(si)) == W // 2 and not set(si) & set(ci)
            c, sx = coords[ci], coords[si]
            q = c - (c - sx) * r["zz"][t, sp][:, None]
            nl = f(q)
            acc = (r["factors"][t, sp] + nl - lp[si]) > r["logu"][t, sp]
            coords[si[acc]] = q[acc]
            lp[si[acc]] = nl[acc]
        chain.append(coords.copy())
    assert np.array_equal(np.array(chain), s.get_chain())


def test_philox_restatement_known_answer_vectors():
    """oracle/philox.py against the Random123 known-answer vectors for philox4x32-10 (zero, all-ones, digits of pi)."""
    from oracle import philox as ph

    def run(c, k):
        return [int(v) for v in ph.philox4x32_10(*[np.uint64(x) for x in c], *[np.uint64(x) for x in k])]

    assert run([0, 0, 0, 0], [0, 0]) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert run([0xffffffff] * 4, [0xffffffff] * 2) == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert run([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0]) == \
        [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]
    r = ph.stretch_randoms(seed=5, step0=3, nsteps=4, nwalkers=12, ndim=3)
    assert r["zz"].shape == (4, 2, 6) and np.all((r["zz"] >= 0.5) & (r["zz"] <= 2.0)) and np.all(r["logu"] < 0)
    assert np.array_equal(r["s_idx"][0, 0], np.arange(0, 12, 2)) and np.array_equal(r["s_idx"][0, 1], np.arange(1, 12, 2))
    assert np.all(r["c_idx"][:, 0] % 2 == 1) and np.all(r["c_idx"][:, 1] % 2 == 0)      # partners from the other half
    later = ph.stretch_randoms(seed=5, step0=5, nsteps=2, nwalkers=12, ndim=3)          # a step does not depend on the block
    assert all(np.array_equal(later[k], r[k][2:]) for k in r)


def test_initial_state_check_is_emcee_3_1_4():
    """The reference pins emcee 3.1.4 (pdm.lock:504-505): ``walkers_independent`` normalises every column before it asks
    for the condition number, so parameters of very different ranges (jet_substructure.yaml:127-137: 0.4 ... 100) and the
    tight cluster of the burn-in re-seed (mcmc.py:99) are accepted, while a truly degenerate ensemble is refused."""
    from bayesian_inference_b200.ensemble import check_initial_state, walkers_independent

    rng = np.random.RandomState(3)
    scales = np.array([1e-5, 0.4, 100.0, 1e6])
    wide = rng.uniform(0, 1, (40, 4)) * scales                     # raw covariance condition number ~1e22
    assert np.linalg.cond(np.cov(wide, rowvar=False)) > 1e8        # what emcee 3.0.x (and round 1 here) tested
    assert walkers_independent(wide)
    base = rng.uniform(0, 1, 4) * scales                           # a strongly correlated, tight cluster around one point
    z = rng.normal(size=(40, 1))
    cluster = base + (z + 1e-3 * rng.normal(size=(40, 4))) * 1e-6 * scales
    assert np.linalg.cond(np.cov(cluster, rowvar=False)) > 1e8
    assert walkers_independent(cluster)
    check_initial_state(cluster)
    s = StretchEnsemble(40, 4, lambda x: -0.5 * np.sum((x / scales) ** 2, axis=1), seed=1)
    s.run_mcmc(wide, 3)                                            # round 1's check raised here
    degenerate = np.repeat(rng.uniform(0, 1, (1, 4)), 40, axis=0)
    assert not walkers_independent(degenerate)                     # a column without spread
    collinear = rng.uniform(0, 1, (40, 1)) * np.ones((1, 4))
    assert not walkers_independent(collinear)                      # exactly dependent columns
    bad = wide.copy()
    bad[3, 1] = np.inf
    assert not walkers_independent(bad)
    with pytest.raises(ValueError, match="condition number"):
        check_initial_state(collinear)
