"""The reference-facing Python modules (emulation.predict, log_posterior.log_posterior) on real sklearn objects."""
from __future__ import annotations

import numpy as np
import pytest
from conftest import fro_err, load_golden, rel_err

from bayesian_inference_b200 import synthetic

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def toy():
    return synthetic.generate("TOY", oob_fraction=0.1)


def test_emulation_predict_signature_and_values(toy):
    from bayesian_inference_b200 import emulation

    _, ref = load_golden("toy")
    th = toy.theta[:24]
    out = emulation.predict(th, toy.emulation_config, emulation_group_results=toy.emulation_results)
    assert set(out) == {"central_value", "cov"}
    assert fro_err(out["central_value"], ref["ref_mean"]) < 1e-12 and fro_err(out["cov"], ref["ref_cov"]) < 1e-12
    per = emulation.predict(th, toy.emulation_config, merge_predictions_over_groups=False,
                            emulation_group_results=toy.emulation_results)
    assert list(per) == ["alpha", "beta"]
    for gi, g in enumerate(per):
        assert fro_err(per[g]["cov"], ref[f"ref_g{gi}_cov"]) < 1e-12
    gcfg = toy.emulation_config.emulation_groups_config["alpha"]
    one = emulation.predict_emulation_group(th, toy.emulation_results["alpha"], gcfg)
    assert fro_err(one["central_value"], ref["ref_g0_mean"]) < 1e-12 and fro_err(one["cov"], ref["ref_g0_cov"]) < 1e-12
    emulation.clear_cache()


def test_log_posterior_module_is_a_drop_in(toy):
    from bayesian_inference_b200 import log_posterior as lp

    _, ref = load_golden("toy")
    lp.initialize_pool_variables(toy.min, toy.max, toy.emulation_config, toy.emulation_results,
                                 toy.experimental_results, None)
    th = toy.theta[:24]
    batch = lp.log_posterior(th)
    fin = np.isfinite(ref["ref_logp_batch"])
    assert np.array_equal(np.isneginf(batch), ~fin) and rel_err(batch[fin], ref["ref_logp_batch"][fin]) < 1e-9
    rows = np.array([lp.log_posterior(x)[0] for x in th])        # emcee without vectorize: (d,) in, (1,) out
    assert rel_err(rows[fin], ref["ref_logp_rows"][fin]) < 1e-9
    assert np.array_equal(lp.log_posterior_walkers(th), rows)
    assert lp.log_posterior(list(th[0])).shape == (1,)           # list input (numpy>=2 breaks the reference here, F13)
    # _loglikelihood(y, cov) on the merged prediction reproduces the same number
    from bayesian_inference_b200 import emulation

    i = int(np.flatnonzero(fin)[0])
    pred = emulation.predict(th[i:i + 1], toy.emulation_config, emulation_group_results=toy.emulation_results)
    y, ye = toy.experimental_results["y"], toy.experimental_results["y_err"]
    val = lp._loglikelihood(pred["central_value"][0] - y, pred["cov"][0] + np.diag(ye ** 2))
    assert abs(val - ref["ref_logp_rows"][i]) < 1e-9 * abs(ref["ref_logp_rows"][i])
    lp.close()
    emulation.clear_cache()


def test_run_mcmc_device_and_host_samplers_agree(toy, tmp_path):
    """mcmc.run_mcmc (mcmc.py:34-134): burn-in -> re-seed -> production with the ensemble on the GPU gives exactly the
    chain of the host-side sampler calling the same log-posterior, and the reference's output keys."""
    import types

    from bayesian_inference_b200 import emulation, log_posterior as lp, mcmc

    d = len(toy.min)
    cfg = types.SimpleNamespace(
        analysis_config={"parameterization": {"p": {"names": [f"p{i}" for i in range(d)], "min": list(toy.min), "max": list(toy.max)}}},
        parameterization="p", analysis_name="toy", config_file="", output_dir=str(tmp_path), mcmc_output_dir=str(tmp_path),
        n_walkers=16, n_burn_steps=40, n_sampling_steps=60, n_logging_steps=50, sampler_outputfile=None)
    kw = dict(emulation_config=toy.emulation_config, emulation_results=toy.emulation_results,
              experimental_results=toy.experimental_results, seed=12, write=False)
    a = mcmc.run_mcmc(cfg, sampler="device", **kw)
    b = mcmc.run_mcmc(cfg, sampler="host", **kw)
    assert a["chain"].shape == (60, 16, d) and a["log_prob"].shape == (60, 16)
    assert np.array_equal(a["chain"], b["chain"]) and np.array_equal(a["log_prob"], b["log_prob"])
    assert np.array_equal(a["acceptance_fraction"], b["acceptance_fraction"])
    assert np.all(np.isfinite(a["log_prob"])) and a["n_log_prob_rows"] == b["n_log_prob_rows"]
    # device-side random numbers (not emcee's stream): reproducible, graph replay changes nothing, burn-in and production
    # do not reuse random numbers
    cfg.n_burn_steps, cfg.n_sampling_steps = 80, 130
    p1 = mcmc.run_mcmc(cfg, sampler="device_philox", **kw)
    p2 = mcmc.run_mcmc(cfg, sampler="device_philox_graph", **kw)
    assert p1["chain"].shape == (130, 16, d) and np.array_equal(p1["chain"], p2["chain"])
    assert not np.array_equal(p1["chain"][:60], a["chain"]) and np.all(np.isfinite(p1["log_prob"]))
    assert 0.1 < p1["acceptance_fraction"].mean() < 0.9
    lp.close()
    emulation.clear_cache()


def test_batched_closure_tests_equal_separate_runs(toy, tmp_path):
    """steer_analysis.py:168-183 runs one MCMC per validation point, each against its own pseudo-data: batching them
    (bi_set_data_sets / bi_log_posterior_sets, one call per half-step for all chains) gives exactly those chains."""
    import types

    from bayesian_inference_b200 import emulation, log_posterior as lp, mcmc

    d = len(toy.min)
    F = toy.experimental_results["y"].shape[0]
    rng = np.random.default_rng(4)
    pseudo = [{"y": toy.experimental_results["y"] * (1.0 + 0.05 * rng.normal(size=F)),
               "y_err": toy.experimental_results["y_err"] * rng.uniform(0.8, 1.5, size=F)} for _ in range(3)]
    cfg = types.SimpleNamespace(
        analysis_config={"parameterization": {"p": {"names": [f"p{i}" for i in range(d)], "min": list(toy.min), "max": list(toy.max)}}},
        parameterization="p", analysis_name="toy", config_file="", output_dir=str(tmp_path), mcmc_output_dir=str(tmp_path),
        n_walkers=16, n_burn_steps=30, n_sampling_steps=40, n_logging_steps=50, sampler_outputfile=None)
    th = toy.theta[:12]
    sid = np.arange(12) % 3
    for mode in ("dense", "structured"):
        batched = mcmc.run_mcmc_closures(cfg, pseudo, emulation_config=toy.emulation_config,
                                         emulation_results=toy.emulation_results, seeds=[11, 12, 13], mode=mode)
        got = lp.device().log_posterior_sets(th, sid, n_div=1.0, mode=mode)      # the handle that holds the three sets
        for i, p in enumerate(pseudo):
            single = mcmc.run_mcmc(cfg, closure_index=i, emulation_config=toy.emulation_config,
                                   emulation_results=toy.emulation_results, experimental_results=p, design_point=None,
                                   seed=11 + i, write=False, mode=mode)
            assert np.array_equal(batched[i]["chain"], single["chain"]), (mode, i)
            assert np.array_equal(batched[i]["log_prob"], single["log_prob"])
            assert np.array_equal(batched[i]["acceptance_fraction"], single["acceptance_fraction"])
        # ... and against the ORACLE: every row of the data-set call is the reference's log_posterior of that walker against
        # that pseudo-data set (data_IO.py:362-372 -> log_posterior.py:42-101), n_div = 1; likewise the log-probabilities the
        # batched chains carry at their last step
        from bayesian_inference_b200 import image
        from oracle import closed_form as cf

        for c, p in enumerate(pseudo):
            z_c = image.pack(toy.emulation_config, toy.emulation_results, min=toy.min, max=toy.max,
                             experimental_results=p).to_flat()
            want = cf.log_posterior(z_c, th[sid == c], n_div=1.0)
            fin = np.isfinite(want)
            assert np.array_equal(np.isneginf(got[sid == c]), ~fin), (mode, c)
            assert rel_err(got[sid == c][fin], want[fin]) < 1e-9, (mode, c)
            last_x, last_lp = batched[c]["chain"][-1], batched[c]["log_prob"][-1]
            assert rel_err(last_lp, cf.log_posterior(z_c, last_x, n_div=1.0)) < 1e-9, (mode, c)
        # the data-set entry point itself against one handle per data set, same mode: bit for bit
        for c, p in enumerate(pseudo):
            lp.initialize_pool_variables(toy.min, toy.max, toy.emulation_config, toy.emulation_results, p, None, mode=mode)
            want = lp.log_posterior_walkers(th[sid == c])
            assert np.array_equal(got[sid == c], want, equal_nan=True), (mode, c)
    lp.close()
    emulation.clear_cache()


def test_emulator_cov_unexplained_is_honoured(toy):
    """emulation.py:445-448 / :487-488: a non-empty dictionary replaces the truncated-PC covariance.  Expected values from
    the oracle on an image packed with the same dictionary (pinned to the reference in tests/test_host_logic.py) and,
    independently, from the default prediction corrected by (U' - U) s s^T / n_div in numpy."""
    from bayesian_inference_b200 import emulation, image, log_posterior as lp
    from oracle import closed_form as cf

    cfgs = toy.emulation_config.emulation_groups_config
    own = emulation.compute_emulator_cov_unexplained(toy.emulation_config, toy.emulation_results)
    other = {g: 2.5 * u + 0.02 * np.diag(np.arange(1, u.shape[0] + 1.0)) for g, u in own.items()}
    th = toy.theta[:10]
    same = emulation.predict(th, toy.emulation_config, emulation_group_results=toy.emulation_results, emulator_cov_unexplained=own)
    base = emulation.predict(th, toy.emulation_config, emulation_group_results=toy.emulation_results)
    assert np.array_equal(same["cov"], base["cov"])                  # the PCA's own matrix: the default image serves it
    got = emulation.predict(th, toy.emulation_config, emulation_group_results=toy.emulation_results, emulator_cov_unexplained=other)
    img = image.pack(toy.emulation_config, toy.emulation_results, min=toy.min, max=toy.max, emulator_cov_unexplained=other)
    mean, cov = cf.predict(img.to_flat(), th)
    assert fro_err(got["central_value"], mean) < 1e-12 and fro_err(got["cov"], cov) < 1e-12
    per = emulation.predict(th, toy.emulation_config, merge_predictions_over_groups=False,
                            emulation_group_results=toy.emulation_results, emulator_cov_unexplained=other)
    per0 = emulation.predict(th, toy.emulation_config, merge_predictions_over_groups=False,
                             emulation_group_results=toy.emulation_results)
    for g in per:
        s = toy.emulation_results[g]["PCA"]["scaler"].scale_
        want = per0[g]["cov"] + ((other[g] - own[g]) * np.outer(s, s) / th.shape[0])[None]
        assert fro_err(per[g]["cov"], want) < 1e-12
        one = emulation.predict_emulation_group(th, toy.emulation_results[g], cfgs[g], emulator_group_cov_unexplained=other[g])
        assert fro_err(one["cov"], want) < 1e-12
    # the same dictionary through initialize_pool_variables (log_posterior.py:80-82)
    lp.initialize_pool_variables(toy.min, toy.max, toy.emulation_config, toy.emulation_results, toy.experimental_results, other)
    z = image.pack(toy.emulation_config, toy.emulation_results, min=toy.min, max=toy.max,
                   experimental_results=toy.experimental_results, emulator_cov_unexplained=other).to_flat()
    want_lp = cf.log_posterior(z, th)
    got_lp = lp.log_posterior(th)
    fin = np.isfinite(want_lp)
    assert np.array_equal(np.isneginf(got_lp), ~fin) and rel_err(got_lp[fin], want_lp[fin]) < 1e-9
    lp.initialize_pool_variables(toy.min, toy.max, toy.emulation_config, toy.emulation_results, toy.experimental_results, None)
    assert rel_err(lp.log_posterior(th)[fin], want_lp[fin]) > 1e-6   # and it changes the answer
    lp.close()
    emulation.clear_cache()
