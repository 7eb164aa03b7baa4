"""Guard mode (BI_DEBUG_GUARD=1): canary bands around every device buffer of a handle and NaN-poisoned payloads.

compute-sanitizer is closed on this pool ("find a bad access with bounds checks and asserts of your own, small cases, and a
comparison with the CPU reference"), so this is the memcheck / initcheck stand-in the test-suite carries: every kernel of
the log-posterior path runs between canaries, results must equal the unguarded run bit for bit (a read of never-written
scratch would surface as NaN), agree with the oracle, and no canary band may change.  Plus a repetition stress test of the
hand-rolled mbarrier / bulk-copy protocol of the fused Cholesky kernel at wave boundaries (W = resident CTAs +- 1)."""
from __future__ import annotations

import numpy as np
import pytest
from conftest import load_golden, rel_err

from bayesian_inference_b200 import image, synthetic
from oracle import closed_form as cf

pytestmark = pytest.mark.gpu


def _run_everything(img, theta, with_sets=True):
    from bayesian_inference_b200.device import DeviceEmulator
    from bayesian_inference_b200.ensemble import DevicePhiloxEnsemble, DeviceStretchEnsemble

    out = {}
    with DeviceEmulator(img, max_batch=16) as dev:            # small max_batch: several chunks per call
        for mode in ("dense", "structured", "woodbury"):
            out[mode] = dev.log_posterior(theta, n_div=1.0, mode=mode)
            out[mode + "_nd"] = dev.log_posterior(theta, n_div=None, mode=mode)
        out["m"], out["v"] = dev.predict_pc(theta)
        out["mean"], out["cov"] = dev.predict(theta[:5])
        out["gmean"], out["gcov"] = dev.predict_group(0, theta[:5])
        out["k"] = dev.kernel_vectors(0, theta[:5])
        F = img.n_features
        rng = np.random.RandomState(0)
        A = rng.normal(size=(2, F, F))
        out["ll"] = dev.loglikelihood(rng.normal(size=(2, F)), A @ A.transpose(0, 2, 1) + F * np.eye(F))
        if with_sets:
            dev.set_data_sets(np.stack([img.y, img.y * 1.01]), np.stack([img.y_err, img.y_err * 1.1]))
            n = min(9, theta.shape[0])
            for mode in ("dense", "structured"):
                out["sets_" + mode] = dev.log_posterior_sets(theta[:n], np.arange(n) % 2, mode=mode)
        W, d = 16, img.d
        x0 = np.random.RandomState(1).uniform(0.3, 0.7, (W, d))
        s = DeviceStretchEnsemble(W, d, dev, seed=3, block=4, graph=True)
        s.run_mcmc(x0, 10)
        out["chain"] = s.get_chain()
        p = DevicePhiloxEnsemble(W, d, dev, seed=3, block=4)
        p.run_mcmc(x0, 6)
        out["pchain"] = p.get_chain()
        out["guards"] = dev.query(32)
    return out


@pytest.mark.parametrize("case", ["toy", "mid", "align"])
def test_every_kernel_between_canaries(case, monkeypatch):
    if case == "align":                    # F = 70: three tiles, ragged observable blocks, walkers outside the prior
        spec = synthetic.ProblemSpec("guard70", d=3, n_design=40, n_walkers=24, seed=5, oob_fraction=0.1, groups=[
            synthetic.GroupSpec("alpha", 4, 5, 7, constant=True), synthetic.GroupSpec("beta", 3, 5, 7, kernel="rbf")])
        a = synthetic.generate(spec)
        img = image.pack(a.emulation_config, a.emulation_results, min=a.min, max=a.max,
                         experimental_results=a.experimental_results)
        theta = a.theta
    else:
        z, ref = load_golden(case)
        img, theta = image.EmulatorImage.from_flat(z), ref["theta"]
    plain = _run_everything(img, theta)
    assert plain["guards"] == -1
    monkeypatch.setenv("BI_DEBUG_GUARD", "1")
    guarded = _run_everything(img, theta)
    assert guarded["guards"] == 0, "a kernel wrote outside one of the handle's buffers"
    for k in plain:
        if k != "guards":
            assert np.array_equal(plain[k], guarded[k], equal_nan=True), k       # nothing read before it was written
    want = cf.log_posterior(img.to_flat(), theta, n_div=1.0)
    fin = np.isfinite(want)
    for mode in ("dense", "structured", "woodbury"):
        assert np.array_equal(np.isneginf(guarded[mode]), np.isneginf(want))
        assert rel_err(guarded[mode][fin], want[fin]) < 1e-9


def test_wave_boundaries_repeated_and_guarded(monkeypatch):
    """C3a shapes (19 tiles, the 4-stage bulk-copy ring sweeps up to 36 half-chunks per round): 60 repetitions each at
    W = resident CTAs - 1, resident CTAs + 1 and one walker per SM + 1 (main + wide-tail launches), guard mode on; every
    repetition must reproduce the first bit for bit, match the oracle on a sample, and leave the canaries intact."""
    from bayesian_inference_b200.device import DeviceEmulator

    monkeypatch.setenv("BI_DEBUG_GUARD", "1")
    a = synthetic.generate("C3A", n_walkers=512, oob_fraction=0.01)
    img = image.pack(a.emulation_config, a.emulation_results, min=a.min, max=a.max,
                     experimental_results=a.experimental_results)
    with DeviceEmulator(img, max_batch=512) as dev:
        slots, sms = dev.query(5), dev.query(4)
        for W in (slots - 1, slots + 1, sms + 1, 1):
            th = a.theta[:W]
            first = dev.log_posterior(th, n_div=1.0, mode="dense")
            for _ in range(60 if W > 1 else 5):
                assert np.array_equal(dev.log_posterior(th, n_div=1.0, mode="dense"), first, equal_nan=True), W
            idx = np.flatnonzero(np.isfinite(first))[:6]
            assert rel_err(first[idx], cf.log_posterior(img.to_flat(), th[idx], n_div=1.0)) < 1e-9
            s1 = dev.log_posterior(th, n_div=1.0, mode="structured")
            assert np.array_equal(dev.log_posterior(th, n_div=1.0, mode="structured"), s1, equal_nan=True)
        assert dev.query(32) == 0
