"""The numpy oracle (oracle/closed_form.py) against golden vectors produced by executing the reference
(tests/golden/make_golden.py), against the live reference when /root/reference is present, and against analytic
known-answer cases.  CPU only."""
from __future__ import annotations

import os

import numpy as np
import pytest
from conftest import fro_err, load_golden, rel_err

from oracle import closed_form as cf
from oracle import ref_harness as rh

# tolerances of SURVEY §8(c)
TOL_K, TOL_PC, TOL_OBS, TOL_LOGP = 1e-13, 1e-11, 1e-12, 1e-9


def _n_groups(z):
    return int(z["n_groups"])


def test_kernel_vectors_match_sklearn(golden):
    _, z, ref = golden
    for g in range(_n_groups(z)):
        assert rel_err(cf.kernel_vectors(z, g, ref["theta"]), ref[f"ref_kvec_g{g}"]) < TOL_K


def test_pc_mean_and_variance_match_sklearn(golden):
    _, z, ref = golden
    m = np.concatenate([cf.predict_pc(z, g, ref["theta"])[0] for g in range(_n_groups(z))], axis=1)
    v = np.concatenate([cf.predict_pc(z, g, ref["theta"])[1] for g in range(_n_groups(z))], axis=1)
    assert np.max(np.abs(m - ref["ref_pc_mean"])) < TOL_PC * max(1.0, np.max(np.abs(ref["ref_pc_mean"])))
    diag = np.concatenate([1.0 + z[f"g{g}_const"] + z[f"g{g}_noise"] for g in range(_n_groups(z))])
    assert np.max(np.abs(v - ref["ref_pc_var"]) / diag) < TOL_PC


def test_predict_matches_reference(golden):
    _, z, ref = golden
    mean, cov = cf.predict(z, ref["theta"])
    assert fro_err(mean, ref["ref_mean"]) < TOL_OBS
    assert fro_err(cov, ref["ref_cov"]) < TOL_OBS
    for g in range(_n_groups(z)):
        mg, cg = cf.predict_group(z, g, ref["theta"])
        assert fro_err(mg, ref[f"ref_g{g}_mean"]) < TOL_OBS
        assert fro_err(cg, ref[f"ref_g{g}_cov"]) < TOL_OBS


def test_merged_covariance_is_block_diagonal_per_observable(golden):
    _, z, ref = golden
    mask = np.zeros((int(z["n_features"]),) * 2, dtype=bool)
    for _, so, _, ln in np.asarray(z["blocks"]).reshape(-1, 4):
        mask[so:so + ln, so:so + ln] = True
    assert np.all(ref["ref_cov"][:, ~mask] == 0.0)      # SURVEY F2, on the reference's own output


def test_log_posterior_matches_reference(golden):
    _, z, ref = golden
    for n_div, key in ((None, "ref_logp_batch"), (1.0, "ref_logp_rows")):
        lp = cf.log_posterior(z, ref["theta"], n_div=n_div)
        want = ref[key]
        assert np.array_equal(np.isinf(lp), np.isinf(want))
        fin = np.isfinite(want)
        assert rel_err(lp[fin], want[fin]) < TOL_LOGP


def test_batch_divisor_changes_the_value():
    z, ref = load_golden("toy")
    fin = np.isfinite(ref["ref_logp_rows"])
    assert np.max(np.abs(ref["ref_logp_rows"][fin] - ref["ref_logp_batch"][fin])) > 1e-6   # SURVEY F3 is real


def test_blockwise_equals_dense():
    for name in ("toy", "mid"):
        z, ref = load_golden(name)
        a = cf.log_posterior(z, ref["theta"], n_div=1.0)
        b = cf.log_posterior_blockwise(z, ref["theta"], n_div=1.0)
        fin = np.isfinite(a)
        assert np.array_equal(np.isinf(a), np.isinf(b))
        assert rel_err(b[fin], a[fin]) < TOL_LOGP


def test_strict_prior_bounds():
    z, ref = load_golden("toy")
    th = np.tile(0.5 * (z["min"] + z["max"]), (4, 1))
    th[1, 0] = z["max"][0]          # on the boundary: outside (strict inequality, log_posterior.py:63)
    th[2, 1] = z["min"][1]
    th[3, 2] = z["max"][2] + 1.0
    lp = cf.log_posterior(z, th, n_div=1.0)
    assert np.isfinite(lp[0]) and np.all(np.isneginf(lp[1:]))


# ---- analytic known-answer tests --------------------------------------------------------------------------------
def _tiny_image(P=2, N=6, F=3, d=2, full_rank=False, seed=0):
    """Hand-built image: RBF kernels, orthonormal S."""
    from scipy.linalg import cholesky

    rng = np.random.default_rng(seed)
    X = rng.uniform(0, 1, (N, d))
    ell = rng.uniform(0.5, 2, (P, d))
    noise = np.full(P, 1e-6)
    Q = F if full_rank else F
    comp = np.linalg.qr(rng.normal(size=(F, F)))[0].T          # rows orthonormal
    yv = rng.normal(size=(P, N))
    L = np.zeros((P, N, N))
    alpha = np.zeros((P, N))
    for p in range(P):
        a, b = X / ell[p], X / ell[p]
        K = np.exp(-0.5 * ((a[:, None, :] - b[None, :, :]) ** 2).sum(-1)) + (noise[p] + 1e-10) * np.eye(N)
        L[p] = cholesky(K, lower=True)
        alpha[p] = np.linalg.solve(K, yv[p])
    z = {
        "d": np.int64(d), "min": np.zeros(d), "max": np.ones(d), "n_groups": np.int64(1), "n_features": np.int64(F),
        "blocks": np.array([[0, 0, 0, F]]), "group_names": np.array(["g"]),
        "g0_X": X, "g0_kind": np.full(P, 3, np.int32), "g0_ell": ell, "g0_const": np.zeros(P), "g0_noise": noise,
        "g0_alpha": alpha, "g0_L": L, "g0_S": comp[:P], "g0_S_rest": comp[P:Q], "g0_ev_rest": np.full(Q - P, 0.25),
        "g0_scale": np.full(F, 2.0), "g0_mean": np.full(F, 1.0), "y": np.zeros(F), "y_err": np.full(F, 0.5),
    }
    return z, yv


def test_kat_interpolates_training_targets_at_design_points():
    z, yv = _tiny_image()
    m, v = cf.predict_pc(z, 0, z["g0_X"])
    assert np.max(np.abs(m - yv.T)) < 1e-4           # eta = 1e-6: GP mean reproduces its training targets
    assert np.max(v) < 1e-5                          # and the predictive variance collapses


def test_kat_full_rank_pca_has_no_unexplained_covariance_and_orthonormal_s_gives_diagonal():
    z, _ = _tiny_image(P=3, F=3)
    assert cf.cov_unexplained(z, 0).shape == (3, 3) and np.all(cf.cov_unexplained(z, 0) == 0.0)
    th = np.array([[0.3, 0.6]])
    m, v = cf.predict_pc(z, 0, th)
    _, cov = cf.predict_group(z, 0, th)
    # S orthonormal and square: cov = s^2 * S^T diag(v) S has eigenvalues s^2 v
    w = np.sort(np.linalg.eigvalsh(cov[0]))
    assert np.allclose(w, np.sort(4.0 * v[0]), rtol=1e-12, atol=1e-18)


def test_kat_single_bin_observables_give_diagonal_likelihood():
    z, _ = _tiny_image(P=2, F=3)
    z["blocks"] = np.array([[0, 0, 0, 1], [0, 1, 1, 1], [0, 2, 2, 1]])
    th = np.array([[0.2, 0.7], [0.8, 0.1]])
    mean, cov = cf.predict(z, th, n_div=1.0)
    lp = cf.log_posterior(z, th, n_div=1.0)
    for i in range(2):
        c = np.diag(cov[i]) + z["y_err"] ** 2
        assert np.count_nonzero(cov[i] - np.diag(np.diag(cov[i]))) == 0
        want = -0.5 * np.sum((mean[i] - z["y"]) ** 2 / c) - 0.5 * np.sum(np.log(c))
        assert abs(lp[i] - want) < 1e-12 * abs(want)


def test_non_positive_definite_gives_nan():
    bad = np.array([[1.0, 2.0], [2.0, 1.0]])
    assert np.isnan(cf.log_likelihood(np.ones(2), bad))


# ---- the live reference, when this is the build container ---------------------------------------------------------
@pytest.mark.skipif(not rh.available(), reason="reference checkout not present (GPU box)")
def test_oracle_against_live_reference():
    from bayesian_inference_b200 import image, synthetic

    a = synthetic.generate("TOY", oob_fraction=0.1)
    img = image.pack(a.emulation_config, a.emulation_results, min=a.min, max=a.max,
                     experimental_results=a.experimental_results)
    z = img.to_flat()
    pred = rh.reference_predict(a, a.theta)
    mean, cov = cf.predict(z, a.theta)
    assert fro_err(mean, pred["central_value"]) < TOL_OBS and fro_err(cov, pred["cov"]) < TOL_OBS
    want = rh.reference_log_posterior(a, a.theta, per_row=True)
    got = cf.log_posterior(z, a.theta, n_div=1.0)
    fin = np.isfinite(want)
    assert np.array_equal(np.isinf(got), np.isinf(want)) and rel_err(got[fin], want[fin]) < TOL_LOGP
    # and the committed fixture is what the generator produces today
    zg, ref = load_golden("toy")
    assert np.array_equal(zg["g0_L"], z["g0_L"]) and np.allclose(ref["ref_logp_rows"][fin[:24]], want[:24][fin[:24]], rtol=1e-12)


def test_reference_pool_runs_the_reference_itself_from_baseline_ref():
    """bench.py's CPU arm: the reference's unmodified modules in baseline/_ref (oracle/make_ref.py), spawn pool initialised
    with the reference's initialize_pool_variables, one theta per task -> the golden per-row values."""
    from bayesian_inference_b200 import synthetic
    from oracle import make_ref, ref_harness as rh
    from oracle.ref_pool import ReferencePool

    if not make_ref.make():
        pytest.skip("no reference checkout and no baseline/_ref")
    old = rh.REFERENCE_SRC
    try:
        assert rh.use_baseline_ref()
        a = synthetic.generate("TOY", oob_fraction=0.1)
        _, ref = load_golden("toy")
        with ReferencePool(a, 2) as pool:
            got = pool.map_rows(a.theta[:10])
            one = pool.one_chunk(a.theta[:3])
        want = ref["ref_logp_rows"][:10]
        assert np.array_equal(np.isneginf(got), np.isneginf(want))
        fin = np.isfinite(want)
        assert np.max(np.abs(got[fin] - want[fin]) / np.abs(want[fin])) < 1e-12
        assert np.array_equal(one, got[:3], equal_nan=True)
    finally:
        rh.REFERENCE_SRC = old
    sums = open(os.path.join(rh.BASELINE_REF, "bayesian_inference", "SHA256SUMS")).read().split()
    assert "log_posterior.py" in sums and "emulation.py" in sums
