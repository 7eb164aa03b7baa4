"""Parity of the CUDA path (through the C ABI) with the reference's golden vectors and with the numpy oracle.

Tolerances (SURVEY §8c / BASELINE north_star): kernel vector <= 1e-13 rel, PC mean <= 1e-11 (scaled), PC variance
<= 1e-11 * kernel diag abs, observable mean / covariance <= 1e-12 rel (Frobenius), log-posterior <= 1e-9 rel."""
from __future__ import annotations

import numpy as np
import pytest
from conftest import fro_err, load_golden, rel_err

from bayesian_inference_b200 import image, synthetic
from oracle import closed_form as cf

pytestmark = pytest.mark.gpu

TOL_K, TOL_PC, TOL_OBS, TOL_LOGP = 1e-13, 1e-11, 1e-12, 1e-9


def _dev(z):
    from bayesian_inference_b200.device import DeviceEmulator

    return DeviceEmulator(image.EmulatorImage.from_flat(z))


def _cmp_logp(got, want, tol=TOL_LOGP):
    assert np.array_equal(np.isneginf(got), np.isneginf(want))
    fin = np.isfinite(want)
    assert np.all(np.isfinite(got[fin]))
    assert rel_err(got[fin], want[fin]) < tol


# ---- golden fixtures (outputs of the reference itself) ------------------------------------------------------------
def test_golden_kernel_vectors(golden):
    _, z, ref = golden
    with _dev(z) as dev:
        for g in range(int(z["n_groups"])):
            assert rel_err(dev.kernel_vectors(g, ref["theta"]), ref[f"ref_kvec_g{g}"]) < TOL_K


def test_golden_pc_mean_variance(golden):
    _, z, ref = golden
    with _dev(z) as dev:
        m, v = dev.predict_pc(ref["theta"])
    assert np.max(np.abs(m - ref["ref_pc_mean"])) < TOL_PC * max(1.0, np.max(np.abs(ref["ref_pc_mean"])))
    diag = np.concatenate([1.0 + z[f"g{g}_const"] + z[f"g{g}_noise"] for g in range(int(z["n_groups"]))])
    assert np.max(np.abs(v - ref["ref_pc_var"]) / diag) < TOL_PC
    assert np.all(v >= 0.0)


def test_golden_predict(golden):
    _, z, ref = golden
    with _dev(z) as dev:
        mean, cov = dev.predict(ref["theta"])                       # divisor = rows in the call, like the reference
        assert fro_err(mean, ref["ref_mean"]) < TOL_OBS
        assert fro_err(cov, ref["ref_cov"]) < TOL_OBS
        assert np.array_equal(cov == 0.0, ref["ref_cov"] == 0.0)    # identical block-diagonal sparsity (F2)
        for g in range(int(z["n_groups"])):
            mg, cg = dev.predict_group(g, ref["theta"])
            assert fro_err(mg, ref[f"ref_g{g}_mean"]) < TOL_OBS
            assert fro_err(cg, ref[f"ref_g{g}_cov"]) < TOL_OBS


def test_golden_log_posterior(golden):
    name, z, ref = golden
    with _dev(z) as dev:
        _cmp_logp(dev.log_posterior(ref["theta"], n_div=None), ref["ref_logp_batch"])
        _cmp_logp(dev.log_posterior(ref["theta"], n_div=1.0), ref["ref_logp_rows"])
        rows = np.array([dev.log_posterior(t, n_div=None)[0] for t in ref["theta"]])      # how emcee drives it
        _cmp_logp(rows, ref["ref_logp_rows"])
        if name != "toy_dense_sigma":
            _cmp_logp(dev.log_posterior(ref["theta"], n_div=1.0, mode="structured"), ref["ref_logp_rows"])
            _cmp_logp(dev.log_posterior(ref["theta"], n_div=1.0, mode="woodbury"), ref["ref_logp_rows"])
            _cmp_logp(dev.log_posterior(ref["theta"], n_div=None, mode="woodbury"), ref["ref_logp_batch"])


def test_structured_mode_refused_when_sigma_couples_observables():
    from bayesian_inference_b200.device import DeviceError

    z, ref = load_golden("toy_dense_sigma")
    F = int(z["n_features"])
    z = dict(z)
    z["sigma_exp"] = z["sigma_exp"] + 1e-3 * np.min(np.diag(z["sigma_exp"])) * np.ones((F, F))
    with _dev(z) as dev:
        with pytest.raises(DeviceError):
            dev.log_posterior(ref["theta"], mode="structured")
        with pytest.raises(DeviceError):
            dev.log_posterior(ref["theta"], mode="woodbury")
        want = cf.log_posterior(z, ref["theta"], n_div=1.0)
        _cmp_logp(dev.log_posterior(ref["theta"], mode="dense"), want)


# ---- generated analyses at the BASELINE shapes vs the oracle ------------------------------------------------------
@pytest.mark.parametrize("name,n_check,over", [
    ("C1", 64, {}),
    ("C3A", 64, {}),
    ("C3B", 64, {}),
    ("C3A", 16, {"dense_sigma_exp": True}),
    ("C4", 32, {"oob_fraction": 0.2}),
])
def test_generated_config_vs_oracle(name, n_check, over):
    a = synthetic.generate(name, **over)
    img = image.pack(a.emulation_config, a.emulation_results, min=a.min, max=a.max,
                     experimental_results=a.experimental_results, sigma_exp=a.sigma_exp)
    z = img.to_flat()
    th = a.theta[:n_check]
    from bayesian_inference_b200.device import DeviceEmulator

    with DeviceEmulator(img) as dev:
        m, v = dev.predict_pc(th)
        mo, vo = [np.concatenate(x, axis=1) for x in zip(*[cf.predict_pc(z, g, th) for g in range(len(img.groups))])]
        assert np.max(np.abs(m - mo)) < TOL_PC * max(1.0, np.max(np.abs(mo)))
        assert np.max(np.abs(v - vo)) < TOL_PC * 4.0
        mean, cov = dev.predict(th[:4], n_div=1.0)
        mo_, co_ = cf.predict(z, th[:4], n_div=1.0)
        assert fro_err(mean, mo_) < TOL_OBS and fro_err(cov, co_) < TOL_OBS
        want = cf.log_posterior(z, th, n_div=1.0)
        _cmp_logp(dev.log_posterior(th, n_div=1.0), want)
        if a.sigma_exp is None:
            _cmp_logp(dev.log_posterior(th, n_div=1.0, mode="structured"), want)
            _cmp_logp(dev.log_posterior(th, n_div=1.0, mode="woodbury"), want)


# ---- every alignment of the merged axis against the 32-row tiles (the dY row is row F of the factorised matrix) -----
@pytest.mark.parametrize("n_obs,n_bins", [(1, 31), (2, 16), (3, 11), (7, 9), (4, 16), (3, 32), (1, 127), (5, 32), (9, 19)])
def test_log_posterior_every_tile_alignment(n_obs, n_bins):
    """F = 31, 32, 33, 63, 64, 96, 127, 160, 171: F % 32 == 0 adds a tile that holds only the augmented row, F % 32 == 31
    puts it in the last row of the last tile, odd F exercises the unaligned covariance stores of predict()."""
    from bayesian_inference_b200.device import DeviceEmulator

    spec = synthetic.ProblemSpec(f"align{n_obs}x{n_bins}", d=3, n_design=40, n_walkers=24, seed=n_obs * 100 + n_bins,
                                 groups=[synthetic.GroupSpec("g0", 5, n_obs, n_bins)], oob_fraction=0.1)
    a = synthetic.generate(spec)
    img = image.pack(a.emulation_config, a.emulation_results, min=a.min, max=a.max,
                     experimental_results=a.experimental_results)
    z = img.to_flat()
    want = cf.log_posterior(z, a.theta, n_div=1.0)
    with DeviceEmulator(img) as dev:
        for mode in ("dense", "structured", "woodbury"):
            _cmp_logp(dev.log_posterior(a.theta, n_div=1.0, mode=mode), want)
        mean, cov = dev.predict(a.theta[:5], n_div=1.0)
        mo, co = cf.predict(z, a.theta[:5], n_div=1.0)
        assert fro_err(mean, mo) < TOL_OBS and fro_err(cov, co) < TOL_OBS
        assert np.array_equal(cov == 0.0, co == 0.0)


# ---- structured mode: block by block in registers (f_b <= 31) vs the tile-skipping kernel vs dense vs the oracle --------
@pytest.mark.parametrize("name,W,over", [("C4", 301, {"oob_fraction": 0.1}), ("C1", 100, {}), ("C3A", 257, {"oob_fraction": 0.02})])
def test_structured_block_direct_equals_tile_skipping_dense_and_oracle(name, W, over, monkeypatch):
    """SURVEY Appendix A (structured variant, f_b <= P: direct f_b x f_b Cholesky per observable block).  The block-direct
    kernel, the fused kernel with tile skipping (forced by BI_STRUCT_IMPL=tiles) and the dense factorisation give the same
    log-posterior; 24 rows are compared with the oracle's block-wise restatement and with its dense evaluation."""
    from bayesian_inference_b200.device import DeviceEmulator

    a = synthetic.generate(name, n_walkers=W, **over)
    img = image.pack(a.emulation_config, a.emulation_results, min=a.min, max=a.max,
                     experimental_results=a.experimental_results)
    z = img.to_flat()
    with DeviceEmulator(img, max_batch=128) as dev:                      # several chunks, the last one ragged
        assert dev.query(18) == 1
        direct = dev.log_posterior(a.theta, n_div=1.0, mode="structured")
        direct_nd, info = dev.log_posterior(a.theta, n_div=None, mode="structured", return_info=True)
        dense = dev.log_posterior(a.theta, n_div=1.0, mode="dense")
        dense_nd = dev.log_posterior(a.theta, n_div=None, mode="dense")
        one = np.array([dev.log_posterior(a.theta[i], n_div=1.0, mode="structured")[0] for i in (0, 5, W - 1)])
        again = dev.log_posterior(a.theta, n_div=1.0, mode="structured")
    monkeypatch.setenv("BI_STRUCT_IMPL", "tiles")
    with DeviceEmulator(img, max_batch=128) as dev2:
        assert dev2.query(18) == 0
        tiles = dev2.log_posterior(a.theta, n_div=1.0, mode="structured")
    monkeypatch.delenv("BI_STRUCT_IMPL")
    monkeypatch.setenv("BI_BD_AUGIN", "0")                                # the augmented row in a fragment row of its own
    with DeviceEmulator(img, max_batch=128) as dev3:
        assert dev3.query(18) == 1
        own_row = dev3.log_posterior(a.theta, n_div=1.0, mode="structured")
    assert np.array_equal(np.isneginf(own_row), np.isneginf(direct))
    fin = np.isfinite(direct)
    assert rel_err(own_row[fin], direct[fin]) < 1e-13                    # same products, same elimination of the real columns
    assert np.array_equal(direct, again, equal_nan=True)                 # deterministic (block terms summed in block order)
    assert np.array_equal(one, direct[[0, 5, W - 1]])                    # a walker's value does not depend on its batch
    inside = np.all((a.theta > a.min) & (a.theta < a.max), axis=1)
    assert np.array_equal(np.isneginf(direct), ~inside) and np.array_equal(info == -1, ~inside) and np.all(info[inside] == 0)
    for got, ref in ((direct, dense), (tiles, dense), (direct_nd, dense_nd)):
        assert np.array_equal(np.isneginf(got), np.isneginf(ref))
        assert rel_err(got[inside], ref[inside]) < TOL_LOGP
    idx = np.flatnonzero(inside)[:24]
    _cmp_logp(direct[idx], cf.log_posterior_blockwise(z, a.theta[idx], n_div=1.0))
    _cmp_logp(direct[idx], cf.log_posterior(z, a.theta[idx], n_div=1.0))


def test_structured_block_direct_reports_a_non_positive_definite_block():
    """A covariance that is not positive definite gives NaN and info = the 1-based merged row of the first bad pivot, as on
    the dense path (the reference ends in NaN too)."""
    z, ref = load_golden("toy")
    z = dict(z)
    z["y_err"] = np.zeros_like(z["y_err"])
    F = int(z["n_features"])
    bad = -5.0 * np.eye(F)
    z["sigma_exp"] = bad                                                  # block diagonal (diagonal) but negative
    with _dev(z) as dev:
        assert dev.query(18) == 1
        th = ref["theta"][np.all((ref["theta"] > z["min"]) & (ref["theta"] < z["max"]), axis=1)][:4]
        s, info_s = dev.log_posterior(th, n_div=1.0, mode="structured", return_info=True)
        d, info_d = dev.log_posterior(th, n_div=1.0, mode="dense", return_info=True)
    assert np.all(np.isnan(s)) and np.all(np.isnan(d))
    assert np.all(info_s >= 1) and np.array_equal(info_s, info_d)


# ---- size-independent properties at the full BASELINE size ----------------------------------------------------------
@pytest.fixture(scope="module")
def c3a():
    from bayesian_inference_b200.device import DeviceEmulator

    a = synthetic.generate("C3A", oob_fraction=0.01)
    img = image.pack(a.emulation_config, a.emulation_results, min=a.min, max=a.max,
                     experimental_results=a.experimental_results)
    dev = DeviceEmulator(img)
    yield a, img, dev
    dev.close()


def test_full_size_batch_equals_rows_and_is_deterministic(c3a):
    a, _, dev = c3a
    full = dev.log_posterior(a.theta, n_div=1.0)                         # 1024 walkers, one call
    again = dev.log_posterior(a.theta, n_div=1.0)
    assert np.array_equal(full, again, equal_nan=True)                   # bit-reproducible
    idx = np.arange(0, a.theta.shape[0], 97)
    rows = np.array([dev.log_posterior(a.theta[i], n_div=1.0)[0] for i in idx])
    assert np.array_equal(rows, full[idx])                               # a walker's value does not depend on its batch
    perm = np.random.default_rng(0).permutation(a.theta.shape[0])
    assert np.array_equal(dev.log_posterior(a.theta[perm], n_div=1.0), full[perm])
    assert np.count_nonzero(np.isneginf(full)) == np.count_nonzero(~np.all((a.theta > a.min) & (a.theta < a.max), axis=1))


def test_full_size_dense_equals_structured_equals_woodbury(c3a):
    a, _, dev = c3a
    d = dev.log_posterior(a.theta, n_div=1.0, mode="dense")
    fin = np.isfinite(d)
    for mode in ("structured", "woodbury"):
        s = dev.log_posterior(a.theta, n_div=1.0, mode=mode)
        assert np.array_equal(np.isneginf(d), np.isneginf(s)) and rel_err(s[fin], d[fin]) < TOL_LOGP


def test_full_size_torch_device_path_matches_host_path(c3a):
    import torch

    a, _, dev = c3a
    th = torch.from_numpy(a.theta).cuda()
    out = dev.log_posterior_tensor(th, n_div=1.0)
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), dev.log_posterior(a.theta, n_div=1.0))


def test_largest_configuration_c5_modes_agree_and_match_oracle():
    """BASELINE configs[4] shapes (n_design=500, n_pc=50, n_obs=2000): oracle on 8 rows, and dense == structured ==
    Woodbury on 96 walkers (the factor scratch of one CTA is 16 MB here, the operand ring sweeps 62 chunks)."""
    from bayesian_inference_b200.device import DeviceEmulator

    a = synthetic.generate("C5A", n_walkers=96, oob_fraction=0.05)
    img = image.pack(a.emulation_config, a.emulation_results, min=a.min, max=a.max,
                     experimental_results=a.experimental_results)
    with DeviceEmulator(img, max_batch=96) as dev:
        d = dev.log_posterior(a.theta, n_div=1.0, mode="dense")
        fin = np.isfinite(d)
        assert fin.sum() >= 80
        for mode in ("structured", "woodbury"):
            s_ = dev.log_posterior(a.theta, n_div=1.0, mode=mode)
            assert np.array_equal(np.isneginf(d), np.isneginf(s_)) and rel_err(s_[fin], d[fin]) < TOL_LOGP
        idx = np.flatnonzero(fin)[:8]
        want = cf.log_posterior(img.to_flat(), a.theta[idx], n_div=1.0)
        assert rel_err(d[idx], want) < TOL_LOGP


# ---- edge cases ---------------------------------------------------------------------------------------------------
def test_edge_cases_empty_single_all_outside():
    z, ref = load_golden("toy")
    with _dev(z) as dev:
        assert dev.log_posterior(np.empty((0, int(z["d"])))).shape == (0,)
        one = dev.log_posterior(ref["theta"][0])
        assert one.shape == (1,) and np.isfinite(one[0])
        out = dev.log_posterior(np.full((5, int(z["d"])), 2.0), n_div=None)
        assert np.all(np.isneginf(out))
        th = np.tile(0.5 * (z["min"] + z["max"]), (3, 1))
        th[1, 0] = z["max"][0]            # exactly on the boundary -> outside
        th[2, 0] = np.nan                 # NaN compares false -> outside, as in the reference
        lp, info = dev.log_posterior(th, return_info=True)
        assert np.isfinite(lp[0]) and np.isneginf(lp[1]) and np.isneginf(lp[2])
        assert list(info) == [0, -1, -1]
        m, c = dev.predict(np.empty((0, int(z["d"]))))
        assert m.shape == (0, int(z["n_features"])) and c.shape[0] == 0


@pytest.mark.parametrize("F", [1, 5, 31, 32, 33, 64, 100, 257])
def test_explicit_loglikelihood_matches_lapack(F):
    z, _ = load_golden("toy")
    rng = np.random.default_rng(F)
    n = 5
    A = rng.normal(size=(n, F, F))
    cov = A @ np.swapaxes(A, 1, 2) + F * np.eye(F)
    dy = rng.normal(size=(n, F))
    with _dev(z) as dev:
        got, info = dev.loglikelihood(dy, cov, return_info=True)
    want = np.array([cf.log_likelihood(dy[i], cov[i]) for i in range(n)])
    assert np.all(info == 0) and rel_err(got, want) < 1e-11


def test_not_positive_definite_reports_pivot_and_nan():
    z, _ = load_golden("toy")
    F = 40
    cov = np.eye(F)[None].repeat(2, 0)
    cov[1, 36, 36] = -1.0
    with _dev(z) as dev:
        got, info = dev.loglikelihood(np.ones((2, F)), cov, return_info=True)
    assert np.isfinite(got[0]) and np.isnan(got[1]) and info[0] == 0 and info[1] == 37


# ---- analytic known-answer tests on the CUDA path itself: answers known in closed form, no oracle -------
def test_kat_cuda_path_closed_form_answers():
    from test_oracle_golden import _tiny_image

    from bayesian_inference_b200.device import DeviceEmulator

    # (1) theta = a design point with eta -> 0: the GP mean reproduces its training targets, the variance collapses
    z, yv = _tiny_image()
    with DeviceEmulator(image.EmulatorImage.from_flat(z)) as dev:
        m, v = dev.predict_pc(z["g0_X"])
    assert np.max(np.abs(m - yv.T)) < 1e-4 and np.max(v) < 1e-5
    # (2) P = Q (nothing truncated) and S square orthonormal: cov = s^2 S^T diag(v) S has eigenvalues s^2 v
    z, _ = _tiny_image(P=3, F=3)
    th = np.array([[0.3, 0.6]])
    with DeviceEmulator(image.EmulatorImage.from_flat(z)) as dev:
        m, v = dev.predict_pc(th)
        _, cov = dev.predict(th, n_div=1.0)
    assert np.allclose(np.sort(np.linalg.eigvalsh(cov[0])), np.sort(4.0 * v[0]), rtol=1e-11, atol=1e-18)
    # (3) one-bin observables: the merged covariance is diagonal and the likelihood is a sum of 1-d Gaussians
    z, _ = _tiny_image(P=2, F=3)
    z["blocks"] = np.array([[0, 0, 0, 1], [0, 1, 1, 1], [0, 2, 2, 1]])
    th = np.array([[0.2, 0.7], [0.8, 0.1], [1.2, 0.5]])          # the last row is outside the prior box
    with DeviceEmulator(image.EmulatorImage.from_flat(z)) as dev:
        mean, cov = dev.predict(th[:2], n_div=1.0)
        lps = {mode: dev.log_posterior(th, n_div=1.0, mode=mode) for mode in ("dense", "structured", "woodbury")}
    for mode, lp in lps.items():
        assert lp[2] == -np.inf, mode
        for i in range(2):
            c = np.diag(cov[i]) + z["y_err"] ** 2
            assert np.count_nonzero(cov[i] - np.diag(np.diag(cov[i]))) == 0
            want = -0.5 * np.sum((mean[i] - z["y"]) ** 2 / c) - 0.5 * np.sum(np.log(c))
            assert abs(lp[i] - want) < 1e-11 * abs(want), (mode, lp[i], want)


def test_wave_tail_instantiation_equals_main_instantiation(c3a):
    # K4 runs the walkers of a last, at most half filled wave on its 16-consumer instantiation (one CTA per SM) and the
    # rest on the 8-consumer one (two CTAs per SM): the same arithmetic in the same order, so a walker's value must not
    # depend on which of the two evaluated it.  Batch sizes on both sides of every split point (148 SMs, 296 slots).
    a, _, dev = c3a
    full = dev.log_posterior(a.theta, n_div=1.0)
    slots, sms = dev.query(5), dev.query(4)          # resident K4 CTAs, SMs (bi_query)
    for W in (1, 5, sms, sms + 1, slots - 1, slots, slots + 1, slots + sms, slots + sms + 1, 2 * slots + 7):
        if W > a.theta.shape[0]:
            continue
        assert np.array_equal(dev.log_posterior(a.theta[:W], n_div=1.0), full[:W]), W
        assert np.array_equal(dev.log_posterior(a.theta[-W:], n_div=1.0), full[-W:]), W


def test_few_tile_instantiation_equals_main_instantiation(monkeypatch):
    # few-tile problems (F <= 287) with more than two walkers per SM on offer run K4 with 4 consumers per CTA and three CTAs
    # per SM instead of 8 and two: the same arithmetic per strip, so the values are bit-identical on both sides of the
    # switch point, for either forced choice (BI_PIPE_SMALL), and equal to the oracle
    from bayesian_inference_b200.device import DeviceEmulator

    a = synthetic.generate("C1", n_walkers=700, oob_fraction=0.02)
    img = image.pack(a.emulation_config, a.emulation_results, min=a.min, max=a.max,
                     experimental_results=a.experimental_results)
    out = {}
    for flag in (None, "0", "1"):
        if flag is None:
            monkeypatch.delenv("BI_PIPE_SMALL", raising=False)
        else:
            monkeypatch.setenv("BI_PIPE_SMALL", flag)
        with DeviceEmulator(img, max_batch=700) as dev:
            sms = dev.query(4)
            out[flag] = dev.log_posterior(a.theta, n_div=1.0)
            if flag is None:
                for W in (1, 2 * sms, 2 * sms + 1, 3 * sms + 5):      # both sides of the per-launch choice
                    assert np.array_equal(dev.log_posterior(a.theta[:W], n_div=1.0), out[None][:W], equal_nan=True), W
                struct = dev.log_posterior(a.theta, n_div=1.0, mode="structured")
    assert np.array_equal(out[None], out["0"], equal_nan=True) and np.array_equal(out[None], out["1"], equal_nan=True)
    idx = np.flatnonzero(np.isfinite(out[None]))[:32]
    _cmp_logp(out[None][idx], cf.log_posterior(img.to_flat(), a.theta[idx], n_div=1.0))
    fin = np.isfinite(out[None])
    assert rel_err(struct[fin], out[None][fin]) < TOL_LOGP


def test_mixed_tile_gp_launch_equals_plain_launch_and_oracle(monkeypatch):
    # K1+K2 with a partly empty last wave runs 32-walker tiles for the whole waves and 16-walker tiles for the rest in one
    # launch (gp_predict_mixed_kernel): the same arithmetic per (walker, PC), so (m, v) are bit-identical to the plain
    # launch for the automatic choice, for forced splits (BI_GP_MIXED=-n) and for ragged batch sizes; and equal to the oracle
    from bayesian_inference_b200.device import DeviceEmulator

    a = synthetic.generate("C1", n_walkers=1100)
    img = image.pack(a.emulation_config, a.emulation_results, min=a.min, max=a.max,
                     experimental_results=a.experimental_results)
    sizes = (333, 512, 700, 1024, 1061, 1100)
    out = {}
    for flag in ("0", None, "-1", "-7", "-1000"):
        if flag is None:
            monkeypatch.delenv("BI_GP_MIXED", raising=False)
        else:
            monkeypatch.setenv("BI_GP_MIXED", flag)
        with DeviceEmulator(img, max_batch=1100) as dev:
            out[flag] = [dev.predict_pc(a.theta[:W]) for W in sizes]
            if flag is None:
                lp = dev.log_posterior(a.theta[:700], n_div=1.0, mode="structured")
    for flag in (None, "-1", "-7", "-1000"):
        for (m0, v0), (m1, v1), W in zip(out["0"], out[flag], sizes):
            assert np.array_equal(m0, m1) and np.array_equal(v0, v1), (flag, W)
    z = img.to_flat()
    th = a.theta[1040:1100]                          # rows that sit in the 16-walker tiles of the automatic split
    mo, vo = [np.concatenate(x, axis=1) for x in zip(*[cf.predict_pc(z, g, th) for g in range(len(img.groups))])]
    m, v = out[None][-1]
    assert np.max(np.abs(m[1040:] - mo)) < TOL_PC * max(1.0, np.max(np.abs(mo)))
    assert np.max(np.abs(v[1040:] - vo)) < TOL_PC * 4.0
    idx = np.flatnonzero(np.isfinite(lp))[:16]
    _cmp_logp(lp[idx], cf.log_posterior(z, a.theta[idx], n_div=1.0))


@pytest.mark.parametrize("name", ["TOY", "C1", "C3B"])
def test_covariance_kernels_agree(name, monkeypatch):
    # predict(): the persistent per-warp kernel (T resident in shared memory) and the (row tile, walker) grid kernel are
    # two launch shapes of the same per-tile arithmetic -- bit-identical covariance, merged and per group
    from bayesian_inference_b200.device import DeviceEmulator

    a = synthetic.generate(name, n_walkers=37) if name != "TOY" else synthetic.generate(name)
    img = image.pack(a.emulation_config, a.emulation_results, min=a.min, max=a.max,
                     experimental_results=a.experimental_results, sigma_exp=a.sigma_exp)
    inside = np.all((a.theta > a.min) & (a.theta < a.max), axis=1)
    th = np.ascontiguousarray(a.theta[inside][:37])
    out = {}
    for flag in ("0", "1"):
        monkeypatch.setenv("BI_K3_PERSIST", flag)
        with DeviceEmulator(img) as dev:
            out[flag] = (dev.predict(th, n_div=float(len(th))), dev.predict_group(0, th, n_div=1.0))
    for k in (0, 1):                    # merged, group 0
        for j in (0, 1):                # mean, covariance
            assert np.array_equal(out["0"][k][j], out["1"][k][j]), (k, j)


def test_handles_of_different_sizes_coexist(c3a):
    # the kernels' dynamic shared-memory limit is a per-function attribute shared by all handles of the process: creating
    # and using a small problem must not break the launches of a larger one created earlier (and vice versa)
    from bayesian_inference_b200.device import DeviceEmulator

    a, _, big = c3a
    before = big.log_posterior(a.theta[:64], n_div=1.0)
    t = synthetic.generate("TOY")
    timg = image.pack(t.emulation_config, t.emulation_results, min=t.min, max=t.max, experimental_results=t.experimental_results)
    with DeviceEmulator(timg) as small:
        s0 = small.log_posterior(t.theta, n_div=1.0)
        assert np.array_equal(big.log_posterior(a.theta[:64], n_div=1.0), before)
        mb, cb = big.predict(a.theta[:3], n_div=1.0)
        assert np.array_equal(small.log_posterior(t.theta, n_div=1.0), s0, equal_nan=True)
    assert np.array_equal(big.log_posterior(a.theta[:64], n_div=1.0), before)
    assert mb.shape == (3, 600) and cb.shape == (3, 600, 600)
