"""Host-side logic: observable ordering / block map, image packing and its flat round trip, synthetic configs.  CPU only."""
from __future__ import annotations

import numpy as np
import pytest
from conftest import load_golden

from bayesian_inference_b200 import image, synthetic
from bayesian_inference_b200.observable_map import ObservableBlockMap, ObservableSelection, sort_labels
from oracle import ref_harness as rh


def test_sort_labels_orders_by_type_observable_sub_centrality_sqrts():
    labels = ["5020__PbPb__jet__b__R0.2__0-10", "2760__PbPb__hadron__a__x__0-10", "5020__PbPb__hadron__a__x__0-10",
              "200__AuAu__hadron__a__x__10-20"]
    got = sort_labels(labels)
    assert got == ["2760__PbPb__hadron__a__x__0-10", "5020__PbPb__hadron__a__x__0-10",
                   "200__AuAu__hadron__a__x__10-20", "5020__PbPb__jet__b__R0.2__0-10"]


def test_observable_selection_matches_substring_and_glob():
    f = ObservableSelection(include_list=["hadron__pt_", "jet__*__R0.2"], exclude_list=["0-10"])
    assert f.accept_observable("5020__PbPb__hadron__pt_ch__x__10-20")
    assert not f.accept_observable("5020__PbPb__hadron__pt_ch__x__0-10")
    assert f.accept_observable("5020__PbPb__jet__mass__R0.2__30-50")
    assert not f.accept_observable("5020__PbPb__jet__mass__R0.4__30-50")


def test_block_map_tiles_the_feature_axis_and_interleaves_groups():
    a = synthetic.generate("TOY")
    bm = a.emulation_config.sort_observables_in_matrix
    tab = bm.block_table([g.name for g in a.spec.groups])
    assert tab[:, 3].sum() == bm.n_features == a.spec.n_features
    assert np.array_equal(tab[:, 1], np.concatenate([[0], np.cumsum(tab[:-1, 3])]))
    assert len(set(tab[:, 0])) == 2 and not np.all(np.diff(tab[:, 0]) >= 0)     # groups interleave in merged order


@pytest.mark.skipif(not rh.available(), reason="reference checkout not present")
def test_block_map_equals_reference_learn_mapping():
    a = synthetic.generate("C4", n_walkers=4)
    ref_cfg = rh.ReferenceConfig(a, "c4map")
    ref_map = ref_cfg.sort_observables_in_matrix
    ours = a.emulation_config.sort_observables_in_matrix
    assert list(ref_map.emulation_group_to_observable_matrix.keys()) == list(ours.entries.keys())
    for k, (g, so, si) in ref_map.emulation_group_to_observable_matrix.items():
        assert ours.entries[k] == (g, so, si)
    assert ref_map.shape[1] == ours.n_features


def test_image_flat_round_trip(tmp_path):
    a = synthetic.generate("TOY")
    img = image.pack(a.emulation_config, a.emulation_results, min=a.min, max=a.max,
                     experimental_results=a.experimental_results)
    p = tmp_path / "img.npz"
    img.save(str(p))
    back = image.EmulatorImage.load(str(p))
    za, zb = img.to_flat(), back.to_flat()
    assert set(za) == set(zb)
    for k in za:
        assert np.array_equal(np.asarray(za[k]), np.asarray(zb[k])), k


def test_golden_image_matches_generator():
    z, _ = load_golden("toy")
    a = synthetic.generate("TOY", oob_fraction=0.1)
    img = image.pack(a.emulation_config, a.emulation_results, min=a.min, max=a.max,
                     experimental_results=a.experimental_results)
    assert np.allclose(img.to_flat()["g0_L"], z["g0_L"], rtol=1e-12, atol=0)
    assert np.array_equal(img.blocks, z["blocks"])


def test_kernel_constants_reads_every_kernel_the_reference_can_build():
    from sklearn.gaussian_process import kernels as sk

    k = sk.Matern(length_scale=[1.0, 2.0], nu=2.5) + sk.ConstantKernel(0.7) + sk.WhiteKernel(0.01)
    kind, ell, c, eta = image.kernel_constants(k, 2)
    assert (kind, list(ell), c, eta) == (image.KIND_MATERN52, [1.0, 2.0], 0.7, 0.01)
    kind, ell, c, eta = image.kernel_constants(sk.RBF(length_scale=3.0), 2)
    assert (kind, list(ell), c, eta) == (image.KIND_RBF, [3.0, 3.0], 0.0, 0.0)
    with pytest.raises(ValueError):
        image.kernel_constants(sk.Matern(nu=1.0), 2)
    with pytest.raises(ValueError):
        image.kernel_constants(sk.RationalQuadratic(), 2)


def test_validate_rejects_overlapping_blocks():
    a = synthetic.generate("TOY")
    img = image.pack(a.emulation_config, a.emulation_results, min=a.min, max=a.max)
    img.blocks = img.blocks.copy()
    img.blocks[1, 1] -= 1
    with pytest.raises(ValueError):
        image.validate(img)


@pytest.mark.parametrize("name,F,P", [("C1", 200, 30), ("C2", 600, 30), ("C3A", 600, 30), ("C3B", 600, 30), ("C4", 600, 40),
                                      ("C5A", 2000, 50), ("C5B", 2000, 50)])
def test_baseline_config_shapes(name, F, P):
    s = synthetic.config(name)
    assert s.n_features == F and s.n_pc_total == P


def test_cov_unexplained_override_is_factored_exactly():
    """A caller-supplied emulator_cov_unexplained (emulation.py:445-448, :487-488) enters the image as an eigen-factor pair
    that multiplies back to the matrix; asymmetric input and wrong shapes are refused."""
    rng = np.random.RandomState(0)
    A = rng.normal(size=(15, 6))
    U = A @ A.T * 1e-3                     # rank 6, PSD
    S_rest, ev = image.factor_cov_unexplained(U, 15)
    assert np.allclose(S_rest.T @ (ev[:, None] * S_rest), U, rtol=0, atol=1e-15 * np.abs(U).max() * 15)
    with pytest.raises(ValueError):
        image.factor_cov_unexplained(U[:14, :14], 15)
    with pytest.raises(ValueError):
        image.factor_cov_unexplained(U + np.triu(np.ones((15, 15)), 1), 15)
    a = synthetic.generate("TOY")
    own = {g: image.pack_group(g, a.emulation_results[g], a.emulation_config.emulation_groups_config[g].n_pc).cov_unexplained()
           for g in a.emulation_results}
    other = {g: 3.0 * u + 0.01 * np.eye(u.shape[0]) for g, u in own.items()}
    img = image.pack(a.emulation_config, a.emulation_results, min=a.min, max=a.max, emulator_cov_unexplained=other)
    for g, gi in zip(img.groups, other):
        assert np.allclose(g.cov_unexplained(), other[gi], rtol=0, atol=1e-14 * np.abs(other[gi]).max())
    with pytest.raises(KeyError):
        image.pack(a.emulation_config, a.emulation_results, min=a.min, max=a.max, emulator_cov_unexplained={"alpha": other["alpha"]})


@pytest.mark.skipif(not rh.available(), reason="reference checkout not present")
def test_cov_unexplained_override_equals_reference_with_the_same_dictionary():
    """The reference executed with a non-empty emulator_cov_unexplained == the oracle on the image packed with it."""
    from oracle import closed_form as cf

    a = synthetic.generate("TOY")
    ref_em, _, _ = rh.import_reference()
    cfg = rh.ReferenceConfig(a, "toy_ucov")
    own = {g: ref_em.compute_emulator_group_cov_unexplained(cfg.emulation_groups_config[g], a.emulation_results[g])
           for g in a.emulation_results}
    other = {g: 2.5 * u + 0.02 * np.diag(np.arange(1, u.shape[0] + 1.0)) for g, u in own.items()}
    th = a.theta[:6]
    want = ref_em.predict(th, cfg, emulation_group_results=a.emulation_results, emulator_cov_unexplained=other)
    img = image.pack(a.emulation_config, a.emulation_results, min=a.min, max=a.max, emulator_cov_unexplained=other)
    mean, cov = cf.predict(img.to_flat(), th)
    assert np.allclose(mean, want["central_value"], rtol=1e-13, atol=0)
    assert np.linalg.norm(cov - want["cov"]) < 1e-13 * np.linalg.norm(want["cov"])
    base = ref_em.predict(th, cfg, emulation_group_results=a.emulation_results)
    assert np.linalg.norm(base["cov"] - want["cov"]) > 1e-3 * np.linalg.norm(want["cov"])      # the override matters
