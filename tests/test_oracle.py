"""CPU: pins the oracle (oracle/melissa_oracle.py) against the golden fixtures, the
reference's one live known-answer test, and oracle-free invariants."""
import math

import numpy as np
import pytest

import melissa_oracle as mo
from conftest import subset


def test_log_sum_exp_reference_bracket():
    # reference tests/testthat/test-filtering.R:37-41
    v = mo.log_sum_exp([0.001, 0.5, 2, 1.4, 1.5])
    assert 2.92 < v < 2.93


def test_rbf_object_matches_bundled_basis(c1, c2):
    # basis objects serialised inside the reference's .rda files
    for z in (c1, c2):
        b = mo.create_rbf_object(z["basis"]["M"])
        assert np.array_equal(b["mus"], z["basis"]["mus"])          # bit-exact centres
        assert b["gamma"] == z["basis"]["gamma"]


def test_vectorised_oracle_matches_golden_c1(c1):
    for tag, K in (("k2", 2), ("k3", 3)):
        o = mo.finalize(mo.melissa_inner(c1, K, c1["basis"], c1[f"w0_{tag}"], c1[f"{tag}_delta_0"],
                                         vb_max_iter=int(c1[f"{tag}_iters"])), K)
        np.testing.assert_allclose(o["lb"][1:], c1[f"{tag}_lb"][1:], rtol=1e-12)
        np.testing.assert_allclose(o["r_nk"], c1[f"{tag}_r_nk"], atol=1e-12)
        np.testing.assert_allclose(o["W"], c1[f"{tag}_W"], atol=1e-10)
        assert np.array_equal(o["labels"], c1[f"{tag}_labels"])
        assert o["flags"] == int(c1[f"{tag}_flags"])


def test_literal_transcription_matches_vectorised(c1):
    # the nested-loop transcription of R/melissa_vb.R:232-529 against the vectorised oracle
    K, iters = 2, 20
    a = mo.melissa_inner(c1, K, c1["basis"], c1["w0_k2"], c1["k2_delta_0"], vb_max_iter=iters)
    b = mo.melissa_inner_literal(c1, K, c1["basis"], c1["w0_k2"], c1["k2_delta_0"], vb_max_iter=iters)
    np.testing.assert_allclose(a["lb"][1:], b["lb"][1:], rtol=1e-12)
    np.testing.assert_allclose(a["W"], b["W"], atol=1e-11)
    np.testing.assert_allclose(a["W_Sigma"], b["W_Sigma"], atol=1e-12)
    np.testing.assert_allclose(a["r_nk"], b["r_nk"], atol=1e-12)
    np.testing.assert_allclose(a["beta"], b["beta"], rtol=1e-12)


def test_ground_truth_recovery_c1(c1):
    # melissa_synth_dt ships its generating labels and mixing proportions
    assert mo.cluster_ari(c1["C_true"], c1["k2_r_nk"]) == 1.0
    frac = c1["C_true"].mean(axis=0)
    est = c1["k2_r_nk"].mean(axis=0)
    np.testing.assert_allclose(est, frac, atol=1e-6)


def test_golden_c2_training_fit_and_imputation(c2):
    train, test = subset(c2, c2["is_train"]), subset(c2, ~c2["is_train"])
    o = mo.melissa_inner(train, 4, c2["basis"], c2["w0"], c2["delta_0"], vb_max_iter=int(c2["iters"]))
    np.testing.assert_allclose(o["lb"][1:], c2["lb"][1:] + math.log(math.factorial(4)), rtol=1e-12)
    assert mo.cluster_ari(c2["C_true"], o["r_nk"]) == 1.0
    act, pred = mo.impute_test_met(o, test)
    perf = mo.eval_imputation_performance(act, pred)
    assert abs(perf["auc"] - float(c2["auc"])) < 1e-12
    assert abs(perf["f_measure"] - float(c2["f_measure"])) < 1e-12
    assert 0.8 < perf["auc"] < 0.9


def test_invariants(c1):
    o = mo.melissa_inner(c1, 2, c1["basis"], c1["w0_k2"], c1["k2_delta_0"], vb_max_iter=60)
    lb = o["lb"][1:]
    assert np.all(np.diff(lb) >= -1e-9)                      # ELBO never decreases
    np.testing.assert_allclose(o["r_nk"].sum(axis=1), 1.0, atol=1e-12)
    np.testing.assert_allclose((o["delta"] - o["delta_0"]).sum(), c1["N"], rtol=1e-12)
    ev = np.linalg.eigvalsh(o["W_Sigma"])
    assert ev.min() > 0                                       # S is SPD


def test_stable_ez_form_agrees_on_bundled_data(c1):
    a = mo.melissa_inner(c1, 2, c1["basis"], c1["w0_k2"], c1["k2_delta_0"], vb_max_iter=30, mimic=True)
    b = mo.melissa_inner(c1, 2, c1["basis"], c1["w0_k2"], c1["k2_delta_0"], vb_max_iter=30, mimic=False)
    np.testing.assert_allclose(a["lb"][1:], b["lb"][1:], rtol=1e-13)


def test_update_Ez_forms_diverge_only_in_the_far_wrong_tail():
    mu = np.linspace(-4, 4, 81)
    for yv in (True, False):
        y = np.full(mu.shape, yv)
        a, b = mo.update_Ez(mu, y, True), mo.update_Ez(mu, y, False)
        np.testing.assert_allclose(a, b, rtol=1e-10)


def test_cell_permutation_invariance(c1):
    # permuting cells permutes r_nk rows and leaves the bound unchanged
    perm = np.random.RandomState(0).permutation(c1["N"])
    inv = np.argsort(perm)
    new_cell = inv[c1["pair_cell"]]
    order = np.lexsort((c1["pair_region"], new_cell))
    pp = c1["pair_ptr"]
    lens = np.diff(pp)[order]
    idx = np.concatenate([np.arange(pp[p], pp[p + 1]) for p in order])
    d = dict(N=c1["N"], M=c1["M"], pair_cell=new_cell[order], pair_region=c1["pair_region"][order],
             pair_ptr=np.concatenate([[0], np.cumsum(lens)]), x=c1["x"][idx], y=c1["y"][idx])
    a = mo.melissa_inner(c1, 2, c1["basis"], c1["w0_k2"], c1["k2_delta_0"], vb_max_iter=20)
    b = mo.melissa_inner(d, 2, c1["basis"], c1["w0_k2"], c1["k2_delta_0"], vb_max_iter=20)
    np.testing.assert_allclose(a["lb"][1:], b["lb"][1:], rtol=1e-12)
    np.testing.assert_allclose(a["r_nk"][perm], b["r_nk"], atol=1e-10)


def test_ari_and_auc_known_answers():
    assert mo.adjusted_rand_index([0, 0, 1, 1], [1, 1, 0, 0]) == 1.0
    assert abs(mo.adjusted_rand_index([0, 0, 1, 1], [0, 1, 0, 1]) - (-0.5)) < 1e-12
    perf = mo.eval_imputation_performance([1, 1, 0, 0], [0.9, 0.4, 0.6, 0.1])
    assert abs(perf["auc"] - 0.75) < 1e-12


def test_c_restatement_matches_numpy_oracle(c1, c2):
    # oracle/melissa_oracle.c (the timed CPU baseline) against the NumPy oracle
    import c_oracle
    for threads in (1, 2):
        o = c_oracle.vb_run(c1, 3, c1["basis"], c1["w0_k3"], c1["k3_delta_0"], vb_max_iter=int(c1["k3_iters"]),
                            n_threads=threads)
        np.testing.assert_allclose(o["lb"][1:], c1["k3_lb"][1:] + math.log(math.factorial(3)), rtol=1e-11)
        np.testing.assert_allclose(o["r_nk"], c1["k3_r_nk"], atol=1e-9)
        np.testing.assert_allclose(o["W"], c1["k3_W"], atol=1e-8)
        np.testing.assert_allclose(o["W_Sigma"], c1["k3_W_Sigma"], atol=1e-9)
        assert o["flags"] == int(c1["k3_flags"]) and o["n_iter"] == int(c1["k3_n_iter"])
    train = subset(c2, c2["is_train"])
    o = c_oracle.vb_run(train, 4, c2["basis"], c2["w0"], c2["delta_0"], vb_max_iter=20)
    np.testing.assert_allclose(o["lb"][1:], (c2["lb"][1:] + math.log(math.factorial(4)))[:2], rtol=1e-11)


@pytest.mark.parametrize("n_basis,K,seed", [(4, 10, 1), (9, 6, 2), (15, 3, 3), (0, 2, 4)])
def test_c_restatement_matches_numpy_oracle_at_the_shapes_it_checks_the_gpu_at(n_basis, K, seed):
    """The C restatement is the checker of the GPU tests at the benchmark recipes (K = 10, D = 5) and the timed CPU
    arm; here it is held against the NumPy oracle at those K / D and at the reference scripts' M = 9 and beyond."""
    import c_oracle
    from melissa_b200 import synth
    d, tr = synth.generate(40, 25, K, n_basis=n_basis, seed=seed, sites_per_region=50, cov_lo=0.1, cov_hi=0.4)
    delta_0 = np.full(K, 0.6)
    ref = mo.melissa_inner(d.as_dict(), K, tr["basis"], tr["w0"], delta_0, vb_max_iter=20)
    got = c_oracle.vb_run(d.as_dict(), K, tr["basis"], tr["w0"], delta_0, vb_max_iter=20, n_threads=2)
    assert got["n_iter"] == ref["n_iter"] and got["flags"] == ref["flags"]
    np.testing.assert_allclose(got["lb"][1:], ref["lb"][1:], rtol=1e-11)
    np.testing.assert_allclose(got["r_nk"], ref["r_nk"], atol=1e-9)
    np.testing.assert_allclose(got["W"], ref["W"], atol=1e-8)
