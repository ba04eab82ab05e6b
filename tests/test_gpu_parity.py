"""GPU: the CUDA path (through the C ABI) against the CPU oracle and the golden fixtures.

Tolerances (north star): lower bound within 1e-9 relative at every checkpoint,
identical hard labels, imputation AUC / F-measure within 1e-6.  Per-pair
H^T E[z] is compared at 1e-12 relative."""
import math
import warnings

import numpy as np
import pytest

import melissa_oracle as mo
import melissa_b200 as mb
from melissa_b200 import synth
from conftest import subset

pytestmark = pytest.mark.gpu

LB_RTOL = 1e-9


@pytest.fixture(autouse=True, params=["sliced-block128", "sliced-block8", "flat"])
def layout(request):
    """Every parity test runs on the sliced layout (default cell block and a tiny one, so
    that several cell blocks and work items are exercised on small data) and on the flat
    cell-major layout."""
    if request.param == "flat":
        mb.set_option("sliced", 0)
    else:
        mb.set_option("sliced", 1)
        mb.set_option("cell_block", int(request.param.split("block")[1]))
    yield request.param
    mb.set_option("sliced", -1)
    mb.set_option("cell_block", 0)


def _engine(d, basis, **kw):
    return mb.Engine(mb.from_dict(d), basis, **kw)


@pytest.mark.parametrize("tag,K", [("k2", 2), ("k3", 3)])
def test_c1_melissa_synth_dt_matches_golden(c1, tag, K):
    with _engine(c1, c1["basis"]) as eng:
        res = mb.melissa_inner(eng, K, c1[f"w0_{tag}"], c1[f"{tag}_delta_0"], vb_max_iter=int(c1[f"{tag}_iters"]),
                               warn=False)
    out = mb.finalize(res, K)
    np.testing.assert_allclose(out["lb"][1:], c1[f"{tag}_lb"][1:], rtol=LB_RTOL)
    assert out["lb"][0] == -np.inf and len(out["lb"]) == len(c1[f"{tag}_lb"])
    assert np.array_equal(out["labels"], c1[f"{tag}_labels"])
    np.testing.assert_allclose(out["r_nk"], c1[f"{tag}_r_nk"], atol=1e-8)
    np.testing.assert_allclose(out["W"], c1[f"{tag}_W"], atol=1e-7)
    np.testing.assert_allclose(out["W_Sigma"], c1[f"{tag}_W_Sigma"], atol=1e-8)
    np.testing.assert_allclose(out["beta"], c1[f"{tag}_beta"], rtol=1e-8)
    np.testing.assert_allclose(out["delta"], c1[f"{tag}_delta"], rtol=1e-8)
    np.testing.assert_allclose(out["pi_k"], c1[f"{tag}_pi_k"], rtol=1e-8)
    assert res["flags"] == int(c1[f"{tag}_flags"]) and res["n_iter"] == int(c1[f"{tag}_n_iter"])


def test_first_sweep_HtEz_per_pair(c1):
    with _engine(c1, c1["basis"]) as eng:
        eng.begin(2, c1["w0_k2"], c1["k2_delta_0"])
        g = eng.fetch_HtEz()
    np.testing.assert_allclose(g, c1["k2_HtEz_init"], rtol=1e-12, atol=1e-13)


def test_stepwise_equals_one_shot(c1):
    with _engine(c1, c1["basis"]) as eng:
        one = eng.run(2, c1["w0_k2"], c1["k2_delta_0"], vb_max_iter=30)
        eng.begin(2, c1["w0_k2"], c1["k2_delta_0"])
        lb, n_iter, flags = mb.vb_loop(eng, 30, 1e-5)
        st = eng.fetch()
    np.testing.assert_allclose(lb[1:], one["lb"][1:], rtol=1e-12)
    np.testing.assert_allclose(st["r_nk"], one["r_nk"], atol=1e-12)
    assert (n_iter, flags) == (one["n_iter"], one["flags"])


def test_c2_encode_holdout_fit_and_imputation(c2):
    train, test = subset(c2, c2["is_train"]), subset(c2, ~c2["is_train"])
    K = 4
    with _engine(train, c2["basis"]) as eng:
        res = mb.melissa_inner(eng, K, c2["w0"], c2["delta_0"], vb_max_iter=int(c2["iters"]), warn=False)
    out = mb.finalize(res, K)
    np.testing.assert_allclose(out["lb"][1:], c2["lb"][1:], rtol=LB_RTOL)
    assert np.array_equal(out["labels"], c2["labels"])                      # ARI = 1 vs the reference restatement
    assert mb.cluster_ari(c2["C_true"], out["r_nk"]) == 1.0                 # and vs the shipped ground truth
    np.testing.assert_allclose(out["W"], c2["W"], atol=1e-7)
    imp = mb.impute_test_met(out, mb.from_dict(test))
    assert np.array_equal(imp["act_obs"], test["y"].astype(float))
    head = imp["pred_obs"][:2000]
    assert np.max(np.abs(head - c2["pred_head"])) <= 1e-3 + 1e-12           # rounding to 3 dp may flip a tie
    assert np.mean(head != c2["pred_head"]) < 0.01
    perf = mb.eval_imputation_performance(out, imp)["imputation"]
    assert abs(perf["auc"] - float(c2["auc"])) < 1e-6
    assert abs(perf["f_measure"] - float(c2["f_measure"])) < 1e-6


@pytest.mark.parametrize("n_basis,K", [(0, 2), (1, 3), (3, 3), (4, 5), (7, 2), (4, 10), (2, 17), (9, 3), (11, 4), (15, 2), (3, 40),
                                       (4, 70)])
def test_ragged_sparse_coverage_vs_oracle(n_basis, K):
    # low coverage => many uncovered (cell, region) pairs, ragged 1..20 CpGs per pair
    d, tr = synth.generate(48, 37, K, n_basis=n_basis, seed=11 + n_basis, sites_per_region=40,
                           cov_lo=0.02, cov_hi=0.25)
    assert d.P < 48 * 37                                                      # some pairs really are missing
    # the oracle needs every region covered by some cell, like the reference
    assert np.unique(d.pair_region).size == 37
    delta_0 = np.full(K, 0.5) + 0.01 * np.arange(K)
    ref = mo.melissa_inner(d.as_dict(), K, tr["basis"], tr["w0"], delta_0, vb_max_iter=25)
    with mb.Engine(d, tr["basis"]) as eng:
        res = eng.run(K, tr["w0"], delta_0, vb_max_iter=25)
    np.testing.assert_allclose(res["lb"][1:], ref["lb"][1:], rtol=LB_RTOL)
    np.testing.assert_allclose(res["r_nk"], ref["r_nk"], atol=1e-8)
    np.testing.assert_allclose(res["W"], ref["W"], atol=1e-7)
    np.testing.assert_allclose(res["W_Sigma"], ref["W_Sigma"], atol=1e-8)
    assert res["flags"] == ref["flags"]


def test_convergence_stops_like_the_reference(c1):
    # loose epsilon: the rule fires at the second bound (i = 20) in both implementations
    ref = mo.melissa_inner(c1, 2, c1["basis"], c1["w0_k2"], c1["k2_delta_0"], vb_max_iter=300, epsilon_conv=10.0)
    with _engine(c1, c1["basis"]) as eng:
        res = eng.run(2, c1["w0_k2"], c1["k2_delta_0"], vb_max_iter=300, epsilon_conv=10.0)
    assert res["n_iter"] == ref["n_iter"] == 20 and res["flags"] == ref["flags"] == 0


def test_two_shards_on_one_gpu_equal_the_whole(c1):
    # the multi-GPU path without NCCL: two handles, buffers summed by hand
    import torch
    d = mb.from_dict(c1)
    shards = [d.shard_cells(0, 9), d.shard_cells(9, 20)]
    engs = [mb.Engine(s, c1["basis"], allow_empty_regions=True) for s in shards]

    class Pair:
        def phase_a(self):
            for e in engs: e.phase_a()
            t = [e.reduce_tensor() for e in engs]
            s = t[0] + t[1]
            for x in t: x.copy_(s)
        def phase_b(self, with_lb):
            for e in engs: e.phase_b(with_lb)
            if with_lb:
                t = [e.lb_tensor() for e in engs]
                s = t[0] + t[1]
                for x in t: x.copy_(s)
        def lower_bound(self):
            v = [e.lower_bound() for e in engs]
            assert v[0] == v[1]
            return v[0]
    for e in engs:
        e.begin(2, c1["w0_k2"], c1["k2_delta_0"], n_cells_total=d.N)
    lb, n_iter, flags = mb.vb_loop(Pair(), 30, 1e-5)
    r = np.concatenate([e.fetch()["r_nk"] for e in engs])
    W = engs[0].fetch()["W"]
    for e in engs: e.close()
    with mb.Engine(d, c1["basis"]) as eng:
        one = eng.run(2, c1["w0_k2"], c1["k2_delta_0"], vb_max_iter=30)
    np.testing.assert_allclose(lb[1:], one["lb"][1:], rtol=1e-12)            # <= 1e-12 drift
    np.testing.assert_allclose(r, one["r_nk"], atol=1e-10)
    np.testing.assert_allclose(W, one["W"], atol=1e-10)


def test_melissa_wrapper_default_initialisation(c1, layout):
    if layout == "flat":
        # the default initialisation needs the per-pair MLE kernel, which lives on the sliced layout; there is no CPU path
        with pytest.raises(Exception, match="sliced layout"):
            mb.melissa(mb.from_dict(c1), K=2, basis=c1["basis"], vb_max_iter=12, vb_init_nstart=2, is_verbose=False, seed=3)
        return
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        out = mb.melissa(mb.to_nested(mb.from_dict(c1)), K=2, basis=c1["basis"], vb_max_iter=60,
                         vb_init_nstart=3, vb_init_max_iter=10, is_verbose=False, seed=3)
    assert mb.cluster_ari(c1["C_true"], out["r_nk"]) == 1.0
    assert out["W"].shape == (30, 4, 2) and out["W_Sigma"].shape == (2, 30, 4, 4)
    assert set(["W", "W_Sigma", "r_nk", "delta", "alpha", "beta", "basis", "pi_k", "lb", "delta_0", "alpha_0",
                "beta_0", "labels"]) <= set(out)
    with pytest.warns(UserWarning, match="did not converge"):
        mb.melissa(mb.from_dict(c1), K=2, basis=c1["basis"], w=c1["w0_k2"], delta_0=c1["k2_delta_0"], vb_max_iter=12)


def test_size_independent_properties_at_scale():
    # 1,000 cells x 400 regions (1e7 CpGs): properties that need no oracle
    K = 4
    d, tr = synth.generate(1000, 400, K, n_basis=4, seed=5)
    delta_0 = np.full(K, 0.51)
    with mb.Engine(d, tr["basis"]) as eng:
        res = eng.run(K, tr["w0"], delta_0, vb_max_iter=40)
        # cell permutation: reverse the cell order
        order = np.argsort(-d.pair_cell.astype(np.int64), kind="stable")
        pp = d.pair_ptr
        lens = np.diff(pp)[order]
        starts = pp[:-1][order]
        idx = np.repeat(starts - np.concatenate([[0], np.cumsum(lens)[:-1]]), lens) + np.arange(int(lens.sum()))
        dp = mb.MelissaData(N=d.N, M=d.M, pair_ptr=np.concatenate([[0], np.cumsum(lens)]).astype(np.int64),
                            pair_cell=(d.N - 1 - d.pair_cell[order]).astype(np.int32),
                            pair_region=d.pair_region[order], x=d.x[idx], y=d.y[idx])
    with mb.Engine(dp, tr["basis"]) as eng:
        res_p = eng.run(K, tr["w0"], delta_0, vb_max_iter=40)
    lb = res["lb"][1:]
    assert np.all(np.diff(lb) >= -1e-9 * np.abs(lb[:-1]))                    # ELBO non-decreasing
    np.testing.assert_allclose(res["r_nk"].sum(axis=1), 1.0, atol=1e-12)
    np.testing.assert_allclose((res["delta"] - delta_0).sum(), d.N, rtol=1e-12)
    assert np.linalg.eigvalsh(res["W_Sigma"]).min() > 0
    assert mb.cluster_ari(tr["C_true"], res["r_nk"]) == 1.0
    np.testing.assert_allclose(res_p["lb"][1:], lb, rtol=1e-12)
    np.testing.assert_allclose(res_p["r_nk"][::-1], res["r_nk"], atol=1e-9)


def test_rare_path_positions_outside_the_table_and_far_tails():
    # x beyond [-1, 1] (outside the RBF table) and |mu| >= 8 (outside the Mills table) take the plain
    # exp / erfcx path; the oracle is run with the cancellation-free E[z] form, since the reference's own
    # expression returns Inf there (deviation 2)
    K = 3
    d, tr = synth.generate(40, 25, K, n_basis=4, seed=21, sites_per_region=60, cov_lo=0.1, cov_hi=0.4)
    d.x[::37] *= 1.35
    w0 = tr["w0"] * 7.0
    delta_0 = np.full(K, 0.6)
    ref = mo.melissa_inner(d.as_dict(), K, tr["basis"], w0, delta_0, vb_max_iter=12, mimic=False)
    with mb.Engine(d, tr["basis"]) as eng:
        eng.begin(K, w0, delta_0)
        g = eng.fetch_HtEz()
        res = eng.run(K, w0, delta_0, vb_max_iter=12)
    assert np.isfinite(g).all()
    np.testing.assert_allclose(res["lb"][1:], ref["lb"][1:], rtol=1e-8)
    np.testing.assert_allclose(res["r_nk"], ref["r_nk"], atol=1e-7)


def test_full_size_c3_properties_and_shard_invariance(layout):
    """BASELINE configs[2] at full size (5,000 x 1,000, K = 4, 1.25e8 CpGs, drawn on the device):
    properties that need no oracle, and the same fit from two cell shards."""
    if layout != "sliced-block128":
        pytest.skip("one layout is enough at this size")
    import torch
    mb.set_option("cell_block", 0)
    K, N, M = 4, 5000, 1000
    data, tr = synth.generate_device(N, M, K, n_basis=4, seed=1, chunk_cells=25)
    delta_0 = np.full(K, 0.51)
    with mb.Engine(data, tr["basis"], K_hint=K) as eng:
        res = eng.run(K, tr["w0"], delta_0, vb_max_iter=30)
    lb = res["lb"][1:]
    assert len(lb) == 3 and np.all(np.diff(lb) > 0)                          # ELBO increases
    np.testing.assert_allclose(res["r_nk"].sum(axis=1), 1.0, atol=1e-12)
    np.testing.assert_allclose((res["delta"] - delta_0).sum(), N, rtol=1e-12)
    assert np.linalg.eigvalsh(res["W_Sigma"]).min() > 0
    assert mb.cluster_ari(tr["C_true"], res["r_nk"]) == 1.0
    # two shards of 2,500 cells, reduction buffers summed by hand
    shards = []
    for lo, hi in ((0, 2500), (2500, 5000)):
        d, _ = synth.generate_device(N, M, K, n_basis=4, seed=1, cell_lo=lo, cell_hi=hi, N_total=N, chunk_cells=25)
        shards.append(mb.Engine(d, tr["basis"], allow_empty_regions=True, K_hint=K))
        del d
    del data
    torch.cuda.empty_cache()

    class Pair:
        def phase_a(self):
            for e in shards: e.phase_a()
            t = [e.reduce_tensor() for e in shards]
            s = t[0] + t[1]
            for x in t: x.copy_(s)
        def phase_b(self, with_lb):
            for e in shards: e.phase_b(with_lb)
            if with_lb:
                t = [e.lb_tensor() for e in shards]
                s = t[0] + t[1]
                for x in t: x.copy_(s)
        def lower_bound(self):
            return [e.lower_bound() for e in shards][0]
    for e in shards:
        e.begin(K, tr["w0"], delta_0, n_cells_total=N)
    lb2, _, _ = mb.vb_loop(Pair(), 30, 1e-5)
    r2 = np.concatenate([e.fetch()["r_nk"] for e in shards])
    for e in shards: e.close()
    np.testing.assert_allclose(lb2[1:], lb, rtol=1e-12)                      # <= 1e-12 drift across shard counts
    assert np.array_equal(r2.argmax(1), res["r_nk"].argmax(1))


def test_profiles_mle_needs_the_sliced_layout(c1, layout):
    d = mb.from_dict(c1)
    with mb.Engine(d, c1["basis"]) as eng:
        if layout == "flat":
            with pytest.raises(Exception, match="sliced layout"):
                eng.profiles_mle(0.5, 15)
        else:
            assert eng.profiles_mle(0.5, 15).shape == (d.P, 4)


def test_argument_errors_through_the_abi(c1):
    d = mb.from_dict(c1)
    with mb.Engine(d, c1["basis"]) as eng:
        with pytest.raises(ValueError, match="shape"):
            eng.run(2, c1["w0_k3"], c1["k2_delta_0"])                       # w has K = 3 clusters
        with pytest.raises(mb._lib.MelissaError, match="K must be in 1..256"):
            eng.run(257, np.zeros((d.M, 4, 257)), np.full(257, 0.5))
        with pytest.raises(mb._lib.MelissaError, match="delta_0 must be positive"):
            eng.run(2, c1["w0_k2"], np.array([0.5, -1.0]))
        with pytest.raises(mb._lib.MelissaError, match="vb_max_iter"):
            eng.run(2, c1["w0_k2"], c1["k2_delta_0"], vb_max_iter=1)
        with pytest.raises(mb._lib.MelissaError, match="positive"):
            eng.run(2, c1["w0_k2"], c1["k2_delta_0"], beta_0=-1.0)
    bad = mb.from_dict(c1)
    bad.y = bad.y.copy(); bad.y[5] = 2
    with pytest.raises(mb._lib.MelissaError, match="not 0/1"):
        mb.Engine(bad, c1["basis"])
    with pytest.raises(mb._lib.MelissaError, match="0..15"):
        mb.Engine(d, mb.create_rbf_object(16))


@pytest.mark.parametrize("case", ["K1", "M1", "single_cpg_pairs", "two_iterations", "one_cell_per_region_block"])
def test_edge_shapes_vs_oracle(case):
    rng = np.random.RandomState(5)
    K, n_basis, iters = 3, 3, 12
    if case == "K1":
        K = 1
    d, tr = synth.generate(40, 1 if case == "M1" else 12, K, n_basis=n_basis, seed=17, sites_per_region=50,
                           cov_lo=0.1, cov_hi=0.5)
    if case == "single_cpg_pairs":                     # every pair keeps only its first CpG
        first = d.pair_ptr[:-1]
        d = mb.MelissaData(N=d.N, M=d.M, pair_ptr=np.arange(d.P + 1, dtype=np.int64), pair_cell=d.pair_cell,
                           pair_region=d.pair_region, x=d.x[first].copy(), y=d.y[first].copy())
    if case == "two_iterations":
        iters = 2                                      # the loop runs i = 2 only; the bound is taken at vb_max_iter
    if case == "one_cell_per_region_block":
        mb.set_option("cell_block", 32)                # 40 cells -> a full block and a block of 8
    w0 = tr["w0"] if case != "M1" else tr["w0"][:1]
    delta_0 = np.full(K, 0.7)
    ref = mo.melissa_inner(d.as_dict(), K, tr["basis"], w0, delta_0, vb_max_iter=iters)
    with mb.Engine(d, tr["basis"], K_hint=K) as eng:
        res = eng.run(K, w0, delta_0, vb_max_iter=iters)
    assert len(res["lb"]) == len(ref["lb"]) and res["n_iter"] == ref["n_iter"] and res["flags"] == ref["flags"]
    np.testing.assert_allclose(res["lb"][1:], ref["lb"][1:], rtol=LB_RTOL)
    np.testing.assert_allclose(res["r_nk"], ref["r_nk"], atol=1e-8)
    np.testing.assert_allclose(res["W"], ref["W"], atol=1e-7)
