"""GPU: parity with the oracle at the BENCHMARK shapes (BASELINE configs[2] and [3]), on slices the C restatement
finishes in seconds; the beta clamp; the per-pair MLE against an independent optimiser; one reference-held value.

Tolerances are the north star's: lower bound 1e-9 relative at every checkpoint, identical hard labels, W 1e-7."""
import numpy as np
import pytest

import c_oracle
import melissa_oracle as mo
import melissa_b200 as mb
import probit_mle
from melissa_b200 import synth

pytestmark = pytest.mark.gpu


def _to_host(t):
    return dict(N=int(t["N"]), M=int(t["M"]), **{k: t[k].cpu().numpy() for k in ("pair_ptr", "pair_cell", "pair_region", "x", "y")})


@pytest.mark.parametrize("name,N,M,K,seed", [("c3 slice", 600, 1000, 4, 1), ("c4 slice", 256, 5000, 10, 2)])
def test_benchmark_recipe_slice_matches_the_c_oracle(name, N, M, K, seed):
    """The recipe of the two benchmark configurations at their real region count and K (so that the cell blocks, the
    chunking and the K > 8 code paths run as in the benchmark), on a few hundred cells, 20 iterations, two bounds."""
    dev, tr = synth.generate_device(N, M, K, n_basis=4, seed=seed, chunk_cells=25 if N % 25 == 0 else 32)
    host = _to_host(dev)
    delta_0 = np.full(K, 0.51)
    ref = c_oracle.vb_run(host, K, tr["basis"], tr["w0"], delta_0, vb_max_iter=20)
    with mb.Engine(dev, tr["basis"], K_hint=K) as eng:
        info = eng.layout_info()
        res = eng.run(K, tr["w0"], delta_0, vb_max_iter=20)
    assert info["sliced"] == 1 and info["cell_block"] == 256
    assert len(res["lb"]) == len(ref["lb"]) == 3 and res["n_iter"] == ref["n_iter"] == 20
    np.testing.assert_allclose(res["lb"][1:], ref["lb"][1:], rtol=1e-9)
    assert np.array_equal(res["r_nk"].argmax(1), ref["r_nk"].argmax(1))
    np.testing.assert_allclose(res["r_nk"], ref["r_nk"], atol=1e-8)
    np.testing.assert_allclose(res["W"], ref["W"], atol=1e-7)
    np.testing.assert_allclose(res["beta"], ref["beta"], rtol=1e-8)


def test_beta_reaches_its_clamp_like_the_reference():
    """beta_k <- min(beta_0 + E_ww / 2, 10 alpha_k) (reference R/melissa_vb.R:386-392): a wide initial w and a small
    alpha_0 drive E_ww / 2 past the cap, so the clamp itself decides beta in at least one cluster."""
    K = 3
    d, tr = synth.generate(60, 6, K, n_basis=3, seed=31, sites_per_region=60, cov_lo=0.2, cov_hi=0.5)
    w0 = tr["w0"] * 20.0
    delta_0 = np.full(K, 0.6)
    alpha_0 = 0.01
    ref = mo.melissa_inner(d.as_dict(), K, tr["basis"], w0, delta_0, alpha_0=alpha_0, vb_max_iter=12, mimic=False)
    cap = 10.0 * (alpha_0 + d.M * 4 / 2.0)
    assert np.any(ref["beta"] == cap), ref["beta"]                      # the fixture really hits the clamp
    with mb.Engine(d, tr["basis"]) as eng:
        res = eng.run(K, w0, delta_0, alpha_0=alpha_0, vb_max_iter=12)
    assert np.array_equal(res["beta"] == cap, ref["beta"] == cap)
    np.testing.assert_allclose(res["beta"], ref["beta"], rtol=1e-8)
    np.testing.assert_allclose(res["lb"][1:], ref["lb"][1:], rtol=1e-8)
    np.testing.assert_allclose(res["W"], ref["W"], atol=1e-6)


def test_profiles_mle_reaches_the_optimum_an_independent_optimiser_finds(c1):
    """melissa_profiles_mle (Newton on the GPU) against oracle/probit_mle.py (BFGS to convergence on the published
    objective): the objective is strictly concave, so both must arrive at the same point."""
    d = mb.from_dict(c1)
    pairs = list(range(0, d.P, 7))
    want = probit_mle.profiles_mle(c1, c1["basis"], lam=0.5, pairs=pairs)
    with mb.Engine(d, c1["basis"]) as eng:
        got = eng.profiles_mle(0.5, 30)[pairs]
    np.testing.assert_allclose(got, want, rtol=1e-6, atol=1e-7)


def test_alpha_of_the_bundled_synthetic_fit_is_below_61(c1):
    """The one value of a melissa() fit the reference's tests still spell out without an RNG:
    expect_lt(melissa_obj$alpha[1], 61) for K = 2 on melissa_synth_dt
    (reference tests/testthat/test-running_melissa.R:29; alpha = alpha_0 + M D / 2 = 0.5 + 30 * 4 / 2 = 60.5)."""
    with mb.Engine(mb.from_dict(c1), c1["basis"]) as eng:
        res = eng.run(2, c1["w0_k2"], c1["k2_delta_0"], vb_max_iter=10)
    assert np.all(res["alpha"] < 61) and np.all(res["alpha"] == 60.5)


def test_device_built_layout_equals_the_host_built_one(c1):
    """melissa_pack builds the slice index on the device (radix sorts and scans, pack_gpu.cuh); round 1's host-thread
    builder stays behind the "pack_host" option.  Same slices in the same order, so the fit must agree to the last bit."""
    K = 5
    d, tr = synth.generate(300, 64, K, n_basis=4, seed=9, sites_per_region=60, cov_lo=0.02, cov_hi=0.3)
    delta_0 = np.full(K, 0.55)
    out = {}
    for host in (0, 1):
        mb.set_option("pack_host", host)
        try:
            with mb.Engine(d, tr["basis"]) as eng:
                out[host] = (eng.layout_info(), eng.run(K, tr["w0"], delta_0, vb_max_iter=20))
        finally:
            mb.set_option("pack_host", 0)
    assert out[0][0] == out[1][0]
    assert np.array_equal(out[0][1]["lb"], out[1][1]["lb"])
    assert np.array_equal(out[0][1]["r_nk"], out[1][1]["r_nk"])
    assert np.array_equal(out[0][1]["W"], out[1][1]["W"])


def test_device_resident_index_lists_are_validated_on_the_device():
    """melissa_pack_device checks the pair list with a kernel: same rules and messages as the host check."""
    import torch
    K = 2
    dev, tr = synth.generate_device(64, 20, K, n_basis=3, seed=4, chunk_cells=32)
    bad = dict(dev)
    pc = dev["pair_cell"].clone()
    pc[pc == 5] = 6                                            # cell 5 loses all its regions
    bad["pair_cell"] = pc
    with pytest.raises(mb._lib.MelissaError, match="cell 5: No coverage across all regions"):
        mb.Engine(bad, tr["basis"])
    bad = dict(dev)
    pr = dev["pair_region"].clone()
    pr[3] = pr[2]                                              # not strictly increasing inside a cell
    bad["pair_region"] = pr
    with pytest.raises(mb._lib.MelissaError, match="strictly increasing"):
        mb.Engine(bad, tr["basis"])
    with mb.Engine(dev, tr["basis"]) as eng:                   # and the untouched lists pack
        assert eng.layout_info()["pairs"] == int(dev["pair_cell"].shape[0])


@pytest.mark.parametrize("G", [2, 3])
def test_several_shards_behind_one_call_equal_the_single_gpu_fit(G):
    """melissa_pack_multi + melissa_vb_run_multi (one process, one host thread and stream per shard, partial sums
    added by the reduce kernel in rank order).  With one GPU in the box all shards land on it -- every line of the
    multi-GPU path except the NVLink hop is exercised -- otherwise they spread over the first GPUs."""
    import torch
    K = 4
    d, tr = synth.generate(150, 40, K, n_basis=4, seed=12, sites_per_region=60, cov_lo=0.05, cov_hi=0.3)
    delta_0 = np.full(K, 0.55)
    with mb.Engine(d, tr["basis"]) as eng:
        one = eng.run(K, tr["w0"], delta_0, vb_max_iter=30)
    ndev = torch.cuda.device_count()
    devices = [g % ndev for g in range(G)]
    with mb.MultiEngine.pack(d, tr["basis"], G, devices=devices) as me:
        many = me.run(K, tr["w0"], delta_0, vb_max_iter=30)
        again = me.run(K, tr["w0"], delta_0, vb_max_iter=30)
    assert many["n_iter"] == one["n_iter"] and many["flags"] == one["flags"] and len(many["lb"]) == len(one["lb"])
    np.testing.assert_allclose(many["lb"][1:], one["lb"][1:], rtol=1e-12)          # <= 1e-12 drift
    assert np.array_equal(many["r_nk"].argmax(1), one["r_nk"].argmax(1))
    np.testing.assert_allclose(many["r_nk"], one["r_nk"], atol=1e-10)
    np.testing.assert_allclose(many["W"], one["W"], atol=1e-10)
    assert np.array_equal(again["lb"], many["lb"]) and np.array_equal(again["r_nk"], many["r_nk"])   # reproducible run to run


def test_tma_staged_contraction_equals_the_plain_load_one():
    """contract_tma_kernel (cp.async.bulk + mbarrier staging) walks the slices in the same order as
    contract_mma_kernel and feeds the same tiles to the same DMMA sequence: the fits must agree to the last bit."""
    K = 10
    d, tr = synth.generate(400, 50, K, n_basis=4, seed=13, sites_per_region=60, cov_lo=0.05, cov_hi=0.3)
    delta_0 = np.full(K, 0.55)
    out = {}
    for tma in (1, 0):
        mb.set_option("contract_tma", tma)
        try:
            with mb.Engine(d, tr["basis"]) as eng:
                out[tma] = eng.run(K, tr["w0"], delta_0, vb_max_iter=20)
        finally:
            mb.set_option("contract_tma", 1)
    assert np.array_equal(out[0]["lb"], out[1]["lb"])
    assert np.array_equal(out[0]["W"], out[1]["W"])
    assert np.array_equal(out[0]["r_nk"], out[1]["r_nk"])


def test_progress_hook_is_called_at_every_bound_and_can_stop_the_run(c1):
    """melissa_set_progress: the reference's progress messages / the place for R_CheckUserInterrupt."""
    seen = []
    with mb.Engine(mb.from_dict(c1), c1["basis"]) as eng:
        with mb.progress(lambda i, lb, diff: seen.append((i, lb)) or False):
            full = eng.run(2, c1["w0_k2"], c1["k2_delta_0"], vb_max_iter=45, epsilon_conv=0.0)
        assert [i for i, _ in seen] == [10, 20, 30, 40, 45] and full["flags"] & 4 == 0
        np.testing.assert_array_equal([lb for _, lb in seen], full["lb"][1:])
        with mb.progress(lambda i, lb, diff: i >= 20):
            cut = eng.run(2, c1["w0_k2"], c1["k2_delta_0"], vb_max_iter=45, epsilon_conv=0.0)
        assert cut["n_iter"] == 20 and cut["flags"] & 4 and np.array_equal(cut["lb"], full["lb"][:3])
        after = eng.run(2, c1["w0_k2"], c1["k2_delta_0"], vb_max_iter=45, epsilon_conv=0.0)      # hook removed again
        assert after["n_iter"] == 45
    d = mb.from_dict(c1)
    with mb.MultiEngine.pack(d, c1["basis"], 2, devices=[0, 0]) as me:
        with mb.progress(lambda i, lb, diff: i >= 20):
            cut2 = me.run(2, c1["w0_k2"], c1["k2_delta_0"], vb_max_iter=45, epsilon_conv=0.0)
    assert cut2["n_iter"] == 20 and cut2["flags"] & 4
    np.testing.assert_allclose(cut2["lb"][1:], cut["lb"][1:], rtol=1e-12)


def test_restarts_on_replicas_pick_the_start_the_sequential_loop_picks():
    """Default initialisation with vb_init_nstart mini-VB runs (R/melissa_vb.R:159-194): handed out over two engines
    that hold the same data (one per GPU when the box has two, else both on GPU 0), the restarts must choose exactly
    the w the one-engine loop chooses -- the runs are independent and bit-reproducible, the winner is taken in restart
    order."""
    import torch
    from melissa_b200.init import initial_w
    K = 3
    d, tr = synth.generate(120, 30, K, n_basis=4, seed=21, sites_per_region=60, cov_lo=0.05, cov_hi=0.3)
    delta_0 = np.full(K, 0.6)
    args = dict(K=K, delta_0=delta_0, alpha_0=0.5, beta_0=None, epsilon_conv=1e-5, is_kmeans=True, vb_init_nstart=5,
                vb_init_max_iter=10)
    dev2 = 1 if torch.cuda.device_count() > 1 else 0
    with mb.Engine(d, tr["basis"]) as a, mb.Engine(d, tr["basis"], device=dev2) as b:
        seq = initial_w(a, d, tr["basis"], rng=np.random.default_rng(4), **args)
        rep = initial_w(a, d, tr["basis"], rng=np.random.default_rng(4), replicas=[b], **args)
    assert seq.shape == (30, 5, K) and np.array_equal(seq, rep)


def test_melissa_on_two_gpus_gives_the_one_gpu_fit():
    """melissa(..., n_gpus = 2): restarts as replicas, final fit on cell shards; same start (same seed), so the same fit
    up to the summation order of the two shards' statistics."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    K = 3
    d, tr = synth.generate(200, 30, K, n_basis=4, seed=22, sites_per_region=60, cov_lo=0.05, cov_hi=0.3)
    kw = dict(K=K, basis=tr["basis"], vb_max_iter=40, vb_init_nstart=4, vb_init_max_iter=10, is_verbose=False, seed=9)
    one = mb.melissa(d, **kw)
    two = mb.melissa(d, n_gpus=2, **kw)
    assert np.array_equal(one["labels"], two["labels"])
    np.testing.assert_allclose(two["lb"][1:], one["lb"][1:], rtol=1e-12)
    np.testing.assert_allclose(two["W"], one["W"], atol=1e-10)
    np.testing.assert_allclose(two["r_nk"], one["r_nk"], atol=1e-10)
