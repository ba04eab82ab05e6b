"""CPU: host-side logic -- input flattening, argument checks, the stopping rule, and the
cell-sharded loop over a world_size-2 gloo group (engine = CPU oracle test double)."""
import os
import socket
import warnings

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import melissa_oracle as mo
from melissa_b200 import data as mdata
from melissa_b200 import evaluate, vb
from conftest import load_c1
from oracle_engine import OracleEngine


def test_nested_roundtrip_and_checks(c1):
    d = mdata.from_dict(c1)
    X = mdata.to_nested(d)
    X[3][5] = None                                   # an uncovered region (NA in the reference)
    X[4][0] = float("nan")
    f = mdata.from_nested(X)
    assert f.N == d.N and f.M == d.M and f.P == d.P - 2
    assert not np.any((f.pair_cell == 3) & (f.pair_region == 5))
    with pytest.raises(TypeError):
        mdata.from_nested("nope")                    # assert_that(is.list(X)), R/melissa_vb.R:91
    with pytest.raises(TypeError):
        mdata.from_nested([np.zeros((2, 2))])        # assert_that(is.list(X[[1]])), :92
    X[7] = [None] * d.M
    with pytest.raises(ValueError, match="No coverage across all regions"):
        mdata.from_nested(X)                         # R/utils.R:69-71


def test_flat_file_roundtrip(tmp_path, c1):
    d = mdata.from_dict(c1)
    d.cell_names = [f"cell_{i + 1}" for i in range(d.N)]
    d.save(tmp_path / "synth.npz")
    e = mdata.MelissaData.load(tmp_path / "synth.npz")
    assert (e.N, e.M, e.P, e.nnz) == (d.N, d.M, d.P, d.nnz) and e.cell_names == d.cell_names
    assert np.array_equal(e.x, d.x) and np.array_equal(e.y, d.y) and np.array_equal(e.pair_ptr, d.pair_ptr)


def test_melissa_argument_checks(c1):
    d = mdata.from_dict(c1)
    with pytest.raises(ValueError):
        vb.melissa(d, K=2, basis=c1["basis"], vb_init_nstart=0)      # :93
    with pytest.raises(ValueError):
        vb.melissa(d, K=2, basis=c1["basis"], vb_init_max_iter=0)    # :94


def test_shard_bounds_balance_by_weight():
    w = np.array([1, 1, 1, 1, 10, 1, 1, 1, 1, 10], float)
    b = vb.shard_bounds(w, 2)
    assert b[0] == 0 and b[-1] == 10 and len(b) == 3
    left, right = w[:b[1]].sum(), w[b[1]:].sum()
    assert abs(left - right) <= 10
    b = vb.shard_bounds(np.ones(5), 5)
    assert b == [0, 1, 2, 3, 4, 5]
    with pytest.raises(ValueError):
        vb.shard_bounds(np.ones(2), 3)


def test_vb_loop_with_oracle_engine_matches_oracle(c1):
    K = 2
    eng = OracleEngine(c1, c1["basis"])
    eng.begin(K, c1["w0_k2"], c1["k2_delta_0"])
    lb, n_iter, flags = vb.vb_loop(eng, 40, 1e-5)
    ref = mo.melissa_inner(c1, K, c1["basis"], c1["w0_k2"], c1["k2_delta_0"], vb_max_iter=40)
    np.testing.assert_allclose(lb[1:], ref["lb"][1:], rtol=1e-11)
    assert (n_iter, flags) == (ref["n_iter"], ref["flags"])
    np.testing.assert_allclose(eng.fetch()["r_nk"], ref["r_nk"], atol=1e-10)


def test_stopping_rule_and_flags():
    class Fake:
        def __init__(self, seq):
            self.seq, self.i = seq, 0
        def phase_a(self): pass
        def phase_b(self, with_lb): pass
        def lower_bound(self):
            v = self.seq[self.i]; self.i += 1; return v
    # converges at the 2nd bound (i = 20): |diff| < eps, no "did not converge"
    lb, n_iter, flags = vb.vb_loop(Fake([-10.0, -10.0 + 1e-7]), 300, 1e-5)
    assert n_iter == 20 and flags == 0 and len(lb) == 3 and lb[0] == -np.inf
    # decreasing bound sets bit 0; hitting vb_max_iter sets bit 1 (R/melissa_vb.R:495-497, :511-513)
    lb, n_iter, flags = vb.vb_loop(Fake([-10.0, -11.0, -12.0]), 25, 1e-5)
    assert n_iter == 25 and flags == 3 and len(lb) == 4
    # bound evaluated at vb_max_iter even when not a multiple of 10 (:458)
    lb, n_iter, flags = vb.vb_loop(Fake([-5.0]), 7, 1e-5)
    assert len(lb) == 2 and flags == 2


def test_finalize_labels_and_lb_shift():
    obj = dict(r_nk=np.array([[0.2, 0.8], [0.5, 0.5], [0.9, 0.1]]), lb=np.array([-np.inf, -3.0]))
    out = vb.finalize(obj, 2)
    assert list(out["labels"]) == [2, 1, 1]                          # first arg-max, 1-based (:212-213)
    assert abs(out["lb"][1] - (-3.0 - np.log(2))) < 1e-15            # :216


def test_eval_metrics_match_oracle():
    rng = np.random.RandomState(0)
    act = (rng.rand(500) < 0.4).astype(float)
    pred = np.clip(0.3 * act + 0.7 * rng.rand(500), 0, 1)
    a = evaluate.eval_imputation_performance({}, dict(act_obs=act, pred_obs=pred))["imputation"]
    b = mo.eval_imputation_performance(act, pred)
    assert abs(a["auc"] - b["auc"]) < 1e-15 and abs(a["f_measure"] - b["f_measure"]) < 1e-15
    lab = rng.randint(0, 3, 50)
    assert evaluate.adjusted_rand_index(lab, (lab + 1) % 3) == 1.0


def test_cluster_error_aligns_labels_first():
    rng = np.random.RandomState(2)
    lab = rng.randint(0, 3, 60)
    C_true = np.eye(3)[lab]
    post = np.eye(3)[(lab + 1) % 3] * 0.9 + 0.05                 # same clustering, labels rotated
    assert evaluate.cluster_error(C_true, post) == 0.0           # R/utils.R:597-607 after .align_cluster
    post2 = post.copy()
    post2[:6] = post2[:6][:, [1, 2, 0]]                          # six cells moved to another cluster
    assert abs(evaluate.cluster_error(C_true, post2) - 2 * 6 / C_true.size) < 1e-12
    out = evaluate.eval_cluster_performance(vb.MelissaResult(r_nk=post), C_true)
    assert out["clustering"]["ari"] == 1.0 and out["clustering"]["error"] == 0.0


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _sharded_worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    c1 = load_c1()
    d = mdata.from_dict(c1)
    cell_w = np.bincount(d.pair_cell, weights=np.diff(d.pair_ptr).astype(float), minlength=d.N)
    b = vb.shard_bounds(cell_w, world)
    shard = d.shard_cells(b[rank], b[rank + 1])
    eng = OracleEngine(shard.as_dict(), c1["basis"])
    eng.begin(2, c1["w0_k2"], c1["k2_delta_0"], n_cells_total=d.N)
    lb, n_iter, flags = vb.vb_loop(eng, 30, 1e-5, group=dist.group.WORLD)
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), lb=lb, r=eng.fetch()["r_nk"], lo=b[rank], hi=b[rank + 1],
             W=eng.fetch()["W"], n_iter=n_iter, flags=flags)
    dist.destroy_process_group()


def test_cell_sharded_loop_world_size_2_gloo(tmp_path, c1):
    port = _free_port()
    mp.spawn(_sharded_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    ref = mo.melissa_inner(c1, 2, c1["basis"], c1["w0_k2"], c1["k2_delta_0"], vb_max_iter=30)
    r = np.zeros_like(ref["r_nk"])
    for rank in range(2):
        z = np.load(tmp_path / f"rank{rank}.npz")
        np.testing.assert_allclose(z["lb"][1:], ref["lb"][1:], rtol=1e-11)      # same bound on every rank
        np.testing.assert_allclose(z["W"], ref["W"], atol=1e-9)
        r[int(z["lo"]):int(z["hi"])] = z["r"]
    np.testing.assert_allclose(r, ref["r_nk"], atol=1e-10)
