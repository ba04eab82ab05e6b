"""Host side of the drop-in: `melissa()` / `melissa_vb()` with the reference's
signature and result object (reference R/melissa_vb.R:86-219), driving the CUDA
library.  The VB arithmetic itself lives in libmelissa_b200.so; this module only
validates arguments, packs the input, runs the stopping rule and names things.
"""
from __future__ import annotations

import math
import warnings

import numpy as np

from .data import MelissaData, create_rbf_object, from_nested
from .engine import Engine

FLAG_LB_DECREASED = 1
FLAG_NOT_CONVERGED = 2
FLAG_INTERRUPTED = 4


class MelissaResult(dict):
    """The reference's S3 "melissa" list (R/melissa_vb.R:523-528, :204-216) as a
    dict with attribute access: W, W_Sigma, r_nk, delta, alpha, beta, basis, pi_k,
    lb, delta_0, alpha_0, beta_0, labels."""

    __getattr__ = dict.__getitem__
    r_class = "melissa"


def shard_bounds(cell_weight, world_size):
    """Contiguous cell blocks balanced by weight (CpGs per cell).  Returns
    world_size + 1 boundaries; every shard gets at least one cell."""
    w = np.asarray(cell_weight, dtype=np.float64)
    N = w.shape[0]
    if world_size > N:
        raise ValueError("more ranks than cells")
    cum = np.concatenate([[0.0], np.cumsum(w)])
    bounds = [0]
    for rnk in range(1, world_size):
        target = cum[-1] * rnk / world_size
        b = int(np.searchsorted(cum, target, side="left"))
        b = max(b, bounds[-1] + 1)
        b = min(b, N - (world_size - rnk))
        bounds.append(b)
    bounds.append(N)
    return bounds


def vb_loop(engine, vb_max_iter, epsilon_conv, lb_every=10, group=None, is_verbose=False):
    """The reference's iteration and stopping rule (R/melissa_vb.R:309, :458,
    :493-513) over an engine exposing phase_a / phase_b / lower_bound.  With a
    torch.distributed `group` the two reduction buffers are all-reduced (SUM)
    between the phases; that is the only inter-GPU traffic."""
    dist = None
    if group is not None:
        import torch.distributed as dist
    lb = [-math.inf]
    flags = 0
    n_iter = 1
    for i in range(2, int(vb_max_iter) + 1):
        n_iter = i
        engine.phase_a()
        if dist is not None:
            dist.all_reduce(engine.reduce_tensor(), group=group)
        with_lb = (i % lb_every == 0) or (i == vb_max_iter)
        engine.phase_b(with_lb)
        if with_lb:
            if dist is not None:
                dist.all_reduce(engine.lb_tensor(), group=group)
            v = engine.lower_bound()
            lb.append(v)
            if v < lb[-2]:
                flags |= FLAG_LB_DECREASED
            if abs(v - lb[-2]) < epsilon_conv:
                break
            if is_verbose and i % 50 == 0:
                print(f"\rIt:\t{i}\tLB:\t{v}\tDiff:\t{v - lb[-2]}")
        if i == vb_max_iter:
            flags |= FLAG_NOT_CONVERGED
    return np.asarray(lb), n_iter, flags


def _raise_warnings(flags):
    if flags & FLAG_LB_DECREASED:
        warnings.warn("Warning: Lower bound decreases!\n")          # R/melissa_vb.R:496
    if flags & FLAG_NOT_CONVERGED:
        warnings.warn("VB did not converge!\n")                     # R/melissa_vb.R:512


def melissa_inner(engine: Engine, K, w, delta_0, alpha_0=0.5, beta_0=None, vb_max_iter=300,
                  epsilon_conv=1e-5, lb_every=10, warn=True):
    """One VB run on one GPU from explicit (w, delta_0): the counterpart of the
    reference's un-exported melissa_inner (R/melissa_vb.R:232-529).  Runs the whole
    loop inside the library (melissa_vb_run)."""
    res = engine.run(K, w, delta_0, alpha_0, beta_0, vb_max_iter, epsilon_conv, lb_every)
    if warn:
        _raise_warnings(res["flags"])
    M, D = engine.M, engine.D
    b0 = math.sqrt(alpha_0 + M * D / 2.0) if beta_0 is None else beta_0
    res.update(basis=engine.basis, delta_0=np.asarray(delta_0, np.float64), alpha_0=alpha_0, beta_0=b0)
    return res


def finalize(obj, K, cell_names=None):
    """R/melissa_vb.R:204-216: names, hard labels (first arg-max, 1-based as in R)
    and lb - log(K!)."""
    r = obj["r_nk"]
    labels = np.array([int(np.flatnonzero(row == np.nanmax(row))[0]) + 1 for row in r], dtype=np.int64)
    out = MelissaResult(obj)
    out["labels"] = labels
    out["lb"] = np.asarray(obj["lb"]) - math.log(math.factorial(K))
    out["cluster_names"] = [f"cluster{k + 1}" for k in range(K)]
    out["cell_names"] = cell_names
    return out


def melissa(X, K=3, basis=None, delta_0=None, w=None, alpha_0=.5, beta_0=None, vb_max_iter=300,
            epsilon_conv=1e-5, is_kmeans=True, vb_init_nstart=10, vb_init_max_iter=20,
            is_parallel=False, no_cores=3, is_verbose=True, device=0, seed=None, n_gpus=1):
    """Drop-in for the reference's `melissa()` (R/melissa_vb.R:86-219).

    X: the nested list of the reference (list of cells -> list of regions ->
    I x 2 array or None) or a MelissaData.  `is_parallel` / `no_cores` are
    accepted and ignored (the GPU replaces mclapply).  Unlike the reference,
    a user-supplied `w` works (the reference leaves `optimal_w` unassigned in
    that case, R/melissa_vb.R:134-198).  `device`, `seed` and `n_gpus`
    are trailing extras.  With n_gpus > 1 (GPUs device ... device + n_gpus - 1 of one box) the restarts of the
    default initialisation run as independent replicas, one complete copy of the data per GPU, and the final fit
    runs on cells sharded over the GPUs (melissa_pack_multi / melissa_vb_run_multi).  Returns a MelissaResult.
    """
    if not (vb_init_nstart > 0):
        raise ValueError("vb_init_nstart not greater than 0")       # assert_that, :93
    if not (vb_init_max_iter > 0):
        raise ValueError("vb_init_max_iter not greater than 0")     # :94
    data = X if isinstance(X, MelissaData) else from_nested(X)
    if basis is None:
        warnings.warn("Basis object not defined. Using as default M = 3 RBFs.\n")   # :97
        basis = create_rbf_object(M=3)
    rng = np.random.default_rng(seed)
    if delta_0 is None:
        delta_0 = np.full(K, .5) + rng.beta(1e-1, 1e1, size=K)      # :110-112
    delta_0 = np.asarray(delta_0, np.float64)
    n_gpus = int(n_gpus)
    if n_gpus < 1:
        raise ValueError("n_gpus must be positive")
    if n_gpus > 1:
        return _melissa_several_gpus(data, K, basis, delta_0, w, alpha_0, beta_0, vb_max_iter, epsilon_conv, is_kmeans,
                                     vb_init_nstart, vb_init_max_iter, is_verbose, list(range(device, device + n_gpus)), rng)
    with Engine(data, basis, device=device, K_hint=K) as eng:
        if w is None:
            from .init import initial_w
            optimal_w = initial_w(eng, data, basis, K, delta_0, alpha_0, beta_0, epsilon_conv, is_kmeans,
                                  vb_init_nstart, vb_init_max_iter, rng)
        else:
            optimal_w = np.asarray(w, np.float64)
        if is_verbose:
            # the reference's message every 50th iteration on which a bound was taken (R/melissa_vb.R:503-507)
            from .engine import progress

            def say(i, lb, diff):
                if i % 50 == 0:
                    print(f"\rIt:\t{i}\tLB:\t{lb}\tDiff:\t{diff}")
                return False
            with progress(say):
                obj = melissa_inner(eng, K, optimal_w, delta_0, alpha_0, beta_0, vb_max_iter, epsilon_conv)
        else:
            obj = melissa_inner(eng, K, optimal_w, delta_0, alpha_0, beta_0, vb_max_iter, epsilon_conv)
    return finalize(obj, K, data.cell_names)


def _melissa_several_gpus(data, K, basis, delta_0, w, alpha_0, beta_0, vb_max_iter, epsilon_conv, is_kmeans,
                          vb_init_nstart, vb_init_max_iter, is_verbose, devices, rng):
    """melissa() on several GPUs of one box: replicas for the restarts, shards for the fit."""
    import threading

    from .engine import MultiEngine, progress
    if w is None:
        from .init import initial_w
        n_rep = min(len(devices), int(vb_init_nstart)) if vb_init_nstart > 1 else 1
        engines, errors = [None] * n_rep, []

        def pack(i):
            try:
                engines[i] = Engine(data, basis, device=devices[i], K_hint=K)
            except Exception as exc:
                errors.append(exc)
        threads = [threading.Thread(target=pack, args=(i,)) for i in range(n_rep)]
        for th in threads:
            th.start()
        for th in threads:
            th.join()
        try:
            if errors:
                raise errors[0]
            optimal_w = initial_w(engines[0], data, basis, K, delta_0, alpha_0, beta_0, epsilon_conv, is_kmeans,
                                  vb_init_nstart, vb_init_max_iter, rng, replicas=engines[1:])
        finally:
            for e in engines:
                if e is not None:
                    e.close()
    else:
        optimal_w = np.asarray(w, np.float64)
    M, D = data.M, int(basis["M"]) + 1
    with MultiEngine.pack(data, basis, len(devices), devices) as me:
        if is_verbose:
            def say(i, lb, diff):
                if i % 50 == 0:
                    print(f"\rIt:\t{i}\tLB:\t{lb}\tDiff:\t{diff}")
                return False
            with progress(say):
                obj = me.run(K, optimal_w, delta_0, alpha_0, beta_0, vb_max_iter, epsilon_conv)
        else:
            obj = me.run(K, optimal_w, delta_0, alpha_0, beta_0, vb_max_iter, epsilon_conv)
    _raise_warnings(obj["flags"])
    b0 = math.sqrt(alpha_0 + M * D / 2.0) if beta_0 is None else beta_0
    obj.update(basis=basis, delta_0=np.asarray(delta_0, np.float64), alpha_0=alpha_0, beta_0=b0)
    return finalize(obj, K, data.cell_names)


#: the reference documents the function under both names (R/melissa_vb.R:3)
melissa_vb = melissa
