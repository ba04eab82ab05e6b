"""Imputation and its scores (reference R/utils.R:225-284, :514-529, :622-634).

`impute_test_met` runs on the GPU (one more CpG pass in the library);
the scores are host NumPy restatements of what the reference delegates to ROCR
and mclust, needed for config 2's acceptance check.
"""
from __future__ import annotations

import numpy as np

from .data import MelissaData, from_nested


def impute_test_met(obj, test, basis=None, is_predictive=True, return_test=False, device=0):
    """Predictive imputation sum_k r_nk Phi(h . W[m,,k]) of held-out CpGs, rounded to
    3 decimals (R/utils.R:225-284 with basis = NULL, is_predictive = TRUE).
    `obj` is the result of melissa(); `test` a nested list (NA / None where a
    region has no held-out CpG) or a MelissaData.  Returns dict(act_obs,
    pred_obs) in the reference's cell-major order (:274-282)."""
    if basis is not None or not is_predictive or return_test:
        raise NotImplementedError("only the joint predictive imputation path is provided")
    import ctypes as C

    from . import _lib
    tdata = test if isinstance(test, MelissaData) else from_nested_test(test)
    lib = _lib.load()
    b = obj["basis"]
    W = np.asarray(obj["W"], np.float64)
    r = np.asarray(obj["r_nk"], np.float64)
    M, D, K = W.shape
    N = r.shape[0]
    if tdata.M != M or tdata.N != N:
        raise ValueError("test data and fit disagree on the number of cells / regions")
    mus = np.ascontiguousarray(b["mus"], np.float64)
    pred = np.empty(tdata.nnz, np.float64)
    Wf = np.asfortranarray(W).ravel(order="F").copy()
    rf = np.asfortranarray(r).ravel(order="F").copy()
    _lib.check(lib.melissa_impute(
        N, M, K, 1, int(b["M"]), _lib.ptr(mus, _lib.c_dbl) if mus.size else None, float(b["gamma"]),
        _lib.ptr(Wf, _lib.c_dbl), _lib.ptr(rf, _lib.c_dbl), tdata.P,
        _lib.ptr(np.ascontiguousarray(tdata.pair_ptr, np.int64), _lib.c_i64),
        _lib.ptr(np.ascontiguousarray(tdata.pair_cell, np.int32), _lib.c_i32),
        _lib.ptr(np.ascontiguousarray(tdata.pair_region, np.int32), _lib.c_i32),
        _lib.ptr(np.ascontiguousarray(tdata.x, np.float64), _lib.c_dbl), _lib.ptr(pred, _lib.c_dbl), int(device)))
    return dict(act_obs=tdata.y.astype(np.float64), pred_obs=pred)


def from_nested_test(X) -> MelissaData:
    """Like data.from_nested but cells without any held-out region are allowed."""
    N, M = len(X), len(X[0])
    pair_cell, pair_region, ptr, xs, ys = [], [], [0], [], []
    for n, cell in enumerate(X):
        for m, reg in enumerate(cell):
            if reg is None or np.ndim(reg) < 2:
                continue
            a = np.asarray(reg, np.float64)
            xs.append(a[:, 0]); ys.append(a[:, 1])
            pair_cell.append(n); pair_region.append(m); ptr.append(ptr[-1] + a.shape[0])
    return MelissaData(N=N, M=M, pair_ptr=np.asarray(ptr, np.int64), pair_cell=np.asarray(pair_cell, np.int32),
                       pair_region=np.asarray(pair_region, np.int32), x=np.concatenate(xs),
                       y=np.concatenate(ys).astype(np.uint8))


def eval_imputation_performance(obj, imputation_obj):
    """AUC and F-measure as ROCR computes them after round(pred, 2)
    (R/utils.R:514-529); adds obj["imputation"]."""
    pred = np.round(np.asarray(imputation_obj["pred_obs"], np.float64), 2)
    act = np.asarray(imputation_obj["act_obs"]) == 1
    npos, nneg = int(act.sum()), int((~act).sum())
    order = np.argsort(-pred, kind="stable")
    ps, as_ = pred[order], act[order]
    idx = np.concatenate([np.flatnonzero(np.diff(ps) != 0), [ps.shape[0] - 1]])
    tp = np.concatenate([[0], np.cumsum(as_)[idx]]).astype(np.float64)
    fp = np.concatenate([[0], np.cumsum(~as_)[idx]]).astype(np.float64)
    cutoffs = np.concatenate([[np.inf], ps[idx]])
    tpr, fpr = tp / npos, fp / nneg
    auc = float(np.sum(0.5 * (tpr[1:] + tpr[:-1]) * np.diff(fpr)))
    with np.errstate(divide="ignore", invalid="ignore"):
        prec, rec = tp / (tp + fp), tp / npos
        f = 1.0 / (0.5 / prec + 0.5 / rec)
    j = int(np.flatnonzero(cutoffs <= 0.5)[0])
    obj = type(obj)(obj)
    obj["imputation"] = dict(auc=auc, f_measure=float(f[j]), tpr=tpr, fpr=fpr, prec=prec, rec=rec)
    return obj


def adjusted_rand_index(a, b) -> float:
    a, b = np.asarray(a), np.asarray(b)
    _, ia = np.unique(a, return_inverse=True)
    _, ib = np.unique(b, return_inverse=True)
    tab = np.zeros((ia.max() + 1, ib.max() + 1), dtype=np.int64)
    np.add.at(tab, (ia, ib), 1)
    if tab.shape == (1, 1):
        return 1.0
    comb = lambda v: v * (v - 1) / 2.0
    sa, sb, st = comb(tab.sum(1)).sum(), comb(tab.sum(0)).sum(), comb(tab).sum()
    expected = sa * sb / comb(a.shape[0])
    return float((st - expected) / (0.5 * (sa + sb) - expected))


def cluster_ari(C_true, r_nk) -> float:
    """R/utils.R:622-634 (first arg-max per row, then mclust::adjustedRandIndex)."""
    first_max = lambda A: np.array([int(np.flatnonzero(r == r.max())[0]) for r in np.asarray(A)])
    return adjusted_rand_index(first_max(C_true), first_max(r_nk))


def _align_cluster(Z1, Z2):
    """R/utils.R:646-719 (`.align_cluster`, type = "mat"): harden both assignment matrices, then greedily
    swap pairs of columns of the estimate while that increases the number of matching entries.
    (R's max.col breaks ties at random; the first maximum is taken here.)"""
    Z1, Z2 = np.asarray(Z1, np.float64), np.asarray(Z2, np.float64)
    N, K = Z1.shape
    L1 = np.zeros((N, K)); Lm = np.zeros((N, K))
    L1[np.arange(N), Z1.argmax(1)] = 1
    Lm[np.arange(N), Z2.argmax(1)] = 1
    for k in range(K):
        for tempk in range(K):
            best = np.sum(L1 == Lm)
            dummy = Lm.copy()
            dummy[:, k], dummy[:, tempk] = Lm[:, tempk], Lm[:, k]
            if np.sum(L1 == dummy) > best:
                Lm = dummy
    return Lm


def cluster_error(C_true, C_post) -> float:
    """R/utils.R:597-607: share of assignment-matrix entries that differ after aligning the labels."""
    C_true = np.asarray(C_true, np.float64)
    aligned = _align_cluster(C_true, C_post)
    return float(1.0 - np.sum(aligned == C_true) / C_true.size)


def eval_cluster_performance(obj, C_true):
    """R/utils.R:572-580: obj["clustering"] = list(error, ari)."""
    obj = type(obj)(obj)
    obj["clustering"] = dict(error=cluster_error(C_true, obj["r_nk"]), ari=cluster_ari(C_true, obj["r_nk"]))
    return obj
