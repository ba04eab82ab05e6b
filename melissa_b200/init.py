"""Default initialisation of the cluster means when `w` is not supplied
(reference R/melissa_vb.R:134-194): per-(cell, region) penalised probit MLE,
k-means on the stacked profiles, `vb_init_nstart` short VB runs to rank the
restarts.  The restarts run on the GPU through the same library call as the
final fit, and so does the MLE (melissa_profiles_mle); k-means is host NumPy
(parity here is statistical, the reference path depends on
R's RNG, `optim` and `stats::kmeans`).
"""
from __future__ import annotations

import math

import numpy as np



def _kmeans(X, K, rng, n_iter=10):
    """Plain Lloyd k-means, one random start (role of stats::kmeans(W_tmp, K, nstart = 1))."""
    N = X.shape[0]
    centres = X[rng.choice(N, K, replace=False)].copy()
    lab = np.zeros(N, dtype=np.int64)
    for _ in range(n_iter):
        d2 = ((X * X).sum(1)[:, None] - 2.0 * X @ centres.T + (centres * centres).sum(1)[None])
        new = d2.argmin(1)
        if np.array_equal(new, lab) and _ > 0:
            break
        lab = new
        for k in range(K):
            sel = lab == k
            if sel.any():
                centres[k] = X[sel].mean(0)
    return lab


def initial_w(engine, data, basis, K, delta_0, alpha_0, beta_0, epsilon_conv, is_kmeans,
              vb_init_nstart, vb_init_max_iter, rng):
    """R/melissa_vb.R:134-194.  Keeps the reference's two quirks: the restart
    whose mini-VB ends with the best lower bound contributes its INITIAL w
    (:187-193), and vb_init_nstart == 1 skips the mini-VB (:174-177)."""
    N, M, D = data.N, data.M, int(basis["M"]) + 1
    w_mle = engine.profiles_mle(0.5, 15)                 # GPU; no CPU path (hand-placed RBF centres: pass `w`)
    ww = rng.normal(0.0, 0.01, size=(M, D, N))                               # :148
    ww[data.pair_region, :, data.pair_cell] = w_mle                          # :151
    W_tmp = ww.transpose(2, 1, 0).reshape(N, D * M)                          # c(ww[,,n]) column-major
    lb_prev = -1e120
    optimal_w = None
    for _t in range(int(vb_init_nstart)):
        w = np.zeros((M, D, K))
        if is_kmeans:
            lab = _kmeans(W_tmp, K, rng)
            for k in range(K):
                members = np.flatnonzero(lab == k)
                if members.size == 0:                                        # the reference leaves this as a TODO (:165)
                    members = np.arange(N)
                w[:, :, k] = ww[:, :, rng.choice(members)]                   # :168
        else:
            w = ww[:, :, rng.choice(N, K, replace=False)].copy()             # :171
        if vb_init_nstart == 1:
            return w
        mini = engine.run(K, w, delta_0, alpha_0, beta_0, max(2, int(vb_init_max_iter)), epsilon_conv)
        lb_cur = mini["lb"][-1]
        if lb_cur > lb_prev:
            optimal_w, lb_prev = w, lb_cur
    if optimal_w is None:
        raise FloatingPointError("no restart of the default initialisation ended with a finite lower bound")
    return optimal_w
