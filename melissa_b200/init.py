"""Default initialisation of the cluster means when `w` is not supplied
(reference R/melissa_vb.R:134-194): per-(cell, region) penalised probit MLE,
k-means on the stacked profiles, `vb_init_nstart` short VB runs to rank the
restarts.  The restarts run on the GPU through the same library call as the
final fit; the MLE and k-means are host NumPy (parity here
is statistical, the reference path depends on R's RNG, `optim` and
`stats::kmeans`).
"""
from __future__ import annotations

import math

import numpy as np
from scipy.special import log_ndtr

SQRT_2PI = math.sqrt(2 * math.pi)


def _design(x, basis):
    mus = np.asarray(basis["mus"], np.float64)
    H = np.ones((x.shape[0], mus.shape[0] + 1))
    if mus.shape[0]:
        d = x[:, None] - mus[None, :]
        H[:, 1:] = np.exp(-basis["gamma"] * d * d)
    return H


def probit_mle(data, basis, lam=0.5, n_iter=15):
    """Batched Newton iterations on the ridge-penalised Bernoulli-probit
    log-likelihood of every covered pair (the role of
    BPRMeth::infer_profiles_mle(model = "bernoulli", lambda = .5, opt_itnmax = 15),
    call site R/melissa_vb.R:136-145).  Returns [P, D]."""
    H = _design(data.x, basis)
    D = H.shape[1]
    P = data.P
    ptr = data.pair_ptr
    pair_of = np.repeat(np.arange(P), np.diff(ptr))
    s = np.where(data.y == 1, 1.0, -1.0)
    w = np.zeros((P, D))
    eye = np.eye(D)[None] * (2.0 * lam)
    for _ in range(n_iter):
        t = s * np.einsum("id,id->i", H, w[pair_of])
        lam_t = np.exp(-0.5 * t * t - log_ndtr(t)) / SQRT_2PI          # phi(t) / Phi(t)
        grad_i = s * lam_t                                             # d/d(eta) of log Phi(s eta)
        hess_i = lam_t * (t + lam_t)                                   # -(d2/d eta2)
        g = np.add.reduceat(H * grad_i[:, None], ptr[:-1], axis=0) - 2.0 * lam * w
        A = np.add.reduceat(H[:, :, None] * H[:, None, :] * hess_i[:, None, None], ptr[:-1], axis=0) + eye
        step = np.linalg.solve(A, g[:, :, None])[:, :, 0]
        w = w + step
    return w


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
    try:
        w_mle = engine.profiles_mle(0.5, 15)             # GPU (sliced layout)
    except Exception as exc:                             # flat layout: the same Newton iterations in NumPy
        if "sliced layout" not in str(exc):
            raise
        w_mle = probit_mle(data, basis)
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
    return optimal_w
