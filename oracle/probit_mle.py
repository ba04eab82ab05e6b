"""CPU checker for melissa_profiles_mle: the penalised probit maximum-likelihood profile of ONE pair, stated
independently of the product code.  TEST INFRASTRUCTURE -- imported by tests/ only.

The reference's default initialisation calls BPRMeth::infer_profiles_mle(X[[n]], model = "bernoulli", basis, H,
lambda = .5, opt_itnmax = 15) (reference R/melissa_vb.R:136-145); BPRMeth is not in the reference tree,
so what is pinned here is its published objective for the Bernoulli model,

    maximise   sum_i [ y_i log Phi(h_i.w) + (1 - y_i) log(1 - Phi(h_i.w)) ]  -  lambda |w|^2,   h_i = [1, rbf_1(x_i), ...]

solved to convergence with SciPy's BFGS on the exact gradient.  The objective is strictly concave (lambda > 0), so the
optimum is unique and any convergent method must land on it: the CUDA kernel's Newton iterations are checked against
this optimum, not against another copy of themselves.
"""
import numpy as np
from scipy.optimize import minimize
from scipy.special import log_ndtr


def design(x, basis):
    mus = np.asarray(basis["mus"], np.float64)
    H = np.ones((x.shape[0], mus.shape[0] + 1))
    if mus.shape[0]:
        H[:, 1:] = np.exp(-float(basis["gamma"]) * (x[:, None] - mus[None, :]) ** 2)
    return H


def pair_mle(x, y, basis, lam=0.5):
    """argmax of the penalised Bernoulli-probit log-likelihood of one (cell, region) pair."""
    H = design(np.asarray(x, np.float64), basis)
    s = np.where(np.asarray(y) == 1, 1.0, -1.0)

    def neg(w):
        t = s * (H @ w)
        ll = log_ndtr(t)
        ratio = np.exp(-0.5 * t * t - ll) / np.sqrt(2.0 * np.pi)          # phi(t) / Phi(t)
        return -(ll.sum() - lam * w @ w), -((s * ratio) @ H - 2.0 * lam * w)

    res = minimize(neg, np.zeros(H.shape[1]), jac=True, method="BFGS", options=dict(gtol=1e-11, maxiter=500))
    return res.x


def profiles_mle(d, basis, lam=0.5, pairs=None):
    """[len(pairs), D] optima for the listed pairs (all pairs if None) of a data dict."""
    pp = d["pair_ptr"]
    pairs = range(pp.shape[0] - 1) if pairs is None else pairs
    return np.stack([pair_mle(d["x"][pp[p]:pp[p + 1]], d["y"][pp[p]:pp[p + 1]], basis, lam) for p in pairs])
