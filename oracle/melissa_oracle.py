"""CPU oracle for the Melissa variational-Bayes hot path.  TEST INFRASTRUCTURE ONLY.

This file is a NumPy/SciPy FP64 restatement of the reference's R algorithm.  It
is the checker that `tests/`, `__graft_entry__.smoke()` and `bench.py`'s
`cpu_baseline` leg use; the product (`melissa_b200/`) never imports it and
fails loudly when its CUDA library is missing.

PARITY UNPINNED: the reference (andreaskapou/Melissa, R package 1.5.3) is pure
interpreted R, R is not installed in the build container, and every numeric
assertion on `melissa()` in the reference's own tests is commented out
(reference `tests/testthat/test-running_melissa.R:13-79`).  The only live golden
value on this path is the `log_sum_exp` bracket in
`tests/testthat/test-filtering.R:37-41`, which `tests/test_oracle.py` checks.
Beyond that the oracle is anchored on (i) a literal nested-loop transcription
(`melissa_inner_literal`) that mirrors the R statement order and is compared
with the vectorised form, (ii) recovery of the ground truth (`C_true`, `pi_k`)
shipped inside the reference's bundled datasets, and (iii) ELBO monotonicity.

Third-party arithmetic: the RBF design matrix comes from BPRMeth (Bioconductor,
version unpinned in reference `DESCRIPTION:11-14`, source not in the tree).
`rbf_design` restates it from the serialised `rbf` objects inside the bundled
`.rda` files (fields M, mus, gamma): H = [1, exp(-gamma (x - mu_j)^2)], with
gamma = M^2/4 and mu_j = j*2/(M+1) - 1.

Every function cites the reference lines it follows (paths relative to the
reference root).
"""
from __future__ import annotations

import math

import numpy as np
from scipy.special import digamma, gammaln, ndtr

SQRT_2PI = math.sqrt(2.0 * math.pi)


# --------------------------------------------------------------------------
# helpers
# --------------------------------------------------------------------------
def log_sum_exp(x):
    """R/utils.R:16-20."""
    x = np.asarray(x, dtype=np.float64)
    offset = np.max(x)
    return math.log(np.sum(np.exp(x - offset))) + offset


def create_rbf_object(M=3):
    """BPRMeth::create_rbf_object as pinned from the bundled basis objects
    (R/melissa_vb.R:98 call site): equally spaced centres, gamma = M^2/4."""
    M = int(M)
    if M == 0:
        return dict(M=0, mus=np.zeros(0), gamma=1.0)
    mus = np.array([j * (2.0 / (M + 1)) - 1.0 for j in range(1, M + 1)], dtype=np.float64)
    return dict(M=M, mus=mus, gamma=M * M / 4.0)


def rbf_design(x, basis):
    """BPRMeth::design_matrix for an rbf object (call site R/utils.R:73).
    Returns H [len(x), D] with D = basis M + 1 and a leading column of ones."""
    x = np.asarray(x, dtype=np.float64)
    mus = np.asarray(basis["mus"], dtype=np.float64)
    H = np.ones((x.shape[0], mus.shape[0] + 1), dtype=np.float64)
    if mus.shape[0]:
        d = x[:, None] - mus[None, :]
        H[:, 1:] = np.exp(-basis["gamma"] * d * d)
    return H


def dnorm(x):
    return np.exp(-0.5 * x * x) / SQRT_2PI


def pnorm(x):
    return ndtr(x)


def update_Ez(mu, y, mimic=True):
    """R/melissa_vb.R:222-226.  `mimic=True` evaluates the reference's own
    expressions; `mimic=False` is the cancellation-free Mills-ratio form the
    CUDA path uses (they differ only where |mu| is large on the wrong side)."""
    mu = np.asarray(mu, dtype=np.float64)
    Ez = np.empty_like(mu)
    y1 = y == 1
    y0 = ~y1
    if mimic:
        Ez[y1] = mu[y1] + dnorm(-mu[y1]) / (1.0 - pnorm(-mu[y1]))
        Ez[y0] = mu[y0] - dnorm(-mu[y0]) / pnorm(-mu[y0])
    else:
        from scipy.special import erfcx
        t = np.where(y1, mu, -mu)
        v = np.abs(t) / math.sqrt(2.0)
        e = erfcx(v)
        # Phi(t)/phi(t) = sqrt(pi/2) erfcx(-t/sqrt2)
        ratio = np.where(t <= 0, math.sqrt(math.pi / 2) * e,
                         SQRT_2PI * np.exp(0.5 * t * t) - math.sqrt(math.pi / 2) * e)
        G = t + 1.0 / ratio
        Ez = np.where(y1, G, -G)
    return Ez


def _segment_sum(values, ptr):
    """Sum rows of `values` over segments [ptr[s], ptr[s+1]); empty segments give 0."""
    nseg = ptr.shape[0] - 1
    out = np.zeros((nseg,) + values.shape[1:], dtype=np.float64)
    nonempty = ptr[1:] > ptr[:-1]
    if values.shape[0] and nonempty.any():
        starts = ptr[:-1][nonempty]
        out[nonempty] = np.add.reduceat(values, starts, axis=0)
    return out


def _group_sum(values, keys, nkeys):
    """Sum rows of `values` grouped by integer key (stable order)."""
    order = np.argsort(keys, kind="stable")
    counts = np.bincount(keys, minlength=nkeys)
    ptr = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    return _segment_sum(values[order], ptr)


# --------------------------------------------------------------------------
# the VB loop, vectorised
# --------------------------------------------------------------------------
def melissa_inner(data, K, basis, w, delta_0, alpha_0=0.5, beta_0=None,
                  vb_max_iter=300, epsilon_conv=1e-5, mimic=True,
                  lb_every=10, trace=None):
    """R/melissa_vb.R:232-529 (`melissa_inner`), vectorised over CpGs and pairs.

    data: dict with N, M, pair_cell, pair_region, pair_ptr, x, y (cell-major CSR).
    w: [M, D, K] initial means; delta_0: [K].
    Returns a dict with the fields of the reference's result list (:523-528)
    plus `flags` (bit 0 lower bound decreased :495-497, bit 1 did not
    converge :511-513) and `n_iter` (last i executed).
    `trace`, if a dict, receives per-pair H^T E[z] after the set-up pass.
    """
    N, M = int(data["N"]), int(data["M"])
    pc = np.asarray(data["pair_cell"], dtype=np.int64)
    pr = np.asarray(data["pair_region"], dtype=np.int64)
    pp = np.asarray(data["pair_ptr"], dtype=np.int64)
    x = np.asarray(data["x"], dtype=np.float64)
    y = np.asarray(data["y"]).astype(np.float64)
    yb = y == 1
    D = int(basis["M"]) + 1
    P = pc.shape[0]
    lens = np.diff(pp)
    cpg_pair = np.repeat(np.arange(P), lens)
    delta_0 = np.asarray(delta_0, dtype=np.float64).copy()
    w = np.asarray(w, dtype=np.float64)
    assert w.shape == (M, D, K), (w.shape, (M, D, K))
    assert N > 0 and M > 0                                           # :240-241

    H = rbf_design(x, basis)                                         # R/utils.R:64-75
    HH = _segment_sum(H[:, :, None] * H[:, None, :], pp)            # :254-255
    LB = [-np.inf]                                                   # :242
    E_ww = np.zeros(K)                                               # :244
    alpha_k = np.full(K, alpha_0 + M * D / 2.0)                     # :264
    m_k = w.copy()                                                   # :266
    S_k = np.tile(np.linalg.inv(np.diag(np.full(D, 50.0)))[None, None], (K, M, 1, 1))  # :268-269
    if beta_0 is None:
        beta_0 = math.sqrt(alpha_k[0])                               # :271-273
    beta_k = np.full(K, float(beta_0))                               # :274
    delta_k = delta_0.copy()                                         # :276
    e_log_pi = digamma(delta_k) - digamma(delta_k.sum())             # :278

    def make_mk_Sk():                                                # :279-280, :450-451
        mm = np.einsum("mdk,mek->kmde", m_k, m_k)
        return mm + S_k                                              # [K, M, D, D]

    mk_Sk = make_mk_Sk()
    wbar = m_k.mean(axis=2)[pr]                                      # :291 rowMeans
    mu = np.einsum("id,id->i", H, wbar[cpg_pair])                    # :283-293
    E_z = update_Ez(mu, yb, mimic)                                   # :295-299
    E_zz = 1.0 + mu * E_z                                            # :300-302
    if trace is not None:
        trace["HtEz_init"] = _segment_sum(H * E_z[:, None], pp)

    cell_ptr = np.concatenate([[0], np.cumsum(np.bincount(pc, minlength=N))]).astype(np.int64)
    assert np.all(np.diff(pc) >= 0), "pairs must be cell-major"
    flags = 0
    n_iter = 1
    pi_k = None
    r_nk = np.zeros((N, K))
    log_r_nk = np.zeros((N, K))
    for i in range(2, vb_max_iter + 1):                              # :309
        n_iter = i
        # ---- E-step :322-327
        HtEz = _segment_sum(H * E_z[:, None], pp)                    # t(H) %*% E_z
        ezz2 = _segment_sum((E_zz * E_zz)[:, None], pp)[:, 0]       # crossprod(E_zz)
        lin = np.einsum("pdk,pd->pk", m_k[pr], HtEz)
        tr = np.einsum("pde,kped->pk", HH, mk_Sk[:, pr])             # tr(HH mk_Sk)
        per_pair = -0.5 * ezz2[:, None] + lin - 0.5 * tr
        log_rho = e_log_pi[None, :] + _segment_sum(per_pair, cell_ptr)
        lse = np.array([log_sum_exp(row) for row in log_rho])        # :330
        log_r_nk = log_rho - lse[:, None]
        r_nk = np.exp(log_r_nk)                                      # :331
        # ---- M-step
        cs = r_nk.sum(axis=0)
        delta_k = delta_0 + cs                                       # :336
        pi_k = (delta_0 + cs) / (K * delta_0 + N)                    # :338
        rp = r_nk[pc]                                                # [P, K]
        for k in range(K):                                           # :341-393
            A = _group_sum(rp[:, k, None, None] * HH, pr, M)
            A = A + np.eye(D)[None] * (alpha_k[k] / beta_k[k] + 1e-7)
            Sk = np.linalg.inv(A)                                    # :368
            b = _group_sum(rp[:, k, None] * HtEz, pr, M)
            mk = np.einsum("mde,me->md", Sk, b)                      # :372
            S_k[k] = Sk                                              # :380
            m_k[:, :, k] = mk                                        # :382
            E_ww[k] = np.sum(np.einsum("md,md->m", mk, mk) + np.trace(Sk, axis1=1, axis2=2))  # :375,:384
            beta_k[k] = beta_0 + 0.5 * E_ww[k]                       # :386
            if beta_k[k] > 10 * alpha_k[k]:                          # :390-392
                beta_k[k] = 10 * alpha_k[k]
        # ---- refresh :418-435
        wbar = np.einsum("pk,pdk->pd", rp, m_k[pr])
        mu = np.einsum("id,id->i", H, wbar[cpg_pair])
        E_z = update_Ez(mu, yb, mimic)
        E_zz = 1.0 + mu * E_z
        e_log_pi = digamma(delta_k) - digamma(delta_k.sum())         # :443
        E_alpha = alpha_k / beta_k                                   # :445
        mk_Sk = make_mk_Sk()                                         # :450-451
        # ---- lower bound :458-509
        if i % lb_every == 0 or i == vb_max_iter:
            pm = pnorm(-mu)
            ll = y * np.log(1.0 - (pm - 1e-10)) + (1.0 - y) * np.log(pm + 1e-10)   # :471-472
            tr_new = np.einsum("pde,kped->pk", HH, mk_Sk[:, pr])
            per_pair_lb = (0.5 * _segment_sum((mu * mu)[:, None], pp)[:, 0]
                           + _segment_sum(ll[:, None], pp)[:, 0]
                           - 0.5 * np.sum(rp * tr_new, axis=1))
            lb_pz_qz = per_pair_lb.sum()
            dg = digamma(alpha_k) - np.log(beta_k)
            lb_p_w = np.sum(-0.5 * M * D * math.log(2 * math.pi) + 0.5 * M * D * dg - 0.5 * E_alpha * E_ww)
            lb_p_c = np.sum(r_nk @ e_log_pi)
            lb_p_pi = np.sum((delta_0 - 1) * e_log_pi) + gammaln(delta_0.sum()) - gammaln(delta_0).sum()
            lb_p_tau = np.sum(alpha_0 * math.log(beta_0) + (alpha_0 - 1) * dg - beta_0 * E_alpha - gammaln(alpha_0))
            lb_q_c = np.sum(r_nk * log_r_nk)
            lb_q_pi = np.sum((delta_k - 1) * e_log_pi) + gammaln(delta_k.sum()) - gammaln(delta_k).sum()
            lb_q_w = np.sum(-0.5 * np.log(np.linalg.det(S_k)) - 0.5 * D * (1 + math.log(2 * math.pi)))
            lb_q_tau = np.sum(-gammaln(alpha_k) + (alpha_k - 1) * digamma(alpha_k) + np.log(beta_k) - alpha_k)
            LB.append(lb_pz_qz + lb_p_c + lb_p_pi + lb_p_w + lb_p_tau - lb_q_c - lb_q_pi - lb_q_w - lb_q_tau)
            if LB[-1] < LB[-2]:                                      # :495-497
                flags |= 1
            if abs(LB[-1] - LB[-2]) < epsilon_conv:                  # :499-501
                break
        if i == vb_max_iter:                                         # :511-513
            flags |= 2
    return dict(W=m_k, W_Sigma=S_k, r_nk=r_nk, delta=delta_k, alpha=alpha_k, beta=beta_k,
                basis=basis, pi_k=pi_k, lb=np.asarray(LB), delta_0=delta_0, alpha_0=alpha_0,
                beta_0=beta_0, flags=flags, n_iter=n_iter, E_ww=E_ww.copy())


# --------------------------------------------------------------------------
# the VB loop, literal nested loops (small inputs only)
# --------------------------------------------------------------------------
def melissa_inner_literal(data, K, basis, w, delta_0, alpha_0=0.5, beta_0=None,
                          vb_max_iter=300, epsilon_conv=1e-5):
    """R/melissa_vb.R:232-529 transcribed statement by statement with the
    reference's own nested cell/region/cluster loops and per-pair list
    structure.  Slow; used to validate `melissa_inner` on small inputs."""
    N, M = int(data["N"]), int(data["M"])
    D = int(basis["M"]) + 1
    pp = data["pair_ptr"]
    H = [[None] * M for _ in range(N)]
    yv = [[None] * M for _ in range(N)]
    for p, (n, m) in enumerate(zip(data["pair_cell"], data["pair_region"])):
        sl = slice(int(pp[p]), int(pp[p + 1]))
        H[n][m] = rbf_design(data["x"][sl], basis)
        yv[n][m] = np.asarray(data["y"][sl]).astype(np.float64)
    region_ind = [[m for m in range(M) if H[n][m] is not None] for n in range(N)]   # :119
    cell_ind = [[n for n in range(N) if H[n][m] is not None] for m in range(M)]     # :121
    delta_0 = np.asarray(delta_0, dtype=np.float64)
    LB = [-np.inf]
    E_ww = np.zeros(K)
    HH = [[None] * M for _ in range(N)]
    for n in range(N):
        for m in region_ind[n]:
            HH[n][m] = H[n][m].T @ H[n][m]                           # :254-255
    alpha_k = np.full(K, alpha_0 + M * D / 2.0)
    m_k = np.array(w, dtype=np.float64)
    S_k = [[np.linalg.inv(np.diag(np.full(D, 50.0))) for _ in range(M)] for _ in range(K)]
    if beta_0 is None:
        beta_0 = math.sqrt(alpha_k[0])
    beta_k = np.full(K, float(beta_0))
    delta_k = delta_0.copy()
    e_log_pi = digamma(delta_k) - digamma(delta_k.sum())
    mk_Sk = [[np.outer(m_k[m, :, k], m_k[m, :, k]) + S_k[k][m] for k in range(K)] for m in range(M)]
    mu = [[None] * M for _ in range(N)]
    E_z = [[None] * M for _ in range(N)]
    E_zz = [[None] * M for _ in range(N)]
    for n in range(N):
        for m in region_ind[n]:
            mu[n][m] = H[n][m] @ (m_k[m].sum(axis=1) / K)            # :291
            E_z[n][m] = update_Ez(mu[n][m], yv[n][m] == 1, True)
            E_zz[n][m] = 1 + mu[n][m] * E_z[n][m]
    flags, pi_k, n_iter = 0, None, 1
    r_nk = np.zeros((N, K))
    log_r_nk = np.zeros((N, K))
    for i in range(2, vb_max_iter + 1):
        n_iter = i
        log_rho = np.zeros((N, K))
        for k in range(K):                                           # :322-327
            for n in range(N):
                s = 0.0
                for m in region_ind[n]:
                    s += (-0.5 * (E_zz[n][m] @ E_zz[n][m])
                          + m_k[m, :, k] @ H[n][m].T @ E_z[n][m]
                          - 0.5 * np.trace(HH[n][m] @ mk_Sk[m][k]))
                log_rho[n, k] = e_log_pi[k] + s
        for n in range(N):
            log_r_nk[n] = log_rho[n] - log_sum_exp(log_rho[n])
        r_nk = np.exp(log_r_nk)
        delta_k = delta_0 + r_nk.sum(axis=0)
        pi_k = (delta_0 + r_nk.sum(axis=0)) / (K * delta_0 + N)
        for k in range(K):
            tot = 0.0
            for m in range(M):
                A = np.diag(np.full(D, alpha_k[k] / beta_k[k] + 1e-7))
                acc = None
                for n in cell_ind[m]:
                    term = HH[n][m] * r_nk[n, k]
                    acc = term if acc is None else acc + term
                Sm = np.linalg.inv(A + acc)
                acc = None
                for n in cell_ind[m]:
                    term = H[n][m].T @ E_z[n][m] * r_nk[n, k]
                    acc = term if acc is None else acc + term
                mm = Sm @ acc
                S_k[k][m] = Sm
                m_k[m, :, k] = mm
                tot += mm @ mm + np.trace(Sm)
            E_ww[k] = tot
            beta_k[k] = beta_0 + 0.5 * E_ww[k]
            if beta_k[k] > 10 * alpha_k[k]:
                beta_k[k] = 10 * alpha_k[k]
        for n in range(N):                                           # :418-435
            for m in region_ind[n]:
                wb = np.zeros(D)
                for k in range(K):
                    wb = wb + r_nk[n, k] * m_k[m, :, k]
                mu[n][m] = H[n][m] @ wb
                E_z[n][m] = update_Ez(mu[n][m], yv[n][m] == 1, True)
                E_zz[n][m] = 1 + mu[n][m] * E_z[n][m]
        e_log_pi = digamma(delta_k) - digamma(delta_k.sum())
        E_alpha = alpha_k / beta_k
        mk_Sk = [[np.outer(m_k[m, :, k], m_k[m, :, k]) + S_k[k][m] for k in range(K)] for m in range(M)]
        if i % 10 == 0 or i == vb_max_iter:
            lb_pz_qz = 0.0
            for n in range(N):
                for m in region_ind[n]:
                    pm = pnorm(-mu[n][m])
                    yy = yv[n][m]
                    lb_pz_qz += (0.5 * (mu[n][m] @ mu[n][m])
                                 + np.sum(yy * np.log(1 - (pm - 1e-10)) + (1 - yy) * np.log(pm + 1e-10))
                                 - 0.5 * sum(r_nk[n, k] * np.trace(HH[n][m] @ mk_Sk[m][k]) for k in range(K)))
            dg = digamma(alpha_k) - np.log(beta_k)
            lb_p_w = np.sum(-0.5 * M * D * math.log(2 * math.pi) + 0.5 * M * D * dg - 0.5 * E_alpha * E_ww)
            lb_p_c = np.sum(r_nk @ e_log_pi)
            lb_p_pi = np.sum((delta_0 - 1) * e_log_pi) + gammaln(delta_0.sum()) - gammaln(delta_0).sum()
            lb_p_tau = np.sum(alpha_0 * math.log(beta_0) + (alpha_0 - 1) * dg - beta_0 * E_alpha - gammaln(alpha_0))
            lb_q_c = np.sum(r_nk * log_r_nk)
            lb_q_pi = np.sum((delta_k - 1) * e_log_pi) + gammaln(delta_k.sum()) - gammaln(delta_k).sum()
            lb_q_w = sum(sum(-0.5 * math.log(np.linalg.det(S_k[k][m])) for k in range(K))
                         - 0.0 for m in range(M)) - 0.5 * D * (1 + math.log(2 * math.pi)) * M * K
            lb_q_tau = np.sum(-gammaln(alpha_k) + (alpha_k - 1) * digamma(alpha_k) + np.log(beta_k) - alpha_k)
            LB.append(lb_pz_qz + lb_p_c + lb_p_pi + lb_p_w + lb_p_tau - lb_q_c - lb_q_pi - lb_q_w - lb_q_tau)
            if LB[-1] < LB[-2]:
                flags |= 1
            if abs(LB[-1] - LB[-2]) < epsilon_conv:
                break
        if i == vb_max_iter:
            flags |= 2
    W_Sigma = np.array([[S_k[k][m] for m in range(M)] for k in range(K)])
    return dict(W=m_k, W_Sigma=W_Sigma, r_nk=r_nk, delta=delta_k, alpha=alpha_k, beta=beta_k,
                pi_k=pi_k, lb=np.asarray(LB), flags=flags, n_iter=n_iter)


def finalize(obj, K):
    """R/melissa_vb.R:204-216: hard labels (first arg-max, 1-based) and
    lb - log(K!)."""
    r = obj["r_nk"]
    labels = np.array([int(np.flatnonzero(row == np.max(row))[0]) + 1 for row in r])
    out = dict(obj)
    out["labels"] = labels
    out["lb"] = obj["lb"] - math.log(math.factorial(K))
    return out


# --------------------------------------------------------------------------
# config-2 tail: imputation and its metrics
# --------------------------------------------------------------------------
def r_round(v, digits):
    """R's round(): IEEE round-half-even on the scaled value (R >= 4.0.0 picks
    the closer representable candidate; for probabilities in [0,1] at 2-3
    digits this agrees with numpy's algorithm)."""
    return np.round(np.asarray(v, dtype=np.float64), digits)


def impute_test_met(obj, test, basis=None):
    """R/utils.R:225-284 with `basis = NULL`, `is_predictive = TRUE`.
    `test`: CSR dict of held-out CpGs.  Returns (act_obs, pred_obs)."""
    basis = obj["basis"] if basis is None else basis
    pc, pr, pp = test["pair_cell"], test["pair_region"], test["pair_ptr"]
    H = rbf_design(test["x"], basis)
    lens = np.diff(pp)
    cp = np.repeat(np.arange(pc.shape[0]), lens)
    K = obj["r_nk"].shape[1]
    pred = np.zeros(H.shape[0])
    for k in range(K):                                               # :246-252
        wk = obj["W"][pr, :, k][cp]
        pred = pred + obj["r_nk"][pc, k][cp] * pnorm(np.einsum("id,id->i", H, wk))
    pred = r_round(pred, 3)                                          # :265
    return np.asarray(test["y"]).astype(np.float64), pred


def eval_imputation_performance(act_obs, pred_obs):
    """R/utils.R:514-529 (ROCR::prediction / performance restated): AUC with
    ties, and the F-measure at the first cutoff <= 0.5."""
    pred = r_round(pred_obs, 2)                                      # :516
    act = np.asarray(act_obs) == 1
    npos, nneg = int(act.sum()), int((~act).sum())
    order = np.argsort(-pred, kind="stable")
    ps, as_ = pred[order], act[order]
    distinct = np.flatnonzero(np.diff(ps) != 0)
    idx = np.concatenate([distinct, [ps.shape[0] - 1]])
    tp = np.concatenate([[0], np.cumsum(as_)[idx]]).astype(np.float64)
    fp = np.concatenate([[0], np.cumsum(~as_)[idx]]).astype(np.float64)
    cutoffs = np.concatenate([[np.inf], ps[idx]])
    tpr, fpr = tp / npos, fp / nneg
    auc = float(np.sum(0.5 * (tpr[1:] + tpr[:-1]) * (fpr[1:] - fpr[:-1])))
    with np.errstate(divide="ignore", invalid="ignore"):
        prec = tp / (tp + fp)
        rec = tp / npos
        f = 1.0 / (0.5 / prec + 0.5 / rec)
    j = int(np.flatnonzero(cutoffs <= 0.5)[0])                       # :523
    return dict(auc=auc, f_measure=float(f[j]))


def adjusted_rand_index(a, b):
    """mclust::adjustedRandIndex (call site R/utils.R:632)."""
    a = np.asarray(a)
    b = np.asarray(b)
    ua, ia = np.unique(a, return_inverse=True)
    ub, ib = np.unique(b, return_inverse=True)
    tab = np.zeros((ua.shape[0], ub.shape[0]), dtype=np.int64)
    np.add.at(tab, (ia, ib), 1)
    if tab.shape == (1, 1):
        return 1.0
    comb = lambda v: v * (v - 1) / 2.0
    sa = comb(tab.sum(axis=1)).sum()
    sb = comb(tab.sum(axis=0)).sum()
    st = comb(tab).sum()
    tot = comb(a.shape[0])
    expected = sa * sb / tot
    return float((st - expected) / (0.5 * (sa + sb) - expected))


def cluster_ari(C_true, r_nk):
    """R/utils.R:622-634."""
    lt = np.array([int(np.flatnonzero(r == r.max())[0]) for r in np.asarray(C_true)])
    le = np.array([int(np.flatnonzero(r == r.max())[0]) for r in np.asarray(r_nk)])
    return adjusted_rand_index(lt, le)
