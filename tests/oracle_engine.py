"""Test double for melissa_b200.engine.Engine built on the CPU oracle, so the host
loop (vb_loop) and the cell-sharded path can be exercised without a GPU.
TEST INFRASTRUCTURE: lives under tests/, imports oracle/."""
import math

import numpy as np
import torch
from scipy.special import digamma, gammaln

import melissa_oracle as mo


class OracleEngine:
    """Same phases as the CUDA library and the same kind of reduction buffers (its own layout:
    red = [acc (M*K*(D + D*D), full H^T H instead of the packed triangle) | colsum (K) | sum r log r],
    lbbuf = [lb_pz_qz, count]); what is shared with the product is vb.vb_loop, not the layout."""

    def __init__(self, data, basis):
        self.d, self.basis = data, basis
        self.N, self.M = int(data["N"]), int(data["M"])
        self.D = int(basis["M"]) + 1
        self.pc = np.asarray(data["pair_cell"], np.int64)
        self.pr = np.asarray(data["pair_region"], np.int64)
        self.pp = np.asarray(data["pair_ptr"], np.int64)
        self.H = mo.rbf_design(data["x"], basis)
        self.y = np.asarray(data["y"]) == 1
        self.HH = mo._segment_sum(self.H[:, :, None] * self.H[:, None, :], self.pp)
        self.cpg_pair = np.repeat(np.arange(self.pc.shape[0]), np.diff(self.pp))
        self.cell_ptr = np.concatenate([[0], np.cumsum(np.bincount(self.pc, minlength=self.N))]).astype(np.int64)

    def begin(self, K, w0, delta_0, alpha_0=0.5, beta_0=None, n_cells_total=None, stream=0):
        M, D = self.M, self.D
        self.K = K
        self.N_total = self.N if n_cells_total is None else n_cells_total
        self.alpha_0 = alpha_0
        self.alpha_k = alpha_0 + M * D / 2.0
        self.beta_0 = math.sqrt(self.alpha_k) if beta_0 is None else beta_0
        self.delta_0 = np.asarray(delta_0, np.float64)
        self.m_k = np.array(w0, np.float64)
        self.S_k = np.tile((np.eye(D) / 50.0)[None, None], (K, M, 1, 1))
        self.beta = np.full(K, self.beta_0)
        self.delta = self.delta_0.copy()
        self.E_ww = np.zeros(K)
        self.e_log_pi = digamma(self.delta) - digamma(self.delta.sum())
        self.red = torch.zeros(M * K * (D + D * D) + K + 1, dtype=torch.float64)
        self.lbbuf = torch.zeros(2, dtype=torch.float64)
        self._sweep(np.full((self.N, K), 1.0 / K), False)

    def _mk_Sk(self):
        return np.einsum("mdk,mek->kmde", self.m_k, self.m_k) + self.S_k

    def _sweep(self, r, with_lb):
        wbar = np.einsum("pk,pdk->pd", r[self.pc], self.m_k[self.pr])
        mu = np.einsum("id,id->i", self.H, wbar[self.cpg_pair])
        Ez = mo.update_Ez(mu, self.y, False)
        self.HtEz = mo._segment_sum(self.H * Ez[:, None], self.pp)
        mkSk = self._mk_Sk()
        lin = np.einsum("pdk,pd->pk", self.m_k[self.pr], self.HtEz)
        tr = np.einsum("pde,kped->pk", self.HH, mkSk[:, self.pr])
        self.log_rho = mo._segment_sum(lin - 0.5 * tr, self.cell_ptr)
        if with_lb:
            pm = mo.pnorm(-mu)
            yf = self.y.astype(np.float64)
            ll = yf * np.log(1 - (pm - 1e-10)) + (1 - yf) * np.log(pm + 1e-10)
            self.lbbuf[0] = float(0.5 * mu @ mu + ll.sum() - 0.5 * np.sum(r[self.pc] * tr))
            self.lbbuf[1] = 0.0

    def phase_a(self):
        K, M, D = self.K, self.M, self.D
        lr = self.log_rho + self.e_log_pi[None]
        lse = np.array([mo.log_sum_exp(v) for v in lr])
        self.log_r = lr - lse[:, None]
        self.r = np.exp(self.log_r)
        rp = self.r[self.pc]
        acc_b = np.stack([mo._group_sum(rp[:, k, None] * self.HtEz, self.pr, M) for k in range(K)], 1)
        acc_A = np.stack([mo._group_sum(rp[:, k, None, None] * self.HH, self.pr, M) for k in range(K)], 1)
        flat = np.concatenate([acc_b.reshape(M * K, D), acc_A.reshape(M * K, D * D)], axis=1).ravel()
        self.red[: flat.shape[0]] = torch.from_numpy(flat)
        self.red[flat.shape[0]: flat.shape[0] + K] = torch.from_numpy(self.r.sum(0))
        self.red[-1] = float(np.sum(self.r * self.log_r))

    def reduce_tensor(self):
        return self.red

    def lb_tensor(self):
        return self.lbbuf

    def phase_b(self, with_lb):
        K, M, D = self.K, self.M, self.D
        red = self.red.numpy()
        rec = red[: M * K * (D + D * D)].reshape(M, K, D + D * D)
        b, A = rec[:, :, :D], rec[:, :, D:].reshape(M, K, D, D)
        self.colsum = red[M * K * (D + D * D): M * K * (D + D * D) + K].copy()
        self.rlogr = float(red[-1])
        for k in range(K):
            Ak = A[:, k] + np.eye(D)[None] * (self.alpha_k / self.beta[k] + 1e-7)
            Sk = np.linalg.inv(Ak)
            mk = np.einsum("mde,me->md", Sk, b[:, k])
            self.S_k[k], self.m_k[:, :, k] = Sk, mk
            self.E_ww[k] = np.sum(np.einsum("md,md->m", mk, mk) + np.trace(Sk, axis1=1, axis2=2))
            self.beta[k] = min(self.beta_0 + 0.5 * self.E_ww[k], 10 * self.alpha_k)
        self.delta = self.delta_0 + self.colsum
        self.pi = (self.delta_0 + self.colsum) / (K * self.delta_0 + self.N_total)
        self.e_log_pi = digamma(self.delta) - digamma(self.delta.sum())
        self._sweep(self.r, with_lb)

    def lower_bound(self):
        K, M, D = self.K, self.M, self.D
        a0, ak, b0 = self.alpha_0, self.alpha_k, self.beta_0
        dg = digamma(ak) - np.log(self.beta)
        Ea = ak / self.beta
        lb_p_w = np.sum(-0.5 * M * D * math.log(2 * math.pi) + 0.5 * M * D * dg - 0.5 * Ea * self.E_ww)
        lb_p_c = float(self.colsum @ self.e_log_pi)
        lb_p_pi = np.sum((self.delta_0 - 1) * self.e_log_pi) + gammaln(self.delta_0.sum()) - gammaln(self.delta_0).sum()
        lb_p_tau = np.sum(a0 * math.log(b0) + (a0 - 1) * dg - b0 * Ea - gammaln(a0))
        lb_q_pi = np.sum((self.delta - 1) * self.e_log_pi) + gammaln(self.delta.sum()) - gammaln(self.delta).sum()
        lb_q_w = np.sum(-0.5 * np.log(np.linalg.det(self.S_k)) - 0.5 * D * (1 + math.log(2 * math.pi)))
        lb_q_tau = np.sum(-gammaln(ak) + (ak - 1) * digamma(ak) + np.log(self.beta) - ak)
        return float(self.lbbuf[0]) + lb_p_c + lb_p_pi + lb_p_w + lb_p_tau - self.rlogr - lb_q_pi - lb_q_w - lb_q_tau

    def fetch(self):
        return dict(W=self.m_k.copy(), W_Sigma=self.S_k.copy(), r_nk=self.r.copy(), delta=self.delta.copy(),
                    beta=self.beta.copy(), pi_k=self.pi.copy())
