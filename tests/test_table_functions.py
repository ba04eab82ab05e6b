"""CPU: the two table-driven special functions of the sweep (rbf_row3 / trunc_mean3, melissa_b200/csrc/sliced.cuh), restated
operation by operation in NumPy and compared with mpmath at 40 digits.  This pins the table spacing
and the series orders chosen there; the device code itself is covered by the GPU parity tests."""
import math

import numpy as np
import pytest
from scipy.special import erfcx

mp = pytest.importorskip("mpmath")
mp.mp.dps = 40
MAGIC = 6755399441055744.0


def trunc_mean_table(t, sp=512, order=4):
    """E[z | z > 0], z ~ N(t, 1): t + 1 / R(t), R = Phi / phi from a table at spacing 1/sp and the
    recurrence c_{n+1} = (c_{n-1} + t0 c_n) / (n + 1) (trunc_mean3 in sliced.cuh: the rounding constant carries the
    scale and the table offset, the reciprocal is a 2^-23 seed plus one cubically convergent step)."""
    magic = MAGIC / sp + 8.0                                            # MILLS_MAGIC
    v = t + magic
    t0 = v - magic
    d = t - t0
    R0 = math.sqrt(math.pi / 2) * erfcx(-t0 / math.sqrt(2))           # the table entry
    c = [R0, 1 + t0 * R0]
    for n in range(1, order):
        c.append((c[n - 1] + t0 * c[n]) / (n + 1))
    R = c[order]
    for n in range(order - 1, -1, -1):
        R = R * d + c[n]
    y = (1.0 / R) * (1 + 2.0 ** -23)                                   # what MUFU.RCP64H guarantees
    e = 1 - R * y
    y = y * (e * e + e) + y
    return t + y


def log_phi_eps_table(t, sp=512):
    """log(Phi(t) + 1e-10) from the (L, L') table at spacing 1/sp and the recurrence of q = L', q' = -q (t + q)
    (log_phi_eps in sliced.cuh), operation by operation."""
    magic = MAGIC / sp + 8.0
    v = t + magic
    t0 = v - magic
    d = t - t0
    Phi0 = mp.ncdf(t0)
    L0 = float(mp.log(Phi0 + mp.mpf("1e-10")))                          # the table entry, correctly rounded
    a0 = float(mp.npdf(t0) / (Phi0 + mp.mpf("1e-10")))
    s2 = 2.0 * a0 + t0
    a1 = -a0 * (t0 + a0)
    a2 = -0.5 * (a1 * s2 + a0)
    a3 = -(1.0 / 3) * (a2 * s2 + (a1 * a1 + a1))
    p = d * (0.25 * a3) + (1.0 / 3) * a2
    p = d * p + 0.5 * a1
    p = d * p + a0
    return d * p + L0


def rbf_table_log2(M):
    """melissa.cu rbf_table_log2: grid spacing 2^-k with gamma s u_max <= 0.02 (series of order 6 / 5), 0.055 from
    D = 10 on (order 8 / 7)."""
    gamma, mu1 = M * M / 4.0, 2.0 / (M + 1) - 1.0
    u_max = max(abs(-1 - mu1), abs(1 - mu1))
    w_max = 0.02 if M + 1 <= 9 else 0.055
    for k in range(8, 13):
        if gamma * 2.0 ** -k * u_max <= w_max:
            return k
    raise AssertionError("gamma too large for the table")


def rbf_row_table(x, M):
    """RBF row from the (E, B) table on the grid x0 = j 2^-k and the short series in d = x - x0 (rbf_row3 in
    sliced.cuh)."""
    gamma, delta = M * M / 4.0, 2.0 / (M + 1)
    mus = np.array([j * (2.0 / (M + 1)) - 1.0 for j in range(1, M + 1)])
    k = rbf_table_log2(M)
    s = 2.0 ** -k
    nw = 6 if M + 1 <= 9 else 8
    magic = (MAGIC + (2 ** k + 1)) * s
    v = x + magic
    x0 = v - magic
    d = x - x0
    u0 = x0 - mus[0]
    q = math.exp(-gamma * delta * delta)
    E, B = np.exp(-gamma * u0 * u0), q * np.exp(2 * gamma * delta * u0)  # the table entries (q folded into B)
    w = d * (-gamma * d + (x0 * (-2 * gamma) + 2 * gamma * mus[0]))
    fact = [1.0 / math.factorial(i) for i in range(10)]
    c = np.full_like(w, fact[nw])
    for i in range(nw - 1, -1, -1):
        c = c * w + fact[i]
    z = 2 * gamma * delta * d
    e = np.full_like(z, fact[nw - 1])
    for i in range(nw - 2, -1, -1):
        e = e * z + fact[i]
    h = [E * c]
    bj = B * e
    for _ in range(1, M):
        h.append(h[-1] * bj)
        bj = bj * (q * q)
    return np.stack(h, 1), mus, gamma


def test_truncated_normal_mean_from_the_table_is_at_the_rounding_floor():
    rng = np.random.RandomState(0)
    for lo, hi, tol in ((-7.9, -4, 5e-14), (-4, 0, 2e-14), (0, 3, 1.5e-15), (3, 7.9, 5e-16)):
        t = rng.uniform(lo, hi, 1500)
        got = trunc_mean_table(t)
        ref = np.array([float(mp.mpf(float(v)) + mp.npdf(float(v)) / mp.ncdf(float(v))) for v in t])
        erf = t + 1.0 / (math.sqrt(math.pi / 2) * erfcx(-t / math.sqrt(2)))     # the erfcx form (v1 kernels)
        rel = np.abs(got - ref) / np.abs(ref)
        rel_erf = np.abs(erf - ref) / np.abs(ref)
        assert rel.max() < tol, (lo, hi, rel.max())
        assert rel.max() < 3 * max(rel_erf.max(), 2.3e-16)          # no worse than the closed form


def test_log_phi_from_the_table_is_at_the_rounding_floor():
    """The bound's per-CpG term log(Phi(t) + 1e-10) (R/melissa_vb.R:471-472): absolute error against 40-digit
    arithmetic, everywhere on the table including the stretch around t = -6.4 where the 1e-10 takes over."""
    rng = np.random.default_rng(5)
    ts = np.concatenate([rng.uniform(-7.99, 7.99, 4000), rng.uniform(-6.8, -6.0, 1000),
                         (np.arange(-4090, 4090, 37) + 0.4999) / 512.0])
    worst = 0.0
    for t in ts:
        want = mp.log(mp.ncdf(mp.mpf(float(t))) + mp.mpf("1e-10"))
        got = log_phi_eps_table(float(t))
        worst = max(worst, abs(float(got - want)) / max(1.0, abs(float(want))))
    assert worst < 1e-15, worst                       # a few ulp of a term that is -3 ... -23 there


@pytest.mark.parametrize("M", [1, 3, 4, 7, 9, 11, 15])
def test_rbf_row_from_the_table_matches_direct_evaluation(M):
    rng = np.random.RandomState(M)
    x = rng.uniform(-1, 1, 1500)
    H, mus, gamma = rbf_row_table(x, M)
    ref = np.array([[float(mp.e ** (-mp.mpf(gamma) * (mp.mpf(float(xx)) - mp.mpf(float(m))) ** 2)) for m in mus]
                    for xx in x])
    direct = np.exp(-gamma * (x[:, None] - mus[None]) ** 2)
    err = (np.abs(H - ref) / ref).max()
    err_direct = (np.abs(direct - ref) / ref).max()
    assert err < 2 * err_direct + 1e-15, (err, err_direct)      # as accurate as calling exp on the rounded argument
    assert err < 1e-14 if M <= 8 else err < 5e-14               # (the error of both grows with gamma (x - mu)^2)
