"""CPU: the two table-driven special functions of the sweep (melissa_b200/csrc/sliced.cuh), restated
operation by operation in NumPy and compared with mpmath at 40 digits.  This pins the table spacing
and the series orders chosen there; the device code itself is covered by the GPU parity tests."""
import math

import numpy as np
import pytest
from scipy.special import erfcx

mp = pytest.importorskip("mpmath")
mp.mp.dps = 40
MAGIC = 6755399441055744.0


def trunc_mean_table(t, sp=128, order=5):
    """E[z | z > 0], z ~ N(t, 1): t + 1 / R(t), R = Phi / phi from a table at spacing 1/sp and the
    recurrence c_{n+1} = (c_{n-1} + t0 c_n) / (n + 1) (trunc_mean_tab in sliced.cuh)."""
    v = t * sp + MAGIC
    t0 = (v - MAGIC) / sp
    d = t - t0
    R0 = math.sqrt(math.pi / 2) * erfcx(-t0 / math.sqrt(2))           # the table entry
    c = [R0, 1 + t0 * R0]
    for n in range(1, order):
        c.append((c[n - 1] + t0 * c[n]) / (n + 1))
    R = c[order]
    for n in range(order - 1, -1, -1):
        R = R * d + c[n]
    return t + 1.0 / R


def rbf_row_table(x, M, sub=128):
    """RBF row from the (E, B) table and the degree-7 / degree-6 series (eval_basis_tab in sliced.cuh)."""
    gamma, delta = M * M / 4.0, 2.0 / (M + 1)
    mus = np.array([j * (2.0 / (M + 1)) - 1.0 for j in range(1, M + 1)])
    s = delta / sub
    u = x - mus[0]
    v = u * (1.0 / s) + MAGIC
    u0 = (v - MAGIC) * s
    d = u - u0
    q = math.exp(-gamma * delta * delta)
    E, B = np.exp(-gamma * u0 * u0), q * np.exp(2 * gamma * delta * u0)  # the table entries (q folded into B)
    w = (-gamma * d) * (2 * u0 + d)
    c = np.full_like(w, 1.0 / 5040)
    for k in (720, 120, 24, 6, 2, 1, 1):
        c = c * w + 1.0 / k
    z = 2 * gamma * delta * d
    e = np.full_like(z, 1.0 / 720)
    for k in (120, 24, 6, 2, 1, 1):
        e = e * z + 1.0 / k
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


@pytest.mark.parametrize("M", [1, 3, 4, 7])
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
    assert err < 1e-14
