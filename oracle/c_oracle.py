"""ctypes wrapper of oracle/liboracle.so (the plain-C restatement).  TEST INFRASTRUCTURE."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "liboracle.so")


def load():
    if not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(os.path.join(HERE, "melissa_oracle.c")):
        subprocess.run(["make", "-C", HERE], check=True, capture_output=True)
    lib = C.CDLL(LIB)
    lib.oracle_num_threads.restype = C.c_int
    lib.oracle_vb_run.restype = C.c_int
    return lib


def vb_run(d, K, basis, w0, delta_0, alpha_0=0.5, beta_0=None, vb_max_iter=300, epsilon_conv=1e-5,
           lb_every=10, n_threads=0):
    lib = load()
    N, M = int(d["N"]), int(d["M"])
    D = int(basis["M"]) + 1
    pp = np.ascontiguousarray(d["pair_ptr"], np.int64)
    pc = np.ascontiguousarray(d["pair_cell"], np.int32)
    pr = np.ascontiguousarray(d["pair_region"], np.int32)
    x = np.ascontiguousarray(d["x"], np.float64)
    y = np.ascontiguousarray(d["y"], np.uint8)
    mus = np.ascontiguousarray(basis["mus"], np.float64)
    w0f = np.asfortranarray(np.asarray(w0, np.float64)).ravel(order="F").copy()
    d0 = np.ascontiguousarray(delta_0, np.float64)
    W = np.empty(M * D * K); WS = np.empty(K * M * D * D); r = np.empty(N * K)
    delta = np.empty(K); beta = np.empty(K); pi = np.empty(K)
    cap = vb_max_iter // lb_every + 3
    lb = np.empty(cap)
    n_iter, flags, loop_s = C.c_int(0), C.c_int(0), C.c_double(0.0)
    P = lambda a, t: a.ctypes.data_as(C.POINTER(t))
    n_lb = lib.oracle_vb_run(
        C.c_int(N), C.c_int(M), C.c_int64(pc.shape[0]), P(pp, C.c_int64), P(pc, C.c_int32), P(pr, C.c_int32),
        P(x, C.c_double), P(y, C.c_uint8), C.c_int(D - 1), P(mus, C.c_double), C.c_double(float(basis["gamma"])),
        C.c_int(K), P(w0f, C.c_double), P(d0, C.c_double), C.c_double(alpha_0),
        C.c_double(float("nan") if beta_0 is None else beta_0), C.c_int(vb_max_iter), C.c_double(epsilon_conv),
        C.c_int(lb_every), C.c_int(n_threads), P(W, C.c_double), P(WS, C.c_double), P(r, C.c_double),
        P(delta, C.c_double), P(beta, C.c_double), P(pi, C.c_double), P(lb, C.c_double), C.c_int(cap),
        C.byref(n_iter), C.byref(flags), C.byref(loop_s))
    if n_lb < 0:
        raise RuntimeError("oracle_vb_run failed")
    return dict(W=W.reshape((M, D, K), order="F"), W_Sigma=WS.reshape(K, M, D, D), r_nk=r.reshape((N, K), order="F"),
                delta=delta, beta=beta, pi_k=pi, lb=lb[:n_lb].copy(), n_iter=n_iter.value, flags=flags.value, loop_seconds=loop_s.value,
                threads=lib.oracle_num_threads() if n_threads <= 0 else n_threads)
