"""Thin object wrapper over the C ABI: one `Engine` = one `melissa_handle`."""
from __future__ import annotations

import ctypes as C
import math

import numpy as np

from . import _lib
from .data import MelissaData


def set_option(key: str, value: int):
    """Library-wide knob for the next pack: "cell_block" (cells per block of the sliced
    layout, default 128) or "sliced" (-1 auto, 0 flat layout, 1 require sliced)."""
    _lib.check(_lib.load().melissa_set_option(key.encode(), int(value)))


class progress:
    """Context manager around melissa_set_progress: `fn(iteration, lower_bound, change)` is called at every bound of
    the runs started inside the block; returning True stops the run (MELISSA_FLAG_INTERRUPTED)."""

    def __init__(self, fn):
        self._cb = _lib.PROGRESS_FN(lambda i, lb, diff, _user: 1 if fn(int(i), float(lb), float(diff)) else 0)

    def __enter__(self):
        _lib.check(_lib.load().melissa_set_progress(self._cb, None))
        return self

    def __exit__(self, *exc):
        _lib.load().melissa_set_progress(_lib.PROGRESS_FN(), None)


class _DeviceBuffer:
    """Exposes a raw device pointer through __cuda_array_interface__ so torch can
    alias it (used only to hand the reduction buffers to torch.distributed)."""

    def __init__(self, ptr: int, n: int):
        self.__cuda_array_interface__ = dict(shape=(int(n),), typestr="<f8", data=(int(ptr), False), version=2)


class Engine:
    """Packed observations on one GPU plus the state of one VB fit.

    Mirrors the phases of the reference's melissa_inner (R/melissa_vb.R:232-529);
    every method is one call into libmelissa_b200.so.
    """

    def __init__(self, data, basis: dict, device: int = 0, allow_empty_regions: bool = False, K_hint=None):
        lib = _lib.load()
        if K_hint is not None:                  # lets the packer size its cell blocks for this K
            _lib.check(lib.melissa_set_option(b"clusters_hint", int(min(64, max(1, K_hint)))))
        if basis.get("kind", "rbf") != "rbf":
            raise ValueError("only RBF basis objects are supported")
        self.lib = lib
        self.data = data
        self.basis = basis
        self.device = int(device)
        self.D = int(basis["M"]) + 1
        self.K = 0
        mus = np.ascontiguousarray(basis["mus"], np.float64)
        h = _lib.handle_t()
        if isinstance(data, dict) and data.get("on_device"):
            # torch CUDA tensors already resident on `device` (melissa_pack_device)
            t = data
            self.N, self.M, self.P = int(t["N"]), int(t["M"]), int(t["pair_cell"].shape[0])
            for k, dt in (("pair_ptr", "int64"), ("pair_cell", "int32"), ("pair_region", "int32"),
                          ("x", "float64"), ("y", "uint8")):
                if str(t[k].dtype) != "torch." + dt or not t[k].is_contiguous() or t[k].device.index != self.device:
                    raise ValueError(f"{k}: need a contiguous {dt} tensor on cuda:{self.device}")
            _lib.check(lib.melissa_pack_device(
                self.N, self.M, self.P, t["pair_ptr"].data_ptr(), t["pair_cell"].data_ptr(),
                t["pair_region"].data_ptr(), t["x"].data_ptr(), t["y"].data_ptr(), 1, int(basis["M"]),
                _lib.ptr(mus, _lib.c_dbl) if mus.size else None, float(basis["gamma"]), self.device,
                int(bool(allow_empty_regions)), C.byref(h)))
            self.h = h
            return
        self.N, self.M, self.P = int(data.N), int(data.M), data.P
        _lib.check(lib.melissa_pack(
            self.N, self.M, self.P, _lib.ptr(np.ascontiguousarray(data.pair_ptr, np.int64), _lib.c_i64),
            _lib.ptr(np.ascontiguousarray(data.pair_cell, np.int32), _lib.c_i32),
            _lib.ptr(np.ascontiguousarray(data.pair_region, np.int32), _lib.c_i32),
            _lib.ptr(np.ascontiguousarray(data.x, np.float64), _lib.c_dbl),
            _lib.ptr(np.ascontiguousarray(data.y, np.uint8), C.c_uint8),
            1, int(basis["M"]), _lib.ptr(mus, _lib.c_dbl) if mus.size else None, float(basis["gamma"]),
            self.device, int(bool(allow_empty_regions)), C.byref(h)))
        self.h = h

    # -- lifetime -----------------------------------------------------------
    def close(self):
        if getattr(self, "h", None):
            self.lib.melissa_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- one-shot fit (the loop runs inside the library) -------------------------
    def run(self, K, w0, delta_0, alpha_0=0.5, beta_0=None, vb_max_iter=300, epsilon_conv=1e-5, lb_every=10):
        K = int(K)
        w0 = self._check_w(w0, K)
        delta_0 = np.ascontiguousarray(delta_0, np.float64)
        if delta_0.shape != (K,):
            raise ValueError("delta_0 must have length K")
        cap = (int(vb_max_iter) // int(lb_every)) + 3
        out = self._alloc_outputs(K)
        lb = np.empty(cap, np.float64)
        n_lb, n_iter, flags = _lib.c_i32(0), _lib.c_i32(0), _lib.c_i32(0)
        _lib.check(self.lib.melissa_vb_run(
            self.h, K, _lib.ptr(w0, _lib.c_dbl), _lib.ptr(delta_0, _lib.c_dbl), float(alpha_0),
            float("nan") if beta_0 is None else float(beta_0), int(vb_max_iter), float(epsilon_conv),
            int(lb_every), *[_lib.ptr(out[k], _lib.c_dbl) for k in ("W", "W_Sigma", "r_nk", "delta", "alpha",
                                                                    "beta", "pi_k")],
            _lib.ptr(lb, _lib.c_dbl), cap, C.byref(n_lb), C.byref(n_iter), C.byref(flags)))
        self.K = K
        res = self._shape_outputs(out, K)
        res.update(lb=lb[:n_lb.value].copy(), n_iter=n_iter.value, flags=flags.value)
        return res

    # -- step-wise fit ---------------------------------------------------------
    def begin(self, K, w0, delta_0, alpha_0=0.5, beta_0=None, n_cells_total=None, stream=0):
        K = int(K)
        w0 = self._check_w(w0, K)
        delta_0 = np.ascontiguousarray(delta_0, np.float64)
        if delta_0.shape != (K,):
            raise ValueError("delta_0 must have length K")
        _lib.check(self.lib.melissa_vb_begin(
            self.h, K, _lib.ptr(w0, _lib.c_dbl), _lib.ptr(delta_0, _lib.c_dbl), float(alpha_0),
            float("nan") if beta_0 is None else float(beta_0),
            int(self.N if n_cells_total is None else n_cells_total), C.c_void_p(int(stream) or None)))
        self.K = K

    def phase_a(self):
        _lib.check(self.lib.melissa_vb_phase_a(self.h))

    def phase_b(self, with_lb: bool):
        _lib.check(self.lib.melissa_vb_phase_b(self.h, int(bool(with_lb))))

    def lower_bound(self) -> float:
        v = _lib.c_dbl(0.0)
        _lib.check(self.lib.melissa_vb_lower_bound(self.h, C.byref(v)))
        return v.value

    def _tensor(self, fn):
        import torch
        p, n = C.c_void_p(), _lib.c_i64(0)
        _lib.check(fn(self.h, C.byref(p), C.byref(n)))
        return torch.as_tensor(_DeviceBuffer(p.value, n.value), device=f"cuda:{self.device}")

    def reduce_tensor(self):
        """torch view of the sufficient-statistics buffer (all-reduce SUM between phase_a and phase_b)."""
        return self._tensor(self.lib.melissa_vb_reduce_buffer)

    def lb_tensor(self):
        """torch view of the lb_pz_qz partial (all-reduce SUM before lower_bound)."""
        return self._tensor(self.lib.melissa_vb_lb_buffer)

    def fetch(self):
        out = self._alloc_outputs(self.K)
        _lib.check(self.lib.melissa_vb_fetch(self.h, *[_lib.ptr(out[k], _lib.c_dbl) for k in
                                                       ("W", "W_Sigma", "r_nk", "delta", "alpha", "beta", "pi_k")]))
        return self._shape_outputs(out, self.K)

    def fetch_HtEz(self):
        out = np.empty((self.P, self.D), np.float64)
        _lib.check(self.lib.melissa_vb_fetch_HtEz(self.h, _lib.ptr(out, _lib.c_dbl)))
        return out

    def profiles_mle(self, lam: float = 0.5, n_iter: int = 15):
        """Per-pair penalised probit MLE on the GPU, [P, D] (melissa_profiles_mle)."""
        out = np.empty((self.P, self.D), np.float64)
        _lib.check(self.lib.melissa_profiles_mle(self.h, float(lam), int(n_iter), _lib.ptr(out, _lib.c_dbl)))
        return out

    def profiles_mle_device(self, lam: float = 0.5, n_iter: int = 15):
        """The same as a torch CUDA tensor [P, D] on the engine's GPU (melissa_profiles_mle_device)."""
        import torch
        out = torch.empty((self.P, self.D), dtype=torch.float64, device=f"cuda:{self.device}")
        _lib.check(self.lib.melissa_profiles_mle_device(self.h, float(lam), int(n_iter), C.c_void_p(out.data_ptr())))
        return out

    def sync(self):
        _lib.check(self.lib.melissa_sync(self.h))

    def launch_count(self) -> int:
        return int(self.lib.melissa_launch_count(self.h))

    def layout_info(self) -> dict:
        """Shape of the packed layout (melissa_layout_info)."""
        v = (C.c_int64 * 8)()
        _lib.check(self.lib.melissa_layout_info(self.h, v, 8))
        keys = ("sliced", "cell_block", "slices", "cpg_slots", "cpgs", "pairs", "chunks", "sweep_warps")
        out = dict(zip(keys, (int(x) for x in v)))
        out["row_padding"] = out["cpg_slots"] / max(1, out["cpgs"])
        return out

    KERNELS = dict(sweep=0, resp=1, contract=2, solve=3, scalars=4, lb=5)

    def profile(self, on: bool):
        _lib.check(self.lib.melissa_profile_enable(self.h, int(bool(on))))

    def profile_read(self):
        """{kernel: (total ms, launches)} since profile(True)."""
        out = {}
        for name, kid in self.KERNELS.items():
            ms, cnt = _lib.c_dbl(0.0), _lib.c_i64(0)
            _lib.check(self.lib.melissa_profile_read(self.h, kid, C.byref(ms), C.byref(cnt)))
            out[name] = (ms.value, cnt.value)
        return out

    # -- helpers ----------------------------------------------------------------
    def _check_w(self, w0, K):
        w0 = np.asarray(w0, np.float64)
        if w0.shape != (self.M, self.D, K):
            raise ValueError(f"w must have shape (M, D, K) = {(self.M, self.D, K)}, got {w0.shape}")
        return np.asfortranarray(w0).ravel(order="F").copy()      # R array layout

    def _alloc_outputs(self, K):
        M, D, N = self.M, self.D, self.N
        return dict(W=np.empty(M * D * K), W_Sigma=np.empty(K * M * D * D), r_nk=np.empty(N * K),
                    delta=np.empty(K), alpha=np.empty(K), beta=np.empty(K), pi_k=np.empty(K))

    def _shape_outputs(self, out, K):
        M, D, N = self.M, self.D, self.N
        return dict(W=out["W"].reshape((M, D, K), order="F"),
                    W_Sigma=out["W_Sigma"].reshape(K, M, D, D),
                    r_nk=out["r_nk"].reshape((N, K), order="F"),
                    delta=out["delta"], alpha=out["alpha"], beta=out["beta"], pi_k=out["pi_k"])


class MultiEngine:
    """Several GPUs of one box behind one object (melissa_multi): either packed from the whole job's host arrays
    (`MultiEngine.pack`) or assembled from Engines the caller packed shard by shard, in cell order
    (`MultiEngine(engines)`; they stay the caller's).  `run` is melissa_vb_run_multi: one call, one process."""

    def __init__(self, engines=None, _handle=None, _shape=None):
        self.lib = _lib.load()
        if _handle is not None:
            self.h, (self.N, self.M, self.D, self.basis) = _handle, _shape
            self.engines = None
            return
        self.engines = list(engines)
        arr = (_lib.handle_t * len(self.engines))(*[e.h for e in self.engines])
        h = _lib.handle_t()
        _lib.check(self.lib.melissa_multi_create(len(self.engines), arr, C.byref(h)))
        self.h = h
        e0 = self.engines[0]
        self.N, self.M, self.D, self.basis = sum(e.N for e in self.engines), e0.M, e0.D, e0.basis

    @classmethod
    def pack(cls, data: MelissaData, basis: dict, n_gpus: int, devices=None):
        lib = _lib.load()
        mus = np.ascontiguousarray(basis["mus"], np.float64)
        dev = None if devices is None else np.ascontiguousarray(devices, np.int32)
        h = _lib.handle_t()
        _lib.check(lib.melissa_pack_multi(
            int(data.N), int(data.M), data.P, _lib.ptr(np.ascontiguousarray(data.pair_ptr, np.int64), _lib.c_i64),
            _lib.ptr(np.ascontiguousarray(data.pair_cell, np.int32), _lib.c_i32),
            _lib.ptr(np.ascontiguousarray(data.pair_region, np.int32), _lib.c_i32),
            _lib.ptr(np.ascontiguousarray(data.x, np.float64), _lib.c_dbl),
            _lib.ptr(np.ascontiguousarray(data.y, np.uint8), C.c_uint8), 1, int(basis["M"]),
            _lib.ptr(mus, _lib.c_dbl) if mus.size else None, float(basis["gamma"]), int(n_gpus),
            _lib.ptr(dev, _lib.c_i32) if dev is not None else None, C.byref(h)))
        return cls(_handle=h, _shape=(int(data.N), int(data.M), int(basis["M"]) + 1, basis))

    def close(self):
        if getattr(self, "h", None):
            self.lib.melissa_multi_free(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def last_loop_ms(self) -> float:
        return float(self.lib.melissa_multi_last_loop_ms(self.h))

    def launch_count(self) -> int:
        return int(self.lib.melissa_multi_launch_count(self.h))

    def run(self, K, w0, delta_0, alpha_0=0.5, beta_0=None, vb_max_iter=300, epsilon_conv=1e-5, lb_every=10):
        K = int(K)
        w0 = np.asarray(w0, np.float64)
        if w0.shape != (self.M, self.D, K):
            raise ValueError(f"w must have shape (M, D, K) = {(self.M, self.D, K)}, got {w0.shape}")
        w0 = np.asfortranarray(w0).ravel(order="F").copy()
        delta_0 = np.ascontiguousarray(delta_0, np.float64)
        if delta_0.shape != (K,):
            raise ValueError("delta_0 must have length K")
        M, D, N = self.M, self.D, self.N
        out = dict(W=np.empty(M * D * K), W_Sigma=np.empty(K * M * D * D), r_nk=np.empty(N * K),
                   delta=np.empty(K), alpha=np.empty(K), beta=np.empty(K), pi_k=np.empty(K))
        cap = (int(vb_max_iter) // int(lb_every)) + 3
        lb = np.empty(cap, np.float64)
        n_lb, n_iter, flags = _lib.c_i32(0), _lib.c_i32(0), _lib.c_i32(0)
        _lib.check(self.lib.melissa_vb_run_multi(
            self.h, K, _lib.ptr(w0, _lib.c_dbl), _lib.ptr(delta_0, _lib.c_dbl), float(alpha_0),
            float("nan") if beta_0 is None else float(beta_0), int(vb_max_iter), float(epsilon_conv),
            int(lb_every), *[_lib.ptr(out[k], _lib.c_dbl) for k in ("W", "W_Sigma", "r_nk", "delta", "alpha",
                                                                    "beta", "pi_k")],
            _lib.ptr(lb, _lib.c_dbl), cap, C.byref(n_lb), C.byref(n_iter), C.byref(flags)))
        res = dict(W=out["W"].reshape((M, D, K), order="F"), W_Sigma=out["W_Sigma"].reshape(K, M, D, D),
                   r_nk=out["r_nk"].reshape((N, K), order="F"), delta=out["delta"], alpha=out["alpha"],
                   beta=out["beta"], pi_k=out["pi_k"])
        res.update(lb=lb[:n_lb.value].copy(), n_iter=n_iter.value, flags=flags.value)
        return res
