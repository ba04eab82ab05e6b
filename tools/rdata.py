"""Minimal reader for R's XDR serialisation (RDX2 / RDS, serialize version 2).

Test/fixture tooling only: it exists so the bundled reference datasets
(`data/melissa_synth_dt.rda`, `data/melissa_encode_dt.rda`, reference
`R/Melissa_package.R:52-113` documents them) can be turned into flat CSR
fixtures without an R installation.  Handles exactly the SEXP types those files
contain.
"""
from __future__ import annotations

import bz2
import gzip
import lzma
import struct

import numpy as np

NA_INT = -2147483648


class RObj:
    """A decoded R value with its attribute pairlist."""

    __slots__ = ("value", "attr")

    def __init__(self, value, attr=None):
        self.value = value
        self.attr = attr or {}

    @property
    def names(self):
        n = self.attr.get("names")
        return None if n is None else list(n.value)

    def __getitem__(self, key):
        if isinstance(key, str):
            return self.value[self.names.index(key)]
        return self.value[key]

    def __len__(self):
        return len(self.value)

    def __repr__(self):
        return f"RObj({type(self.value).__name__}, attr={list(self.attr)})"


class _Reader:
    def __init__(self, buf: bytes):
        self.b = buf
        self.o = 0
        self.refs = []

    def i32(self) -> int:
        (v,) = struct.unpack_from(">i", self.b, self.o)
        self.o += 4
        return v

    def take(self, n: int) -> bytes:
        v = self.b[self.o:self.o + n]
        self.o += n
        return v

    def item(self):
        flags = self.i32()
        t = flags & 0xFF
        has_attr = bool(flags & (1 << 9))
        has_tag = bool(flags & (1 << 10))
        if t == 254:                      # NILVALUE_SXP
            return None
        if t == 255:                      # REFSXP
            idx = flags >> 8
            if idx == 0:
                idx = self.i32()
            return self.refs[idx - 1]
        if t in (253, 242, 241, 251, 250):  # global/empty/base env, missing, unbound
            return None
        if t == 1:                        # SYMSXP
            name = self.item()
            self.refs.append(name)
            return name
        if t == 2:                        # LISTSXP (pairlist) -> ordered dict
            out = {}
            pos = 0
            while True:
                attr = self.item() if has_attr else None
                tag = self.item() if has_tag else None
                car = self.item()
                out[tag if tag is not None else pos] = car
                pos += 1
                flags = self.i32()
                t2 = flags & 0xFF
                if t2 == 254:
                    return out
                if t2 != 2:
                    raise ValueError(f"pairlist cdr of type {t2}")
                has_attr = bool(flags & (1 << 9))
                has_tag = bool(flags & (1 << 10))
        if t == 9:                        # CHARSXP
            n = self.i32()
            return None if n == -1 else self.take(n).decode("utf-8", "replace")
        if t == 10 or t == 13:            # LGLSXP / INTSXP
            n = self.i32()
            v = np.frombuffer(self.take(4 * n), dtype=">i4").astype(np.int32)
        elif t == 14:                     # REALSXP
            n = self.i32()
            v = np.frombuffer(self.take(8 * n), dtype=">f8").astype(np.float64)
        elif t == 16:                     # STRSXP
            n = self.i32()
            v = [self.item() for _ in range(n)]
        elif t == 19:                     # VECSXP
            n = self.i32()
            v = [self.item() for _ in range(n)]
        else:
            raise ValueError(f"unsupported SEXP type {t} at offset {self.o}")
        attr = self.item() if has_attr else None
        obj = RObj(v, attr)
        obj_is_logical = (t == 10)
        if obj_is_logical:
            obj.attr = dict(obj.attr)
            obj.attr["__logical__"] = True
        return obj


def _decompress(raw: bytes) -> bytes:
    if raw[:3] == b"BZh":
        return bz2.decompress(raw)
    if raw[:2] == b"\x1f\x8b":
        return gzip.decompress(raw)
    if raw[:6] == b"\xfd7zXZ\x00":
        return lzma.decompress(raw)
    return raw


def read_rda(path: str) -> dict:
    """Return {object name: RObj} for an .rda / .RData file."""
    buf = _decompress(open(path, "rb").read())
    if buf[:5] != b"RDX2\n":
        raise ValueError("not an RDX2 file")
    r = _Reader(buf[5:])
    if r.take(2) != b"X\n":
        raise ValueError("only XDR serialisation is supported")
    r.i32(), r.i32(), r.i32()             # serialize version, writer, min reader
    return r.item()


def read_rds(path: str):
    """Return the single object held by an .rds file."""
    buf = _decompress(open(path, "rb").read())
    r = _Reader(buf)
    if r.take(2) != b"X\n":
        raise ValueError("only XDR serialisation is supported")
    r.i32(), r.i32(), r.i32()
    return r.item()


def is_na_scalar(obj) -> bool:
    """True for R's logical NA of length one (an uncovered region)."""
    return (isinstance(obj, RObj) and obj.attr.get("__logical__")
            and len(obj.value) == 1 and int(obj.value[0]) == NA_INT)


def melissa_data_to_csr(dt: RObj) -> dict:
    """Flatten a `melissa_data_obj` (reference R/process_data.R:266-267) to CSR.

    Pairs are emitted cell-major (cell, then region), 0-based.
    """
    met = dt["met"]
    N = len(met)
    M = len(met[0])
    pair_cell, pair_region, pair_ptr = [], [], [0]
    xs, ys = [], []
    for n in range(N):
        cell = met[n]
        if len(cell) != M:
            raise ValueError("ragged region count")
        for m in range(M):
            reg = cell[m]
            if is_na_scalar(reg):
                continue
            dim = reg.attr["dim"].value
            I = int(dim[0])
            v = reg.value
            xs.append(v[:I])
            ys.append(v[I:2 * I])
            pair_cell.append(n)
            pair_region.append(m)
            pair_ptr.append(pair_ptr[-1] + I)
    x = np.concatenate(xs)
    y = np.concatenate(ys)
    if not np.all((y == 0) | (y == 1)):
        raise ValueError("non-binary methylation state")
    out = dict(N=N, M=M,
               pair_cell=np.asarray(pair_cell, np.int32),
               pair_region=np.asarray(pair_region, np.int32),
               pair_ptr=np.asarray(pair_ptr, np.int64),
               x=x.astype(np.float64), y=y.astype(np.uint8))
    opts = dt["opts"]
    if opts is not None:
        names = opts.names
        if "C_true" in names:
            c = opts["C_true"]
            d = c.attr["dim"].value
            out["C_true"] = np.asarray(c.value, np.float64).reshape(int(d[1]), int(d[0])).T.copy()
        if "W" in names:
            w = opts["W"]
            d = [int(v) for v in w.attr["dim"].value]
            out["W_true"] = np.asarray(w.value, np.float64).reshape(d[::-1]).transpose(2, 1, 0).copy()
        if "pi_k" in names:
            out["pi_true"] = np.asarray(opts["pi_k"].value, np.float64)
        if "basis_obj" in names:
            b = opts["basis_obj"]
            out["basis_M"] = int(np.asarray(b["M"].value)[0])
            out["basis_mus"] = np.asarray(b["mus"].value, np.float64)
            out["basis_gamma"] = float(np.asarray(b["gamma"].value)[0])
    return out
