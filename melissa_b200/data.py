"""Flat (CSR-style) form of Melissa's nested-list input.

The reference's `X` is a list over cells of lists over regions of `I x 2`
matrices (column 1: CpG location scaled to [-1, 1], column 2: methylation state)
or NA where the cell does not cover the region (reference R/melissa_vb.R:11-17).
`MelissaData` is the same information as five flat arrays, cell-major -- the
form `melissa_pack` takes (include/melissa_b200.h).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Optional, Sequence

import numpy as np


@dataclass
class MelissaData:
    N: int
    M: int
    pair_ptr: np.ndarray      # int64 [P+1]
    pair_cell: np.ndarray     # int32 [P]
    pair_region: np.ndarray   # int32 [P]
    x: np.ndarray             # float64 [nnz]
    y: np.ndarray             # uint8 [nnz]
    cell_names: Optional[Sequence[str]] = None
    extras: dict = field(default_factory=dict)

    @property
    def P(self) -> int:
        return int(self.pair_cell.shape[0])

    @property
    def nnz(self) -> int:
        return int(self.x.shape[0])

    def as_dict(self) -> dict:
        return dict(N=self.N, M=self.M, pair_ptr=self.pair_ptr, pair_cell=self.pair_cell,
                    pair_region=self.pair_region, x=self.x, y=self.y)

    def save(self, path: str):
        """Flat on-disk form (a compressed .npz of the five arrays + N, M): the format that lets a
        real data set reach the GPU path without ever becoming a nested R list."""
        np.savez_compressed(path, N=self.N, M=self.M, pair_ptr=self.pair_ptr, pair_cell=self.pair_cell,
                            pair_region=self.pair_region, x=self.x, y=self.y,
                            cell_names=np.asarray([] if self.cell_names is None else list(self.cell_names)))

    @staticmethod
    def load(path: str) -> "MelissaData":
        z = np.load(path, allow_pickle=False)
        names = [str(v) for v in z["cell_names"]] if "cell_names" in z.files and z["cell_names"].size else None
        return MelissaData(N=int(z["N"]), M=int(z["M"]), pair_ptr=z["pair_ptr"].astype(np.int64),
                           pair_cell=z["pair_cell"].astype(np.int32), pair_region=z["pair_region"].astype(np.int32),
                           x=z["x"].astype(np.float64), y=z["y"].astype(np.uint8), cell_names=names)

    def shard_cells(self, lo: int, hi: int) -> "MelissaData":
        """Cells [lo, hi) as a stand-alone data set (cell indices re-based to 0)."""
        sel = np.flatnonzero((self.pair_cell >= lo) & (self.pair_cell < hi))
        if sel.size == 0:
            raise ValueError("shard without any covered pair")
        p0, p1 = int(sel[0]), int(sel[-1]) + 1
        c0, c1 = int(self.pair_ptr[p0]), int(self.pair_ptr[p1])
        return MelissaData(N=hi - lo, M=self.M, pair_ptr=(self.pair_ptr[p0:p1 + 1] - c0).astype(np.int64),
                           pair_cell=(self.pair_cell[p0:p1] - lo).astype(np.int32),
                           pair_region=self.pair_region[p0:p1].copy(), x=self.x[c0:c1].copy(),
                           y=self.y[c0:c1].copy(),
                           cell_names=None if self.cell_names is None else list(self.cell_names[lo:hi]))


def _is_na(v) -> bool:
    if v is None:
        return True
    a = np.asarray(v)
    return a.ndim < 2 and (a.size == 0 or (a.size == 1 and (a.dtype.kind in "fc") and np.isnan(a.ravel()[0])))


def from_nested(X, cell_names=None) -> MelissaData:
    """Flatten the reference's nested list (list of cells -> list of regions ->
    [I x 2] array or None / NaN).  Mirrors the checks `melissa()` makes
    (R/melissa_vb.R:91-92) and the NA handling of :119-121."""
    if not isinstance(X, (list, tuple)):
        raise TypeError("X is not a list")                      # assert_that(is.list(X))
    if len(X) == 0 or not isinstance(X[0], (list, tuple)):
        raise TypeError("X[[1]] is not a list")                 # assert_that(is.list(X[[1]]))
    N, M = len(X), len(X[0])
    pair_cell, pair_region, ptr, xs, ys = [], [], [0], [], []
    for n, cell in enumerate(X):
        if len(cell) != M:
            raise ValueError(f"cell {n} has {len(cell)} regions, expected {M}")
        covered = 0
        for m, reg in enumerate(cell):
            if _is_na(reg):
                continue
            a = np.asarray(reg, dtype=np.float64)
            if a.ndim != 2 or a.shape[1] < 2:
                raise ValueError(f"X[[{n + 1}]][[{m + 1}]] is not an I x 2 matrix")
            xs.append(a[:, 0])
            ys.append(a[:, 1])
            pair_cell.append(n)
            pair_region.append(m)
            ptr.append(ptr[-1] + a.shape[0])
            covered += 1
        if covered == 0:
            raise ValueError("No coverage across all regions. This cell should be discarded!")   # R/utils.R:69-71
    y = np.concatenate(ys)
    if not np.all((y == 0) | (y == 1)):
        raise ValueError("methylation states must be 0 or 1")
    return MelissaData(N=N, M=M, pair_ptr=np.asarray(ptr, np.int64), pair_cell=np.asarray(pair_cell, np.int32),
                       pair_region=np.asarray(pair_region, np.int32), x=np.concatenate(xs).astype(np.float64),
                       y=y.astype(np.uint8), cell_names=cell_names)


def to_nested(d: MelissaData):
    """Inverse of from_nested (small data only)."""
    X = [[None] * d.M for _ in range(d.N)]
    for p in range(d.P):
        sl = slice(int(d.pair_ptr[p]), int(d.pair_ptr[p + 1]))
        X[int(d.pair_cell[p])][int(d.pair_region[p])] = np.stack([d.x[sl], d.y[sl].astype(np.float64)], axis=1)
    return X


def from_dict(d: dict) -> MelissaData:
    return MelissaData(N=int(d["N"]), M=int(d["M"]), pair_ptr=np.ascontiguousarray(d["pair_ptr"], np.int64),
                       pair_cell=np.ascontiguousarray(d["pair_cell"], np.int32),
                       pair_region=np.ascontiguousarray(d["pair_region"], np.int32),
                       x=np.ascontiguousarray(d["x"], np.float64), y=np.ascontiguousarray(d["y"], np.uint8))


def create_rbf_object(M: int = 3) -> dict:
    """BPRMeth::create_rbf_object(M) as pinned from the basis objects inside the
    reference's bundled data: equally spaced centres
    mu_j = j * 2/(M+1) - 1 and gamma = M^2 / 4."""
    M = int(M)
    if M < 0:
        raise ValueError("M must be >= 0")
    mus = np.array([j * (2.0 / (M + 1)) - 1.0 for j in range(1, M + 1)], dtype=np.float64)
    return dict(kind="rbf", M=M, mus=mus, gamma=(M * M / 4.0) if M > 0 else 1.0)
