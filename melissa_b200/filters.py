"""Region filters on the flat pair list (reference R/utils.R:762-847, `filter_regions`).

Each is a segmented reduction over `MelissaData`; they run on the host (NumPy), once,
before packing.  Semantics follow the reference function by function:
an uncovered (cell, region) pair is the reference's NA list element.
"""
from __future__ import annotations

import numpy as np

from .data import MelissaData


def _select_pairs(d: MelissaData, keep_pair: np.ndarray, region_map=None, M=None) -> MelissaData:
    lens = np.diff(d.pair_ptr)
    keep_cpg = np.repeat(keep_pair, lens)
    new_ptr = np.concatenate([[0], np.cumsum(lens[keep_pair])]).astype(np.int64)
    region = d.pair_region[keep_pair]
    if region_map is not None:
        region = region_map[region].astype(np.int32)
    return MelissaData(N=d.N, M=d.M if M is None else M, pair_ptr=new_ptr,
                       pair_cell=d.pair_cell[keep_pair].copy(), pair_region=region.copy(),
                       x=d.x[keep_cpg].copy(), y=d.y[keep_cpg].copy(), cell_names=d.cell_names)


def filter_by_cpg_coverage(d: MelissaData, min_cpgcov: int = 10) -> MelissaData:
    """R/utils.R:762-771: pairs with fewer than `min_cpgcov` CpGs become uncovered (NA);
    the number of regions does not change."""
    if not min_cpgcov >= 0:
        raise ValueError("min_cpgcov >= 0 is not TRUE")             # assert_that, :763
    return _select_pairs(d, np.diff(d.pair_ptr) >= min_cpgcov)


def filter_by_coverage_across_cells(d: MelissaData, min_cell_cov_prcg: float = 0.5) -> MelissaData:
    """R/utils.R:785-814: drop regions covered by less than `min_cell_cov_prcg` of the cells;
    the remaining regions are renumbered."""
    if not (0 <= min_cell_cov_prcg <= 1):
        raise ValueError("min_cell_cov_prcg >= 0 & min_cell_cov_prcg <= 1 is not TRUE")   # :786
    cov_cells = np.bincount(d.pair_region, minlength=d.M)
    keep_region = ~(cov_cells / d.N < min_cell_cov_prcg)            # :798
    return _drop_regions(d, keep_region)


def filter_by_variability(d: MelissaData, min_var: float = 0.1) -> MelissaData:
    """R/utils.R:826-847: keep regions whose mean methylation rate varies across cells with a
    standard deviation (stats::sd, n - 1 denominator) above `min_var`."""
    if not min_var >= 0:
        raise ValueError("min_var >= 0 is not TRUE")                # :827
    rate = np.add.reduceat(d.y.astype(np.float64), d.pair_ptr[:-1]) / np.diff(d.pair_ptr)   # mean(x[,2]) per pair
    n = np.bincount(d.pair_region, minlength=d.M).astype(np.float64)
    s1 = np.bincount(d.pair_region, weights=rate, minlength=d.M)
    with np.errstate(invalid="ignore", divide="ignore"):
        mean = s1 / n
        dev2 = np.bincount(d.pair_region, weights=(rate - mean[d.pair_region]) ** 2, minlength=d.M)
        sd = np.sqrt(dev2 / (n - 1.0))                              # NaN for a single cell, like R's sd()
    sd[n == 0] = 0.0                                                # :836-837
    return _drop_regions(d, sd > min_var)                           # which(NA > x) drops NA as well


def _drop_regions(d: MelissaData, keep_region: np.ndarray) -> MelissaData:
    region_map = np.cumsum(keep_region) - 1
    return _select_pairs(d, keep_region[d.pair_region], region_map, int(keep_region.sum()))
