"""CPU: region filters against the reference's own live test values
(reference tests/testthat/test-filtering.R:3-35, run on melissa_encode_dt)."""
import numpy as np
import pytest

import melissa_b200 as mb


@pytest.fixture(scope="module")
def encode(c2):
    return mb.from_dict(c2)


def _na_regions_of_cell(d, n):
    return d.M - int(np.sum(d.pair_cell == n))


def test_filter_by_cpg_coverage(encode):
    f = mb.filter_by_cpg_coverage(encode, min_cpgcov=20)
    assert _na_regions_of_cell(f, 0) == 31                                 # test-filtering.R:6
    assert not np.any((f.pair_cell == 0) & (f.pair_region == 93))          # met[[1]][[94]] is NA, :7
    assert f.M == encode.M
    with pytest.raises(ValueError):
        mb.filter_by_cpg_coverage(encode, min_cpgcov=-1)                   # :8
    f = mb.filter_by_cpg_coverage(encode, min_cpgcov=1000)
    assert f.P == 0                                                        # every region NA, :10-12


def test_filter_by_coverage_across_cells(encode):
    f = mb.filter_by_coverage_across_cells(encode, min_cell_cov_prcg=0.2)
    assert _na_regions_of_cell(f, 0) == 0 and f.M == encode.M              # :17-18
    with pytest.raises(ValueError):
        mb.filter_by_coverage_across_cells(encode, min_cell_cov_prcg=-1)   # :19
    with pytest.raises(ValueError):
        mb.filter_by_coverage_across_cells(encode, min_cell_cov_prcg=2)    # :20
    # after thinning, sparsely covered regions go and the rest are renumbered
    thin = mb.filter_by_cpg_coverage(encode, min_cpgcov=22)
    cov = np.bincount(thin.pair_region, minlength=thin.M) / thin.N
    thr = float(np.median(cov))
    f = mb.filter_by_coverage_across_cells(thin, min_cell_cov_prcg=thr)
    assert 0 < f.M < thin.M
    assert f.M == int(np.sum(cov >= thr)) and f.pair_region.max() == f.M - 1
    assert np.all(np.bincount(f.pair_region, minlength=f.M) / f.N >= thr)


def test_filter_by_variability(encode):
    assert mb.filter_by_variability(encode, min_var=0.1).M == 99           # :25-26
    assert mb.filter_by_variability(encode, min_var=0.15).M == 3           # :28-29
    with pytest.raises(ValueError):
        mb.filter_by_variability(encode, min_var=-1)                       # :31
    assert mb.filter_by_variability(encode, min_var=1000).M == 0           # :33-34
