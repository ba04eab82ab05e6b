"""CPU: the C-ABI library builds, loads and exports every symbol the header declares."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    src = open(os.path.join(ROOT, "include", "melissa_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(melissa_[a-zA-Z0-9_]+)\s*\(", src)))


def test_header_declares_the_boundary():
    names = header_functions()
    for must in ("melissa_pack", "melissa_vb_run", "melissa_free", "melissa_last_error", "melissa_impute"):
        assert must in names


def test_library_exports_every_declared_symbol():
    from melissa_b200 import _lib
    from melissa_b200.build import build_library
    build_library()
    lib = _lib.load()
    names = header_functions()
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"
    assert sorted(_lib.SIGNATURES) == names, "ctypes signature table and header disagree"
    assert lib.melissa_abi_version() == 1


def test_argument_errors_do_not_need_a_gpu():
    # validation happens before the first CUDA call, and mirrors the reference's checks
    import numpy as np
    from melissa_b200 import _lib
    lib = _lib.load()
    h = _lib.handle_t()
    ptr = np.array([0, 2, 4], np.int64)
    cell = np.array([0, 2], np.int32)          # cell 1 has no covered region
    reg = np.array([0, 0], np.int32)
    x = np.zeros(4)
    y = np.zeros(4, np.uint8)
    mus = np.array([-0.5, 0, 0.5])
    rc = lib.melissa_pack(3, 1, 2, _lib.ptr(ptr, _lib.c_i64), _lib.ptr(cell, _lib.c_i32), _lib.ptr(reg, _lib.c_i32),
                          _lib.ptr(x, _lib.c_dbl), _lib.ptr(y, ctypes.c_uint8), 1, 3, _lib.ptr(mus, _lib.c_dbl), 2.25,
                          0, 0, ctypes.byref(h))
    assert rc == 1
    assert b"No coverage across all regions" in lib.melissa_last_error()   # R/utils.R:69-71
    rc = lib.melissa_pack(3, 1, 2, _lib.ptr(ptr, _lib.c_i64), _lib.ptr(cell, _lib.c_i32), _lib.ptr(reg, _lib.c_i32),
                          _lib.ptr(x, _lib.c_dbl), _lib.ptr(y, ctypes.c_uint8), 2, 3, _lib.ptr(mus, _lib.c_dbl), 2.25,
                          0, 0, ctypes.byref(h))
    assert rc == 1 and b"RBF" in lib.melissa_last_error()
    assert lib.melissa_vb_phase_a(None) == 3


def test_a_fault_in_one_shard_of_melissa_pack_multi_is_named_without_a_gpu():
    # every shard validates its window of the caller's lists on the host before its first CUDA call; the fault is
    # reported with the shard and the shard's first cell / pair, ahead of any CUDA failure of another shard
    import numpy as np
    from melissa_b200 import _lib
    lib = _lib.load()
    n_cells, n_regions = 8, 3
    cell = np.repeat(np.arange(n_cells, dtype=np.int32), n_regions)
    reg = np.tile(np.arange(n_regions, dtype=np.int32), n_cells)
    reg[-2], reg[-1] = 2, 1                                   # cell 7: regions 0, 2, 1 -- not increasing
    P = cell.size
    ptr = np.arange(P + 1, dtype=np.int64) * 2
    x = np.zeros(2 * P)
    y = np.zeros(2 * P, np.uint8)
    mus = np.array([-0.5, 0, 0.5])
    mm = _lib.handle_t()
    rc = lib.melissa_pack_multi(n_cells, n_regions, P, _lib.ptr(ptr, _lib.c_i64), _lib.ptr(cell, _lib.c_i32),
                                _lib.ptr(reg, _lib.c_i32), _lib.ptr(x, _lib.c_dbl), _lib.ptr(y, ctypes.c_uint8), 1, 3,
                                _lib.ptr(mus, _lib.c_dbl), 2.25, 2, None, ctypes.byref(mm))
    msg = lib.melissa_last_error()
    assert rc == 1, msg
    assert b"shard 1 (cells from 4, pairs from 12)" in msg and b"regions must be strictly increasing" in msg and b"pair 11" in msg


def test_every_option_the_header_documents_is_accepted_and_every_accepted_one_is_documented():
    import re
    from melissa_b200 import _lib
    lib = _lib.load()
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "melissa_b200.h")).read()
    doc = hdr[hdr.index("Library-wide tuning and diagnostic knobs"):hdr.index("int melissa_set_option")]
    documented = set(re.findall(r'^ \*   "([a-z_]+)"', doc, re.M))
    src = open(os.path.join(root, "melissa_b200", "csrc", "melissa.cu")).read()
    accepted = set(re.findall(r'std::strcmp\(key, "([a-z_]+)"\)', src))
    assert documented == accepted, (documented ^ accepted)
    neutral = {"sliced": -1}                                   # the value that means "default" where 0 does not
    for key in sorted(documented):
        assert lib.melissa_set_option(key.encode(), neutral.get(key, 0)) == 0, (key, lib.melissa_last_error())
    lib.melissa_set_option(b"contract_mma", 1)                 # 0 is not the default of these two
    lib.melissa_set_option(b"contract_tma", 1)
    assert lib.melissa_set_option(b"no_such_knob", 1) == 1 and b"unknown option" in lib.melissa_last_error()
