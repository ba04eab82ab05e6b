"""Parity against the reference ITSELF, for the day R is available.

tools/rebaseline_with_R.R runs Melissa:::melissa_inner (reference R/melissa_vb.R:232) on the two bundled data sets
with the fixtures' initialisation and leaves raw doubles under tests/golden/r_dump/.  No machine used so far has R
(SURVEY.md 0.2), so these tests skip with the reason "parity unpinned"; the moment the dump exists they compare the
CPU oracle, and on a GPU the CUDA path, with the reference's own numbers at the north-star tolerances."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, subset

DUMP = os.path.join(GOLDEN, "r_dump")
RUNS = [("c1_k2", 2), ("c1_k3", 3), ("c2", 4)]
needs_dump = pytest.mark.skipif(not os.path.isdir(DUMP), reason="parity unpinned: no R dump (run tools/rebaseline_with_R.R "
                                                                "on a machine with R + BPRMeth + Melissa)")


def _load(tag, N, M, D, K):
    rd = lambda name: np.fromfile(os.path.join(DUMP, f"{tag}_{name}.bin"), dtype="<f8")
    return dict(lb=rd("lb"), r_nk=rd("r_nk").reshape((N, K), order="F"), W=rd("W").reshape((M, D, K), order="F"))


def _case(tag, c1, c2):
    if tag.startswith("c1"):
        k = tag.split("_")[1]
        return c1, c1["basis"], c1[f"w0_{k}"], c1[f"{k}_delta_0"], int(c1[f"{k}_iters"])
    train = subset(c2, c2["is_train"])
    return train, c2["basis"], c2["w0"], c2["delta_0"], int(c2["iters"])


@needs_dump
@pytest.mark.parametrize("tag,K", RUNS)
def test_oracle_matches_the_reference_dump(tag, K, c1, c2):
    import melissa_oracle as mo
    d, basis, w0, delta_0, iters = _case(tag, c1, c2)
    ref = _load(tag, d["N"], d["M"], int(basis["M"]) + 1, K)
    o = mo.melissa_inner(d, K, basis, w0, delta_0, vb_max_iter=iters)
    np.testing.assert_allclose(o["lb"][1:], ref["lb"][1:], rtol=1e-9)
    assert np.array_equal(o["r_nk"].argmax(1), ref["r_nk"].argmax(1))
    np.testing.assert_allclose(o["W"], ref["W"], atol=1e-7)


@needs_dump
@pytest.mark.gpu
@pytest.mark.parametrize("tag,K", RUNS)
def test_cuda_path_matches_the_reference_dump(tag, K, c1, c2):
    import melissa_b200 as mb
    d, basis, w0, delta_0, iters = _case(tag, c1, c2)
    ref = _load(tag, d["N"], d["M"], int(basis["M"]) + 1, K)
    with mb.Engine(mb.from_dict(d), basis) as eng:
        res = eng.run(K, w0, delta_0, vb_max_iter=iters)
    np.testing.assert_allclose(res["lb"][1:], ref["lb"][1:], rtol=1e-9)
    assert np.array_equal(res["r_nk"].argmax(1), ref["r_nk"].argmax(1))
    np.testing.assert_allclose(res["W"], ref["W"], atol=1e-7)


def test_the_inputs_the_R_script_reads_are_the_fixtures(c1, c2):
    """tests/golden/r_inputs/ (tools/export_fixture_inputs.py) must stay in step with the .npz fixtures."""
    inp = os.path.join(GOLDEN, "r_inputs")
    w = np.fromfile(os.path.join(inp, "c1_k2_w0.bin"), dtype="<f8").reshape(c1["w0_k2"].shape, order="F")
    assert np.array_equal(w, c1["w0_k2"])
    assert np.array_equal(np.fromfile(os.path.join(inp, "c2_delta_0.bin"), dtype="<f8"), c2["delta_0"])
    mask = np.fromfile(os.path.join(inp, "c2_train_mask.bin"), dtype=np.int8).astype(bool)
    assert np.array_equal(mask, c2["is_train"])
