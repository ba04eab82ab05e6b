import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _positions_to_x(pos):
    return (pos.astype(np.float64) + 1000.0) / 2000.0 * 2.0 - 1.0


def load_c1():
    """melissa_synth_dt (20 cells x 30 regions) + the oracle's results on it."""
    z = dict(np.load(os.path.join(GOLDEN, "synth_c1.npz")))
    z["x"] = _positions_to_x(z["pos"])
    z["basis"] = dict(kind="rbf", M=int(z["basis_M"]), mus=z["basis_mus"], gamma=float(z["basis_gamma"]))
    z["N"], z["M"] = int(z["N"]), int(z["M"])
    return z


def load_c2():
    """melissa_encode_dt (200 x 100) with the frozen 50 % CpG hold-out split."""
    z = dict(np.load(os.path.join(GOLDEN, "encode_c2.npz")))
    z["x"] = _positions_to_x(z["pos"])
    z["basis"] = dict(kind="rbf", M=int(z["basis_M"]), mus=z["basis_mus"], gamma=float(z["basis_gamma"]))
    z["N"], z["M"] = int(z["N"]), int(z["M"])
    z["is_train"] = np.unpackbits(z["is_train"])[: z["x"].shape[0]].astype(bool)
    return z


def subset(d, mask):
    """CpGs selected by `mask` as a data dict (pairs that lose all CpGs vanish)."""
    pp = d["pair_ptr"]
    lens = np.add.reduceat(mask.astype(np.int64), pp[:-1])
    keep = lens > 0
    return dict(N=d["N"], M=d["M"], pair_cell=d["pair_cell"][keep], pair_region=d["pair_region"][keep],
                pair_ptr=np.concatenate([[0], np.cumsum(lens[keep])]).astype(np.int64),
                x=d["x"][mask], y=d["y"][mask])


@pytest.fixture(scope="session")
def c1():
    return load_c1()


@pytest.fixture(scope="session")
def c2():
    return load_c2()
