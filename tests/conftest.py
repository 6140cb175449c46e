import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def etihad():
    m = np.load(os.path.join(GOLDEN, "etihad_model.npz"))
    return m["xyz"], m["viewpoint"]


@pytest.fixture(scope="session")
def space_2p5():
    return np.load(os.path.join(GOLDEN, "etihad_searchspace_2p5m.npz"))


@pytest.fixture(scope="session")
def space_1p5():
    return np.load(os.path.join(GOLDEN, "etihad_searchspace_1p5m.npz"))


@pytest.fixture(scope="session")
def oracle_mod():
    from oracle import oracle as O
    O.build()
    return O


@pytest.fixture(scope="session")
def lib_built():
    """Build the C-ABI library once per session when the toolchain is present (CPU box and GPU box)."""
    import subprocess
    from asscpp_b200 import _abi
    if not os.path.exists(_abi.LIB_PATH):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "asscpp_b200", "csrc")])
    return _abi.LIB_PATH


# (result field of AsscppViewResult, oracle stats column)
def field_pairs(O):
    return [("survivors", O.ST_SURVIVORS), ("visible", O.ST_VISIBLE), ("carry", O.ST_CARRY),
            ("unique_rays", O.ST_UNIQUE), ("cells", O.ST_CELLS), ("steps", O.ST_STEPS),
            ("gain_pts", O.ST_GAIN_PTS), ("gain_vox", O.ST_GAIN_VOX), ("vis_hash", O.ST_HASH),
            ("surv_hash", O.ST_SURV_HASH)]


def assert_results_equal(O, res, st, label=""):
    """Bit-exact bar: every integer field of every view equals the oracle's."""
    for name, col in field_pairs(O):
        a = res[name].astype(np.uint64)
        b = st[:, col].astype(np.uint64)
        bad = np.nonzero(a != b)[0]
        assert bad.size == 0, "%s: %s differs on %d views, first view %d: gpu=%d oracle=%d" % (
            label, name, bad.size, bad[0], a[bad[0]], b[bad[0]])
