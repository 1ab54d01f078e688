import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLD = os.path.join(ROOT, "tests", "golden")
CONFIGS = [(0, 1), (0, 0), (1, 1), (1, 0), (2, 1), (2, 0)]   # (nfix_crc, fix_df)
GOLDEN_NAMES = ["c2_sparse", "c3_dense", "boundary"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle_mod():
    from oracle import oracle1090 as O
    O.build() if not os.path.exists(O.LIB_PATH) else None
    return O


@pytest.fixture(scope="session")
def golden():
    out = {}
    for name in GOLDEN_NAMES:
        with np.load(os.path.join(GOLD, name + ".npz")) as z:
            out[name] = {k: z[k] for k in z.files}
    return out


@pytest.fixture(scope="session")
def adsbcu_mod():
    """The CUDA library binding.  Built in-tree if missing; never replaced by a CPU path."""
    if os.environ.get("ADSBCU_TEST_EMU") == "1":
        # CPU test suite only (tests/test_emu_kernels.py runs selected -m gpu tests in a subprocess with this set):
        # the same kernel sources, compiled by g++ against the SIMT emulator of tests/emu.  Test infrastructure;
        # the product binding has no such switch.
        sys.path.insert(0, os.path.join(ROOT, "tests", "emu"))
        import emu_load
        return emu_load.load_emulated()
    from libadsb_b200 import adsbcu, build
    build.build()
    adsbcu.load()
    return adsbcu
