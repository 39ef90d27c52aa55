import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    # tools/cpu_emu: DRM_EMU_LIB=<libdrm_emu*.so> points the Python mirror at the host-thread emulation, so that -m gpu test
    # files whose kernels are emulated can be run where there is no GPU (tests/test_emu_kernels.py does this in a child)
    emu = os.environ.get("DRM_EMU_LIB")
    if emu:
        import deepreadmapper_b200.api as api
        api._LIB_PATH = emu
        api._lib = None


@pytest.fixture(scope="session")
def golden():
    with open(os.path.join(ROOT, "tests", "golden", "sw_golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def fixtures(golden):
    """The reference's small test inputs (genome, reads, window table) rebuilt from the golden file."""
    fx = dict(golden["fixtures"])
    g = fx["genome"]
    comp = {"A": "T", "T": "A", "C": "G", "G": "C", "N": "N"}
    wins = []
    for p in range(len(g) - fx["ref_len"] + 1):
        w = g[p:p + fx["ref_len"]]
        wins.append(w)
        wins.append("".join(comp[c] for c in reversed(w)))
    assert len(wins) == fx["n_windows"]
    fx["windows"] = wins
    return fx


@pytest.fixture(scope="session")
def ob():
    from oracle import bindings
    bindings.oracle()
    return bindings
