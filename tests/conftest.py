import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def raw_table():
    from mitsuba_customization_b200 import synth
    return synth.synthetic_table(0)


@pytest.fixture(scope="session")
def oracle_table(raw_table):
    from oracle import merl_oracle as O
    t = O.OracleTable(raw_table)
    t.build_cdf()
    return t


@pytest.fixture(scope="session")
def built_lib():
    """Build libmerl_b200.so once (nvcc cross-compiles without a GPU)."""
    import __graft_entry__ as g
    g.build()
    from mitsuba_customization_b200 import _cabi
    return _cabi.lib()
