import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (runs on the B200 box)")


@pytest.fixture(scope="session")
def built():
    """Both shared objects exist (built in-tree by __graft_entry__.build())."""
    import __graft_entry__ as G
    G.build()
    return True


@pytest.fixture(scope="session")
def api(built):
    from raytracingoneweekend_b200 import api as a
    return a


@pytest.fixture(scope="session")
def A(built):
    from raytracingoneweekend_b200 import _abi
    return _abi


@pytest.fixture(scope="session")
def O(built):
    from oracle import binding
    binding.lib()
    return binding


def oracle_scene_from(O, frozen):
    o = O.OracleScene()
    for k, arr in frozen.lists().items():
        o.add(k, arr)
    for pod, code in frozen.dyn_marched():
        o.add_dyn_marched(pod, code)
    return o


def bits(a):
    import numpy as np
    return a.view(np.uint32) if a.dtype == np.float32 else a
