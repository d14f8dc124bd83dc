import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on a B200)")


@pytest.fixture(scope="session")
def native_libs():
    """Build the native libraries if they are not there yet (on the GPU box they arrive prebuilt)."""
    from spark_druid_olap_b200 import ENGINE_LIB
    from spark_druid_olap_b200.tpch import GEN_LIB
    if not (os.path.exists(ENGINE_LIB) and os.path.exists(GEN_LIB)):
        from spark_druid_olap_b200 import build
        build.build_all()
    return ENGINE_LIB, GEN_LIB


@pytest.fixture(scope="session")
def gpu_engine(native_libs):
    import spark_druid_olap_b200 as sdo
    eng = sdo.Engine(0)  # raises NoDevice without a GPU: there is no fallback
    yield eng
    eng.close()
