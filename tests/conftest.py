import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run by the driver with -m gpu)")


@pytest.fixture(scope="session")
def port_oracle():
    import oracle
    oracle.build()
    return oracle.Oracle("port")


@pytest.fixture(scope="session")
def ref_oracle():
    import oracle
    if not oracle.have_reference():
        pytest.skip("oracle/_ref/libadacom_ref.so not built (needs /root/reference)")
    return oracle.Oracle("reference")


@pytest.fixture(scope="session")
def gpu_ctx():
    import duckdb_adaptive_compression_b200 as adacom
    ctx = adacom.Context(0)  # raises loudly when there is no sm_100 device
    yield ctx
    ctx.close()
