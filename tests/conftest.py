import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def port_oracle():
    from oracle.pyoracle import Oracle
    return Oracle("port")


@pytest.fixture(scope="session")
def ref_oracle():
    """The reference's own code (oracle/_ref); present when built where /root/reference exists."""
    from oracle.pyoracle import Oracle, REF_SO
    if not os.path.exists(REF_SO):
        pytest.skip("oracle/_ref not built (reference sources absent)")
    return Oracle("reference")


@pytest.fixture(scope="session")
def engine_cli():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from dindel_tgi_b200 import engine
    e = engine.Engine(engine.cli_params(), 0)
    yield e
    e.close()
