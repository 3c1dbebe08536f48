import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def ref():
    """The compiled reference (oracle/_ref/libcpprl_ref.so)."""
    from oracle import ref_binding
    if not ref_binding.available():
        pytest.skip("oracle/_ref/libcpprl_ref.so not built (run oracle/build_ref.sh where /root/reference exists)")
    return ref_binding
