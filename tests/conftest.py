import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

GOLDEN = ROOT / "tests" / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (sm_100a); run with -m gpu on the B200 box")


@pytest.fixture(scope="session")
def built_library():
    """The in-tree CUDA library; compiled here if absent (nvcc cross-compiles without a GPU)."""
    from yoho_b200 import build
    return build.build()


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN
