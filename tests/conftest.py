import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def built():
    """Makes sure the in-tree native artefacts exist (they are prebuilt by __graft_entry__.build())."""
    import avrecode_b200.build as b
    if not (os.path.exists(os.path.join(ROOT, "avrecode_b200", "libavrecode_b200.so"))
            and os.path.exists(os.path.join(ROOT, "oracle", "_build", "libavr_oracle.so"))
            and os.path.exists(os.path.join(ROOT, "build", "h264gen"))
            and os.path.exists(os.path.join(ROOT, "build", "libavr_emu.so"))):
        b.build()
    return True
