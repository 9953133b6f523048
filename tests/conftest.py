import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run by the driver with -m gpu)")
    config.addinivalue_line("markers", "needs_ref: needs oracle/_ref/libvkt_ref.so (the compiled reference)")


def pytest_collection_modifyitems(config, items):
    from oracle import pyoracle
    if pyoracle.have_reference():
        return
    skip = pytest.mark.skip(reason="oracle/_ref/libvkt_ref.so not built (reference sources absent)")
    for it in items:
        if "needs_ref" in it.keywords:
            it.add_marker(skip)
