import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

REF_DIRS = [os.path.join(ROOT, "baseline", "_ref"), "/root/reference"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def _ref_dir():
    for d in REF_DIRS:
        if os.path.isfile(os.path.join(d, "cnn.py")):
            return d
    return None


@pytest.fixture(scope="session")
def ref_modules():
    """The reference's own modules (cnn, splitdataset), imported unmodified as the oracle."""
    d = _ref_dir()
    if d is None:
        pytest.skip("reference sources not available")
    sys.path.insert(0, d)
    try:
        import cnn
        import splitdataset
    finally:
        sys.path.remove(d)
    return {"cnn": cnn, "splitdataset": splitdataset, "dir": d}
