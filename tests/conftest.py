import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(ROOT, "tests", "golden")


def pytest_collection_modifyitems(config, items):
    """`gpu` tests are skipped (not failed) on a box without a CUDA device or without the built library."""
    try:
        import torch
        have = torch.cuda.is_available()
    except Exception:
        have = False
    have = have and os.path.exists(os.path.join(ROOT, "alpha-zero_b200", "libaz_b200.so"))
    if have:
        return
    skip = pytest.mark.skip(reason="needs a CUDA device and libaz_b200.so")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)
