import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


def _has_gpu() -> bool:
    try:
        import h2b200
        return h2b200.load().h2b_device_count() > 0
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def ctx():
    import h2b200
    c = h2b200.Context()
    yield c
    c.close()


@pytest.fixture(scope="session")
def engine(ctx):
    import h2b200
    return h2b200.B200Engine(ctx)


@pytest.fixture
def knob(ctx):
    """Set tuning knobs of the shared context for one test (h2b_set_option) and restore the defaults afterwards."""
    touched = []

    def set_(name, value):
        ctx.set_option(name, value)
        touched.append(name)

    yield set_
    for name in touched:
        ctx.set_option(name, None)
