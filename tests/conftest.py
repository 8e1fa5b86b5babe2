import os
import sys

import pytest

# Several tests run a handful of library contexts (3 streams each) side by side on ONE GPU, and the sharded-MSM test lets their
# exchange kernels wait for each other.  With the default of 8 hardware queues, streams of different contexts alias one queue and
# a waiting kernel can sit in front of the work it waits for; give every stream its own queue.  (Test set-up only: in production
# every rank owns its GPU.  Must be set before the CUDA runtime initialises.)
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

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
