import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _has_gpu() -> bool:
    try:
        import torch

        return bool(torch.cuda.is_available())
    except Exception:
        return False


@pytest.hookimpl(tryfirst=True)
def pytest_cmdline_main(config):
    """The CPU suite starts real masters, agents and harness processes and takes over half an hour one test at a time; without
    an explicit -n it is spread over four pytest-xdist workers (DET_TEST_WORKERS=0 keeps it serial, =N picks another count).
    Runs that see a GPU stay serial: device tests time kernels and must not share the device with each other."""
    if os.environ.get("PYTEST_XDIST_WORKER") or not config.pluginmanager.hasplugin("xdist"):
        return None
    if getattr(config.option, "numprocesses", None) is not None or getattr(config.option, "collectonly", False):
        return None
    workers = int(os.environ.get("DET_TEST_WORKERS", "4"))
    if workers > 0 and not _has_gpu():
        config.option.numprocesses = workers
    return None


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "slow: longer-running CPU test")


def pytest_collection_modifyitems(config, items):
    try:
        import torch

        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)
