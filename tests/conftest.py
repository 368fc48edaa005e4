import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_runtest_setup(item):
    """-m gpu tests compare with the compiled reference (oracle/_ref, shipped to the GPU box with the snapshot): a box
    without it must FAIL those tests, not skip them (a silent skip would leave the reference comparison gone)."""
    if item.get_closest_marker("gpu") is not None:
        item.config._leon_gpu_item = True


@pytest.hookimpl(hookwrapper=True)
def pytest_runtest_makereport(item, call):
    outcome = yield
    rep = outcome.get_result()
    if item.get_closest_marker("gpu") is not None and rep.skipped and call.excinfo is not None \
            and "oracle/_ref not built" in str(call.excinfo.value):
        rep.outcome = "failed"
        rep.longrepr = "oracle/_ref is missing on this box: a GPU parity test may not skip its reference comparison (%s)" % str(call.excinfo.value)
