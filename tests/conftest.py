import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:  # pragma: no cover
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device here")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


def pytest_terminal_summary(terminalreporter, exitstatus, config):
    """One line per checked greedy sequence (oracle/parity.py) so the counts survive `-q`."""
    try:
        from oracle.parity import REPORT
    except Exception:  # pragma: no cover
        return
    if REPORT:
        terminalreporter.write_line("")
        terminalreporter.write_line("greedy parity vs the fp32 oracle (teacher-forced, near-ties re-aligned):")
        for line in REPORT:
            terminalreporter.write_line("  " + line)
