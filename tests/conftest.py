import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_terminal_summary(terminalreporter):
    """Achieved parity errors of every CUDA-vs-oracle comparison of this run (tests/parity.py)."""
    import parity
    if not parity.LOG:
        return
    terminalreporter.write_line("")
    terminalreporter.write_line("achieved parity errors (tolerances: 2e-2 relative, 1e-3 LSE, <= 2x the 16-bit-pipeline reference):")
    for line in parity.summary_lines():
        terminalreporter.write_line("  " + line)
    parity.write_log(ROOT)
