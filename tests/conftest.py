import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle_mod():
    from oracle import oracle as om
    om.build()
    return om


def pytest_terminal_summary(terminalreporter):
    """Largest relative error each 1e-12 parity assertion achieved in this session (tests/dsl.py assert_close)."""
    try:
        from dsl import ACHIEVED
    except Exception:
        return
    if ACHIEVED:
        terminalreporter.write_line("achieved relative errors (bar: 1e-12):")
        for k in sorted(ACHIEVED):
            terminalreporter.write_line(f"  {k}: {ACHIEVED[k]:.3e}")
