import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "slow: takes more than ~20 s on CPU")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def load_golden(name):
    return torch.load(os.path.join(GOLDEN_DIR, name + ".pt"), weights_only=False)


def pytest_terminal_summary(terminalreporter):
    """measured parity deviations (tests/helpers.record), so they land in the GPU box's test log"""
    try:
        from helpers import DEVIATIONS
    except Exception:
        return
    if not DEVIATIONS:
        return
    terminalreporter.write_sep("-", "measured parity deviations (max relative; bound)")
    seen = set()
    for case, what, dev, bound in DEVIATIONS:
        key = (case, what)
        if key in seen:
            continue
        seen.add(key)
        flag = "" if dev <= bound else "   <-- over"
        terminalreporter.write_line(f"{case:28s} {what:72s} {dev:9.2e}  ({bound:.1e}){flag}")
