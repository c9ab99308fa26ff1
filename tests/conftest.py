"""pytest configuration: `gpu` marker, library / oracle fixtures.

`-m "not gpu"` runs here without a GPU: oracle vs golden vectors (and vs the compiled reference
when oracle/_ref exists), host-side mesh logic, C-ABI symbol table.  `-m gpu` are the parity
tests proper; they call the CUDA path through the C-ABI and fail (not skip) if the extension is
missing on a GPU box.
"""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
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
def lib():
    from minicfd_b200 import abi
    if not os.path.exists(abi.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    return abi.load()


@pytest.fixture(scope="session")
def oracle_lib():
    from oracle import oracle
    return oracle.load()


def pytest_terminal_summary(terminalreporter):
    """the asserted measure is per component; the worst per-DOF figure seen is reported."""
    try:
        from tests import common
    except Exception:
        return
    w = common._worst
    if w["per_component"] or w["per_dof"]:
        terminalreporter.write_line(
            f"parity: worst per-component max|d|/max|ref_c| = {w['per_component']:.3e}; "
            f"worst per-DOF |d|/|ref| (|ref| > 1e-12) = {w['per_dof']:.3e}")
    if common._relaxed:
        terminalreporter.write_line(f"parity: {len(common._relaxed)} assert_parity component checks passed above 1e-10 (bound in brackets):")
        worst = {}
        for test, what, err, bound in common._relaxed:
            key = (test, what)
            if key not in worst or err > worst[key][0]:
                worst[key] = (err, bound)
        for (test, what), (err, bound) in sorted(worst.items()):
            terminalreporter.write_line(f"  {test}  {what}: {err:.3e} [{bound:g}]")
