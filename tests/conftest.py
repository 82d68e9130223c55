import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "pytorch-dppo_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


def pytest_sessionfinish(session, exitstatus):
    """Table of every comparison made through util.assert_close_scaled: per-tensor relative error and the
    element-wise atol it would need (written where gpurun brings it back; nothing is written on a box without it)."""
    try:
        import json
        import util
        out = os.path.join(ROOT, "gpurun_out")
        if util.CLOSENESS and os.path.isdir(out):
            worst = {}
            for c in util.CLOSENESS:
                w = worst.get(c["name"])
                if w is None or c["elem_atol_needed"] > w["elem_atol_needed"]:
                    worst[c["name"]] = c
            with open(os.path.join(out, "closeness_report.json"), "w") as f:
                json.dump(sorted(worst.values(), key=lambda c: -c["elem_atol_needed"]), f, indent=1)
    except Exception:                                            # noqa: BLE001  (a report must never fail the run)
        pass
