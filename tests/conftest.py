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
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def pytest_sessionfinish(session, exitstatus):
    """Worst fp32 errors seen by the north_star checks (tests/golden_cases.py:assert_fp32_north_star), for the record."""
    try:
        from tests import golden_cases as gc
        if gc.MARGINS:
            import json
            out = os.path.join(ROOT, "gpurun_out")
            os.makedirs(out, exist_ok=True)
            with open(os.path.join(out, "fp32_margins.json"), "w") as f:
                json.dump({k: dict(pos_m=v[0], ang_rad=v[1], other_rel=v[2]) for k, v in sorted(gc.MARGINS.items())}, f, indent=1)
    except Exception:  # noqa: BLE001
        pass
