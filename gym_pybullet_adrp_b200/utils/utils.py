"""Helpers of the reference's ``utils/utils.py`` that the evaluation driver and user controllers rely on."""
from __future__ import annotations

import importlib.util
import sys
import time
from pathlib import Path

import numpy as np

from ..configs import load_config  # noqa: F401  (utils/utils.py:95-108)


def sync(i, start_time, timestep):
    """utils/utils.py:47-64: sleep so that iteration ``i`` of period ``timestep`` does not run ahead of the wall clock."""
    if timestep > .04 or i % (int(1 / (24 * timestep))) == 0:
        elapsed = time.time() - start_time
        if elapsed < (i * timestep):
            time.sleep(timestep * i - elapsed)


def map2pi(angle):
    """utils/utils.py: wrap an angle to [-pi, pi]."""
    return (np.asarray(angle) + np.pi) % (2 * np.pi) - np.pi


def load_controller(path, class_name: str = None):
    """utils/utils.py:112-135: import a controller module from a file path and return the class named like the file."""
    path = Path(path)
    assert path.exists(), f"Controller file not found: {path}"
    assert path.is_file(), f"Controller path is not a file: {path}"
    class_name = class_name or path.name.split(".")[0]
    spec = importlib.util.spec_from_file_location("controller", path)
    module = importlib.util.module_from_spec(spec)
    sys.modules["controller"] = module
    spec.loader.exec_module(module)
    assert hasattr(module, class_name), f"{path} does not define {class_name}"
    return getattr(module, class_name)


def str2bool(val):
    """utils/utils.py:70-91: argparse helper, the strings the reference's example scripts accept for a Boolean."""
    if isinstance(val, bool):
        return val
    if str(val).lower() in ("yes", "true", "t", "y", "1"):
        return True
    if str(val).lower() in ("no", "false", "f", "n", "0"):
        return False
    import argparse
    raise argparse.ArgumentTypeError("[ERROR] in str2bool(), a Boolean value is expected")
