"""CPU: the C-ABI library builds, loads, and exports every symbol include/adrp.h declares."""
import ctypes as C
import os
import re

import pytest

from gym_pybullet_adrp_b200 import _abi, _lib, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    build.build()
    return _lib.load()


def test_header_symbols_exported(lib):
    hdr = open(os.path.join(ROOT, "include", "adrp.h")).read()
    declared = set(re.findall(r"\b(adrp_[a-z_0-9]+)\s*\(", hdr))
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), name


def test_struct_layout_matches(lib):
    assert lib.adrp_abi_version() == _abi.ABI_VERSION
    assert lib.adrp_config_size() == C.sizeof(_abi.ConfigC)


def test_create_rejects_bad_config_without_gpu(lib):
    cfg = _abi.make_config(num_envs=4)
    cfg.abi_version = 1
    h = C.c_void_p()
    assert lib.adrp_create(C.byref(cfg), C.byref(h)) == -3          # ADRP_EABI
    cfg = _abi.make_config(num_envs=4)
    cfg.ctrl_freq = 7
    assert lib.adrp_create(C.byref(cfg), C.byref(h)) == -1
    assert b"divisible" in lib.adrp_last_error(None)


def test_make_config_errors_like_reference():
    with pytest.raises(ValueError):                                 # BaseAviary.py:79-80
        _abi.make_config(num_envs=1, pyb_freq=240, ctrl_freq=7)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "gym_pybullet_adrp_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert not re.search(r"import[^\n]*oracle|from[^\n]*oracle|liboracle|oracle_lib|oracle/", src), f


def test_no_gpu_fails_loudly_and_null_handles_are_safe(lib):
    """No CPU fallback: a valid config on a machine without a usable device is an error with a message, and the
    query entry points tolerate a NULL handle."""
    import torch
    assert lib.adrp_obs_dim(None) == 0 and lib.adrp_action_dim(None) == 0 and lib.adrp_substeps(None) == 0
    assert lib.adrp_launch_count(None) == 0 and lib.adrp_destroy(None) == -1
    assert lib.adrp_step_host(None, None, None, None, None, None, None) == -1
    cfg = _abi.make_config(num_envs=4)
    cfg.device = 9999 if torch.cuda.is_available() else 0
    h = C.c_void_p()
    assert lib.adrp_create(C.byref(cfg), C.byref(h)) == -2           # ADRP_ECUDA
    assert b"cudaSetDevice" in lib.adrp_last_error(None)
    assert not h.value
