"""Test-side loader of the CPU oracle (oracle/liboracle.so).  Only tests, smoke() and bench.py's
cpu_baseline / --impl reference legs import this."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
_LIB = None


def build_oracle():
    subprocess.run(["make", "-s", "-C", ORACLE_DIR], check=True)


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(ORACLE_DIR, "liboracle.so")
        srcs = [os.path.join(ORACLE_DIR, f) for f in os.listdir(ORACLE_DIR) if f.endswith((".c", ".h"))]
        srcs.append(os.path.join(ROOT, "include", "adrp.h"))
        if not os.path.exists(path) or any(os.path.getmtime(s) > os.path.getmtime(path) for s in srcs):
            build_oracle()
        _LIB = C.CDLL(path)
        _LIB.orc_create.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
        _LIB.orc_step.argtypes = [C.c_void_p] * 8 + [C.c_int]
        for f in ("orc_destroy", "orc_obs_dim", "orc_action_dim", "orc_substeps", "orc_reset_draws_dim"):
            getattr(_LIB, f).argtypes = [C.c_void_p]
        _LIB.orc_reset.argtypes = [C.c_void_p] * 4
        _LIB.orc_set_state.argtypes = [C.c_void_p] * 2
        _LIB.orc_get_state.argtypes = [C.c_void_p] * 2
        _LIB.orc_race_post.argtypes = [C.c_void_p] * 6
    return _LIB


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class OracleEnv:
    """Host-buffer twin of gym_pybullet_adrp_b200._lib.Handle."""

    def __init__(self, cfg):
        self.cfg = cfg
        self.h = C.c_void_p()
        rc = lib().orc_create(C.byref(cfg), C.byref(self.h))
        if rc != 0:
            raise ValueError(f"orc_create failed: {rc}")
        self.N, self.M = cfg.num_envs, cfg.drones_per_env
        self.dtype = np.float64 if cfg.precision == 1 else np.float32
        self.O, self.A, self.S = lib().orc_obs_dim(self.h), lib().orc_action_dim(self.h), lib().orc_substeps(self.h)
        self.R = lib().orc_reset_draws_dim(self.h)

    def close(self):
        if self.h:
            lib().orc_destroy(self.h)
            self.h = None

    __del__ = close

    def reset(self, env_mask=None, draws=None):
        obs = np.zeros((self.N, self.M, self.O), self.dtype)
        m = None if env_mask is None else np.ascontiguousarray(env_mask, np.uint8)
        dr = None if draws is None else np.ascontiguousarray(draws, self.dtype)
        lib().orc_reset(self.h, _ptr(m), _ptr(dr), _ptr(obs))
        return obs

    def set_state(self, state):
        s = np.ascontiguousarray(state, self.dtype).reshape(self.N, self.M, 20)
        lib().orc_set_state(self.h, _ptr(s))

    def get_state(self):
        s = np.zeros((self.N, self.M, 20), self.dtype)
        lib().orc_get_state(self.h, _ptr(s))
        return s

    def step(self, action, noise=None, nthreads=1):
        a = np.ascontiguousarray(action, self.dtype).reshape(self.N, self.M, self.A)
        nz = None if noise is None else np.ascontiguousarray(noise, self.dtype)
        obs = np.zeros((self.N, self.M, self.O), self.dtype)
        rew = np.zeros(self.N, self.dtype)
        te, tr = np.zeros(self.N, np.uint8), np.zeros(self.N, np.uint8)
        fl = np.zeros((self.N, self.M), np.uint8)
        lib().orc_step(self.h, _ptr(a), _ptr(nz), _ptr(obs), _ptr(rew), _ptr(te), _ptr(tr), _ptr(fl), nthreads)
        return obs, rew, te, tr, fl

    def race_post(self):
        """Race env: gate progress, observation and flags of the state currently held (no substeps)."""
        obs = np.zeros((self.N, self.M, self.O), self.dtype)
        rew = np.zeros(self.N, self.dtype)
        te, tr = np.zeros(self.N, np.uint8), np.zeros(self.N, np.uint8)
        fl = np.zeros((self.N, self.M), np.uint8)
        rc = lib().orc_race_post(self.h, _ptr(obs), _ptr(rew), _ptr(te), _ptr(tr), _ptr(fl))
        assert rc == 0
        return obs, rew, te, tr, fl


def dslpid_rows(precision, model_id, drone_c, n, dt, pos, quat, vel, tpos, trpy, tvel, trates, ctrl):
    dt_ = np.float64 if precision == 1 else np.float32
    arrs = [np.ascontiguousarray(x, dt_) for x in (pos, quat, vel, tpos, trpy, tvel, trates)]
    ctrl = np.ascontiguousarray(ctrl, dt_).copy()
    rpm, pos_e, yaw_e = np.zeros((n, 4), dt_), np.zeros((n, 3), dt_), np.zeros(n, dt_)
    f = lib().orc_dslpid_control
    f.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_double, C.c_int, C.c_double] + [C.c_void_p] * 11
    f(precision, model_id, C.byref(drone_c), 9.8, n, dt, *[_ptr(a) for a in arrs], _ptr(ctrl), _ptr(rpm),
      _ptr(pos_e), _ptr(yaw_e))
    return rpm, pos_e, yaw_e, ctrl
