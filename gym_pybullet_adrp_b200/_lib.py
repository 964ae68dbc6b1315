"""Loader of the CUDA library (libadrp_b200.so, built in-tree by build.py) and a thin handle class.

There is NO CPU fallback: if the library is missing or a call fails, this raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import _abi

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
# ADRP_LIB_PATH: another build of the same library (tools/variant_build.sh: launch-bound / block-size experiments)
LIB_PATH = os.environ.get("ADRP_LIB_PATH") or os.path.join(PKG_DIR, "libadrp_b200.so")

EXPORTS = ["adrp_create", "adrp_destroy", "adrp_obs_dim", "adrp_action_dim", "adrp_noise_dim",
           "adrp_reset_draws_dim", "adrp_substeps", "adrp_reset", "adrp_set_state", "adrp_get_state",
           "adrp_step", "adrp_step_n", "adrp_set_rollout_step", "adrp_step_host", "adrp_dslpid_control", "adrp_launch_count", "adrp_set_profiling",
           "adrp_step_kernel_ms", "adrp_fp32_peak_tflops", "adrp_last_error", "adrp_abi_version", "adrp_config_size",
           "adrp_step_rows_static"]

_lib = None


class AdrpError(RuntimeError):
    pass


def load():
    """dlopen the CUDA library and declare the C signatures of include/adrp.h."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise AdrpError(f"{LIB_PATH} not found: build it with `python -m gym_pybullet_adrp_b200.build` "
                        "(nvcc, sm_100a). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    vp, u8p = C.c_void_p, C.c_void_p
    lib.adrp_create.argtypes = [C.POINTER(_abi.ConfigC), C.POINTER(vp)]
    lib.adrp_destroy.argtypes = [vp]
    for f in ("adrp_obs_dim", "adrp_action_dim", "adrp_noise_dim", "adrp_reset_draws_dim", "adrp_substeps"):
        getattr(lib, f).argtypes = [vp]
    lib.adrp_reset.argtypes = [vp, u8p, vp, vp, vp]
    lib.adrp_set_state.argtypes = [vp, vp, vp]
    lib.adrp_get_state.argtypes = [vp, vp, vp]
    lib.adrp_step.argtypes = [vp, vp, vp, vp, vp, u8p, u8p, u8p, vp]
    lib.adrp_step_n.argtypes = [vp, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, u8p, u8p, u8p, vp]
    lib.adrp_set_rollout_step.argtypes = [vp, C.c_uint32]
    lib.adrp_step_host.argtypes = [vp, vp, vp, vp, u8p, u8p, u8p]
    lib.adrp_dslpid_control.argtypes = [C.c_int, C.c_int, C.POINTER(_abi.DroneParamsC), C.c_double, C.c_int,
                                        C.c_double] + [vp] * 12
    lib.adrp_step_rows_static.argtypes = [C.POINTER(_abi.ConfigC), vp, vp, vp, vp, vp, u8p, u8p, vp]
    lib.adrp_launch_count.argtypes = [vp]
    lib.adrp_launch_count.restype = C.c_int64
    lib.adrp_set_profiling.argtypes = [vp, C.c_int]
    lib.adrp_step_kernel_ms.argtypes = [vp, C.c_int]
    lib.adrp_step_kernel_ms.restype = C.c_double
    lib.adrp_fp32_peak_tflops.argtypes = [C.c_int]
    lib.adrp_fp32_peak_tflops.restype = C.c_double
    lib.adrp_last_error.argtypes = [vp]
    lib.adrp_last_error.restype = C.c_char_p
    if lib.adrp_abi_version() != _abi.ABI_VERSION or lib.adrp_config_size() != C.sizeof(_abi.ConfigC):
        raise AdrpError("libadrp_b200.so does not match include/adrp.h as mirrored in _abi.py (rebuild)")
    _lib = lib
    return lib


def _torch():
    import torch
    return torch


class Handle:
    """Owns one adrp_handle; all tensors are torch CUDA tensors on the handle's device."""

    def __init__(self, cfg: _abi.ConfigC):
        self.lib = load()
        self.cfg = cfg
        self._h = C.c_void_p()
        rc = self.lib.adrp_create(C.byref(cfg), C.byref(self._h))
        if rc != 0:
            msg = self.lib.adrp_last_error(None).decode()
            if rc == _abi_EINVAL and "divisible" in msg:
                raise ValueError(msg)
            raise AdrpError(f"adrp_create failed ({rc}): {msg}")
        torch = _torch()
        self.device = torch.device("cuda", cfg.device)
        self.dtype = torch.float64 if cfg.precision == _abi.F64 else torch.float32
        self.np_dtype = np.float64 if cfg.precision == _abi.F64 else np.float32
        self.N, self.M = cfg.num_envs, cfg.drones_per_env
        self.O = self.lib.adrp_obs_dim(self._h)
        self.A = self.lib.adrp_action_dim(self._h)
        self.S = self.lib.adrp_substeps(self._h)
        self.R = self.lib.adrp_reset_draws_dim(self._h)
        self.obs = torch.zeros((self.N, self.M, self.O), dtype=self.dtype, device=self.device)
        self.reward = torch.zeros(self.N, dtype=self.dtype, device=self.device)
        self.terminated = torch.zeros(self.N, dtype=torch.uint8, device=self.device)
        self.truncated = torch.zeros(self.N, dtype=torch.uint8, device=self.device)
        self.flags = torch.zeros((self.N, self.M), dtype=torch.uint8, device=self.device)
        # (the outputs are persistent tensors: their addresses are looked up once, not on every step)
        self._out_ptrs = tuple(t.data_ptr() for t in (self.obs, self.reward, self.terminated, self.truncated, self.flags))
        self._act_numel = self.N * self.M * self.A

    # ------------------------------------------------------------------ helpers
    def _check(self, rc, what):
        if rc != 0:
            raise AdrpError(f"{what} failed ({rc}): {self.lib.adrp_last_error(self._h).decode()}")

    def _stream(self):
        return C.c_void_p(_torch().cuda.current_stream(self.device).cuda_stream)

    def _dev(self, x, shape):
        torch = _torch()
        t = torch.as_tensor(x, dtype=self.dtype, device=self.device).contiguous()
        if tuple(t.shape) != tuple(shape):
            t = t.reshape(shape)
        return t

    def close(self):
        if getattr(self, "_h", None):
            self.lib.adrp_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ API
    def reset(self, env_mask=None, draws=None):
        torch = _torch()
        m = None if env_mask is None else torch.as_tensor(env_mask, dtype=torch.uint8, device=self.device).contiguous()
        dr = None if draws is None else self._dev(draws, (self.N, self.R))
        self._check(self.lib.adrp_reset(self._h, None if m is None else m.data_ptr(),
                                        None if dr is None else dr.data_ptr(), self.obs.data_ptr(), self._stream()),
                    "adrp_reset")
        return self.obs

    def set_state(self, state):
        s = self._dev(state, (self.N, self.M, _abi.STATE_DIM))
        self._check(self.lib.adrp_set_state(self._h, s.data_ptr(), self._stream()), "adrp_set_state")

    def get_state(self):
        torch = _torch()
        s = torch.empty((self.N, self.M, _abi.STATE_DIM), dtype=self.dtype, device=self.device)
        self._check(self.lib.adrp_get_state(self._h, s.data_ptr(), self._stream()), "adrp_get_state")
        return s

    def step(self, action, noise=None):
        """Device path: ``action`` is (or is copied to) a CUDA tensor; outputs are the handle's tensors.

        ALIASING: ``obs, reward, terminated, truncated, flags`` are the handle's PERSISTENT output tensors; the next ``step``
        / ``reset`` overwrites them in place.  Rollout code that keeps a reference across steps (``obs_list.append(obs)``,
        ``next_obs`` vs ``obs``) must ``clone()``; ``step_n(..., keep_outputs=True)`` and ``vec_env.TorchVecEnv`` return
        fresh tensors."""
        torch = _torch()
        if torch.is_tensor(action) and action.dtype == self.dtype and action.device == self.device \
                and action.is_contiguous() and action.numel() == self._act_numel:
            a = action                                     # already what the kernel reads: no conversion, no reshape
        else:
            a = self._dev(action, (self.N, self.M, self.A))
        nz = None if noise is None else self._dev(noise, (self.N, self.M, self.S, 7))
        rc = self.lib.adrp_step(self._h, a.data_ptr(), None if nz is None else nz.data_ptr(), *self._out_ptrs, self._stream())
        if rc != 0:
            self._check(rc, "adrp_step")
        return self.obs, self.reward, self.terminated, self.truncated, self.flags

    def step_n(self, actions, n_steps=None, noise=None, keep_outputs=False):
        """``n_steps`` control steps in ONE launch (adrp_step_n), bit-identical to that many ``step`` calls.

        ``actions``: (K, N, M, A) -- one slice per step -- or (N, M, A), held for all ``n_steps`` (a fixed DSLPID target, a
        hover setpoint).  ``keep_outputs``: return every step's outputs, shaped (K, N, M, O) / (K, N) / (K, N, M), in
        fresh tensors; otherwise the handle's own tensors end up holding the last step's.  ``actions=None``: a random
        policy -- every component drawn U[-1, 1) inside the kernel (ADRP_ACTIONS_PHILOX: Philox4x32-10 keyed by the config's
        seed, counter (global env, 192 + drone, rollout step, 0); ``set_rollout_step`` replays)."""
        torch = _torch()
        if actions is None:
            a, per_step, K = None, 2, int(n_steps)
        else:
            a = torch.as_tensor(actions, dtype=self.dtype, device=self.device).contiguous()
            per_step = a.dim() == 4
            K = int(a.shape[0]) if per_step else int(n_steps)
            if n_steps is not None and per_step and int(n_steps) != K:
                raise ValueError("n_steps does not match the leading dimension of actions")
            a = a.reshape((K, self.N, self.M, self.A) if per_step else (self.N, self.M, self.A))
        nz = None
        if noise is not None:
            nz = torch.as_tensor(noise, dtype=self.dtype, device=self.device).contiguous()
            nz = nz.reshape((K, self.N, self.M, self.S, 7) if per_step else (self.N, self.M, self.S, 7))
        if keep_outputs:
            out = (torch.empty((K, self.N, self.M, self.O), dtype=self.dtype, device=self.device),
                   torch.empty((K, self.N), dtype=self.dtype, device=self.device),
                   torch.empty((K, self.N), dtype=torch.uint8, device=self.device),
                   torch.empty((K, self.N), dtype=torch.uint8, device=self.device),
                   torch.empty((K, self.N, self.M), dtype=torch.uint8, device=self.device))
        else:
            out = (self.obs, self.reward, self.terminated, self.truncated, self.flags)
        self._check(self.lib.adrp_step_n(self._h, K, int(per_step), int(keep_outputs), None if a is None else a.data_ptr(),
                                         None if nz is None else nz.data_ptr(), *[t.data_ptr() for t in out],
                                         self._stream()), "adrp_step_n")
        if keep_outputs:                                   # the handle's tensors mirror the last step, as after step()
            for dst, src in zip((self.obs, self.reward, self.terminated, self.truncated, self.flags), out):
                dst.copy_(src[-1])
        return out

    def set_rollout_step(self, step: int):
        """Third Philox counter word of the next ``step_n(None, ...)`` draw (adrp_set_rollout_step)."""
        self._check(self.lib.adrp_set_rollout_step(self._h, int(step)), "adrp_set_rollout_step")

    def step_host(self, action: np.ndarray):
        """Host path (NumPy in, NumPy out): H2D, launch, D2H inside the C call."""
        a = np.ascontiguousarray(action, self.np_dtype).reshape(self.N, self.M, self.A)
        if not hasattr(self, "_hbuf"):
            # the observation buffer is page-locked, so the library DMAs into it directly (no staging copy)
            # every output buffer is page-locked, so the library DMAs into them directly (no staging copy)
            torch = _torch()
            u8 = lambda shape: torch.empty(tuple(shape), dtype=torch.uint8, pin_memory=True).numpy()   # noqa: E731
            self._hbuf = (self.pinned_empty((self.N, self.M, self.O)), self.pinned_empty((self.N,)),
                          u8((self.N,)), u8((self.N,)), u8((self.N, self.M)))
        o, r, te, tr, fl = self._hbuf
        self._check(self.lib.adrp_step_host(self._h, a.ctypes.data, o.ctypes.data, r.ctypes.data, te.ctypes.data,
                                            tr.ctypes.data, fl.ctypes.data), "adrp_step_host")
        return o, r, te, tr, fl

    def pinned_empty(self, shape):
        """Page-locked NumPy array in the handle's dtype (for actions handed to ``step_host`` without a staging copy)."""
        torch = _torch()
        return torch.empty(tuple(shape), dtype=self.dtype, pin_memory=True).numpy()

    def set_profiling(self, on=True):
        self._check(self.lib.adrp_set_profiling(self._h, int(on)), "adrp_set_profiling")

    def step_kernel_ms(self, reset=True):
        return self.lib.adrp_step_kernel_ms(self._h, int(reset))

    @property
    def launch_count(self):
        return int(self.lib.adrp_launch_count(self._h))


_abi_EINVAL = -1


def fp32_peak_tflops(device=0):
    """Measured FP32 FMA peak of the device (TFLOP/s)."""
    v = load().adrp_fp32_peak_tflops(int(device))
    if v <= 0:
        raise AdrpError("adrp_fp32_peak_tflops failed")
    return float(v)


def dslpid_control(precision, drone_model, n, dt, pos, quat, vel, tpos, trpy, tvel, trates, ctrl, device=0):
    """Batched DSLPIDControl.computeControl on the GPU (torch tensors in / out; ``ctrl`` updated in place)."""
    from .assets import drone_params
    torch = _torch()
    lib = load()
    dtype = torch.float64 if precision == _abi.F64 else torch.float32
    dev = torch.device("cuda", device)
    ts = [torch.as_tensor(x, dtype=dtype, device=dev).contiguous() for x in (pos, quat, vel, tpos, trpy, tvel, trates)]
    rpm = torch.empty((n, 4), dtype=dtype, device=dev)
    pos_e = torch.empty((n, 3), dtype=dtype, device=dev)
    yaw_e = torch.empty(n, dtype=dtype, device=dev)
    dp = _abi.drone_params_c(drone_params(drone_model))
    rc = lib.adrp_dslpid_control(precision, _abi.MODEL_ID[drone_model], C.byref(dp), 9.8, n, float(dt),
                                 *[t.data_ptr() for t in ts], ctrl.data_ptr(), rpm.data_ptr(), pos_e.data_ptr(),
                                 yaw_e.data_ptr(), C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
    if rc != 0:
        raise AdrpError(f"adrp_dslpid_control failed ({rc})")
    return rpm, pos_e, yaw_e
