"""Batched ``BaseAviary``: the reference's env base class (gym_pybullet_adrp/envs/BaseAviary.py:19-1154) over the
CUDA step kernel.  Same constructor arguments, attributes and ``reset`` / ``step`` / ``close`` conventions; the
extra keyword-only arguments (``num_envs``, ``device``, ``precision``, ...) select the batch.

* ``num_envs == 1`` (default): NumPy in, NumPy out, shapes and types exactly as the reference returns them
  (obs ``(NUM_DRONES, O)``, scalar reward, bool flags) -- a drop-in for one env.
* ``num_envs > 1``: torch CUDA tensors with a leading env axis; ``action`` may be NumPy or torch.

Everything graphical (``gui``, ``record``, ``vision_attributes``) is unavailable on this box and raises.
"""
from __future__ import annotations

import time

import numpy as np

from .. import _abi
from ..assets import derived_constants, drone_params
from ..utils.enums import DroneModel, Physics


class BaseAviary:
    ENV_KIND = _abi.ENV_CTRL
    ACTION_TYPE = _abi.ACT_RPM_RAW

    def __init__(self, drone_model: DroneModel = DroneModel.CF2X, num_drones: int = 1, neighbourhood_radius: float = np.inf,
                 initial_xyzs=None, initial_rpys=None, physics: Physics = Physics.PYB, pyb_freq: int = 240,
                 ctrl_freq: int = 240, gui=False, record=False, obstacles=False, user_debug_gui=True,
                 vision_attributes=False, output_folder='results', *, num_envs: int = 1, device: int = 0,
                 precision: str = "fp32", auto_reset: bool | None = None, seed: int = 0, env_offset: int = 0,
                 obs_columns: int = 0):
        if gui or record or vision_attributes:
            raise NotImplementedError("GUI / video / camera observations need a graphics stack: unavailable on this box")
        #### Constants (BaseAviary.py:74-128) ######################
        self.G = 9.8
        self.RAD2DEG, self.DEG2RAD = 180 / np.pi, np.pi / 180
        self.CTRL_FREQ, self.PYB_FREQ = ctrl_freq, pyb_freq
        if self.PYB_FREQ % self.CTRL_FREQ != 0:
            raise ValueError('[ERROR] in BaseAviary.__init__(), pyb_freq is not divisible by env_freq.')
        self.PYB_STEPS_PER_CTRL = int(self.PYB_FREQ / self.CTRL_FREQ)
        self.CTRL_TIMESTEP, self.PYB_TIMESTEP = 1. / self.CTRL_FREQ, 1. / self.PYB_FREQ
        self.NUM_DRONES, self.NEIGHBOURHOOD_RADIUS = num_drones, neighbourhood_radius
        self.DRONE_MODEL, self.PHYSICS = drone_model, physics
        self.GUI, self.RECORD, self.OBSTACLES, self.USER_DEBUG = False, False, obstacles, user_debug_gui
        self.URDF, self.OUTPUT_FOLDER = drone_model.value + ".urdf", output_folder
        p = drone_params(drone_model)
        self.M, self.L, self.THRUST2WEIGHT_RATIO = p.mass, p.arm, p.thrust2weight
        self.J = np.diag([p.ixx, p.iyy, p.izz])
        self.J_INV = np.linalg.inv(self.J)
        self.KF, self.KM = p.kf, p.km
        self.COLLISION_H, self.COLLISION_R, self.COLLISION_Z_OFFSET = p.collision_h, p.collision_r, p.collision_z_offset
        self.MAX_SPEED_KMH, self.GND_EFF_COEFF, self.PROP_RADIUS = p.max_speed_kmh, p.gnd_eff_coeff, p.prop_radius
        self.DRAG_COEFF = np.array([p.drag_coeff_xy, p.drag_coeff_xy, p.drag_coeff_z])
        self.DW_COEFF_1, self.DW_COEFF_2, self.DW_COEFF_3 = p.dw_coeff_1, p.dw_coeff_2, p.dw_coeff_3
        dc = derived_constants(p, drone_model, self.G)
        self.GRAVITY, self.HOVER_RPM, self.MAX_RPM, self.MAX_THRUST = dc.GRAVITY, dc.HOVER_RPM, dc.MAX_RPM, dc.MAX_THRUST
        self.MAX_XY_TORQUE, self.MAX_Z_TORQUE, self.GND_EFF_H_CLIP = dc.MAX_XY_TORQUE, dc.MAX_Z_TORQUE, dc.GND_EFF_H_CLIP
        #### Initial poses (BaseAviary.py:193-206) #################
        if initial_xyzs is None:
            self.INIT_XYZS = np.array(_abi.default_init_xyz(p, num_drones), dtype=float)
        elif np.array(initial_xyzs).shape == (self.NUM_DRONES, 3):
            self.INIT_XYZS = np.array(initial_xyzs, dtype=float)
        else:
            raise ValueError("[ERROR] invalid initial_xyzs in BaseAviary.__init__(), try initial_xyzs.reshape(NUM_DRONES,3)")
        if initial_rpys is None:
            self.INIT_RPYS = np.zeros((self.NUM_DRONES, 3))
        elif np.array(initial_rpys).shape == (self.NUM_DRONES, 3):
            self.INIT_RPYS = np.array(initial_rpys, dtype=float)
        else:
            raise ValueError("[ERROR] invalid initial_rpys in BaseAviary.__init__(), try initial_rpys.reshape(NUM_DRONES,3)")
        #### Batch ##################################################
        self.NUM_ENVS, self.DEVICE = int(num_envs), int(device)
        self.PRECISION = {"fp32": _abi.F32, "fp64": _abi.F64}[precision]
        self._auto_reset = (self.NUM_ENVS > 1) if auto_reset is None else bool(auto_reset)
        self._seed, self._env_offset = int(seed), int(env_offset)
        self._obs_columns = int(obs_columns)               # CtrlAviary-kind: 12 = the kinematic columns only (adrp_config.obs_columns)
        self._single = self.NUM_ENVS == 1
        self._handle = None
        self.CLIENT = -1                                   # there is no PyBullet client
        self.step_counter = 0
        self.RESET_TIME = time.time()
        self.action_space = self._actionSpace()
        self.observation_space = self._observationSpace()

    # ------------------------------------------------------------------ config / handle
    def _make_config(self):
        return _abi.make_config(num_envs=self.NUM_ENVS, num_drones=self.NUM_DRONES, drone_model=self.DRONE_MODEL,
                                physics=self.PHYSICS, pyb_freq=self.PYB_FREQ, ctrl_freq=self.CTRL_FREQ,
                                action_type=self.ACTION_TYPE, env_kind=self.ENV_KIND, precision=self.PRECISION,
                                device=self.DEVICE, initial_xyzs=self.INIT_XYZS.tolist(), initial_rpys=self.INIT_RPYS.tolist(),
                                target_pos=self._target_pos(), episode_len_sec=getattr(self, "EPISODE_LEN_SEC", 8.0),
                                auto_reset=self._auto_reset, seed=self._seed, env_offset=self._env_offset,
                                obs_columns=self._obs_columns)

    def _target_pos(self):
        return None

    @property
    def handle(self):
        """The C-ABI handle (created on first use; raises if the CUDA library or a GPU is missing)."""
        if self._handle is None:
            from .. import _lib
            self._handle = _lib.Handle(self._make_config())
        return self._handle

    # ------------------------------------------------------------------ gymnasium API
    def reset(self, seed: int = None, options: dict = None):
        """BaseAviary.reset (:223-258): like the reference, ``seed`` is accepted and ignored (its RNG is unseeded)."""
        obs = self.handle.reset()
        self.step_counter = 0
        self.RESET_TIME = time.time()
        return self._out_obs(obs), self._computeInfo()

    def step(self, action):
        """BaseAviary.step (:262-387): one control step = PYB_STEPS_PER_CTRL fused physics substeps, one launch."""
        h = self.handle
        a = self._in_action(action)
        obs, reward, terminated, truncated, self._last_flags = h.step(a)
        self.step_counter += self.PYB_STEPS_PER_CTRL
        if self._single:
            return (self._out_obs(obs), float(reward[0]), bool(terminated[0]), bool(truncated[0]), self._computeInfo())
        return obs, reward, terminated.bool(), truncated.bool(), self._computeInfo()

    def step_n(self, actions=None, n_steps=None, keep_outputs=False):
        """``n_steps`` calls of ``step`` in ONE launch (adrp_step_n; no counterpart in the reference, whose loop is one Python
        call per control step -- examples/pid.py:138-147, examples/learn.py's rollouts).  ``actions``: ``(K, ...)`` slices in
        the shape ``step`` takes, one per control step; one such slice, held for ``n_steps``; or ``None``: a random policy, every
        component drawn U[-1, 1) inside the kernel (``action_space.sample()`` without the host).  Returns torch tensors
        ``(obs, reward, terminated, truncated, info)`` of the last step, or of every step ``(K, num_envs, ...)`` with
        ``keep_outputs`` (envs that finish inside the launch restart there when ``auto_reset`` is on)."""
        import torch
        h = self.handle
        a = None
        if actions is not None:
            a = actions.to(device=h.device, dtype=h.dtype) if isinstance(actions, torch.Tensor) else \
                torch.tensor(np.asarray(actions), dtype=h.dtype, device=h.device)
            per_step = a.numel() != h.N * h.M * h.A
            a = a.reshape((-1, h.N, h.M, h.A) if per_step else (h.N, h.M, h.A))
            if per_step and n_steps is None:
                n_steps = int(a.shape[0])
        obs, reward, terminated, truncated, flags = h.step_n(a, n_steps=n_steps, keep_outputs=keep_outputs)
        self._last_flags = flags[-1] if keep_outputs else flags
        self.step_counter += int(n_steps) * self.PYB_STEPS_PER_CTRL
        return obs, reward, terminated.bool(), truncated.bool(), self._computeInfo()

    def render(self, mode='human', close=False):
        """BaseAviary.render (:391-416): text only."""
        t = self.step_counter * self.PYB_TIMESTEP
        wall = time.time() - self.RESET_TIME
        print("\n[INFO] BaseAviary.render() ——— it {:04d}".format(self.step_counter),
              "——— wall-clock time {:.1f}s,".format(wall),
              "simulation time {:.1f}s@{:d}Hz ({:.2f}x)".format(t, self.PYB_FREQ, t / max(wall, 1e-9)))

    def close(self):
        if self._handle is not None:
            self._handle.close()
            self._handle = None

    def getPyBulletClient(self):
        return self.CLIENT

    def getDroneIds(self):
        return np.arange(1, self.NUM_DRONES + 1)

    # ------------------------------------------------------------------ helpers
    def _in_action(self, action):
        import torch
        h = self.handle
        if isinstance(action, torch.Tensor):
            a = action.to(device=h.device, dtype=h.dtype)
        else:
            a = torch.tensor(np.asarray(action), dtype=h.dtype, device=h.device)      # (copies: read-only arrays are fine)
        return a.reshape(h.N, h.M, h.A)

    def _out_obs(self, obs):
        if not self._single:
            return obs
        return obs[0].cpu().numpy().astype(self.observation_space.dtype, copy=False)

    def _getDroneStateVector(self, nth_drone, env: int = 0):
        """20-vector [pos3 quat4 rpy3 vel3 ang_v3 last_clipped_action4] (BaseAviary.py:545-565)."""
        import torch
        s = self.handle.get_state()[env, nth_drone].double().cpu().numpy()
        x, y, z, w = s[3:7]
        sarg = -2 * (x * z - w * y)
        if sarg <= -0.99999:
            rpy = [0, -np.pi / 2, 2 * np.arctan2(x, -y)]
        elif sarg >= 0.99999:
            rpy = [0, np.pi / 2, 2 * np.arctan2(-x, y)]
        else:
            rpy = [np.arctan2(2 * (y * z + w * x), w * w - x * x - y * y + z * z), np.arcsin(sarg),
                   np.arctan2(2 * (x * y + w * z), w * w + x * x - y * y - z * z)]
        del torch
        return np.hstack([s[0:3], s[3:7], rpy, s[7:10], s[10:13], s[16:20]])

    def _computeInfo(self):
        return {"answer": 42}

    def _actionSpace(self):
        raise NotImplementedError

    def _observationSpace(self):
        raise NotImplementedError
