"""``MultiRaceAviary`` (gym_pybullet_adrp/envs/MultiRaceAviary.py:25-725): M drones race through the gates of a
``level*.yaml`` track; full-state setpoints ``[x, y, z, yaw]`` per drone go to a Mellinger controller that runs
every physics substep (500 Hz) inside the step kernel, next to the gate / obstacle / drone-drone contact tests.

The reference spawns one OS process per drone for the firmware (MultiRaceAviary.py:107-115); here the firmware
state of every drone of every env lives in HBM and the controller is a device function.
"""
from __future__ import annotations

import numpy as np

from .. import _abi
from ..utils.constants import CTRL_FREQ, FIRMWARE_FREQ
from ..utils.enums import ActionType, Command, DroneModel, ObservationType, Physics, RaceMode
from ._spaces import Box
from .BaseAviary import BaseAviary


class MultiRaceAviary(BaseAviary):
    ENV_KIND = _abi.ENV_RACE
    ACTION_TYPE = _abi.ACT_MEL

    def __init__(self, race_config, drone_model: DroneModel = DroneModel.CF2X, num_drones: int = 2,
                 physics: Physics = Physics.PYB, pyb_freq: int = FIRMWARE_FREQ, ctrl_freq: int = CTRL_FREQ,
                 gui: bool = False, record: bool = False, racemode: RaceMode = RaceMode.COMPARE,
                 obs: ObservationType = ObservationType.KIN, act: ActionType = ActionType.PID, **batch):
        if obs != ObservationType.KIN:
            raise NotImplementedError("ObservationType.RGB needs a renderer: unavailable on this box")
        self.config = race_config
        self.racemode = racemode
        self.observation_type, self.action_type = obs, act
        self.gates_nominal = np.array(race_config["gates"], dtype=float)          # :82
        self.obstacles_nominal = np.array(race_config["obstacles"], dtype=float)
        self.bounds = np.array(race_config["bounds"], dtype=float)                # :84
        self.EPISODE_LEN_SEC = race_config["episode_len_sec"]
        names = list(race_config["init_states"].keys())[:num_drones]
        init_xyzs = np.array([race_config["init_states"][d]["pos"] for d in names], dtype=float)
        init_rpys = np.array([race_config["init_states"][d]["rpy"] for d in names], dtype=float) * np.pi / 180   # :95
        self._physics_mask = batch.pop("physics_mask", None)
        self._wrappers = int(batch.pop("wrappers", 0))       # _abi.WRAP_* (utils/wrapper.py fused into the step kernel)
        # (Command, args) tuple actions are 16-float command rows (ADRP_ACT_MEL_CMD: the firmware's high-level commander runs
        # on the device) -- the default for a single env, like the reference; full_state_actions: 13-float FULLSTATE rows
        # (ADRP_ACT_MEL_FULL); the batched default is the 4-float [x, y, z, yaw] row of the reference's ndarray path (:190-195)
        full, cmd = batch.pop("full_state_actions", None), batch.pop("command_actions", None)
        self._cmd = (batch.get("num_envs", 1) == 1 and not full) if cmd is None else bool(cmd)
        self._full = bool(full) and not self._cmd
        super().__init__(drone_model=drone_model, num_drones=num_drones, initial_xyzs=init_xyzs, initial_rpys=init_rpys,
                         physics=physics, pyb_freq=pyb_freq, ctrl_freq=ctrl_freq, gui=gui, record=record, **batch)

    def _make_config(self):
        return _abi.make_race_config(self.config, num_envs=self.NUM_ENVS, num_drones=self.NUM_DRONES,
                                     drone_model=self.DRONE_MODEL, physics=self.PHYSICS, pyb_freq=self.PYB_FREQ,
                                     ctrl_freq=self.CTRL_FREQ, racemode=self.racemode, precision=self.PRECISION,
                                     device=self.DEVICE, seed=self._seed or None, env_offset=self._env_offset,
                                     auto_reset=self._auto_reset, physics_mask=self._physics_mask,
                                     wrappers=self._wrappers, full_state_actions=self._full, command_actions=self._cmd)

    # ------------------------------------------------------------------ spaces (:284-343)
    def _actionSpace(self):
        lim = np.ones((self.NUM_DRONES, 4))
        return Box(low=-1 * lim, high=lim, dtype=np.float64)

    def _observationSpace(self):
        G, NO = len(self.config["gates"]), len(self.config["obstacles"])
        lo = [-5] * 3 + [-np.pi] * 3 + [-10] * 6 + [-5, -5, -5, -np.pi] * G + [-1] * G + [-5] * 3 * NO + [-1] * NO + [-1]
        hi = [5] * 3 + [np.pi] * 3 + [10] * 6 + [5, 5, 5, np.pi] * G + [1] * G + [5] * 3 * NO + [1] * NO + [4]
        if self.racemode == RaceMode.COMPETE:
            lo += ([-5] * 3 + [-np.pi] * 3) * (self.NUM_DRONES - 1)
            hi += ([5] * 3 + [np.pi] * 3) * (self.NUM_DRONES - 1)
        lo, hi = np.array(lo, dtype=float), np.array(hi, dtype=float)
        return Box(low=np.vstack([lo] * self.NUM_DRONES), high=np.vstack([hi] * self.NUM_DRONES),
                   dtype=np.float64 if self.PRECISION == _abi.F64 else np.float32)

    # ------------------------------------------------------------------ API
    def reset(self, seed: int = None, options: dict = None, draws=None):
        """MultiRaceAviary.reset (:127-167).  ``draws``: optional ``(num_envs, R)`` uniform numbers replacing the Philox
        stream (the reference's RNG is unseeded, so tests inject its recorded draws)."""
        obs = self.handle.reset(draws=draws)
        self.step_counter = 0
        return self._out_obs(obs), self._computeInfo()

    def _rows_from_commands(self, action):
        """[(Command, args)] per drone -> (NUM_DRONES, 16) command rows (MultiRaceAviary.py:197-205): every Command of
        utils/enums.py:58-70 -- FULLSTATE, the TAKEOFF / LAND variants, GOTO, STOP, NOTIFY, NONE."""
        if not self._cmd:
            raise ValueError("(Command, args) actions need command_actions=True (16-float ADRP_ACT_MEL_CMD rows)")
        return np.stack([_abi.command_row(cmd, args) for cmd, args in action])

    def step(self, action, noise=None):
        """MultiRaceAviary.step (:171-270).  ``action``: ``(NUM_DRONES, 4)`` full-state setpoints [x, y, z, yaw]
        (``(num_envs, NUM_DRONES, 4)`` when batched); ``(..., 13)`` rows [pos, vel, acc, yaw, rpy_rates] with
        ``full_state_actions=True``; ``(..., 16)`` command rows (``_abi.command_row``) with ``command_actions=True``; or --
        single env -- the reference's list of ``(Command, args)`` tuples, one per drone."""
        if isinstance(action, (list, tuple)) and len(action) and isinstance(action[0], (list, tuple)) \
                and len(action[0]) == 2 and isinstance(action[0][0], Command):
            if not self._single:
                raise ValueError("(Command, args) tuples are the single-env interface; pass (num_envs, NUM_DRONES, 16) rows")
            action = self._rows_from_commands(action)
        elif self._full or self._cmd:
            a = np.asarray(action.cpu() if hasattr(action, "cpu") else action, dtype=float)
            if a.shape[-1] == 4:                                   # ndarray path: FULLSTATE(pos, 0, 0, yaw, 0, step_counter) (:190-195)
                full = np.zeros(a.shape[:-1] + ((16,) if self._cmd else (13,)))
                o = 1 if self._cmd else 0
                full[..., o:o + 3], full[..., o + 9] = a[..., 0:3], a[..., 3]
                if self._cmd:
                    full[..., 0], full[..., 14] = _abi.CMD_ID["fst"], self.step_counter
                action = full
        h = self.handle
        obs, reward, terminated, truncated, self._last_flags = h.step(self._in_action(action), noise=noise)
        self.step_counter += self.PYB_STEPS_PER_CTRL
        if self._single:
            return self._out_obs(obs), float(reward[0]), bool(terminated[0]), bool(truncated[0]), self._computeInfo()
        return obs, reward, terminated.bool(), truncated.bool(), self._computeInfo()

    # state the reference exposes as attributes
    def _flags(self):
        f = getattr(self, "_last_flags", None)
        if f is None:
            return np.zeros((self.NUM_ENVS, self.NUM_DRONES), np.uint8)
        return f.cpu().numpy()

    @property
    def drones_eliminated(self):
        e = (self._flags() & _abi.FLAG_ELIMINATED) != 0
        return e[0] if self._single else e

    @property
    def drones_finished(self):
        e = (self._flags() & _abi.FLAG_FINISHED) != 0
        return e[0] if self._single else e

    @property
    def current_gate(self):
        G, NO = len(self.config["gates"]), len(self.config["obstacles"])
        g = self.handle.obs[..., 12 + 5 * G + 4 * NO].cpu().numpy().astype(float)
        return g[0] if self._single else g
