"""``BaseRLAviary`` (gym_pybullet_adrp/envs/BaseRLAviary.py:15-322): normalised actions (RPM, ONE_D_RPM, PID,
ONE_D_PID, VEL), 12 kinematic floats + the action buffer of the last half second as observation."""
from __future__ import annotations

import numpy as np

from .. import _abi
from ..utils.enums import ActionType, DroneModel, ObservationType, Physics
from ._spaces import Box
from .BaseAviary import BaseAviary


class BaseRLAviary(BaseAviary):
    def __init__(self, drone_model: DroneModel = DroneModel.CF2X, num_drones: int = 1, neighbourhood_radius: float = np.inf,
                 initial_xyzs=None, initial_rpys=None, physics: Physics = Physics.PYB, pyb_freq: int = 240,
                 ctrl_freq: int = 240, gui=False, record=False, obs: ObservationType = ObservationType.KIN,
                 act: ActionType = ActionType.RPM, **batch):
        if obs != ObservationType.KIN:
            raise NotImplementedError("ObservationType.RGB needs a renderer: unavailable on this box")
        if act not in _abi.ACTION_ID or act == ActionType.MEL:
            print("[ERROR] in BaseRLAviary._actionSpace()")
            raise ValueError(f"unsupported ActionType {act}")
        self.ACTION_BUFFER_SIZE = int(ctrl_freq // 2)       # BaseRLAviary.py:66
        self.OBS_TYPE, self.ACT_TYPE = obs, act
        self.ACTION_TYPE = _abi.ACTION_ID[act]
        if act in (ActionType.PID, ActionType.VEL, ActionType.ONE_D_PID) and drone_model not in (DroneModel.CF2X, DroneModel.CF2P):
            print("[ERROR] in BaseRLAviary.__init()__, no controller is available for the specified drone_model")
        super().__init__(drone_model=drone_model, num_drones=num_drones, neighbourhood_radius=neighbourhood_radius,
                         initial_xyzs=initial_xyzs, initial_rpys=initial_rpys, physics=physics, pyb_freq=pyb_freq,
                         ctrl_freq=ctrl_freq, gui=gui, record=record, obstacles=False, user_debug_gui=False, **batch)
        if act == ActionType.VEL:
            self.SPEED_LIMIT = 0.03 * self.MAX_SPEED_KMH * (1000 / 3600)   # :94-95

    def _actionSpace(self):                                # BaseRLAviary.py:122-156
        size = _abi.ACTION_DIM[self.ACTION_TYPE]
        return Box(low=-np.ones((self.NUM_DRONES, size)), high=np.ones((self.NUM_DRONES, size)), dtype=np.float32)

    def _observationSpace(self):                           # BaseRLAviary.py:243-280
        lo = np.array([[-np.inf, -np.inf, 0, -np.inf, -np.inf, -np.inf, -np.inf, -np.inf, -np.inf, -np.inf, -np.inf, -np.inf]
                       for _ in range(self.NUM_DRONES)])
        hi = np.full((self.NUM_DRONES, 12), np.inf)
        size = _abi.ACTION_DIM[self.ACTION_TYPE]
        for _ in range(self.ACTION_BUFFER_SIZE):
            lo = np.hstack([lo, -np.ones((self.NUM_DRONES, size))])
            hi = np.hstack([hi, np.ones((self.NUM_DRONES, size))])
        return Box(low=lo, high=hi, dtype=np.float32)
