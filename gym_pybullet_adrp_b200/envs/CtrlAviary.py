"""``CtrlAviary`` (gym_pybullet_adrp/envs/CtrlAviary.py:8-200): raw RPM commands in, 20-float state rows out."""
from __future__ import annotations

import numpy as np

from .. import _abi
from ..utils.enums import DroneModel, Physics
from ._spaces import Box
from .BaseAviary import BaseAviary


class CtrlAviary(BaseAviary):
    ENV_KIND = _abi.ENV_CTRL
    ACTION_TYPE = _abi.ACT_RPM_RAW

    def __init__(self, drone_model: DroneModel = DroneModel.CF2X, num_drones: int = 1, neighbourhood_radius: float = np.inf,
                 initial_xyzs=None, initial_rpys=None, physics: Physics = Physics.PYB, pyb_freq: int = 240,
                 ctrl_freq: int = 240, gui=False, record=False, obstacles=False, user_debug_gui=True,
                 output_folder='results', **batch):
        super().__init__(drone_model=drone_model, num_drones=num_drones, neighbourhood_radius=neighbourhood_radius,
                         initial_xyzs=initial_xyzs, initial_rpys=initial_rpys, physics=physics, pyb_freq=pyb_freq,
                         ctrl_freq=ctrl_freq, gui=gui, record=record, obstacles=obstacles, user_debug_gui=user_debug_gui,
                         output_folder=output_folder, **batch)

    def _actionSpace(self):                                # CtrlAviary.py:74-87
        lo = np.zeros((self.NUM_DRONES, 4))
        hi = np.full((self.NUM_DRONES, 4), self.MAX_RPM)
        return Box(low=lo, high=hi, dtype=np.float32)

    def _observationSpace(self):                           # CtrlAviary.py:91-102
        lo = np.array([[-np.inf, -np.inf, 0., -1., -1., -1., -1., -np.pi, -np.pi, -np.pi, -np.inf, -np.inf, -np.inf,
                        -np.inf, -np.inf, -np.inf, 0., 0., 0., 0.] for _ in range(self.NUM_DRONES)])
        hi = np.array([[np.inf, np.inf, np.inf, 1., 1., 1., 1., np.pi, np.pi, np.pi, np.inf, np.inf, np.inf,
                        np.inf, np.inf, np.inf, self.MAX_RPM, self.MAX_RPM, self.MAX_RPM, self.MAX_RPM]
                       for _ in range(self.NUM_DRONES)])
        if self._obs_columns == 12:                        # [pos3, rpy3, vel3, ang_vel3] (not in the reference: obs_columns=12)
            keep = [0, 1, 2, 7, 8, 9, 10, 11, 12, 13, 14, 15]
            lo, hi = lo[:, keep], hi[:, keep]
        return Box(low=lo, high=hi, dtype=np.float64 if self.PRECISION == _abi.F64 else np.float32)
