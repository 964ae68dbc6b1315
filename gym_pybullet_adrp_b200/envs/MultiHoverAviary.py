"""``MultiHoverAviary`` (gym_pybullet_adrp/envs/MultiHoverAviary.py:8-145): M drones, summed hover reward."""
from __future__ import annotations

import numpy as np

from .. import _abi
from ..utils.enums import ActionType, DroneModel, ObservationType, Physics
from .BaseRLAviary import BaseRLAviary


class MultiHoverAviary(BaseRLAviary):
    ENV_KIND = _abi.ENV_MULTIHOVER

    def __init__(self, drone_model: DroneModel = DroneModel.CF2X, num_drones: int = 2, neighbourhood_radius: float = np.inf,
                 initial_xyzs=None, initial_rpys=None, physics: Physics = Physics.PYB, pyb_freq: int = 240,
                 ctrl_freq: int = 30, gui=False, record=False, obs: ObservationType = ObservationType.KIN,
                 act: ActionType = ActionType.RPM, **batch):
        self.EPISODE_LEN_SEC = 8                           # :52
        super().__init__(drone_model=drone_model, num_drones=num_drones, neighbourhood_radius=neighbourhood_radius,
                         initial_xyzs=initial_xyzs, initial_rpys=initial_rpys, physics=physics, pyb_freq=pyb_freq,
                         ctrl_freq=ctrl_freq, gui=gui, record=record, obs=obs, act=act, **batch)
        self.TARGET_POS = self.INIT_XYZS + np.array([[0, 0, 1 / (i + 1)] for i in range(num_drones)])   # :71

    def _target_pos(self):
        return (self.INIT_XYZS + np.array([[0, 0, 1 / (i + 1)] for i in range(self.NUM_DRONES)])).tolist()
