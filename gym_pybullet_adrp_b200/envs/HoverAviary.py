"""``HoverAviary`` (gym_pybullet_adrp/envs/HoverAviary.py:8-132): one drone, reward max(0, 2 - |target - pos|^4),
truncation on leaving the 1.5 m box / tilting / 8 s."""
from __future__ import annotations

import numpy as np

from .. import _abi
from ..utils.enums import ActionType, DroneModel, ObservationType, Physics
from .BaseRLAviary import BaseRLAviary


class HoverAviary(BaseRLAviary):
    ENV_KIND = _abi.ENV_HOVER

    def __init__(self, drone_model: DroneModel = DroneModel.CF2X, initial_xyzs=None, initial_rpys=None,
                 physics: Physics = Physics.PYB, pyb_freq: int = 240, ctrl_freq: int = 30, gui=False, record=False,
                 obs: ObservationType = ObservationType.KIN, act: ActionType = ActionType.RPM, **batch):
        self.TARGET_POS = np.array([0, 0, 1])              # :51
        self.EPISODE_LEN_SEC = 8                           # :52
        super().__init__(drone_model=drone_model, num_drones=1, initial_xyzs=initial_xyzs, initial_rpys=initial_rpys,
                         physics=physics, pyb_freq=pyb_freq, ctrl_freq=ctrl_freq, gui=gui, record=record, obs=obs, act=act,
                         **batch)

    def _target_pos(self):
        return [self.TARGET_POS.tolist()]
