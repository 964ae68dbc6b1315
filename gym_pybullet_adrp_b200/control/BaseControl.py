"""``BaseControl`` (gym_pybullet_adrp/control/BaseControl.py:9-224): controller interface."""
from __future__ import annotations

import numpy as np

from ..assets import drone_params
from ..utils.enums import DroneModel


class BaseControl:
    def __init__(self, drone_model: DroneModel, g: float = 9.8):
        self.DRONE_MODEL = drone_model
        p = drone_params(drone_model)                       # _getURDFParameter (:185-224)
        self.GRAVITY = g * p.mass
        self.KF, self.KM = p.kf, p.km
        self.reset()

    def reset(self):
        self.control_counter = 0

    def computeControlFromState(self, control_timestep, state, target_pos, target_rpy=np.zeros(3),
                                target_vel=np.zeros(3), target_rpy_rates=np.zeros(3)):
        """BaseControl.py:55-95: slices the 20-float state vector."""
        state = np.asarray(state)
        return self.computeControl(control_timestep=control_timestep, cur_pos=state[..., 0:3], cur_quat=state[..., 3:7],
                                   cur_vel=state[..., 10:13], cur_ang_vel=state[..., 13:16], target_pos=target_pos,
                                   target_rpy=target_rpy, target_vel=target_vel, target_rpy_rates=target_rpy_rates)

    def computeControl(self, control_timestep, cur_pos, cur_quat, cur_vel, cur_ang_vel, target_pos,
                       target_rpy=np.zeros(3), target_vel=np.zeros(3), target_rpy_rates=np.zeros(3)):
        raise NotImplementedError
