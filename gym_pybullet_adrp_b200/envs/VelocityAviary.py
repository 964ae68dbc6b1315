"""``VelocityAviary`` (gym_pybullet_adrp/envs/VelocityAviary.py:9-200): velocity commands in, 20-float state rows out.

The action is ``(vx, vy, vz, speed_fraction)`` per drone; the step kernel turns it into a DSLPID call with the
current position as position target, the current yaw as yaw target and ``SPEED_LIMIT * |a[3]| * a[:3]/|a[:3]|`` as
velocity target (VelocityAviary.py:129-170), inside the same launch as the physics substeps
(``ADRP_ACT_VEL`` in ``csrc/adrp_step.cuh``).  The controller is always the CF2X DSLPID (:68-69).
"""
from __future__ import annotations

import numpy as np

from .. import _abi
from ..utils.enums import DroneModel, Physics
from ._spaces import Box
from .CtrlAviary import CtrlAviary


class VelocityAviary(CtrlAviary):
    ENV_KIND = _abi.ENV_CTRL
    ACTION_TYPE = _abi.ACT_VEL

    def __init__(self, drone_model: DroneModel = DroneModel.CF2X, num_drones: int = 1, neighbourhood_radius: float = np.inf,
                 initial_xyzs=None, initial_rpys=None, physics: Physics = Physics.PYB, pyb_freq: int = 240,
                 ctrl_freq: int = 240, gui=False, record=False, obstacles=False, user_debug_gui=True,
                 output_folder='results', **batch):
        if drone_model not in (DroneModel.CF2X, DroneModel.CF2P):
            # the reference leaves self.ctrl undefined for other models and fails on the first step (:68-69,158)
            raise ValueError("[ERROR] in VelocityAviary.__init__(), DSLPIDControl requires DroneModel.CF2X or DroneModel.CF2P")
        super().__init__(drone_model=drone_model, num_drones=num_drones, neighbourhood_radius=neighbourhood_radius,
                         initial_xyzs=initial_xyzs, initial_rpys=initial_rpys, physics=physics, pyb_freq=pyb_freq,
                         ctrl_freq=ctrl_freq, gui=gui, record=record, obstacles=obstacles, user_debug_gui=user_debug_gui,
                         output_folder=output_folder, **batch)
        self.SPEED_LIMIT = 0.03 * self.MAX_SPEED_KMH * (1000 / 3600)          # :87

    def _actionSpace(self):                                # VelocityAviary.py:91-104
        lo = np.array([[-1, -1, -1, 0] for _ in range(self.NUM_DRONES)])
        hi = np.array([[1, 1, 1, 1] for _ in range(self.NUM_DRONES)])
        return Box(low=lo, high=hi, dtype=np.float32)
