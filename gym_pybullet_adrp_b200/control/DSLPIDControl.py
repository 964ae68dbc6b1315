"""``DSLPIDControl`` (gym_pybullet_adrp/control/DSLPIDControl.py:9-287) on the GPU: one controller object holds
the integrators of ``n`` drones (``n = 1`` by default, exactly the reference's call pattern in examples/pid.py)."""
from __future__ import annotations

import numpy as np

from .. import _abi
from ..utils.enums import DroneModel
from .BaseControl import BaseControl


class DSLPIDControl(BaseControl):
    def __init__(self, drone_model: DroneModel, g: float = 9.8, *, n: int = 1, precision: str = "fp64", device: int = 0):
        if drone_model not in (DroneModel.CF2X, DroneModel.CF2P):
            print("[ERROR] in DSLPIDControl.__init__(), DSLPIDControl requires DroneModel.CF2X or DroneModel.CF2P")
            raise SystemExit                                  # the reference calls exit() (:34-36)
        self.n, self.device = int(n), int(device)
        self.precision = {"fp32": _abi.F32, "fp64": _abi.F64}[precision]
        self._ctrl = None
        super().__init__(drone_model=drone_model, g=g)
        self.P_COEFF_FOR, self.I_COEFF_FOR, self.D_COEFF_FOR = np.array([.4, .4, 1.25]), np.array([.05, .05, .05]), np.array([.2, .2, .5])
        self.P_COEFF_TOR, self.I_COEFF_TOR, self.D_COEFF_TOR = np.array([70000., 70000., 60000.]), np.array([.0, .0, 500.]), np.array([20000., 20000., 12000.])
        self.PWM2RPM_SCALE, self.PWM2RPM_CONST, self.MIN_PWM, self.MAX_PWM = 0.2685, 4070.3, 20000, 65535

    def reset(self):
        super().reset()
        if self._ctrl is not None:
            self._ctrl.zero_()                                # last_rpy, integral_pos_e, integral_rpy_e (:65-78)

    def _rows(self, x):
        return np.array(np.broadcast_to(np.asarray(x, dtype=np.float64), (self.n, 3))) if np.ndim(x) < 2 else np.asarray(x)

    def computeControl(self, control_timestep, cur_pos, cur_quat, cur_vel, cur_ang_vel, target_pos,
                       target_rpy=np.zeros(3), target_vel=np.zeros(3), target_rpy_rates=np.zeros(3)):
        """Returns ``(rpm[4], pos_e[3], yaw_e)`` like the reference (:82-145); ``(n, .)`` arrays when ``n > 1``."""
        import torch
        from .. import _lib
        if self._ctrl is None:
            dt = torch.float64 if self.precision == _abi.F64 else torch.float32
            self._ctrl = torch.zeros((self.n, 9), dtype=dt, device=torch.device("cuda", self.device))
        self.control_counter += 1
        quat = np.asarray(cur_quat, dtype=np.float64).reshape(self.n, 4)
        rpm, pos_e, yaw_e = _lib.dslpid_control(self.precision, self.DRONE_MODEL, self.n, control_timestep, self._rows(cur_pos),
                                                quat, self._rows(cur_vel), self._rows(target_pos), self._rows(target_rpy),
                                                self._rows(target_vel), self._rows(target_rpy_rates), self._ctrl, device=self.device)
        rpm, pos_e, yaw_e = rpm.cpu().numpy(), pos_e.cpu().numpy(), yaw_e.cpu().numpy()
        if self.n == 1:
            return rpm[0], pos_e[0], float(yaw_e[0])
        return rpm, pos_e, yaw_e
