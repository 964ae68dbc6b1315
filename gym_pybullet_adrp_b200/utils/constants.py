"""Module constants (values of gym_pybullet_adrp/utils/constants.py:9-44)."""
import math

import numpy as np

RAD_TO_DEG = 180 / math.pi
DEG_TO_RAD = math.pi / 180
VEC3_UP = np.array([0, 0, 1], dtype=np.float64)
ZERO3 = np.zeros(3)

# lsy-drone-racing track constants (constants.py:23-25)
Z_LOW, Z_HIGH, VISIBILITY_RANGE = 0.3, 0.775, 0.45

# crazyflie firmware constants (constants.py:29-44)
FIRMWARE_FREQ, CTRL_FREQ = 500, 25
FIRMWARE_DT, CTRL_DT = 1.0 / FIRMWARE_FREQ, 1.0 / CTRL_FREQ
MIN_PWM, MAX_PWM = 20000, 65535
PWM2RPM_SCALE, PWM2RPM_CONST = 0.2685, 4070.3
ACTION_DELAY = SENSOR_DELAY = 0
GYRO_LPF_CUTOFF_FREQ, ACCEL_LPF_CUTOFF_FREQ = 80, 30
SUPPLY_VOLTAGE = 3
MOTOR_SET_ENABLE = QUAD_FORMATION_X = True
SPEED_LIMIT = 10
