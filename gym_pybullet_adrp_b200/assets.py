"""Physical constants of the drone models and of the race-track bodies.

The reference reads these from URDF files at construction time
(envs/BaseAviary.py:989-1021, control/BaseControl.py:185-224). Here they are a plain table --
the numbers of assets/cf2x_IROS.urdf:5,12-13,33-35,43-79, assets/cf2x.urdf:5,11-12,42-78,
assets/cf2p.urdf:5,11-12,42-78 and assets/racer.urdf:5,11-12,28,36-72 -- so that the C ABI can
take them as one POD struct and the kernel can keep them in constant memory.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field, replace

from .utils.enums import DroneModel

_CF_PROPS = dict(arm=0.0397, kf=3.16e-10, km=7.94e-12, thrust2weight=2.25, max_speed_kmh=30.0,
                 gnd_eff_coeff=11.36859, prop_radius=2.31348e-2, drag_coeff_xy=9.1785e-7,
                 drag_coeff_z=10.311e-7, dw_coeff_1=2267.18, dw_coeff_2=0.16, dw_coeff_3=-0.11)


@dataclass(frozen=True)
class DroneParams:
    """URDF-level parameters of one drone model."""
    urdf: str
    mass: float
    ixx: float
    iyy: float
    izz: float
    arm: float
    kf: float
    km: float
    thrust2weight: float
    max_speed_kmh: float
    gnd_eff_coeff: float
    prop_radius: float
    drag_coeff_xy: float
    drag_coeff_z: float
    dw_coeff_1: float
    dw_coeff_2: float
    dw_coeff_3: float
    collision_r: float = 0.06
    collision_h: float = 0.025
    collision_z_offset: float = 0.0
    #: inertial origins of prop links 0..3 in the base frame (PYB-mode wrench arms)
    prop_xy: tuple = field(default=((0.0, 0.0),) * 4)


DRONES = {
    "cf2x_IROS": DroneParams(
        urdf="cf2x_IROS", mass=0.03454, ixx=1.4e-5, iyy=1.4e-5, izz=2.17e-5,
        prop_xy=((0.028, 0.028), (-0.028, 0.028), (-0.028, -0.028), (0.028, -0.028)), **_CF_PROPS),
    # legacy urdf: only its mass/inertia are used (race drones' Bullet body)
    "cf2x": DroneParams(
        urdf="cf2x", mass=0.027, ixx=1.4e-5, iyy=1.4e-5, izz=2.17e-5,
        prop_xy=((0.028, -0.028), (-0.028, -0.028), (-0.028, 0.028), (0.028, 0.028)), **_CF_PROPS),
    "cf2p": DroneParams(
        urdf="cf2p", mass=0.027, ixx=2.3951e-5, iyy=2.3951e-5, izz=3.2347e-5,
        prop_xy=((0.0397, 0.0), (0.0, 0.0397), (-0.0397, 0.0), (0.0, -0.0397)), **_CF_PROPS),
    "racer": DroneParams(
        urdf="racer", mass=0.830, ixx=0.003113, iyy=0.003113, izz=0.003113,
        arm=0.109, kf=8.47e-9, km=2.13e-11, thrust2weight=4.17, max_speed_kmh=200.0,
        gnd_eff_coeff=11.36859, prop_radius=12.7e-2, drag_coeff_xy=9.1785e-7,
        drag_coeff_z=10.311e-7, dw_coeff_1=2267.18, dw_coeff_2=0.16, dw_coeff_3=-0.11,
        prop_xy=((0.085, 0.0675), (-0.085, 0.0675), (-0.085, -0.0675), (0.085, -0.0675))),
}


def drone_params(model: DroneModel | str) -> DroneParams:
    key = model.value if isinstance(model, DroneModel) else str(model)
    return DRONES[key]


@dataclass(frozen=True)
class DerivedConstants:
    """BaseAviary.__init__ derived constants (envs/BaseAviary.py:116-128)."""
    G: float
    GRAVITY: float
    HOVER_RPM: float
    MAX_RPM: float
    MAX_THRUST: float
    MAX_XY_TORQUE: float
    MAX_Z_TORQUE: float
    GND_EFF_H_CLIP: float


def derived_constants(p: DroneParams, model: DroneModel, g: float = 9.8) -> DerivedConstants:
    gravity = g * p.mass
    hover = math.sqrt(gravity / (4 * p.kf))
    max_rpm = math.sqrt((p.thrust2weight * gravity) / (4 * p.kf))
    max_thrust = 4 * p.kf * max_rpm ** 2
    if model == DroneModel.CF2P:
        max_xy = p.arm * p.kf * max_rpm ** 2
    else:
        max_xy = (2 * p.arm * p.kf * max_rpm ** 2) / math.sqrt(2)
    max_z = 2 * p.km * max_rpm ** 2
    h_clip = 0.25 * p.prop_radius * math.sqrt((15 * max_rpm ** 2 * p.kf * p.gnd_eff_coeff) / max_thrust)
    return DerivedConstants(g, gravity, hover, max_rpm, max_thrust, max_xy, max_z, h_clip)


# ---------------------------------------------------------------------------------------------
# Race-track bodies (assets/portal.urdf:23-167, low_portal.urdf:138-166, obstacle.urdf:5-58).
# All shapes are given in the body frame of the gate / obstacle as
# (kind, centre xyz, half-extents or (radius, half-length), pitch about body y).
# ---------------------------------------------------------------------------------------------
GATE_EDGE_HALF = (0.25, 0.025, 0.025)      # box .5 x .05 x .05
GATE_EDGE_OFFSET = 0.225                   # edge centres at +-0.225 from the gate centre
GATE_SIDE_PITCH = 1.57                     # "rpy=0 1.57 0": not pi/2
TALL_SUPPORT = dict(radius=0.05, half_len=0.4, z=-0.6)      # portal.urdf:141-159
LOW_SUPPORT = dict(half=(0.075, 0.075, 0.125), z=-0.4)      # low_portal.urdf:141-159
OBSTACLE_CYL = dict(radius=0.05, half_len=0.4, z=0.0)       # obstacle.urdf:8-26
OBSTACLE_BOX = dict(half=(0.075, 0.075, 0.125), z=-0.4)     # obstacle.urdf:34-50
GATE_RAY_HALF_LEN = 0.1875                                   # MultiRaceAviary.py:486
GATE_RAY_STEP = 0.05                                         # MultiRaceAviary.py:487-488
GATE_HEIGHT_TALL, GATE_HEIGHT_LOW = 1.0, 0.525               # MultiRaceAviary.py:479-482


def with_body(p: DroneParams, **kw) -> DroneParams:
    return replace(p, **kw)
