"""Enumerations of the env/controller API surface.

Member names and values are the reference's (gym_pybullet_adrp/utils/enums.py:8-87) so
that user code written against it (``DroneModel.CF2X``, ``Physics("pyb_gnd")`` ...) keeps
working. ``ImageType`` / ``ObservationType.RGB`` exist for API completeness only: there is
no renderer on this path ("out of scope").
"""
from enum import Enum

# enums.py:8-14 -- CF2X points at the IROS urdf (different mass and prop order)
DroneModel = Enum("DroneModel", {"CF2X": "cf2x_IROS", "CF2P": "cf2p", "RACE": "racer"})

# enums.py:18-26
Physics = Enum("Physics", {
    "PYB": "pyb", "DYN": "dyn", "PYB_GND": "pyb_gnd", "PYB_DRAG": "pyb_drag",
    "PYB_DW": "pyb_dw", "PYB_GND_DRAG_DW": "pyb_gnd_drag_dw",
    # extension: the analytic integrator with the three force models added
    "DYN_GND_DRAG_DW": "dyn_gnd_drag_dw",
})

# enums.py:30-36
ImageType = Enum("ImageType", {"RGB": 0, "DEP": 1, "SEG": 2, "BW": 3})

# enums.py:40-47
ActionType = Enum("ActionType", {
    "MEL": "mel", "RPM": "rpm", "PID": "pid", "VEL": "vel",
    "ONE_D_RPM": "one_d_rpm", "ONE_D_PID": "one_d_pid",
})

# enums.py:51-54
ObservationType = Enum("ObservationType", {"KIN": "kin", "RGB": "rgb"})

# enums.py:58-70
Command = Enum("Command", {
    "FULLSTATE": "fst", "TAKEOFF": "tko", "TAKEOFFYAW": "toy", "TAKEOFFVEL": "tov",
    "LAND": "lnd", "LANDYAW": "ldy", "LANDVEL": "ldv", "STOP": "stp", "GOTO": "gto",
    "NOTIFY": "ntf", "NONE": "non",
})

# enums.py:74-80
State = Enum("State", {"INIT": 0, "TAKEOFF": 1, "CONTROL": 2, "LAND": 3, "END": 4})

# enums.py:84-87
RaceMode = Enum("RaceMode", {"COMPARE": 0, "COMPETE": 1})
