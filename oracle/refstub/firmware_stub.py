"""TEST INFRASTRUCTURE -- stand-in for the SWIG module ``pycffirmware`` (utiasDSL, un-vendored, no
version pin): the names control/MellingerControl.py touches, the high-level commander included.  Arithmetic is delegated to the
declared C restatement in oracle/adrp_oracle_race.h through liboracle.so.  One ``Firmware`` object
per controller mirrors the reference loading a private copy of the module per drone
(MellingerControl.py:85-86) because the real firmware keeps its integrators in file-scope statics.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_LIB = C.CDLL(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "liboracle.so"))
_F = C.c_float
_LIB.fw_lpf2p_apply.restype = _F
_LIB.fw_lpf2p_apply.argtypes = [C.POINTER(_F), C.POINTER(_F), _F]
_LIB.fw_lpf2p_init.argtypes = [C.POINTER(_F), _F, _F]
_V3 = _F * 3
_LIB.fw_hl_tell_state.argtypes = [C.c_void_p, _V3, _V3, _F]
_LIB.fw_hl_update_time.argtypes = [C.c_void_p, _F]
_LIB.fw_hl_takeoff.argtypes = _LIB.fw_hl_land.argtypes = [C.c_void_p, _F, _F, C.c_int, _F]
_LIB.fw_hl_takeoff_vel.argtypes = _LIB.fw_hl_land_vel.argtypes = [C.c_void_p, _F, _F, C.c_int]
_LIB.fw_hl_go_to.argtypes = [C.c_void_p, _F, _F, _F, _F, _F, C.c_int]
_LIB.fw_hl_get_setpoint.argtypes = [C.c_void_p, _V3, _V3, _F, _F * 13]
_LIB.fw_controller_mellinger_yaw.argtypes = [_F * 10, _F * 16, C.c_int, _F, _V3, _V3, _F * 4, _V3, _F, _F * 4]

MEL_MASS = 0.027


class _Struct:
    """Attribute bag whose numeric fields round like C ``float`` / integer members."""
    _ints = ()

    def __setattr__(self, k, v):
        if isinstance(v, _Struct):
            object.__setattr__(self, k, v)
        elif k in self._ints or k == "timestamp" or k == "interruptTimestamp":
            object.__setattr__(self, k, int(v))
        else:
            object.__setattr__(self, k, float(np.float32(v)))


def _vec(names):
    s = _Struct()
    for n in names:
        setattr(s, n, 0.0)
    return s


class _Mode(_Struct):
    _ints = ("x", "y", "z", "roll", "pitch", "yaw", "quat")


class Firmware:
    modeDisable, modeAbs, modeVelocity = 0, 1, 2

    def __init__(self):
        self._statics = (_F * 10)()
        self._plan = C.create_string_buffer(_LIB.fw_hl_sizeof())      # struct planner + the commander's statics

    # ---- structs
    def lpf2pData(self):
        s = _Struct()
        object.__setattr__(s, "coef", (_F * 5)())
        object.__setattr__(s, "delay", (_F * 2)())
        return s

    def control_t(self):
        c = _Struct()
        object.__setattr__(c, "_ints", ("roll", "pitch", "yaw"))
        c.roll = c.pitch = c.yaw = 0
        c.thrust = 0.0
        return c

    def setpoint_t(self):
        s = _Struct()
        s.position, s.velocity, s.acceleration = (_vec("xyz") for _ in range(3))
        s.attitudeRate, s.attitude = _vec(("roll", "pitch", "yaw")), _vec(("roll", "pitch", "yaw"))
        s.attitudeQuaternion = _vec("xyzw")
        m = _Mode()
        for n in _Mode._ints:
            setattr(m, n, 0)
        s.mode = m
        s.timestamp = 0
        return s

    def sensorData_t(self):
        s = _Struct()
        s.acc, s.gyro = _vec("xyz"), _vec("xyz")
        s.interruptTimestamp = 0
        return s

    def state_t(self):
        s = _Struct()
        s.attitude = _vec(("roll", "pitch", "yaw"))
        s.attitudeQuaternion = _vec("xyzw")
        s.position, s.velocity, s.acc = _vec("xyz"), _vec("xyz"), _vec("xyz")
        return s

    # ---- filter.c
    def lpf2pInit(self, d, fs, fc):
        _LIB.fw_lpf2p_init(d.coef, fs, fc)
        d.delay[0] = d.delay[1] = 0.0

    def lpf2pApply(self, d, v):
        return float(_LIB.fw_lpf2p_apply(d.coef, d.delay, float(np.float32(v))))

    # ---- controller_mellinger.c
    def controllerMellingerInit(self):
        for i in range(10):
            self._statics[i] = 0.0

    def controllerMellinger(self, control, setpoint, sensors, state, tick):
        if tick % 2 != 0:            # RATE_DO_EXECUTE(ATTITUDE_RATE, tick): 1000 Hz loop / 500 Hz rate
            return
        yaw_abs = setpoint.mode.yaw == self.modeAbs
        assert setpoint.mode.x == self.modeAbs and (yaw_abs or setpoint.mode.quat == self.modeAbs), \
            "position setpoints with a yaw angle or a quaternion only"
        sp = (_F * 16)(setpoint.position.x, setpoint.position.y, setpoint.position.z,
                       setpoint.velocity.x, setpoint.velocity.y, setpoint.velocity.z,
                       setpoint.acceleration.x, setpoint.acceleration.y, setpoint.acceleration.z,
                       setpoint.attitudeRate.roll, setpoint.attitudeRate.pitch, setpoint.attitudeRate.yaw,
                       setpoint.attitudeQuaternion.x, setpoint.attitudeQuaternion.y, setpoint.attitudeQuaternion.z,
                       setpoint.attitudeQuaternion.w)
        pos = (_F * 3)(state.position.x, state.position.y, state.position.z)
        vel = (_F * 3)(state.velocity.x, state.velocity.y, state.velocity.z)
        q = (_F * 4)(state.attitudeQuaternion.x, state.attitudeQuaternion.y, state.attitudeQuaternion.z,
                     state.attitudeQuaternion.w)
        gyro = (_F * 3)(sensors.gyro.x, sensors.gyro.y, sensors.gyro.z)
        out = (_F * 4)()
        _LIB.fw_controller_mellinger_yaw(self._statics, sp, int(yaw_abs), setpoint.attitude.yaw, pos, vel, q, gyro,
                                         MEL_MASS, out)
        control.roll, control.pitch, control.yaw, control.thrust = int(out[0]), int(out[1]), int(out[2]), out[3]

    # ---- crtp_commander_high_level.c (declared restatement: oracle/adrp_oracle_planner.h)
    @staticmethod
    def _state_args(state):
        return (_V3(state.position.x, state.position.y, state.position.z),
                _V3(state.velocity.x, state.velocity.y, state.velocity.z), state.attitude.yaw)

    def crtpCommanderHighLevelInit(self):
        _LIB.fw_hl_init(self._plan)

    def crtpCommanderHighLevelTellState(self, state):
        _LIB.fw_hl_tell_state(self._plan, *self._state_args(state))

    def crtpCommanderHighLevelStop(self):
        _LIB.fw_hl_stop(self._plan)

    def crtpCommanderHighLevelUpdateTime(self, t):
        _LIB.fw_hl_update_time(self._plan, float(t))

    def crtpCommanderHighLevelTakeoff(self, height, duration):
        return _LIB.fw_hl_takeoff(self._plan, height, duration, 1, 0.0)

    def crtpCommanderHighLevelTakeoffYaw(self, height, duration, yaw):
        return _LIB.fw_hl_takeoff(self._plan, height, duration, 0, yaw)

    def crtpCommanderHighLevelTakeoffWithVelocity(self, height, velocity, relative):
        return _LIB.fw_hl_takeoff_vel(self._plan, height, velocity, int(bool(relative)))

    def crtpCommanderHighLevelLand(self, height, duration):
        return _LIB.fw_hl_land(self._plan, height, duration, 1, 0.0)

    def crtpCommanderHighLevelLandYaw(self, height, duration, yaw):
        return _LIB.fw_hl_land(self._plan, height, duration, 0, yaw)

    def crtpCommanderHighLevelLandWithVelocity(self, height, velocity, relative):
        return _LIB.fw_hl_land_vel(self._plan, height, velocity, int(bool(relative)))

    def crtpCommanderHighLevelGoTo(self, x, y, z, yaw, duration, relative):
        return _LIB.fw_hl_go_to(self._plan, x, y, z, yaw, duration, int(bool(relative)))

    def crtpCommanderHighLevelGetSetpoint(self, setpoint, state):
        out = (_F * 13)()
        if not _LIB.fw_hl_get_setpoint(self._plan, *self._state_args(state), out):
            return                                                   # planner stopped: the setpoint is left as it is
        sp = setpoint
        sp.position.x, sp.position.y, sp.position.z = out[0], out[1], out[2]
        sp.velocity.x, sp.velocity.y, sp.velocity.z = out[3], out[4], out[5]
        sp.acceleration.x, sp.acceleration.y, sp.acceleration.z = out[6], out[7], out[8]
        sp.attitudeRate.roll, sp.attitudeRate.pitch, sp.attitudeRate.yaw = out[9], out[10], out[11]
        sp.attitude.yaw = out[12]
        sp.mode.x = sp.mode.y = sp.mode.z = sp.mode.yaw = self.modeAbs
        sp.mode.roll = sp.mode.pitch = sp.mode.quat = self.modeDisable
