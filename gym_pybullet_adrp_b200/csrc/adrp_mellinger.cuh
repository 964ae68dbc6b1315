// adrp_mellinger.cuh -- the race env's low-level controller as device functions, one call per drone
// per physics SUBSTEP (500 Hz), state in registers.
//
//   mellinger_tick   MellingerControl.computeControl (control/MellingerControl.py:154-262): sensor
//                    synthesis by finite differences (:206-214), _update_state (:446-471),
//                    _update_sensorData gyro low-pass (:347-365), _step_controller tumble check and
//                    sub-rate schedule (:378-419), _compute_pwms power distribution + battery curve
//                    (:423-442), PWM -> thrust -> prop reorder -> noise -> _thr2pwm -> RPM (:246-262,
//                    :307-343).  Wrapper arithmetic in T (the reference: float64).
//   fw_mellinger     firmware core controllerMellinger() (pycffirmware, un-vendored, float32 C):
//                    DECLARED restatement, statement order fixed so that the CPU checker can match it.
//                    float32 in both builds; the fp64 (parity) translation unit is compiled with
//                    -fmad=false so that no multiply-add is contracted, like the C build of the firmware.
#pragma once
#include "adrp_math.cuh"
#include "adrp_params.cuh"

#ifndef ADRP_PACKED_PWM
#define ADRP_PACKED_PWM 1
#endif
namespace adrp {

// libm float functions of the firmware.  Parity build (T = double): correctly rounded via the
// double-precision function; fp32 build: CUDA's float functions.
template <typename T> struct FwMath;
template <> struct FwMath<double> {
    static __device__ __forceinline__ float atan2(float y, float x) { return (float)::atan2((double)y, (double)x); }
    static __device__ __forceinline__ void sincos(float a, float* s, float* c) { *s = (float)::sin((double)a); *c = (float)::cos((double)a); }
};
template <> struct FwMath<float> {
    static __device__ __forceinline__ float atan2(float y, float x) { return atan2f(y, x); }
    static __device__ __forceinline__ void sincos(float a, float* s, float* c) { sincosf(a, s, c); }
};

struct MelFw {                       // controller_mellinger.c file-scope statics + control_t
    float i_error_x, i_error_y, i_error_z, i_error_m_x, i_error_m_y, i_error_m_z;
    float prev_omega_roll, prev_omega_pitch, prev_setpoint_omega_roll, prev_setpoint_omega_pitch;
    float thrust;
    float roll, pitch, yaw;          // int16 values (held as floats: exact, and the PWM mix reads them as such)
};

template <typename T>
struct Mel {
    V3<T> prev_rpy, prev_vel;        // wrapper state (MellingerControl.reset, :99-150)
    T rpms[4];                       // MultiRaceAviary.rpms[i]: the command the physics applies next substep
    float gd1[3], gd2[3];            // gyro lpf2p delay elements
    MelFw fw;
    float acc_z;                     // state_t.acc.z as stored (tumble check, :382)
    int tick, tumble, last_att, error;
};

// Full-state command (MellingerControl._sendFullStateCmd, :510-543): yaw-only quaternion; vel / acc / attitude rates
// (deg/s) are zero for ndarray actions (MultiRaceAviary.py:190-195) and literal zeros in those kernels.
// yaw_abs / yaw_deg: the setpoint as the high-level commander leaves it (crtpCommanderHighLevelGetSetpoint: mode.yaw = modeAbs
// with attitude.yaw in degrees, mode.quat = modeDisable); only the ADRP_ACT_MEL_CMD instances (HL) read them.
struct MelSetpoint { float pos[3], qz, qw, vel[3], acc[3], rate_deg[3], xc0, xc1, yaw_deg; int yaw_abs; };
struct HlState;                      // adrp_planner.cuh
template <typename T> __device__ inline void hl_update_setpoint(HlState& h, const float pos[3], const float vel[3], float yaw_deg, int tick);
__device__ inline MelSetpoint& hl_setpoint(HlState& h);

// fast build: heading of the desired x axis, constant over the control step.  The yaw-only setpoint quaternion is
// (0, 0, sin(yaw/2), cos(yaw/2)): the firmware's atan2 arguments (dy_s, dy_c) ARE (sin yaw, cos yaw) up to their norm, so
// its atan2 -> degrees -> radians -> sincos round trip reduces to a normalisation.
template <typename T> __device__ __forceinline__ void mel_setpoint_heading(MelSetpoint& sp)
{
    if (!M<T>::fast) return;
    const float dy_s = 2.0f * (sp.qw * sp.qz), dy_c = 1 - 2 * (sp.qz * sp.qz);
    const float in = M<float>::frsqrt(dy_c * dy_c + dy_s * dy_s);
    sp.xc0 = dy_c * in; sp.xc1 = dy_s * in;
}

__device__ __forceinline__ float fw_clamp(float v, float lo, float hi) { return v < lo ? lo : (v > hi ? hi : v); }
template <typename T> __device__ __forceinline__ float fw_clampT(float v, float lo, float hi)
{
    return M<T>::fast ? fminf(fmaxf(v, lo), hi) : fw_clamp(v, lo, hi);
}
__device__ __forceinline__ float fw_radians(float d) { return (3.14159265358979323846f / 180.0f) * d; }
__device__ __forceinline__ float fw_degrees(float r) { return (180.0f / 3.14159265358979323846f) * r; }

// filter.c lpf2pApply
__device__ __forceinline__ float fw_lpf2p(float b0, float b1, float b2, float a1, float a2, float& d1, float& d2, float sample)
{
    float d0 = sample - d1 * a1 - d2 * a2;
    if (!isfinite(d0)) d0 = sample;
    const float out = d0 * b0 + d1 * b1 + d2 * b2;
    d2 = d1;
    d1 = d0;
    return out;
}

// controllerMellinger() with mode.x/y/z/quat = modeAbs, roll/pitch/yaw = modeDisable.
template <typename T, bool HL = false>
__device__ __forceinline__ void fw_mellinger(MelFw& m, const MelSetpoint& sp, float mass, const float pos[3],
                                             const float vel[3], const float q[4], const Mat3<T>& Rq, const float gyro_deg[3])
{
    const float massThrust = 132000.0f;
    const float kp_xy = 0.4f, kd_xy = 0.2f, ki_xy = 0.05f, i_range_xy = 2.0f;
    const float kp_z = 1.25f, kd_z = 0.4f, ki_z = 0.05f, i_range_z = 0.4f;
    const float kR_xy = 70000.0f, kw_xy = 20000.0f, ki_m_xy = 0.0f, i_range_m_xy = 1.0f;
    const float kR_z = 60000.0f, kw_z = 12000.0f, ki_m_z = 500.0f, i_range_m_z = 1500.0f;
    const float kd_omega_rp = 200.0f, GRAVITY_MAGNITUDE = 9.81f;
    const float dt = (float)(1.0f / 500.0f);
    float r_error[3], v_error[3], tt[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) { r_error[k] = sp.pos[k] - pos[k]; v_error[k] = sp.vel[k] - vel[k]; }
    m.i_error_z += r_error[2] * dt; m.i_error_z = fw_clampT<T>(m.i_error_z, -i_range_z, i_range_z);
    m.i_error_x += r_error[0] * dt; m.i_error_x = fw_clampT<T>(m.i_error_x, -i_range_xy, i_range_xy);
    m.i_error_y += r_error[1] * dt; m.i_error_y = fw_clampT<T>(m.i_error_y, -i_range_xy, i_range_xy);
    tt[0] = mass * sp.acc[0] + kp_xy * r_error[0] + kd_xy * v_error[0] + ki_xy * m.i_error_x;
    tt[1] = mass * sp.acc[1] + kp_xy * r_error[1] + kd_xy * v_error[1] + ki_xy * m.i_error_y;
    tt[2] = mass * (sp.acc[2] + GRAVITY_MAGNITUDE) + kp_z * r_error[2] + kd_z * v_error[2] + ki_z * m.i_error_z;
    // desired yaw from the setpoint quaternion (0, 0, qz, qw), in degrees
    const float sx = 0.0f, sy = 0.0f, sz = sp.qz, sw = sp.qw;
    // (fp32 product build: cos / sin of that angle are the atan2 arguments themselves, see below)
    const float dy_s = 2.0f * (sw * sz + sx * sy), dy_c = 1 - 2 * (sy * sy + sz * sz);
    float R00, R01, R02, R10, R11, R12, R20, R21, R22;
    if (M<T>::fast) {
        // the substep's rotation matrix (shared with the force models and the Euler angles): the firmware's
        // quat2rotmat of a unit quaternion is the same matrix
        R00 = (float)Rq.m00; R01 = (float)Rq.m01; R02 = (float)Rq.m02; R10 = (float)Rq.m10; R11 = (float)Rq.m11;
        R12 = (float)Rq.m12; R20 = (float)Rq.m20; R21 = (float)Rq.m21; R22 = (float)Rq.m22;
    } else {
        const float x = q[0], y = q[1], z = q[2], w = q[3];
        R00 = 1 - 2 * y * y - 2 * z * z; R01 = 2 * x * y - 2 * z * w; R02 = 2 * x * z + 2 * y * w;
        R10 = 2 * x * y + 2 * z * w; R11 = 1 - 2 * x * x - 2 * z * z; R12 = 2 * y * z - 2 * x * w;
        R20 = 2 * x * z - 2 * y * w; R21 = 2 * y * z + 2 * x * w; R22 = 1 - 2 * x * x - 2 * y * y;
    }
    const float current_thrust = tt[0] * R02 + tt[1] * R12 + tt[2] * R22;
    const float nt = M<T>::fast ? M<float>::fsqrt(tt[0] * tt[0] + tt[1] * tt[1] + tt[2] * tt[2]) : sqrtf(tt[0] * tt[0] + tt[1] * tt[1] + tt[2] * tt[2]);
    float zd0, zd1, zd2;
    if (M<T>::fast) { const float in = M<float>::rcp(nt); zd0 = tt[0] * in; zd1 = tt[1] * in; zd2 = tt[2] * in; }
    else { zd0 = tt[0] / nt; zd1 = tt[1] / nt; zd2 = tt[2] / nt; }
    float xc0, xc1;
    if (M<T>::fast && !HL) {
        xc0 = sp.xc0; xc1 = sp.xc1;                           // mel_setpoint_heading, once per control step
    } else {
        float desiredYaw = fw_degrees(FwMath<T>::atan2(dy_s, dy_c));
        if (HL && sp.yaw_abs) desiredYaw = sp.yaw_deg;        // mode.yaw == modeAbs: setpoint->attitude.yaw (the planner's, per tick)
        FwMath<T>::sincos(fw_radians(desiredYaw), &xc1, &xc0);
    }
    const float xc2 = 0.0f;
    float yd0 = zd1 * xc2 - zd2 * xc1, yd1 = zd2 * xc0 - zd0 * xc2, yd2 = zd0 * xc1 - zd1 * xc0;
    if (M<T>::fast) { const float in = M<float>::frsqrt(yd0 * yd0 + yd1 * yd1 + yd2 * yd2); yd0 *= in; yd1 *= in; yd2 *= in; }
    else { const float ny = sqrtf(yd0 * yd0 + yd1 * yd1 + yd2 * yd2); yd0 /= ny; yd1 /= ny; yd2 /= ny; }
    const float xd0 = yd1 * zd2 - yd2 * zd1, xd1 = yd2 * zd0 - yd0 * zd2, xd2 = yd0 * zd1 - yd1 * zd0;
    // eRM = Rdes^T R - R^T Rdes with Rdes = [xd yd zd] (columns): Rd[k][0] = xd_k, Rd[k][1] = yd_k, Rd[k][2] = zd_k
    float e21 = 0, e02 = 0, e10 = 0;
    e21 += zd0 * R01 - R02 * yd0; e02 += xd0 * R02 - R00 * zd0; e10 += yd0 * R00 - R01 * xd0;
    e21 += zd1 * R11 - R12 * yd1; e02 += xd1 * R12 - R10 * zd1; e10 += yd1 * R10 - R11 * xd1;
    e21 += zd2 * R21 - R22 * yd2; e02 += xd2 * R22 - R20 * zd2; e10 += yd2 * R20 - R21 * xd2;
    const float eR0 = e21, eR1 = -e02, eR2 = e10;
    const float rate_roll = fw_radians(gyro_deg[0]), rate_pitch = -fw_radians(gyro_deg[1]), rate_yaw = fw_radians(gyro_deg[2]);
    const float ew0 = fw_radians(sp.rate_deg[0]) - rate_roll, ew1 = -fw_radians(sp.rate_deg[1]) - rate_pitch,
                ew2 = fw_radians(sp.rate_deg[2]) - rate_yaw;
    float err_d_roll = (fw_radians(sp.rate_deg[0]) - m.prev_setpoint_omega_roll) - (rate_roll - m.prev_omega_roll);
    float err_d_pitch = -(fw_radians(sp.rate_deg[1]) - m.prev_setpoint_omega_pitch) - (rate_pitch - m.prev_omega_pitch);
    if (M<T>::fast) { err_d_roll *= 500.0f; err_d_pitch *= 500.0f; }
    else { err_d_roll = err_d_roll / dt; err_d_pitch = err_d_pitch / dt; }
    m.prev_omega_roll = rate_roll; m.prev_omega_pitch = rate_pitch;
    m.prev_setpoint_omega_roll = fw_radians(sp.rate_deg[0]);
    m.prev_setpoint_omega_pitch = fw_radians(sp.rate_deg[1]);
    m.i_error_m_x += (-eR0) * dt; m.i_error_m_x = fw_clampT<T>(m.i_error_m_x, -i_range_m_xy, i_range_m_xy);
    m.i_error_m_y += (-eR1) * dt; m.i_error_m_y = fw_clampT<T>(m.i_error_m_y, -i_range_m_xy, i_range_m_xy);
    m.i_error_m_z += (-eR2) * dt; m.i_error_m_z = fw_clampT<T>(m.i_error_m_z, -i_range_m_z, i_range_m_z);
    // (ki_m_xy == 0: the fast build leaves the term out)
    const float Mx = M<T>::fast ? -kR_xy * eR0 + kw_xy * ew0 + kd_omega_rp * err_d_roll
                                : -kR_xy * eR0 + kw_xy * ew0 + ki_m_xy * m.i_error_m_x + kd_omega_rp * err_d_roll;
    const float My = M<T>::fast ? -kR_xy * eR1 + kw_xy * ew1 + kd_omega_rp * err_d_pitch
                                : -kR_xy * eR1 + kw_xy * ew1 + ki_m_xy * m.i_error_m_y + kd_omega_rp * err_d_pitch;
    const float Mz = -kR_z * eR2 + kw_z * ew2 + ki_m_z * m.i_error_m_z;
    m.thrust = massThrust * current_thrust;
    if (m.thrust > 0) {
        // control->roll = clamp(M.x, -32000, 32000) into an int16: truncation toward zero
        if (M<T>::fast) {
            m.roll = truncf(fw_clampT<T>(Mx, -32000, 32000)); m.pitch = truncf(fw_clampT<T>(My, -32000, 32000));
            m.yaw = truncf(fw_clampT<T>(-Mz, -32000, 32000));
        } else {
            m.roll = (float)(int)(short)fw_clampT<T>(Mx, -32000, 32000);
            m.pitch = (float)(int)(short)fw_clampT<T>(My, -32000, 32000);
            m.yaw = (float)(int)(short)fw_clampT<T>(-Mz, -32000, 32000);
        }
    } else {
        m.roll = m.pitch = m.yaw = 0;
        m.i_error_x = m.i_error_y = m.i_error_z = 0;          // controllerMellingerReset()
        m.i_error_m_x = m.i_error_m_y = m.i_error_m_z = 0;
    }
}

// (tick/500 - last/500 > 0.002) exactly as Python evaluates it in float64 (:395-411): always true for a gap
// of two or more ticks, never for a gap of zero; for a gap of one it depends on the rounding of tick/500,
// which the host tabulates once (bit `tick` of lut).
// Fast build: the S schedule bits of one control step, fetched once before the substep loop (the tick advances by
// exactly one per substep).  ok == false (window not inside the table): per-tick evaluation as below.
struct MelWindow { uint32_t bits; bool ok; };
template <typename T>
__device__ __forceinline__ MelWindow mel_schedule_window(const RaceParams<T>& RP, const uint32_t* __restrict__ lut, int tick0, int S)
{
    MelWindow w = {0u, false};
    if (S <= 32 && tick0 >= 0 && tick0 + 64 <= RP.lut_ticks) {
        const uint32_t lo = __ldg(lut + (tick0 >> 5)), hi = __ldg(lut + (tick0 >> 5) + 1);
        w.bits = __funnelshift_r(lo, hi, tick0 & 31);
        w.ok = true;
    }
    return w;
}

// (beyond the table: the float64 test itself, out of line -- its double-precision arithmetic would otherwise sit in the substep loop)
static __device__ __noinline__ bool mel_schedule_fires_f64(int tick, int last)
{
    return (double)tick / 500.0 - (double)last / 500.0 > 0.002;
}
template <typename T>
__device__ __forceinline__ bool mel_schedule_fires(const RaceParams<T>& RP, const uint32_t* __restrict__ lut, int tick, int last)
{
    const int gap = tick - last;
    if (gap != 1) return gap > 1;
    if (tick < RP.lut_ticks) return (__ldg(lut + (tick >> 5)) >> (tick & 31)) & 1u;
    return mel_schedule_fires_f64(tick, last);
}

// One 500 Hz tick.  p, q, v: the drone's state AFTER this substep's physics.  noise4: motor thrust noise (N).
// HL (ADRP_ACT_MEL_CMD): the high-level commander's _update_setpoint runs before the controller and the setpoint is hl's.
template <typename T, bool HL = false>
__device__ __forceinline__ void mellinger_tick(const RaceParams<T>& RP, const uint32_t* __restrict__ lut, Mel<T>& m,
                                               const MelSetpoint& sp_in, const V3<T>& p, const Q4<T>& q, const V3<T>& v,
                                               const T noise4[4], const Mat3<T>& Rq, const MelWindow win, int s,
                                               HlState* hl = nullptr)
{
    const T FIRMWARE_DT = (T)(1.0 / 500), RAD_TO_DEG = (T)(180.0 / 3.14159265358979323846), DEG_TO_RAD = (T)(3.14159265358979323846 / 180.0);
    // obs[i, 3:6] (:242); Rq: rotmat_loop(q), valid in the fast build only
    const V3<T> rpy = M<T>::fast ? euler_rpy_rot(q, Rq) : euler_rpy_loop(q);
    T rates[3], acc_z;
    float qf[4];
    if (M<T>::fast) {
        // (rates in deg/s straight away: * 500 * 180 / pi; the conversion below is then the identity)
        const T K = T(500.0 * 180.0 / 3.14159265358979323846);
        rates[0] = (rpy.x - m.prev_rpy.x) * K; rates[1] = (rpy.y - m.prev_rpy.y) * K; rates[2] = (rpy.z - m.prev_rpy.z) * K;
        acc_z = (v.z - m.prev_vel.z) * T(500.0 / 9.8) + T(1);
        // euler -> degrees -> radians -> quaternion (:446-471) reproduces +-q up to rounding, and the firmware only
        // reads q through its rotation matrix: hand it the state quaternion
        qf[0] = (float)q.x; qf[1] = (float)q.y; qf[2] = (float)q.z; qf[3] = (float)q.w;
    } else {
        rates[0] = (rpy.x - m.prev_rpy.x) / FIRMWARE_DT; rates[1] = (rpy.y - m.prev_rpy.y) / FIRMWARE_DT;
        rates[2] = (rpy.z - m.prev_rpy.z) / FIRMWARE_DT;                         // :209
        acc_z = (v.z - m.prev_vel.z) / FIRMWARE_DT / T(9.8) + T(1);              // :213 (only acc.z is read downstream)
        // _update_state (:446-471): degrees and back, then utils.get_quaternion_from_euler (utils/utils.py:20-43)
        const T rdx = rpy.x * RAD_TO_DEG * DEG_TO_RAD, rdy = rpy.y * RAD_TO_DEG * DEG_TO_RAD, rdz = rpy.z * RAD_TO_DEG * DEG_TO_RAD;
        T sr, cr, sp_, cp, sy, cy;
        M<T>::sincos(rdx / T(2), &sr, &cr);
        M<T>::sincos(rdy / T(2), &sp_, &cp);
        M<T>::sincos(rdz / T(2), &sy, &cy);
        qf[0] = (float)(sr * cp * cy - cr * sp_ * sy); qf[1] = (float)(cr * sp_ * cy + sr * cp * sy);
        qf[2] = (float)(cr * cp * sy - sr * sp_ * cy); qf[3] = (float)(cr * cp * cy + sr * sp_ * sy);
    }
    m.prev_rpy = rpy;
    m.prev_vel = v;
    const float pos[3] = {(float)p.x, (float)p.y, (float)p.z}, vel[3] = {(float)v.x, (float)v.y, (float)v.z};
    m.acc_z = (float)acc_z;
    // _update_sensorData (:347-365): gyro in deg/s through lpf2p
    float gyro[3];
#pragma unroll
    for (int k = 0; k < 3; ++k)
        gyro[k] = fw_lpf2p(RP.lpf_b0, RP.lpf_b1, RP.lpf_b2, RP.lpf_a1, RP.lpf_a2, m.gd1[k], m.gd2[k], (float)(M<T>::fast ? rates[k] : rates[k] * RAD_TO_DEG));
    // _update_setpoint (:369-374): while no full-state command overrides it, the high-level commander writes the setpoint
    // (state_t fields are C floats: position, velocity, attitude.yaw in degrees)
    if (HL) hl_update_setpoint<T>(*hl, pos, vel, (float)(rpy.z * RAD_TO_DEG), m.tick);
    const MelSetpoint& sp = HL ? hl_setpoint(*hl) : sp_in;
    // _step_controller (:378-419)
    if (m.acc_z < -0.5f) m.tumble += 1; else m.tumble = 0;                       // :382-385
    const bool stop = m.tumble >= 30;
    if (stop) {
        m.tick += 1; m.error = 1;                                                // :387-391: zero PWM
    } else {
        const int gap = m.tick - m.last_att;
        const bool fires = (M<T>::fast && win.ok) ? (gap != 1 ? gap > 1 : ((win.bits >> s) & 1u) != 0u)
                                                  : mel_schedule_fires(RP, lut, m.tick, m.last_att);
        if (fires) {
            m.last_att = m.tick;
            fw_mellinger<T, HL>(m.fw, sp, RP.mel_mass, pos, vel, qf, Rq, gyro);
        }
        m.tick += 1;
    }
    // _compute_pwms (:423-442)
    const T r = (T)m.fw.roll * T(0.5), pp = (T)m.fw.pitch * T(0.5), y = (T)m.fw.yaw, t = (T)m.fw.thrust;
    const T th[4] = {t - r + pp + y, t - r - pp - y, t + r - pp + y, t + r + pp - y};
    const T KF = T(3.16e-10), SCALE = T(0.2685), CONST = T(4070.3);
    if (M<T>::fast) {
        // Same chain with the constants folded: battery curve volts / 3 * 65535 = x (A x + B); min(pct, 1) * 65535 and the
        // PWM clamp of :246 are one clamp; thrust -> (+ noise) -> _thr2pwm -> RPM is sqrt(rp^2 + noise / KF) clamped to the
        // RPM images of the PWM limits (SCALE > 0: the maps are monotone).
        const T A = T(-0.0006239 * 65535.0 / 3.0), B = T(0.088 * 65535.0 / 3.0);
        const T RPM_LO = T(0.2685 * 20000.0 + 4070.3), RPM_HI = T(0.2685 * 65535.0 + 4070.3);
        T rp[4];
        if (stop) {
#pragma unroll
            for (int k = 0; k < 4; ++k) rp[k] = RPM_LO;                          // zero PWM, clamped to 20000 (:246)
        } else if constexpr (ADRP_PACKED_PWM && sizeof(T) == 4) {
            // motors (0, 1) and (2, 3) in the packed fp32 forms (two results per issue slot; clamps stay scalar)
#pragma unroll
            for (int k = 0; k < 4; k += 2) {
                const float2 x = __fmul2_rn(make_float2(fminf(fmaxf((float)th[k], 0.0f), 65535.0f), fminf(fmaxf((float)th[k + 1], 0.0f), 65535.0f)),
                                            make_float2((float)(60.0 / 65535.0), (float)(60.0 / 65535.0)));
                const float2 v = __fmul2_rn(x, __ffma2_rn(make_float2((float)A, (float)A), x, make_float2((float)B, (float)B)));
                const float2 r = __ffma2_rn(make_float2((float)SCALE, (float)SCALE),
                                            make_float2(fminf(fmaxf(v.x, 20000.0f), 65535.0f), fminf(fmaxf(v.y, 20000.0f), 65535.0f)),
                                            make_float2((float)CONST, (float)CONST));
                rp[k] = (T)r.x; rp[k + 1] = (T)r.y;
            }
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const T x = M<T>::clamp(th[k], T(0), T(65535)) * T(60.0 / 65535.0);
                rp[k] = SCALE * M<T>::clamp(x * (A * x + B), T(20000), T(65535)) + CONST;
            }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {                                            // thrust[[3,2,1,0]] + disturbance (:250-253)
            if (noise4) {
                const T r2 = rp[3 - k] * rp[3 - k] + noise4[k] * T(1.0 / 3.16e-10);
                m.rpms[k] = M<T>::clamp(M<T>::fsqrt(M<T>::fmax(r2, T(0))), RPM_LO, RPM_HI);
            } else {
                m.rpms[k] = rp[3 - k];                                           // no draw: sqrt(rp^2) = rp, already inside the limits
            }
        }
        return;
    }
    T pwms[4] = {T(0), T(0), T(0), T(0)};
    if (!stop) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const T x = clampT(th[k], T(0), T(65535)) / T(65535) * T(60);
            const T volts = T(-0.0006239) * (x * x) + T(0.088) * x;
            const T v3 = volts / T(3);
            const T pct = v3 < T(1) ? v3 : T(1);
            pwms[k] = pct * T(65535);
        }
    }
    // :246-262
    T thrust[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const T c = M<T>::clamp(pwms[k], T(20000), T(65535));
        const T rp = SCALE * c + CONST;
        thrust[k] = KF * (rp * rp);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        T tk = thrust[3 - k] + noise4[k];                                        // thrust[[3,2,1,0]] + disturbance (:250-253)
        if (tk < T(0)) tk = T(0);                                                // _thr2pwm (:307-343)
        T pwm = (M<T>::sqrt(tk / T(1) / KF) - CONST) / SCALE;
        pwm = M<T>::clamp(pwm, T(20000), T(65535));
        m.rpms[k] = SCALE * pwm + CONST;
    }
}

}  // namespace adrp
