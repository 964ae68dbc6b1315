// adrp_dslpid.cuh -- DSL PID controller as a device function (control/DSLPIDControl.py:149-259).
//
// One call per drone per CONTROL step.  The controller state (integral_pos_e, integral_rpy_e,
// last_rpy) lives in registers for the call and in the SoA `pid` buffer between launches.
// scipy's Rotation.from_matrix(..).as_euler('XYZ') / from_euler('XYZ') (DSLPIDControl.py:205,242)
// are the closed forms of the intrinsic X-Y-Z sequence R = Rx(a) Ry(b) Rz(c).
#pragma once
#include "../../include/adrp.h"
#include "adrp_math.cuh"

namespace adrp {

template <typename T>
struct PidState { V3<T> ipos, irpy, last_rpy; };

template <typename T>
__device__ __forceinline__ void dslpid_control(T pid_gravity, T pid_kf, int pid_model, T dt,
                                               const V3<T>& pos, const Q4<T>& quat, const V3<T>& vel,
                                               const V3<T>& tpos, const V3<T>& trpy, const V3<T>& tvel,
                                               const V3<T>& trates, PidState<T>& c, T rpm[4], V3<T>& pos_e, T& yaw_e)
{
    const Mat3<T> R = rotmat(quat);                                             // :187
    pos_e = {tpos.x - pos.x, tpos.y - pos.y, tpos.z - pos.z};                   // :188
    const V3<T> vel_e = {tvel.x - vel.x, tvel.y - vel.y, tvel.z - vel.z};       // :189
    c.ipos.x = clampT(c.ipos.x + pos_e.x * dt, T(-2), T(2));                    // :190-191
    c.ipos.y = clampT(c.ipos.y + pos_e.y * dt, T(-2), T(2));
    c.ipos.z = clampT(clampT(c.ipos.z + pos_e.z * dt, T(-2), T(2)), T(-0.15), T(0.15));   // :192
    const V3<T> tt = {T(.4) * pos_e.x + T(.05) * c.ipos.x + T(.2) * vel_e.x,    // :194-196
                      T(.4) * pos_e.y + T(.05) * c.ipos.y + T(.2) * vel_e.y,
                      T(1.25) * pos_e.z + T(.05) * c.ipos.z + T(.5) * vel_e.z + pid_gravity};
    const T scalar_thrust = M<T>::fmax(T(0), tt.x * R.m02 + tt.y * R.m12 + tt.z * R.m22);   // :197
    const T thrust = (M<T>::sqrt(scalar_thrust / (T(4) * pid_kf)) - T(4070.3)) / T(0.2685); // :198
    const T nt = M<T>::sqrt(dot(tt, tt));
    const V3<T> z_ax = {tt.x / nt, tt.y / nt, tt.z / nt};                       // :199
    T sy, cy;
    M<T>::sincos(trpy.z, &sy, &cy);
    const V3<T> x_c = {cy, sy, T(0)};                                           // :200
    V3<T> y_ax = cross(z_ax, x_c);
    const T ny = M<T>::sqrt(dot(y_ax, y_ax));
    y_ax = {y_ax.x / ny, y_ax.y / ny, y_ax.z / ny};                             // :201
    const V3<T> x_ax = cross(y_ax, z_ax);                                       // :202
    // target_rotation = [x_ax y_ax z_ax] (columns); as_euler('XYZ')               :203-205
    const T eb = M<T>::asin(clampT(z_ax.x, T(-1), T(1)));
    const T ea = M<T>::atan2(-z_ax.y, z_ax.z);
    const T ec = M<T>::atan2(-y_ax.x, x_ax.x);
    // ---- attitude loop (:212-259)
    const V3<T> cur_rpy = euler_rpy(quat);                                      // :241
    T sa, ca, sb, cb, sc, cc;
    M<T>::sincos(ea, &sa, &ca);
    M<T>::sincos(eb, &sb, &cb);
    M<T>::sincos(ec, &sc, &cc);
    const Mat3<T> Rd = {cb * cc, -cb * sc, sb,                                  // :242-244
                        ca * sc + sa * sb * cc, ca * cc - sa * sb * sc, -sa * cb,
                        sa * sc - ca * sb * cc, sa * cc + ca * sb * sc, ca * cb};
    // vee(Rd^T R - R^T Rd) = [e21, e02, e10]                                      :245-246
    const T e21 = (Rd.m02 * R.m01 - R.m02 * Rd.m01) + (Rd.m12 * R.m11 - R.m12 * Rd.m11) + (Rd.m22 * R.m21 - R.m22 * Rd.m21);
    const T e02 = (Rd.m00 * R.m02 - R.m00 * Rd.m02) + (Rd.m10 * R.m12 - R.m10 * Rd.m12) + (Rd.m20 * R.m22 - R.m20 * Rd.m22);
    const T e10 = (Rd.m01 * R.m00 - R.m01 * Rd.m00) + (Rd.m11 * R.m10 - R.m11 * Rd.m10) + (Rd.m21 * R.m20 - R.m21 * Rd.m20);
    const V3<T> rate_e = {trates.x - (cur_rpy.x - c.last_rpy.x) / dt,           // :247
                          trates.y - (cur_rpy.y - c.last_rpy.y) / dt,
                          trates.z - (cur_rpy.z - c.last_rpy.z) / dt};
    c.last_rpy = cur_rpy;                                                       // :248
    c.irpy.x = clampT(clampT(c.irpy.x - e21 * dt, T(-1500), T(1500)), T(-1), T(1));   // :249-251
    c.irpy.y = clampT(clampT(c.irpy.y - e02 * dt, T(-1500), T(1500)), T(-1), T(1));
    c.irpy.z = clampT(c.irpy.z - e10 * dt, T(-1500), T(1500));
    const T tx = clampT(T(-70000) * e21 + T(20000) * rate_e.x + T(0) * c.irpy.x, T(-3200), T(3200));    // :253-256
    const T ty = clampT(T(-70000) * e02 + T(20000) * rate_e.y + T(0) * c.irpy.y, T(-3200), T(3200));
    const T tz = clampT(T(-60000) * e10 + T(12000) * rate_e.z + T(500) * c.irpy.z, T(-3200), T(3200));
    T mix[4];
    if (pid_model == ADRP_CF2P) {                                               // :55-60
        mix[0] = -ty - tz; mix[1] = tx + tz; mix[2] = ty - tz; mix[3] = -tx + tz;
    } else {                                                                    // :48-53
        mix[0] = T(-.5) * tx + T(-.5) * ty - tz; mix[1] = T(-.5) * tx + T(.5) * ty + tz;
        mix[2] = T(.5) * tx + T(.5) * ty - tz;   mix[3] = T(.5) * tx + T(-.5) * ty + tz;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
        rpm[i] = T(0.2685) * clampT(thrust + mix[i], T(20000), T(65535)) + T(4070.3);   // :257-259
    yaw_e = ec - cur_rpy.z;                                                     // :145
}

}  // namespace adrp
