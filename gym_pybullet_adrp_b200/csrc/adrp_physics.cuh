// adrp_physics.cuh -- one physics substep of one drone, state in registers.
//
//   DYN:  BaseAviary._dynamics + _integrateQ (envs/BaseAviary.py:822-896), followed by the pose
//         round trip the reference makes through PyBullet every substep (:869-873 -> :521).
//   PYB:  BaseAviary._physics (:683-718) + p.stepSimulation (:373-374) for ONE unconstrained body:
//         Bullet's btMultiBody base update (default 0.04 damping, gyroscopic term, exponential-map
//         orientation update, renormalisation) -- declared restatement, DESIGN.md section 4.
//   Force models: _groundEffect (:722-757), _drag (:761-788), _downwash (:792-818).
//
// Two arithmetic flavours, selected by M<T>::fast: the fp64 parity build executes the reference's
// statements one IEEE operation at a time; the fp32 product build replaces divisions and square roots
// by MUFU approximations and precomputed reciprocals (same formulas, <= 2 ulp per operation).
#pragma once
#include "adrp_math.cuh"
#include "adrp_params.cuh"

namespace adrp {

template <typename T>
struct Drone {
    V3<T> p, v, w, rt;     // position, velocity, world angular velocity, DYN body rates
    Q4<T> q;
    T last_rpm[4];         // BaseAviary.last_clipped_action
};

// Extra thrust of each propeller near the ground (:739-749).
template <typename T>
__device__ __forceinline__ void ground_effect(const KParams<T>& P, const Drone<T>& d, const Mat3<T>& R,
                                              const T rpm2kf[4], T gnd[4])
{
    const bool on = level_enough(d.q);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        T h = d.p.z + R.m20 * P.prop_x[i] + R.m21 * P.prop_y[i];
        h = M<T>::fmax(h, P.gnd_eff_h_clip);
        const T r = M<T>::fdiv(P.prop_radius, T(4) * h);
        gnd[i] = on ? mul_rn(rpm2kf[i] * P.gnd_eff_coeff, r * r) : T(0);
    }
}

// World-frame drag force (:761-788): R^T (c o v) applied in LINK_FRAME == c o v.
template <typename T>
__device__ __forceinline__ V3<T> drag_force(const KParams<T>& P, const Drone<T>& d)
{
    T sum;
    if (M<T>::fast) {
        sum = T(0.10471975511965977) * ((d.last_rpm[0] + d.last_rpm[1]) + (d.last_rpm[2] + d.last_rpm[3]));   // 2 pi / 60
    } else {
        sum = (T(2) * T(3.141592653589793) * d.last_rpm[0] / T(60) + T(2) * T(3.141592653589793) * d.last_rpm[1] / T(60)) +
              T(2) * T(3.141592653589793) * d.last_rpm[2] / T(60) + T(2) * T(3.141592653589793) * d.last_rpm[3] / T(60);
    }
    return {-P.drag_xy * sum * d.v.x, -P.drag_xy * sum * d.v.y, -P.drag_z * sum * d.v.z};
}

// Downwash from the drones of the same env that fly above this one (:805-811).  The env's M drones
// sit in MP consecutive lanes of one warp, so the neighbour positions come over warp shuffles.
// MP > 0: lanes per env known at compile time (unrolled); MP == 0: read from the params.
template <typename T, int MP>
__device__ __forceinline__ T downwash(const KParams<T>& P, const V3<T>& p, int lane_base, bool valid)
{
    T fz = T(0);
    const T pz_pub = valid ? p.z : T(-1e30);          // padding lanes never count as "above"
    const int mpad = MP > 0 ? MP : P.Mpad;
#pragma unroll
    for (int j = 0; j < mpad; ++j) {
        const int src = lane_base + j;
        const T ox = __shfl_sync(0xffffffffu, p.x, src);
        const T oy = __shfl_sync(0xffffffffu, p.y, src);
        const T oz = __shfl_sync(0xffffffffu, pz_pub, src);
        const T dz = oz - p.z;
        const T dx = ox - p.x, dy = oy - p.y;
        if (M<T>::fast) {
            // alpha = DW1 (R / 4dz)^2, exp(-(dxy / beta)^2 / 2) without the square root; no branch
            const T dxy2 = dx * dx + dy * dy;
            const T a = (T(0.25) * P.prop_radius) * M<T>::rcp(dz);
            const T ib = M<T>::rcp(P.dw2 * dz + P.dw3);
            const T term = (P.dw1 * (a * a)) * M<T>::fexp(T(-0.5) * (dxy2 * (ib * ib)));
            fz -= (dz > T(0) && dxy2 < T(100)) ? term : T(0);
        } else {
            const T dxy = M<T>::sqrt(dx * dx + dy * dy);
            if (dz > T(0) && dxy < T(10)) {
                const T a = P.prop_radius / (T(4) * dz);
                const T alpha = P.dw1 * (a * a);
                const T beta = P.dw2 * dz + P.dw3;
                const T r = dxy / beta;
                fz -= alpha * M<T>::exp(T(-0.5) * (r * r));
            }
        }
    }
    return fz;
}

// Explicit model.  f[i] = KF rpm_i^2 (+ ground effect), tz_km[i] = KM rpm_i^2, extra_z body-z force
// (downwash), fext world force (drag).
template <typename T>
__device__ __forceinline__ void dyn_substep(const KParams<T>& P, Drone<T>& d, const Mat3<T>& R, const T f[4],
                                            const T tq[4], T extra_z, const V3<T>& fext)
{
    const T thrust = (f[0] + f[1] + f[2] + f[3]) + extra_z;
    const V3<T> fw = {R.m02 * thrust + fext.x, R.m12 * thrust + fext.y, R.m22 * thrust - P.gravity + fext.z};   // :847-848
    T zt = -tq[0] + tq[1] - tq[2] + tq[3];                                       // :852
    if (P.drone_model == ADRP_RACE) zt = -zt;                                    // :850-851
    T xt, yt;
    if (P.drone_model == ADRP_CF2P) {
        xt = (f[1] - f[3]) * P.arm_x;                                            // :857-858  (arm_x = L)
        yt = (-f[0] + f[2]) * P.arm_x;
    } else {
        xt = (f[0] + f[1] - f[2] - f[3]) * P.arm_x;                              // :854-855  (arm_x = L/sqrt 2)
        yt = (-f[0] + f[1] + f[2] - f[3]) * P.arm_x;
    }
    const V3<T> Jw = {P.jx * d.rt.x, P.jy * d.rt.y, P.jz * d.rt.z};
    const V3<T> gy = cross(d.rt, Jw);
    const V3<T> deriv = {P.jinvx * (xt - gy.x), P.jinvy * (yt - gy.y), P.jinvz * (zt - gy.z)};   // :860-861
    if (M<T>::fast) {
        const T k = P.dt * P.inv_mass;
        d.v = {d.v.x + k * fw.x, d.v.y + k * fw.y, d.v.z + k * fw.z};
    } else {
        d.v = {d.v.x + P.dt * (fw.x / P.mass), d.v.y + P.dt * (fw.y / P.mass), d.v.z + P.dt * (fw.z / P.mass)};   // :862-864
    }
    d.rt = {d.rt.x + P.dt * deriv.x, d.rt.y + P.dt * deriv.y, d.rt.z + P.dt * deriv.z};          // :865
    d.p = {d.p.x + P.dt * d.v.x, d.p.y + P.dt * d.v.y, d.p.z + P.dt * d.v.z};                    // :866
    // _integrateQ (:883-896): q <- [cos(th) I + (2/|w|) sin(th) Lambda/2] q,  th = |w| dt / 2
    const T wn2 = dot(d.rt, d.rt);
    const T wn = M<T>::fsqrt(wn2);
    if (!(M<T>::abs(wn) <= T(1e-8))) {
        T s, c;
        M<T>::sincos_small(wn * P.dt / T(2), &s, &c);
        const T k = M<T>::fast ? s * M<T>::rcp(wn) : T(2) / wn * s * T(0.5);
        const T px = d.rt.x, qy = d.rt.y, rz = d.rt.z;
        const Q4<T> q = d.q;
        d.q = {c * q.x + k * (rz * q.y - qy * q.z + px * q.w), c * q.y + k * (-rz * q.x + px * q.z + qy * q.w),
               c * q.z + k * (qy * q.x - px * q.y + rz * q.w), c * q.w + k * (-px * q.x - qy * q.y - rz * q.z)};
    }
    d.q = canon_quat_loop(d.q);
    d.w = mul(R, d.rt);                                                          // :875-879 (pre-update R)
}

// Bullet free body.  fz: total body-z force at the props + downwash; tb: body torque; fext: world force;
// m, j*: the body's mass and principal inertia (per drone in the race env, changeDynamics).
template <typename T>
__device__ __forceinline__ void pyb_substep(const KParams<T>& P, int phy, Drone<T>& d, const Mat3<T>& R, T fz,
                                            const V3<T>& tb, const V3<T>& fext, T m, T jx, T jy, T jz)
{
    // world -> base frame
    const V3<T> fw = {R.m02 * fz + fext.x, R.m12 * fz + fext.y, R.m22 * fz + fext.z - m * P.g};
    const V3<T> f_l = mulT(R, fw), v_l = mulT(R, d.v), w_l = mulT(R, d.w);
    const T vn = M<T>::fsqrt(dot(v_l, v_l)), wn = M<T>::fsqrt(dot(w_l, w_l));
    const V3<T> Jw = {jx * w_l.x, jy * w_l.y, jz * w_l.z};
    const V3<T> gy = cross(w_l, Jw);
    const T kl = P.lin_damp + P.lin_damp * vn, ka = P.ang_damp + P.ang_damp * wn;
    V3<T> a_l, al_l;
    if (M<T>::fast) {
        const T im = M<T>::rcp(m);
        a_l = {f_l.x * im - v_l.x * kl, f_l.y * im - v_l.y * kl, f_l.z * im - v_l.z * kl};
        al_l = {(tb.x - Jw.x * ka - gy.x) * M<T>::rcp(jx), (tb.y - Jw.y * ka - gy.y) * M<T>::rcp(jy),
                (tb.z - Jw.z * ka - gy.z) * M<T>::rcp(jz)};
    } else {
        a_l = {(f_l.x - m * v_l.x * kl) / m, (f_l.y - m * v_l.y * kl) / m, (f_l.z - m * v_l.z * kl) / m};
        al_l = {(tb.x - Jw.x * ka - gy.x) / jx, (tb.y - Jw.y * ka - gy.y) / jy, (tb.z - Jw.z * ka - gy.z) / jz};
    }
    const V3<T> a_w = mul(R, a_l), al_w = mul(R, al_l);
    d.v = {d.v.x + P.dt * a_w.x, d.v.y + P.dt * a_w.y, d.v.z + P.dt * a_w.z};
    d.w = {d.w.x + P.dt * al_w.x, d.w.y + P.dt * al_w.y, d.w.z + P.dt * al_w.z};
    d.p = {d.p.x + P.dt * d.v.x, d.p.y + P.dt * d.v.y, d.p.z + P.dt * d.v.z};
    // exponential map (btMultiBody::stepPositionsMultiDof)
    T ang = M<T>::fsqrt(dot(d.w, d.w));
    if (ang * P.dt > T(0.7853981633974483)) ang = T(0.7853981633974483) / P.dt;
    T k, cw;
    if (ang < T(0.001)) {
        k = T(0.5) * P.dt - (P.dt * P.dt * P.dt) * T(0.020833333333) * ang * ang;
        cw = M<T>::fast ? T(1) - T(0.125) * (ang * P.dt) * (ang * P.dt) : M<T>::cos(ang * P.dt * T(0.5));
    } else {
        T s;
        M<T>::sincos_small(T(0.5) * ang * P.dt, &s, &cw);
        k = M<T>::fdiv(s, ang);
    }
    const V3<T> ax = {d.w.x * k, d.w.y * k, d.w.z * k};
    const Q4<T> q = d.q;
    d.q = {cw * q.x + ax.x * q.w + ax.y * q.z - ax.z * q.y, cw * q.y + ax.y * q.w + ax.z * q.x - ax.x * q.z,
           cw * q.z + ax.z * q.w + ax.x * q.y - ax.y * q.x, cw * q.w - ax.x * q.x - ax.y * q.y - ax.z * q.z};
    d.q = canon_quat_loop(d.q);
    if (phy & ADRP_PHY_GROUND) {
        // declared plane response: project the lowest point of the collision cylinder back to z = 0,
        // cancel the approaching normal velocity, Coulomb friction mu = 0.5 on the tangential part
        const T r22 = T(1) - T(2) * (d.q.x * d.q.x + d.q.y * d.q.y);
        const T s2 = T(1) - r22 * r22;
        const T hz = T(0.5) * P.coll_h * M<T>::abs(r22) + P.coll_r * M<T>::fsqrt(M<T>::fmax(s2, T(0)));
        const T pen = hz - d.p.z;
        if (pen > T(0)) {
            d.p.z += pen;
            if (d.v.z < T(0)) {
                const T jn = -d.v.z;
                d.v.z = T(0);
                const T vt = M<T>::sqrt(d.v.x * d.v.x + d.v.y * d.v.y);
                if (vt > T(0)) {
                    const T kf = M<T>::fmax(T(0), T(1) - T(0.5) * jn / vt);
                    d.v.x *= kf; d.v.y *= kf;
                }
            }
        }
    }
}

}  // namespace adrp
