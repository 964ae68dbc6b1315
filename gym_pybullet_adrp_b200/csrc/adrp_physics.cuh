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
    if (M<T>::fast) {
        // |roll|, |pitch| < pi/2 read off the rotation matrix of this substep: for a unit quaternion the
        // cosine term of the roll atan2 is m22 and the pitch asin argument is -m20
        const bool on = R.m22 > T(0) && M<T>::abs(R.m20) < T(0.99999);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            T h = d.p.z + R.m20 * P.prop_x[i] + R.m21 * P.prop_y[i];
            h = M<T>::fmax(h, P.gnd_eff_h_clip);
            gnd[i] = on ? mul_rn(rpm2kf[i] * (P.gnd_eff_coeff * P.qr4 * P.qr4), M<T>::rcp(h * h)) : T(0);
        }
        return;
    }
    const bool on = level_enough(d.q);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        T h = d.p.z + R.m20 * P.prop_x[i] + R.m21 * P.prop_y[i];
        h = M<T>::fmax(h, P.gnd_eff_h_clip);
        const T r = P.prop_radius / (T(4) * h);
        gnd[i] = on ? mul_rn(rpm2kf[i] * P.gnd_eff_coeff, r * r) : T(0);
    }
}

// Product build: thrust INCLUDING the ground effect, f_i (1 + c / h_i^2) -- one select for the tilt gate, one multiply-add
// per propeller.  Equal commands at equal heights still give bit-equal thrusts (the torque sums cancel exactly).
template <typename T>
__device__ __forceinline__ void ground_effect_total(const KParams<T>& P, const Drone<T>& d, const Mat3<T>& R,
                                                    const T rpm2kf[4], T out[4])
{
    const bool on = R.m22 > T(0) && M<T>::abs(R.m20) < T(0.99999);
    const T c = on ? P.gnd_eff_coeff * P.qr4 * P.qr4 : T(0);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        T h = d.p.z + R.m20 * P.prop_x[i] + R.m21 * P.prop_y[i];
        h = M<T>::fmax(h, P.gnd_eff_h_clip);
        out[i] = mul_rn(rpm2kf[i], c * M<T>::rcp(h * h) + T(1));
    }
}

// The same for props (0, 1) and (2, 3) at once in the packed fp32 forms (see downwash_pair_sym2): 20 scalar FP instructions -> 10.
__device__ __forceinline__ void ground_effect_total2(const KParams<float>& P, const Drone<float>& d, const Mat3<float>& R,
                                                     const float rpm2kf[4], float out[4])
{
    const bool on = R.m22 > 0.0f && fabsf(R.m20) < 0.99999f;
    const float c = on ? P.gnd_eff_coeff * P.qr4 * P.qr4 : 0.0f;
    const float2 r20 = make_float2(R.m20, R.m20), r21 = make_float2(R.m21, R.m21), pz = make_float2(d.p.z, d.p.z);
    const float2 c2 = make_float2(c, c), one = make_float2(1.0f, 1.0f);
#pragma unroll
    for (int i = 0; i < 4; i += 2) {
        float2 h = __ffma2_rn(r21, make_float2(P.prop_y[i], P.prop_y[i + 1]), __ffma2_rn(r20, make_float2(P.prop_x[i], P.prop_x[i + 1]), pz));
        h = make_float2(fmaxf(h.x, P.gnd_eff_h_clip), fmaxf(h.y, P.gnd_eff_h_clip));
        const float2 hh = __fmul2_rn(h, h);
        const float2 g = __ffma2_rn(c2, make_float2(M<float>::rcp(hh.x), M<float>::rcp(hh.y)), one);
        const float2 o = __fmul2_rn(make_float2(rpm2kf[i], rpm2kf[i + 1]), g);
        out[i] = o.x; out[i + 1] = o.y;
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

template <typename T>
__device__ __forceinline__ V3<T> drag_force_sum(const KParams<T>& P, const Drone<T>& d, T sum)
{
    return {-P.drag_xy * sum * d.v.x, -P.drag_xy * sum * d.v.y, -P.drag_z * sum * d.v.z};
}

// Downwash from the drones of the same env that fly above this one (:805-811).  The env's M drones
// sit in MP consecutive lanes of one warp, so the neighbour positions come over warp shuffles.
// MP > 0: lanes per env known at compile time (unrolled); MP == 0: read from the params.
//
// fp32 product build: alpha exp(-(dxy/beta)^2 / 2) = 2^(l2c - dxy^2 / (a dz + b)^2) / dz^2 (constants folded on the
// host, KParams::dwf_*): two MUFU.RCP and one MUFU.EX2 per pair, no square root, no branch.  The partners come
// over butterfly shuffles (lane ^ k, k = 1 .. MP-1), which visit every other drone of the env exactly once.
template <typename T>
__device__ __forceinline__ T downwash_pair_fast(const KParams<T>& P, T dx, T dy, T dz)
{
    const T dxy2 = dx * dx + dy * dy;
    const T rz = M<T>::rcp(dz);
    const T ib = M<T>::rcp(P.dwf_a * dz + P.dwf_b);
    const T e = M<T>::fexp2(P.dwf_l2c - (dxy2 * ib) * ib);
    const T term = (e * rz) * rz;
    return (dz > T(0) && dxy2 < T(100)) ? term : T(0);
}

// Both directions of one pair from a single evaluation (the formula depends on |dz| and dxy only): `mine` is the
// force on this drone when the partner flies above it, `theirs` the force on the partner when it flies below.
template <typename T>
__device__ __forceinline__ void downwash_pair_sym(const KParams<T>& P, T dx, T dy, T dz, T& mine, T& theirs)
{
    const T dxy2 = dx * dx + dy * dy;
    const T az = M<T>::abs(dz);
    const T rz = M<T>::rcp(az);
    const T ib = M<T>::rcp(P.dwf_a * az + P.dwf_b);
    const T e = M<T>::fexp2(P.dwf_l2c - (dxy2 * ib) * ib);
    const T term = dxy2 < T(100) ? (e * rz) * rz : T(0);
    mine = dz > T(0) ? term : T(0);
    theirs = dz < T(0) ? term : T(0);
}

#ifndef ADRP_PACKED_DW
#define ADRP_PACKED_DW 1
#endif
// Two evaluations of downwash_pair_sym at once in the packed fp32 forms of sm_100 (FFMA2 / FMUL2 / FADD2: two results per
// issue slot; the FMA pipe is as busy as before, but the step kernel is bound by instruction issue, not by that pipe).
// dx, dy, dz: the two partners' offsets; mine / theirs as in downwash_pair_sym, per partner.
__device__ __forceinline__ void downwash_pair_sym2(const KParams<float>& P, float2 dx, float2 dy, float2 dz, float2& mine, float2& theirs)
{
    const float2 dxy2 = __ffma2_rn(dy, dy, __fmul2_rn(dx, dx));
    const float2 az = make_float2(fabsf(dz.x), fabsf(dz.y));
    const float2 den = __ffma2_rn(make_float2(P.dwf_a, P.dwf_a), az, make_float2(P.dwf_b, P.dwf_b));
    const float2 rz = make_float2(M<float>::rcp(az.x), M<float>::rcp(az.y));
    const float2 ib = make_float2(M<float>::rcp(den.x), M<float>::rcp(den.y));
    const float2 u = __fmul2_rn(dxy2, ib);
    const float2 arg = __ffma2_rn(make_float2(-u.x, -u.y), ib, make_float2(P.dwf_l2c, P.dwf_l2c));
    const float2 e = make_float2(M<float>::fexp2(arg.x), M<float>::fexp2(arg.y));
    const float2 t = __fmul2_rn(__fmul2_rn(e, rz), rz);
    const float tx = dxy2.x < 100.0f ? t.x : 0.0f, ty = dxy2.y < 100.0f ? t.y : 0.0f;
    mine = make_float2(dz.x > 0.0f ? tx : 0.0f, dz.y > 0.0f ? ty : 0.0f);
    theirs = make_float2(dz.x < 0.0f ? tx : 0.0f, dz.y < 0.0f ? ty : 0.0f);
}

// PACKED: the caller's choice per kernel instance (measured, profiles/r02zm_*): it pays in the n-step instances (C5 +4.2 %, C4
// +2.8 %) and in C5's single-step one (+0.8 %), not in C4's single-step one (-0.6 %) nor under the race kernel's register cap
// (C5M -1.2 %).
template <typename T, int MP, bool PACKED = false>
__device__ __forceinline__ T downwash(const KParams<T>& P, const V3<T>& p, int lane_base, bool valid)
{
    T fz = T(0);
    const T pz_pub = valid ? p.z : T(-1e30);          // padding lanes never count as "above"
    const int mpad = MP > 0 ? MP : P.Mpad;
    if constexpr (M<T>::fast && MP == 8 && PACKED && ADRP_PACKED_DW) {
        // the four partners of the all-pairs scheme below (k = 1, 2, 3 and the antipodal k = 4), two per packed evaluation
        const int i = (threadIdx.x & 31) - lane_base;
        float2 ox[2], oy[2], oz[2];
#pragma unroll
        for (int k = 1; k <= 4; ++k) {
            const int fwd = lane_base + ((i + k) & 7);
            const float x = __shfl_sync(0xffffffffu, p.x, fwd), y = __shfl_sync(0xffffffffu, p.y, fwd), z = __shfl_sync(0xffffffffu, pz_pub, fwd);
            if ((k - 1) & 1) { ox[(k - 1) >> 1].y = x; oy[(k - 1) >> 1].y = y; oz[(k - 1) >> 1].y = z; }
            else { ox[(k - 1) >> 1].x = x; oy[(k - 1) >> 1].x = y; oz[(k - 1) >> 1].x = z; }
        }
        const float2 npx = make_float2(-p.x, -p.x), npy = make_float2(-p.y, -p.y), npz = make_float2(-pz_pub, -pz_pub);
        float2 m0, t0, m1, t1;
        downwash_pair_sym2(P, __fadd2_rn(ox[0], npx), __fadd2_rn(oy[0], npy), __fadd2_rn(oz[0], npz), m0, t0);
        downwash_pair_sym2(P, __fadd2_rn(ox[1], npx), __fadd2_rn(oy[1], npy), __fadd2_rn(oz[1], npz), m1, t1);
        fz -= m0.x; fz -= __shfl_sync(0xffffffffu, t0.x, lane_base + ((i - 1) & 7));
        fz -= m0.y; fz -= __shfl_sync(0xffffffffu, t0.y, lane_base + ((i - 2) & 7));
        fz -= m1.x; fz -= __shfl_sync(0xffffffffu, t1.x, lane_base + ((i - 3) & 7));
        fz -= m1.y;                                        // antipodal pair: evaluated from both ends, nothing handed back
        return fz;
    }
    if (M<T>::fast && MP >= 4) {
        // every unordered pair once: lane i evaluates (i, i+k mod MP) for k = 1 .. MP/2-1 and hands the partner's
        // share back over one more shuffle; the antipodal pairs (k = MP/2) are evaluated from both ends
        const int i = (threadIdx.x & 31) - lane_base;
#pragma unroll
        for (int k = 1; k < MP / 2; ++k) {
            const int fwd = lane_base + ((i + k) & (MP - 1)), back = lane_base + ((i - k) & (MP - 1));
            const T ox = __shfl_sync(0xffffffffu, p.x, fwd);
            const T oy = __shfl_sync(0xffffffffu, p.y, fwd);
            const T oz = __shfl_sync(0xffffffffu, pz_pub, fwd);
            T mine, theirs;
            downwash_pair_sym(P, ox - p.x, oy - p.y, oz - pz_pub, mine, theirs);
            fz -= mine;
            fz -= __shfl_sync(0xffffffffu, theirs, back);
        }
        const T ox = __shfl_xor_sync(0xffffffffu, p.x, MP / 2);
        const T oy = __shfl_xor_sync(0xffffffffu, p.y, MP / 2);
        const T oz = __shfl_xor_sync(0xffffffffu, pz_pub, MP / 2);
        fz -= downwash_pair_fast(P, ox - p.x, oy - p.y, oz - pz_pub);
        return fz;
    }
    if (M<T>::fast) {
#pragma unroll
        for (int k = 1; k < mpad; ++k) {
            const T ox = __shfl_xor_sync(0xffffffffu, p.x, k);
            const T oy = __shfl_xor_sync(0xffffffffu, p.y, k);
            const T oz = __shfl_xor_sync(0xffffffffu, pz_pub, k);
            fz -= downwash_pair_fast(P, ox - p.x, oy - p.y, oz - p.z);
        }
        return fz;
    }
#pragma unroll
    for (int j = 0; j < mpad; ++j) {
        const int src = lane_base + j;
        const T ox = __shfl_sync(0xffffffffu, p.x, src);
        const T oy = __shfl_sync(0xffffffffu, p.y, src);
        const T oz = __shfl_sync(0xffffffffu, pz_pub, src);
        const T dz = oz - p.z;
        const T dx = ox - p.x, dy = oy - p.y;
        const T dxy = M<T>::sqrt(dx * dx + dy * dy);
        if (dz > T(0) && dxy < T(10)) {
            const T a = P.prop_radius / (T(4) * dz);
            const T alpha = P.dw1 * (a * a);
            const T beta = P.dw2 * dz + P.dw3;
            const T r = dxy / beta;
            fz -= alpha * M<T>::exp(T(-0.5) * (r * r));
        }
    }
    return fz;
}

// Explicit model.  f[i] = KF rpm_i^2 (+ ground effect), tz_km[i] = KM rpm_i^2, extra_z body-z force
// (downwash), fext world force (drag).
template <typename T>
__device__ __forceinline__ void dyn_substep(const KParams<T>& P, Drone<T>& d, const Mat3<T>& R, const T f[4],
                                            const T tq[4], T extra_z, const V3<T>& fext, bool last)
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
    const T px = d.rt.x, qy = d.rt.y, rz = d.rt.z;
    const Q4<T> q = d.q;
    if (M<T>::fast) {
        // cos(th) and sin(th)/|w| = (dt/2) sinc(th) are even in th: polynomials in th^2, no sqrt / division
        const T z = wn2 * (P.half_dt * P.half_dt);
        T c, k;
        if (z <= T(0.6168502750680849)) {
            M<T>::cos_sinc_sq(z, &c, &k);
            k *= P.half_dt;
        } else {
            const T wn = M<T>::fsqrt(wn2);
            T s;
            M<T>::sincos_small(wn * P.half_dt, &s, &c);
            k = s * M<T>::rcp(wn);
        }
        d.q = {c * q.x + k * (rz * q.y - qy * q.z + px * q.w), c * q.y + k * (-rz * q.x + px * q.z + qy * q.w),
               c * q.z + k * (qy * q.x - px * q.y + rz * q.w), c * q.w + k * (-px * q.x - qy * q.y - rz * q.z)};
        // the pose round trip through Bullet only renormalises and fixes the sign: _integrateQ is linear in q and
        // every consumer inside the loop normalises by |q|^2, so the product build applies it once per control
        // step (step kernel epilogue); the world rates are needed after the last substep only
        if (last) {                                        // (the loop itself reads the third row / column of R only)
            Q4<T> ql = q;
            opaque(ql.x); opaque(ql.y); opaque(ql.z); opaque(ql.w);
            d.w = mul(rotmat_loop(ql), d.rt);
        }
        return;
    }
    const T wn = M<T>::sqrt(wn2);
    if (!(M<T>::abs(wn) <= T(1e-8))) {
        T s, c;
        M<T>::sincos(wn * P.dt / T(2), &s, &c);
        const T k = T(2) / wn * s * T(0.5);
        d.q = {c * q.x + k * (rz * q.y - qy * q.z + px * q.w), c * q.y + k * (-rz * q.x + px * q.z + qy * q.w),
               c * q.z + k * (qy * q.x - px * q.y + rz * q.w), c * q.w + k * (-px * q.x - qy * q.y - rz * q.z)};
    }
    d.q = canon_quat(d.q);
    d.w = mul(R, d.rt);                                                          // :875-879 (pre-update R)
}

// Explicit model, fp32 product build: the same update with the constants folded on the host (KParams::k_*): the body-rate
// derivative written per axis -- J^-1 (tau - w x J w) = J^-1 tau - J^-1 (J_c - J_b) w_b w_c for a diagonal inertia -- and the
// drag force -c sum_rpm v folded into a per-axis gain on the old velocity (axy, az: 1 - dt/m c sum_rpm, the caller keeps the
// value for last_clipped_action's RPMs of this substep).
// DM: prop layout at compile time -- 0: X (CF2X, RACE), 1: + (CF2P), -1: read from the params.
template <typename T, int DM = -1>
__device__ __forceinline__ void dyn_substep_fast(const KParams<T>& P, Drone<T>& d, const Mat3<T>& R, const T f[4],
                                                 const T tq[4], T extra_z, T axy, T az, bool last)
{
    const T kt = ((f[0] + f[1]) + (f[2] + f[3]) + extra_z) * P.k_m;                // dt / m * thrust (:847-848,:862-864)
    const T zt = (tq[1] - tq[0]) + (tq[3] - tq[2]);                              // :852 (RACE: sign in k_tz)
    T sx, sy;
    const bool plus = DM >= 0 ? DM == 1 : P.drone_model == ADRP_CF2P;
    if (plus) { sx = f[1] - f[3]; sy = f[2] - f[0]; }                            // :857-858
    else { sx = (f[0] + f[1]) - (f[2] + f[3]); sy = (f[1] + f[2]) - (f[0] + f[3]); }   // :854-855
    const T px = d.rt.x, qy = d.rt.y, rz = d.rt.z;
    d.rt = {px + P.k_tx * sx - P.k_gx * (qy * rz), qy + P.k_ty * sy - P.k_gy * (rz * px), rz + P.k_tz * zt - P.k_gz * (px * qy)};   // :860-861,:865
    d.v = {d.v.x * axy + kt * R.m02, d.v.y * axy + kt * R.m12, d.v.z * az + kt * R.m22 - P.k_grav};
    d.p = {d.p.x + P.dt * d.v.x, d.p.y + P.dt * d.v.y, d.p.z + P.dt * d.v.z};    // :866
    // _integrateQ (:883-896) with the NEW rates: cos(th) and sin(th)/|w| = (dt/2) sinc(th) as polynomials in th^2
    const T nx = d.rt.x, ny = d.rt.y, nz = d.rt.z;
    const T wn2 = nx * nx + ny * ny + nz * nz;
    const Q4<T> q = d.q;
    const T z = wn2 * (P.half_dt * P.half_dt);
    T c, k;
    if (z <= T(0.6168502750680849)) {
        M<T>::cos_sinc_sq(z, &c, &k);
        k *= P.half_dt;
    } else {
        const T wn = M<T>::fsqrt(wn2);
        T s;
        M<T>::sincos_small(wn * P.half_dt, &s, &c);
        k = s * M<T>::rcp(wn);
    }
    // (the four rows as two packed pairs -- 20 FP instructions -> 10 -- measured slower: C5 adrp_step_n 1.569e10 -> 1.529e10; this update is
    // on the substep's dependency chain, where the packed forms' latency counts: profiles/r02zq_packed_quaternion_ab.txt)
    d.q = {c * q.x + k * (nz * q.y - ny * q.z + nx * q.w), c * q.y + k * (-nz * q.x + nx * q.z + ny * q.w),
           c * q.z + k * (ny * q.x - nx * q.y + nz * q.w), c * q.w + k * (-nx * q.x - ny * q.y - nz * q.z)};
    if (last) {                                            // world rates: R of the pose BEFORE this update (:875-879)
        Q4<T> ql = q;
        opaque(ql.x); opaque(ql.y); opaque(ql.z); opaque(ql.w);
        d.w = mul(rotmat_unit(ql), d.rt);
    }
}

// Base-frame angular velocity for pyb_substep's fast path, and back (R: rotmat_loop of the pose at that moment).
template <typename T> __device__ __forceinline__ V3<T> pyb_body_rates(const Mat3<T>& R, const V3<T>& w) { return mulT(R, w); }
template <typename T> __device__ __forceinline__ V3<T> pyb_world_rates(const Mat3<T>& R, const V3<T>& wl) { return mul(R, wl); }

// Declared plane response of the Bullet free body once the collision cylinder has sunk `pen` into z = 0: project back,
// cancel the approaching normal velocity, Coulomb friction mu = 0.5 on the tangential part.  Product build: out of line
// (by value; returns the new p.z, v.x, v.y, v.z) -- a drone in flight never executes it, and its IEEE division / square
// root would otherwise sit inside the substep loop.
template <typename T> struct PlaneHit { T pz, vx, vy, vz; };
template <typename T> __device__ __noinline__ PlaneHit<T> plane_response(T pen, T pz, T vx, T vy, T vz)
{
    PlaneHit<T> r = {pz + pen, vx, vy, vz};
    if (vz < T(0)) {
        const T jn = -vz;
        r.vz = T(0);
        const T vt = M<T>::sqrt(vx * vx + vy * vy);
        if (vt > T(0)) {
            const T kf = M<T>::fmax(T(0), T(1) - T(0.5) * jn / vt);
            r.vx = vx * kf; r.vy = vy * kf;
        }
    }
    return r;
}

// Per-drone constants of the Bullet free body folded once per launch (product build): dt / m, dt / J and the gyroscopic
// coefficients dt / J_a (J_c - J_b) of a diagonal inertia (race env: per-drone mass / inertia, changeDynamics).
template <typename T> struct PybFold { T dim, kx, ky, kz, gx, gy, gz; };
template <typename T> __device__ __forceinline__ PybFold<T> pyb_fold(const KParams<T>& P, T m, T jx, T jy, T jz)
{
    const T ix = P.dt * M<T>::rcp(jx), iy = P.dt * M<T>::rcp(jy), iz = P.dt * M<T>::rcp(jz);
    return {P.dt * M<T>::rcp(m), ix, iy, iz, ix * (jz - jy), iy * (jx - jz), iz * (jy - jx)};
}

// Bullet free body.  fz: total body-z force at the props + downwash; tb: body torque; fext: world force;
// m, j*: the body's mass and principal inertia (per drone in the race env, changeDynamics).
// wl (fast build only): the angular velocity in the BASE frame, carried by the caller across the substeps of one
// control step (pyb_body_rates before the loop, pyb_world_rates after it); d.w is not touched in between.
template <typename T>
__device__ __forceinline__ void pyb_substep(const KParams<T>& P, int phy, Drone<T>& d, V3<T>& wl, const Mat3<T>& R, T fz,
                                            const V3<T>& tb, const V3<T>& fext, T m, T jx, T jy, T jz)
{
    const V3<T> fw = {R.m02 * fz + fext.x, R.m12 * fz + fext.y, R.m22 * fz + fext.z - m * P.g};
    if (M<T>::fast) {
        // Bullet updates the base in its own frame: a_w = R (R^T F / m - kl R^T v).  R is orthonormal (rotmat_loop
        // divides by |q|^2), so the linear part is F / m - kl v with |R^T v| = |v|.  Angular part: with u = w_l + dt al_l
        // the new world rate is R u and the pose update is q' = dq(R u dt) q = q dq(u dt), hence R' = R Exp(u dt) and the
        // next substep's base-frame rate R'^T R u = Exp(-u dt) u = u: the base-frame rate integrates without any
        // rotation, and the world rate is only needed once the control step is over.
        // w + dt J^-1 (tau - ka J w - w x J w) per axis = w (1 - dt ka) + (dt / J) tau - (dt / J (J_c - J_b)) w_b w_c
        const T vn = M<T>::fsqrt(dot(d.v, d.v)), wn = M<T>::fsqrt(dot(wl, wl));
        const PybFold<T> K = pyb_fold(P, m, jx, jy, jz);                 // loop-invariant: hoisted by the compiler
        const T dkl = (T(1) - P.k_lin) - P.k_lin * vn, dka = (T(1) - P.k_ang) - P.k_ang * wn;
        const T yz = wl.y * wl.z, zx = wl.z * wl.x, xy = wl.x * wl.y;
        wl = {wl.x * dka + K.kx * tb.x - K.gx * yz, wl.y * dka + K.ky * tb.y - K.gy * zx, wl.z * dka + K.kz * tb.z - K.gz * xy};
        const T dim = K.dim;
        d.v = {d.v.x * dkl + dim * fw.x, d.v.y * dkl + dim * fw.y, d.v.z * dkl + dim * fw.z};
        d.p = {d.p.x + P.dt * d.v.x, d.p.y + P.dt * d.v.y, d.p.z + P.dt * d.v.z};
        // exponential map (btMultiBody::stepPositionsMultiDof) in the base frame: q <- q (k u, cos)
        const T w2 = dot(wl, wl);
        T k, cw;
        if (w2 * (P.dt * P.dt) <= T(0.6168502750680849)) {
            // |w| dt <= pi/4 (no clamp): cos(|w| dt/2) and sin(|w| dt/2)/|w| as polynomials in (|w| dt/2)^2, which also
            // covers Bullet's own Taylor branch below 1e-3 rad/s
            M<T>::cos_sinc_sq(w2 * (P.half_dt * P.half_dt), &cw, &k);
            k *= P.half_dt;
        } else {
            // |w| dt > pi/4: Bullet clamps the angle, not the axis -- (w sin(pi/8) / (pi / 4 dt), cos(pi/8)), normalised later
            cw = T(0.9238795325112867);
            k = T(0.3826834323650898 / 0.7853981633974483) * P.dt;
        }
        const V3<T> ax = {wl.x * k, wl.y * k, wl.z * k};
        const Q4<T> q = d.q;
        d.q = {q.w * ax.x + q.x * cw + q.y * ax.z - q.z * ax.y, q.w * ax.y - q.x * ax.z + q.y * cw + q.z * ax.x,
               q.w * ax.z + q.x * ax.y - q.y * ax.x + q.z * cw, q.w * cw - q.x * ax.x - q.y * ax.y - q.z * ax.z};
    } else {
        // world -> base frame
        const V3<T> f_l = mulT(R, fw), v_l = mulT(R, d.v), w_l = mulT(R, d.w);
        const T vn = M<T>::fsqrt(dot(v_l, v_l)), wn = M<T>::fsqrt(dot(w_l, w_l));
        const V3<T> Jw = {jx * w_l.x, jy * w_l.y, jz * w_l.z};
        const V3<T> gy = cross(w_l, Jw);
        const T kl = P.lin_damp + P.lin_damp * vn, ka = P.ang_damp + P.ang_damp * wn;
        const V3<T> a_l = {(f_l.x - m * v_l.x * kl) / m, (f_l.y - m * v_l.y * kl) / m, (f_l.z - m * v_l.z * kl) / m};
        const V3<T> al_l = {(tb.x - Jw.x * ka - gy.x) / jx, (tb.y - Jw.y * ka - gy.y) / jy, (tb.z - Jw.z * ka - gy.z) / jz};
        const V3<T> a_w = mul(R, a_l), al_w = mul(R, al_l);
        d.v = {d.v.x + P.dt * a_w.x, d.v.y + P.dt * a_w.y, d.v.z + P.dt * a_w.z};
        d.w = {d.w.x + P.dt * al_w.x, d.w.y + P.dt * al_w.y, d.w.z + P.dt * al_w.z};
        d.p = {d.p.x + P.dt * d.v.x, d.p.y + P.dt * d.v.y, d.p.z + P.dt * d.v.z};
        // exponential map (btMultiBody::stepPositionsMultiDof)
        T ang = M<T>::sqrt(dot(d.w, d.w));
        T k, cw;
        if (ang * P.dt > T(0.7853981633974483)) ang = T(0.7853981633974483) / P.dt;
        if (ang < T(0.001)) {
            k = T(0.5) * P.dt - (P.dt * P.dt * P.dt) * T(0.020833333333) * ang * ang;
            cw = M<T>::cos(ang * P.dt * T(0.5));
        } else {
            T sn;
            M<T>::sincos(T(0.5) * ang * P.dt, &sn, &cw);
            k = sn / ang;
        }
        const V3<T> ax = {d.w.x * k, d.w.y * k, d.w.z * k};
        const Q4<T> q = d.q;
        d.q = {cw * q.x + ax.x * q.w + ax.y * q.z - ax.z * q.y, cw * q.y + ax.y * q.w + ax.z * q.x - ax.x * q.z,
               cw * q.z + ax.z * q.w + ax.x * q.y - ax.y * q.x, cw * q.w - ax.x * q.x - ax.y * q.y - ax.z * q.z};
        d.q = canon_quat(d.q);                                                   // Bullet renormalises every step
    }
    // (the product build renormalises once per control step, in the kernel epilogue: the update is linear in q and
    // rotmat_loop divides by |q|^2)
    if ((phy & ADRP_PHY_GROUND) && (!M<T>::fast || d.p.z < P.plane_reach)) {   // (product build: exact culling, KParams::plane_reach)
        // declared plane response: project the lowest point of the collision cylinder back to z = 0,
        // cancel the approaching normal velocity, Coulomb friction mu = 0.5 on the tangential part
        const T r22 = T(1) - T(2) * (d.q.x * d.q.x + d.q.y * d.q.y);
        const T s2 = T(1) - r22 * r22;
        const T hz = T(0.5) * P.coll_h * M<T>::abs(r22) + P.coll_r * M<T>::fsqrt(M<T>::fmax(s2, T(0)));
        const T pen = hz - d.p.z;
        if (M<T>::fast) {
            if (pen > T(0)) {
                const PlaneHit<T> r = plane_response<T>(pen, d.p.z, d.v.x, d.v.y, d.v.z);
                d.p.z = r.pz; d.v.x = r.vx; d.v.y = r.vy; d.v.z = r.vz;
            }
        } else if (pen > T(0)) {
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
