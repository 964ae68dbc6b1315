// adrp_math.cuh -- scalar helpers of the fused step kernel (float and double instantiations).
//
// Quaternion order is xyzw (PyBullet).  The closed forms follow the call sites of the reference:
//   rotmat       p.getMatrixFromQuaternion  (envs/BaseAviary.py:778,843; control/DSLPIDControl.py:187,240)
//   euler_rpy    p.getEulerFromQuaternion   (envs/BaseAviary.py:522; control/DSLPIDControl.py:144,241)
//   quat_from_rpy p.getQuaternionFromEuler  (envs/BaseAviary.py:492)
//   canon_quat   the reset->get pose round trip of a Bullet body (envs/BaseAviary.py:869-873 then :521)
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

namespace adrp {

template <typename T> struct M;
template <> struct M<float> {
    static __device__ __forceinline__ float sqrt(float x) { return sqrtf(x); }
    static __device__ __forceinline__ float rsqrt(float x) { return rsqrtf(x); }
    static __device__ __forceinline__ float sin(float x) { return sinf(x); }
    static __device__ __forceinline__ float cos(float x) { return cosf(x); }
    static __device__ __forceinline__ void sincos(float x, float* s, float* c) { sincosf(x, s, c); }
    static __device__ __forceinline__ float atan2(float y, float x) { return atan2f(y, x); }
    static __device__ __forceinline__ float asin(float x) { return asinf(x); }
    static __device__ __forceinline__ float exp(float x) { return expf(x); }
    static __device__ __forceinline__ float log(float x) { return logf(x); }
    static __device__ __forceinline__ float abs(float x) { return fabsf(x); }
    static __device__ __forceinline__ float tan(float x) { return tanf(x); }
    static __device__ __forceinline__ float fmax(float a, float b) { return fmaxf(a, b); }
    static __device__ __forceinline__ float fmin(float a, float b) { return fminf(a, b); }
    // ---- fp32 product build only: MUFU-based approximations (<= 2 ulp) for the substep loop.  The epilogue
    // (flags, observation) keeps the IEEE forms above, so flags stay a bit-exact function of the stored state.
    static constexpr bool fast = true;
    static __device__ __forceinline__ float rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
    static __device__ __forceinline__ float fdiv(float a, float b) { return a * rcp(b); }
    static __device__ __forceinline__ float fsqrt(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
    static __device__ __forceinline__ float frsqrt(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
    static __device__ __forceinline__ float fexp2(float x) { float r; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
    static __device__ __forceinline__ float fexp(float x) { return fexp2(x * 1.4426950408889634f); }
    // sin / cos of the half rotation angle of one substep: minimax polynomials on |x| <= pi/4 (Cephes sinf / cosf
    // kernels, ~1 ulp), the libm call beyond
    static __device__ __forceinline__ void sincos_small(float x, float* s, float* c)
    {
        if (fabsf(x) <= 0.7853981633974483f) {
            const float z = x * x;
            *s = x + x * z * (-1.6666654611e-1f + z * (8.3321608736e-3f + z * -1.9515295891e-4f));
            *c = 1.0f - 0.5f * z + z * z * (4.166664568298827e-2f + z * (-1.388731625493765e-3f + z * 2.443315711809948e-5f));
        } else {
            sincosf(x, s, c);
        }
    }
    static __device__ __forceinline__ float clamp(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }
    // atan2 for the substep loop: octant reduction through one MUFU.RCP, then atan(a) = a + a^3 q(a^2) on [0, 1]
    // (degree-7 minimax q, 7e-8 absolute error = 1 ulp at pi/4); signed zeros / infinities are not special-cased
    static __device__ __forceinline__ float atan2_loop(float y, float x)
    {
        const float ax = fabsf(x), ay = fabsf(y);
        const float mx = fmaxf(ax, ay), mn = fminf(ax, ay);
        const float a = mn * rcp(fmaxf(mx, 1e-37f));
        const float z = a * a;
        float q = 2.6222428084e-03f;
        q = q * z + -1.5132528995e-02f;
        q = q * z + 4.1121847573e-02f;
        q = q * z + -7.3667049965e-02f;
        q = q * z + 1.0573931535e-01f;
        q = q * z + -1.4185975150e-01f;
        q = q * z + 1.9990396606e-01f;
        q = q * z + -3.3332987058e-01f;
        float r = (a * z) * q + a;
        if (ay > ax) r = 1.5707963267948966f - r;
        if (x < 0.0f) r = 3.141592653589793f - r;
        return copysignf(r, y);
    }
    // Box-Muller pair for the substep loop's disturbance draws: radius sqrt(-2 ln u), u in (0, 1], through lg2.approx / sqrt.approx;
    // sin / cos of 2 pi v, v in [0, 1), through sin.approx / cos.approx on the angle shifted into [-pi, pi) (absolute error
    // ~5e-7 of a unit normal; the draws scale a noise of 1e-3 N)
    static __device__ __forceinline__ void box_muller(float u, float v, float* n0, float* n1)
    {
        float l, sn, cs;
        asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l) : "f"(u));
        const float r = fsqrt(l * -1.3862943611198906f);            // -2 ln 2 * log2(u)
        const float a = v * 6.283185307179586f - 3.141592653589793f;
        asm("sin.approx.ftz.f32 %0, %1;" : "=f"(sn) : "f"(a));
        asm("cos.approx.ftz.f32 %0, %1;" : "=f"(cs) : "f"(a));
        *n0 = -(r * cs); *n1 = -(r * sn);                           // cos(a + pi) = -cos a, sin(a + pi) = -sin a
    }
    // asin for the substep loop: Cephes asinf kernel, x + x z p(z) on |x| <= 0.5 (the usual case: pitch below 30 degrees),
    // pi/2 - 2 asin(sqrt((1 - |x|) / 2)) beyond; ~1 ulp
    static __device__ __forceinline__ float asin_poly(float x, float z)
    {
        float p = 4.2163199048e-2f;
        p = p * z + 2.4181311049e-2f;
        p = p * z + 4.5470025998e-2f;
        p = p * z + 7.4953002686e-2f;
        p = p * z + 1.6666752422e-1f;
        return x + (x * z) * p;
    }
    static __device__ __forceinline__ float asin_loop(float x)
    {
        const float a = fabsf(x);
        if (a <= 0.5f) return asin_poly(x, x * x);
        const float z = 0.5f * (1.0f - a);
        const float r = 1.5707963267948966f - 2.0f * asin_poly(fsqrt(z), z);
        return copysignf(r, x);
    }
    // cos(th) and sin(th)/th from z = th^2 (both even in th: no square root, no division); valid for z <= (pi/4)^2
    static __device__ __forceinline__ void cos_sinc_sq(float z, float* c, float* sinc)
    {
        *sinc = 1.0f + z * (-1.6666654611e-1f + z * (8.3321608736e-3f + z * -1.9515295891e-4f));
        *c = 1.0f - 0.5f * z + z * z * (4.166664568298827e-2f + z * (-1.388731625493765e-3f + z * 2.443315711809948e-5f));
    }
};
template <> struct M<double> {
    static __device__ __forceinline__ double sqrt(double x) { return ::sqrt(x); }
    static __device__ __forceinline__ double rsqrt(double x) { return 1.0 / ::sqrt(x); }
    static __device__ __forceinline__ double sin(double x) { return ::sin(x); }
    static __device__ __forceinline__ double cos(double x) { return ::cos(x); }
    static __device__ __forceinline__ void sincos(double x, double* s, double* c) { ::sincos(x, s, c); }
    static __device__ __forceinline__ double atan2(double y, double x) { return ::atan2(y, x); }
    static __device__ __forceinline__ double asin(double x) { return ::asin(x); }
    static __device__ __forceinline__ double exp(double x) { return ::exp(x); }
    static __device__ __forceinline__ double log(double x) { return ::log(x); }
    static __device__ __forceinline__ double abs(double x) { return ::fabs(x); }
    static __device__ __forceinline__ double tan(double x) { return ::tan(x); }
    static __device__ __forceinline__ double fmax(double a, double b) { return ::fmax(a, b); }
    static __device__ __forceinline__ double fmin(double a, double b) { return ::fmin(a, b); }
    // fp64 parity build: every operation stays the IEEE statement the reference executes
    static constexpr bool fast = false;
    static __device__ __forceinline__ double rcp(double x) { return 1.0 / x; }
    static __device__ __forceinline__ double fdiv(double a, double b) { return a / b; }
    static __device__ __forceinline__ double fsqrt(double x) { return ::sqrt(x); }
    static __device__ __forceinline__ double frsqrt(double x) { return 1.0 / ::sqrt(x); }
    static __device__ __forceinline__ double fexp2(double x) { return ::exp2(x); }
    static __device__ __forceinline__ double fexp(double x) { return ::exp(x); }
    static __device__ __forceinline__ void sincos_small(double x, double* s, double* c) { ::sincos(x, s, c); }
    static __device__ __forceinline__ double clamp(double v, double lo, double hi) { return v < lo ? lo : (v > hi ? hi : v); }
    static __device__ __forceinline__ double atan2_loop(double y, double x) { return ::atan2(y, x); }
    static __device__ __forceinline__ double asin_loop(double x) { return ::asin(x); }
    static __device__ __forceinline__ void box_muller(double u, double v, double* n0, double* n1)   // (fast build only; unused here)
    {
        const double r = ::sqrt(-2.0 * ::log(u));
        *n0 = r * ::cos(6.283185307179586 * v); *n1 = r * ::sin(6.283185307179586 * v);
    }
    static __device__ __forceinline__ void cos_sinc_sq(double z, double* c, double* sinc)
    {
        const double th = ::sqrt(z);
        *c = ::cos(th);
        *sinc = th > 0.0 ? ::sin(th) / th : 1.0;
    }
};

// Product that is never contracted into a following add.  The motor thrusts / torques of a symmetric command
// must cancel EXACTLY in the torque sums (as they do in the reference's float64 numpy): a fused multiply-add
// would leave the rounding error of one product behind, and DSLPID's roll / yaw loops amplify any such seed
// (SURVEY F5).  Everything else in the fp32 build is free to use FMA.
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }

// Value barrier: the compiler may not assume anything about v afterwards (keeps a rarely needed recomputation inside
// its branch instead of being merged with, and hoisted next to, the hot path's common subexpressions).
__device__ __forceinline__ void opaque(float& v) { asm volatile("" : "+f"(v)); }
__device__ __forceinline__ void opaque(double& v) { asm volatile("" : "+d"(v)); }

template <typename T> __device__ __forceinline__ T clampT(T v, T lo, T hi) { return v < lo ? lo : (v > hi ? hi : v); }

template <typename T> struct V3 { T x, y, z; };
template <typename T> struct Q4 { T x, y, z, w; };
template <typename T> struct Mat3 { T m00, m01, m02, m10, m11, m12, m20, m21, m22; };

template <typename T> __device__ __forceinline__ V3<T> cross(const V3<T>& a, const V3<T>& b)
{
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
template <typename T> __device__ __forceinline__ T dot(const V3<T>& a, const V3<T>& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
template <typename T> __device__ __forceinline__ V3<T> mul(const Mat3<T>& R, const V3<T>& v)
{
    return {R.m00 * v.x + R.m01 * v.y + R.m02 * v.z, R.m10 * v.x + R.m11 * v.y + R.m12 * v.z,
            R.m20 * v.x + R.m21 * v.y + R.m22 * v.z};
}
template <typename T> __device__ __forceinline__ V3<T> mulT(const Mat3<T>& R, const V3<T>& v)
{
    return {R.m00 * v.x + R.m10 * v.y + R.m20 * v.z, R.m01 * v.x + R.m11 * v.y + R.m21 * v.z,
            R.m02 * v.x + R.m12 * v.y + R.m22 * v.z};
}

// btMatrix3x3::setRotation: normalises by 2/|q|^2.
template <typename T> __device__ __forceinline__ Mat3<T> rotmat(const Q4<T>& q)
{
    const T d = q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w;
    const T s = T(2) / d;
    const T xs = q.x * s, ys = q.y * s, zs = q.z * s;
    const T wx = q.w * xs, wy = q.w * ys, wz = q.w * zs;
    const T xx = q.x * xs, xy = q.x * ys, xz = q.x * zs;
    const T yy = q.y * ys, yz = q.y * zs, zz = q.z * zs;
    return {T(1) - (yy + zz), xy - wz, xz + wy, xy + wz, T(1) - (xx + zz), yz - wx, xz - wy, yz + wx, T(1) - (xx + yy)};
}

// rotmat for the substep loop: same formula, the normalisation through the build's fdiv.
template <typename T> __device__ __forceinline__ Mat3<T> rotmat_loop(const Q4<T>& q)
{
    const T d = q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w;
    const T s = M<T>::fdiv(T(2), d);
    const T xs = q.x * s, ys = q.y * s, zs = q.z * s;
    const T wx = q.w * xs, wy = q.w * ys, wz = q.w * zs;
    const T xx = q.x * xs, xy = q.x * ys, xz = q.x * zs;
    const T yy = q.y * ys, yz = q.y * zs, zz = q.z * zs;
    return {T(1) - (yy + zz), xy - wz, xz + wy, xy + wz, T(1) - (xx + zz), yz - wx, xz - wy, yz + wx, T(1) - (xx + yy)};
}

// Rotation matrix of a UNIT quaternion (no division by |q|^2): the product build normalises the pose when it loads it and
// once per control step afterwards; inside the substep loop the polynomial cos / sinc updates keep |q| = 1 to ~1e-7 per
// substep, which is the rounding level of the matrix entries themselves.
template <typename T> __device__ __forceinline__ Mat3<T> rotmat_unit(const Q4<T>& q)
{
    const T xs = q.x + q.x, ys = q.y + q.y, zs = q.z + q.z;
    const T wx = q.w * xs, wy = q.w * ys, wz = q.w * zs;
    const T xx = q.x * xs, xy = q.x * ys, xz = q.x * zs;
    const T yy = q.y * ys, yz = q.y * zs, zz = q.z * zs;
    return {T(1) - (yy + zz), xy - wz, xz + wy, xy + wz, T(1) - (xx + zz), yz - wx, xz - wy, yz + wx, T(1) - (xx + yy)};
}
template <typename T> __device__ __forceinline__ Q4<T> unit_quat_loop(const Q4<T>& q)
{
    const T inv = M<T>::frsqrt(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
    return {q.x * inv, q.y * inv, q.z * inv, q.w * inv};
}

// canon_quat for the substep loop (normalisation through the build's frsqrt)
template <typename T> __device__ __forceinline__ Q4<T> canon_quat_loop(const Q4<T>& q);

// pybullet.getEulerFromQuaternion, including its gimbal-lock branches.
template <typename T> __device__ __forceinline__ V3<T> euler_rpy(const Q4<T>& q)
{
    const T sqx = q.x * q.x, sqy = q.y * q.y, sqz = q.z * q.z, squ = q.w * q.w;
    const T sarg = T(-2) * (q.x * q.z - q.w * q.y);
    const T HALF_PI = T(1.5707963267948966);
    if (sarg <= T(-0.99999)) return {T(0), -HALF_PI, T(2) * M<T>::atan2(q.x, -q.y)};
    if (sarg >= T(0.99999)) return {T(0), HALF_PI, T(2) * M<T>::atan2(-q.x, q.y)};
    return {M<T>::atan2(T(2) * (q.y * q.z + q.w * q.x), squ - sqx - sqy + sqz), M<T>::asin(sarg),
            M<T>::atan2(T(2) * (q.x * q.y + q.w * q.z), squ + sqx - sqy - sqz)};
}

// euler_rpy for the substep loop (the Mellinger wrapper differentiates it every tick): same branches, the
// build's loop atan2.
template <typename T> __device__ __forceinline__ V3<T> euler_rpy_loop(const Q4<T>& q)
{
    const T sqx = q.x * q.x, sqy = q.y * q.y, sqz = q.z * q.z, squ = q.w * q.w;
    const T sarg = T(-2) * (q.x * q.z - q.w * q.y);
    const T HALF_PI = T(1.5707963267948966);
    if (sarg <= T(-0.99999)) return {T(0), -HALF_PI, T(2) * M<T>::atan2(q.x, -q.y)};
    if (sarg >= T(0.99999)) return {T(0), HALF_PI, T(2) * M<T>::atan2(-q.x, q.y)};
    return {M<T>::atan2_loop(T(2) * (q.y * q.z + q.w * q.x), squ - sqx - sqy + sqz), M<T>::asin_loop(sarg),
            M<T>::atan2_loop(T(2) * (q.x * q.y + q.w * q.z), squ + sqx - sqy - sqz)};
}

// euler_rpy for the fast build's substep loop, read off R = rotmat_loop(q) (already there for the force models):
// m21 : m22 = 2(yz + wx) : (w^2 - x^2 - y^2 + z^2), -m20 = sarg / |q|^2, m10 : m00 = 2(xy + wz) : (w^2 + x^2 - y^2 - z^2).
// The gimbal-lock branches (|pitch| within 0.26 degrees of 90) out of line: they call libm's atan2 and would sit, never
// executed, inside the substep loop, whose footprint in the instruction cache is what the loop's `no_instruction` stalls
// pay for (C5M: 0.22 per issued instruction with 975 static instructions in the loop).  Arguments and result by value, so
// that no live state has its address taken.
template <typename T> __device__ __noinline__ T euler_gimbal_yaw(T qx, T qy, T sarg)
{
    return sarg < T(0) ? T(2) * M<T>::atan2(qx, -qy) : T(2) * M<T>::atan2(-qx, qy);
}
template <typename T> __device__ __forceinline__ V3<T> euler_rpy_rot(const Q4<T>& q, const Mat3<T>& R)
{
    const T sarg = -R.m20;
    const T HALF_PI = T(1.5707963267948966);
    if (M<T>::abs(sarg) >= T(0.99999)) return {T(0), sarg < T(0) ? -HALF_PI : HALF_PI, euler_gimbal_yaw<T>(q.x, q.y, sarg)};
    return {M<T>::atan2_loop(R.m21, R.m22), M<T>::asin_loop(sarg), M<T>::atan2_loop(R.m10, R.m00)};
}

// |roll| < pi/2 and |pitch| < pi/2 (the _groundEffect gate, envs/BaseAviary.py:749) without
// evaluating atan2/asin: roll = atan2(s, c) lies in (-pi/2, pi/2) iff c > 0 or s == c == 0;
// the gimbal branches return pitch = +-pi/2 exactly, which fails the strict inequality.
template <typename T> __device__ __forceinline__ bool level_enough(const Q4<T>& q)
{
    const T sarg = T(-2) * (q.x * q.z - q.w * q.y);
    if (sarg <= T(-0.99999) || sarg >= T(0.99999)) return false;
    const T c = q.w * q.w - q.x * q.x - q.y * q.y + q.z * q.z;
    const T s = T(2) * (q.y * q.z + q.w * q.x);
    return c > T(0) || (c == T(0) && s == T(0));
}

template <typename T> __device__ __forceinline__ Q4<T> quat_from_rpy(const V3<T>& e)
{
    T sp, cp, st, ct, ss, cs;
    M<T>::sincos(e.x / T(2), &sp, &cp);
    M<T>::sincos(e.y / T(2), &st, &ct);
    M<T>::sincos(e.z / T(2), &ss, &cs);
    Q4<T> q = {sp * ct * cs - cp * st * ss, cp * st * cs + sp * ct * ss, cp * ct * ss - sp * st * cs,
               cp * ct * cs + sp * st * ss};
    const T len = M<T>::sqrt(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
    return {q.x / len, q.y / len, q.z / len, q.w / len};
}

// Normalised, sign-canonical quaternion: equals btMatrix3x3(q).getRotation() up to rounding.
// trace = 4w^2/|q|^2 - 1 > 0 keeps w positive; otherwise the component picked by Bullet's
// diagonal comparison chain (m00 < m11 <=> x^2 < y^2 ...) is made positive.
template <typename T> __device__ __forceinline__ Q4<T> canon_quat(const Q4<T>& q)
{
    const T xx = q.x * q.x, yy = q.y * q.y, zz = q.z * q.z, ww = q.w * q.w;
    const T d = xx + yy + zz + ww;
    T inv = T(1) / M<T>::sqrt(d);
    T pivot;
    if (T(4) * ww > d) pivot = q.w;                                   // trace > 0
    else pivot = xx < yy ? (yy < zz ? q.z : q.y) : (xx < zz ? q.z : q.x);
    if (pivot < T(0)) inv = -inv;
    return {q.x * inv, q.y * inv, q.z * inv, q.w * inv};
}

template <typename T> __device__ __forceinline__ Q4<T> canon_quat_loop(const Q4<T>& q)
{
    const T xx = q.x * q.x, yy = q.y * q.y, zz = q.z * q.z, ww = q.w * q.w;
    const T d = xx + yy + zz + ww;
    T inv = M<T>::frsqrt(d);
    T pivot;
    if (T(4) * ww > d) pivot = q.w;
    else pivot = xx < yy ? (yy < zz ? q.z : q.y) : (xx < zz ? q.z : q.x);
    if (pivot < T(0)) inv = -inv;
    return {q.x * inv, q.y * inv, q.z * inv, q.w * inv};
}

// ---------------------------------------------------------------- Philox4x32-10 (Salmon et al. 2011)
__device__ __forceinline__ void philox4x32(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t out[4])
{
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
template <typename T> __device__ __forceinline__ T u01(uint32_t x) { return (T)(x >> 8) * (T)5.9604644775390625e-08; }              // [0,1)
template <typename T> __device__ __forceinline__ T u01_open(uint32_t x) { return (T)((x >> 8) + 1) * (T)5.9604644775390625e-08; }   // (0,1]
// [-1, 1): the action draw of adrp_step_n's ADRP_ACTIONS_PHILOX (exact in fp32 and fp64: 24 bits)
template <typename T> __device__ __forceinline__ T u11(uint32_t x) { return (T)(int)((x >> 8) - 0x800000u) * (T)1.1920928955078125e-07; }

}  // namespace adrp
