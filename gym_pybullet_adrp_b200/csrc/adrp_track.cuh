// adrp_track.cuh -- analytic stand-ins for the Bullet queries of the race env.
//
//   gate_distance / obstacle_distance   p.getClosestPoints(body, drone, 0.45) (envs/MultiRaceAviary.py:615,632)
//                                       and p.getContactPoints (:555): distance between the drone's collision
//                                       cylinder (cf2x_IROS.urdf:33-35) and the URDF primitives of portal.urdf /
//                                       low_portal.urdf / obstacle.urdf, all treated as vertical prisms.
//   ray_cylinder                        one ray of p.rayTestBatch (:496) against a drone's oriented cylinder.
// DECLARED geometric model (gates and obstacles are static here); same formulas as the
// CPU checker used by the tests.
#pragma once
#include "adrp_math.cuh"

namespace adrp {

// gap between a circle (p, r) and an oriented rectangle (centre c, half extents h, cos/sin of its yaw)
template <typename T>
__device__ __forceinline__ T gap_circle_rect(T px, T py, T r, T cx, T cy, T c, T s, T hx, T hy)
{
    const T dx = px - cx, dy = py - cy;
    const T lx = c * dx + s * dy, ly = -s * dx + c * dy;
    const T ex = M<T>::abs(lx) - hx, ey = M<T>::abs(ly) - hy;
    if (ex > T(0) || ey > T(0)) {
        const T ox = ex > T(0) ? ex : T(0), oy = ey > T(0) ? ey : T(0);
        return M<T>::sqrt(ox * ox + oy * oy) - r;
    }
    return (ex > ey ? ex : ey) - r;
}

template <typename T>
__device__ __forceinline__ T gap_circle_circle(T px, T py, T r, T cx, T cy, T r2)
{
    const T dx = px - cx, dy = py - cy;
    return M<T>::sqrt(dx * dx + dy * dy) - r - r2;
}

// horizontal gap and vertical gap -> distance (negative = penetration depth along the shallower axis)
template <typename T>
__device__ __forceinline__ T combine_gaps(T dxy, T dz)
{
    const T a = dxy > T(0) ? dxy : T(0), b = dz > T(0) ? dz : T(0);
    if (a == T(0) && b == T(0)) return dxy > dz ? dxy : dz;
    return M<T>::sqrt(a * a + b * b);
}

template <typename T>
__device__ __forceinline__ T prism_rect(const V3<T>& p, T hz, T r, T cx, T cy, T c, T s, T hx, T hy, T cz, T hz2)
{
    return combine_gaps(gap_circle_rect(p.x, p.y, r, cx, cy, c, s, hx, hy), M<T>::abs(p.z - cz) - hz - hz2);
}

template <typename T>
__device__ __forceinline__ T prism_circ(const V3<T>& p, T hz, T r, T cx, T cy, T r2, T cz, T hz2)
{
    return combine_gaps(gap_circle_circle(p.x, p.y, r, cx, cy, r2), M<T>::abs(p.z - cz) - hz - hz2);
}

// vertical half extent of the tilted collision cylinder (r22 = R[2][2] of the drone)
template <typename T>
__device__ __forceinline__ T drone_half_height(T coll_r, T coll_h, T r22)
{
    const T s2 = T(1) - r22 * r22;
    return T(0.5) * coll_h * M<T>::abs(r22) + coll_r * M<T>::sqrt(s2 > T(0) ? s2 : T(0));
}

// portal.urdf: two horizontal bars at z -+ 0.225, two posts at +-0.225 along the gate, a support cylinder;
// low_portal.urdf: the same frame on a 0.15 x 0.15 x 0.25 box.
template <typename T>
__device__ __forceinline__ T gate_distance(const V3<T>& p, T hz, T r, T x, T y, T z, T c, T s, int low)
{
    T d = prism_rect(p, hz, r, x, y, c, s, T(0.25), T(0.025), z - T(0.225), T(0.025));
    T e = prism_rect(p, hz, r, x, y, c, s, T(0.25), T(0.025), z + T(0.225), T(0.025));
    if (e < d) d = e;
    e = prism_rect(p, hz, r, x + T(0.225) * c, y + T(0.225) * s, c, s, T(0.025), T(0.025), z, T(0.25));
    if (e < d) d = e;
    e = prism_rect(p, hz, r, x - T(0.225) * c, y - T(0.225) * s, c, s, T(0.025), T(0.025), z, T(0.25));
    if (e < d) d = e;
    if (low) e = prism_rect(p, hz, r, x, y, c, s, T(0.075), T(0.075), z - T(0.4), T(0.125));
    else e = prism_circ(p, hz, r, x, y, T(0.05), z - T(0.6), T(0.4));
    return e < d ? e : d;
}

// obstacle.urdf: cylinder r = 0.05, l = 0.8 on a 0.15 x 0.15 x 0.25 box (yaw 0: cos 1, sin 0)
template <typename T>
__device__ __forceinline__ T obstacle_distance(const V3<T>& p, T hz, T r, T x, T y, T z)
{
    const T d = prism_circ(p, hz, r, x, y, T(0.05), z, T(0.4));
    const T e = prism_rect(p, hz, r, x, y, T(1), T(0), T(0.075), T(0.075), z - T(0.4), T(0.125));
    return e < d ? e : d;
}

// First hit fraction of the segment o + t d (t in [0, 1]) with the solid cylinder (centre c, unit axis ax,
// radius r, half height hh); -1 = no hit, also when the segment starts inside (Bullet's convex ray cast
// reports nothing from inside).
template <typename T>
__device__ __forceinline__ T ray_cylinder(const V3<T>& o, const V3<T>& d, const V3<T>& c, const V3<T>& ax, T r, T hh)
{
    const V3<T> oc = {o.x - c.x, o.y - c.y, o.z - c.z};
    const T oa = dot(oc, ax), da = dot(d, ax);
    const V3<T> op = {oc.x - oa * ax.x, oc.y - oa * ax.y, oc.z - oa * ax.z};
    const V3<T> dp = {d.x - da * ax.x, d.y - da * ax.y, d.z - da * ax.z};
    const T rad2 = dot(op, op);
    if (M<T>::abs(oa) <= hh && rad2 <= r * r) return T(-1);
    T t0 = T(0), t1 = T(1);
    if (da == T(0)) {
        if (M<T>::abs(oa) > hh) return T(-1);
    } else {
        T ta = (-hh - oa) / da, tb = (hh - oa) / da;
        if (ta > tb) { const T tmp = ta; ta = tb; tb = tmp; }
        if (ta > t0) t0 = ta;
        if (tb < t1) t1 = tb;
        if (t0 > t1) return T(-1);
    }
    const T A = dot(dp, dp), B = dot(op, dp), C = rad2 - r * r;
    if (A == T(0)) {
        if (C > T(0)) return T(-1);
    } else {
        const T disc = B * B - A * C;
        if (disc < T(0)) return T(-1);
        const T sq = M<T>::sqrt(disc), ta = (-B - sq) / A, tb = (-B + sq) / A;
        if (ta > t0) t0 = ta;
        if (tb < t1) t1 = tb;
        if (t0 > t1) return T(-1);
    }
    return t0;
}

}  // namespace adrp
