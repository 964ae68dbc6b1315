"""TEST INFRASTRUCTURE -- declared analytic stand-in for Bullet's collision queries on the path.

Replaces, for the race env, ``p.rayTestBatch`` (MultiRaceAviary.py:496), ``p.getContactPoints``
(:555) and ``p.getClosestPoints`` (:615,632).  PARITY UNPINNED (third-party GJK/ray-cast code,
no reference test pins a value).  The model, identical in the CUDA kernel (csrc/track.cuh):

* Track bodies are static (deviation F8) and every collision shape is a *vertical prism*: a
  yaw-rotated rectangle or a circle in the xy-plane times a z-interval.  The gate's side edges
  (``rpy="0 1.57 0"``, portal.urdf:83,112) are taken as exactly upright (E20 not reproduced).
* The drone's collision cylinder (r=.06, h=.025, cf2x_IROS.urdf:33-35) is, for distance queries,
  the vertical cylinder of radius r and half-height ``hz = h/2*|R22| + r*sqrt(1-R22^2)`` (the
  exact z-extent of the tilted cylinder).  For the gate ray test the exact oriented cylinder is used.
* distance(A,B) = sqrt(max(d_xy,0)^2 + max(d_z,0)^2) with d_xy the footprint gap and d_z the
  z-interval gap (negative = overlapping in that direction).
* A contact is reported when distance < CONTACT_BREAKING_THRESHOLD (Bullet keeps manifold points
  up to 0.02 m apart and getContactPoints returns them all); visibility when distance < 0.45.
"""
from __future__ import annotations

import math


def drone_half_height(R22, r, h):
    s2 = 1.0 - R22 * R22
    return 0.5 * h * abs(R22) + r * math.sqrt(s2 if s2 > 0.0 else 0.0)


def gap_circle_rect(px, py, r, cx, cy, yaw, hx, hy):
    """Footprint gap between a circle (centre p, radius r) and a rectangle (centre c, yaw, half hx, hy)."""
    c, s = math.cos(yaw), math.sin(yaw)
    dx, dy = px - cx, py - cy
    lx = c * dx + s * dy
    ly = -s * dx + c * dy
    ex, ey = abs(lx) - hx, abs(ly) - hy
    if ex > 0.0 or ey > 0.0:
        ox = ex if ex > 0.0 else 0.0
        oy = ey if ey > 0.0 else 0.0
        return math.sqrt(ox * ox + oy * oy) - r
    return (ex if ex > ey else ey) - r        # inside: negative penetration depth


def gap_circle_circle(px, py, r, cx, cy, r2):
    dx, dy = px - cx, py - cy
    return math.sqrt(dx * dx + dy * dy) - r - r2


def gap_interval(z, hz, cz, hz2):
    return abs(z - cz) - hz - hz2


def combine(dxy, dz):
    a = dxy if dxy > 0.0 else 0.0
    b = dz if dz > 0.0 else 0.0
    if a == 0.0 and b == 0.0:
        return dxy if dxy > dz else dz        # overlapping: negative
    return math.sqrt(a * a + b * b)


def gate_shapes(x, y, z, yaw, low):
    """Prisms of one gate: list of ('rect', cx, cy, yaw, hx, hy, cz, hz) / ('circ', cx, cy, r, cz, hz)."""
    c, s = math.cos(yaw), math.sin(yaw)
    shapes = [
        ("rect", x, y, yaw, 0.25, 0.025, z - 0.225, 0.025),     # grey edge (bottom bar)
        ("rect", x, y, yaw, 0.25, 0.025, z + 0.225, 0.025),     # blue edge (top bar)
        ("rect", x + 0.225 * c, y + 0.225 * s, yaw, 0.025, 0.025, z, 0.25),   # green edge (upright)
        ("rect", x - 0.225 * c, y - 0.225 * s, yaw, 0.025, 0.025, z, 0.25),   # red edge (upright)
    ]
    if low:
        shapes.append(("rect", x, y, yaw, 0.075, 0.075, z - 0.4, 0.125))
    else:
        shapes.append(("circ", x, y, 0.05, z - 0.6, 0.4))
    return shapes


def obstacle_shapes(x, y, z):
    return [("circ", x, y, 0.05, z, 0.4), ("rect", x, y, 0.0, 0.075, 0.075, z - 0.4, 0.125)]


def drone_shape_distance(px, py, pz, hz, r, shape):
    if shape[0] == "rect":
        _, cx, cy, yaw, hx, hy, cz, hz2 = shape
        dxy = gap_circle_rect(px, py, r, cx, cy, yaw, hx, hy)
    else:
        _, cx, cy, r2, cz, hz2 = shape
        dxy = gap_circle_circle(px, py, r, cx, cy, r2)
    return combine(dxy, gap_interval(pz, hz, cz, hz2))


def drone_body_distance(px, py, pz, hz, r, shapes):
    return min(drone_shape_distance(px, py, pz, hz, r, s) for s in shapes)


def drone_drone_distance(p1, hz1, p2, hz2, r):
    dxy = gap_circle_circle(p1[0], p1[1], r, p2[0], p2[1], r)
    return combine(dxy, gap_interval(p1[2], hz1, p2[2], hz2))


def drone_plane_distance(pz, hz):
    return pz - hz


def ray_cylinder(o, d, c, axis, r, hh):
    """First hit fraction t in [0,1] of segment o + t d with the solid cylinder (centre c, unit axis,
    radius r, half-height hh); None if no hit or if the segment starts inside (Bullet's convex ray
    cast reports nothing from inside)."""
    oc = [o[0] - c[0], o[1] - c[1], o[2] - c[2]]
    oa = oc[0] * axis[0] + oc[1] * axis[1] + oc[2] * axis[2]
    da = d[0] * axis[0] + d[1] * axis[1] + d[2] * axis[2]
    # components perpendicular to the axis
    op = [oc[i] - oa * axis[i] for i in range(3)]
    dp = [d[i] - da * axis[i] for i in range(3)]
    rad2 = op[0] ** 2 + op[1] ** 2 + op[2] ** 2
    if abs(oa) <= hh and rad2 <= r * r:
        return None
    t0, t1 = 0.0, 1.0
    # slab |oa + t da| <= hh
    if da == 0.0:
        if abs(oa) > hh:
            return None
    else:
        ta, tb = (-hh - oa) / da, (hh - oa) / da
        if ta > tb:
            ta, tb = tb, ta
        t0, t1 = max(t0, ta), min(t1, tb)
        if t0 > t1:
            return None
    # infinite cylinder |op + t dp|^2 <= r^2
    A = dp[0] ** 2 + dp[1] ** 2 + dp[2] ** 2
    B = op[0] * dp[0] + op[1] * dp[1] + op[2] * dp[2]
    C = rad2 - r * r
    if A == 0.0:
        if C > 0.0:
            return None
    else:
        disc = B * B - A * C
        if disc < 0.0:
            return None
        sq = math.sqrt(disc)
        ta, tb = (-B - sq) / A, (-B + sq) / A
        t0, t1 = max(t0, ta), min(t1, tb)
        if t0 > t1:
            return None
    return t0
