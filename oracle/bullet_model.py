"""TEST INFRASTRUCTURE -- declared CPU restatement of the PyBullet primitives on the path.

PyBullet (PyPI ``pybullet ^3.2.5``, pyproject.toml:20 of the reference) is an un-vendored
third-party dependency that is absent from /root/reference and not installable here, so the
arithmetic below restates Bullet's *published* algorithms (pybullet.c helpers, btMatrix3x3,
btMultiBody ABA for a floating base without moving links, btMultiBody::stepPositionsMultiDof)
and is anchored on the reference's call sites:

* getMatrixFromQuaternion  -- BaseAviary.py:778,843; DSLPIDControl.py:187,240
* getEulerFromQuaternion   -- BaseAviary.py:522; DSLPIDControl.py:144,241
* getQuaternionFromEuler   -- BaseAviary.py:492; MultiRaceAviary.py:392,401,459
* pose round trip reset->get (btTransform(q) -> btMatrix3x3::getRotation) -- BaseAviary.py:869-879,521
* applyExternalForce/Torque + stepSimulation for ONE free body -- BaseAviary.py:706-718,373-374

PARITY UNPINNED for everything in this file: no reference test holds a number for it
(SURVEY.md section 4). It is used by ``oracle/refstub`` (to run the reference's own Python code
without PyBullet) and by the oracle port; never by the product path.
"""
from __future__ import annotations

import math

BULLET_LINEAR_DAMPING = 0.04    # btMultiBody default, not disabled by the reference (BaseAviary.py:496-498)
BULLET_ANGULAR_DAMPING = 0.04
ANGULAR_MOTION_THRESHOLD = 0.5 * (math.pi / 2)
CONTACT_BREAKING_THRESHOLD = 0.02  # gContactBreakingThreshold: getContactPoints reports points closer than this


def matrix_from_quaternion(q):
    """pybullet.getMatrixFromQuaternion == btMatrix3x3::setRotation (row-major 3x3 nested list)."""
    x, y, z, w = q
    d = x * x + y * y + z * z + w * w
    s = 2.0 / d
    xs, ys, zs = x * s, y * s, z * s
    wx, wy, wz = w * xs, w * ys, w * zs
    xx, xy, xz = x * xs, x * ys, x * zs
    yy, yz, zz = y * ys, y * zs, z * zs
    return [[1.0 - (yy + zz), xy - wz, xz + wy],
            [xy + wz, 1.0 - (xx + zz), yz - wx],
            [xz - wy, yz + wx, 1.0 - (xx + yy)]]


def euler_from_quaternion(q):
    """pybullet.getEulerFromQuaternion (roll, pitch, yaw about fixed x, y, z)."""
    x, y, z, w = q
    sqx, sqy, sqz, squ = x * x, y * y, z * z, w * w
    sarg = -2.0 * (x * z - w * y)
    if sarg <= -0.99999:
        return (0.0, -0.5 * math.pi, 2.0 * math.atan2(x, -y))
    if sarg >= 0.99999:
        return (0.0, 0.5 * math.pi, 2.0 * math.atan2(-x, y))
    return (math.atan2(2.0 * (y * z + w * x), squ - sqx - sqy + sqz),
            math.asin(sarg),
            math.atan2(2.0 * (x * y + w * z), squ + sqx - sqy - sqz))


def quaternion_from_euler(rpy):
    """pybullet.getQuaternionFromEuler (normalised, xyzw)."""
    phi, the, psi = rpy[0] / 2.0, rpy[1] / 2.0, rpy[2] / 2.0
    q = [math.sin(phi) * math.cos(the) * math.cos(psi) - math.cos(phi) * math.sin(the) * math.sin(psi),
         math.cos(phi) * math.sin(the) * math.cos(psi) + math.sin(phi) * math.cos(the) * math.sin(psi),
         math.cos(phi) * math.cos(the) * math.sin(psi) - math.sin(phi) * math.sin(the) * math.cos(psi),
         math.cos(phi) * math.cos(the) * math.cos(psi) + math.sin(phi) * math.sin(the) * math.sin(psi)]
    n = math.sqrt(q[0] ** 2 + q[1] ** 2 + q[2] ** 2 + q[3] ** 2)
    return [c / n for c in q]


def quaternion_from_matrix(m):
    """btMatrix3x3::getRotation."""
    trace = m[0][0] + m[1][1] + m[2][2]
    t = [0.0] * 4
    if trace > 0.0:
        s = math.sqrt(trace + 1.0)
        t[3] = s * 0.5
        s = 0.5 / s
        t[0] = (m[2][1] - m[1][2]) * s
        t[1] = (m[0][2] - m[2][0]) * s
        t[2] = (m[1][0] - m[0][1]) * s
    else:
        i = (2 if m[1][1] < m[2][2] else 1) if m[0][0] < m[1][1] else (2 if m[0][0] < m[2][2] else 0)
        j, k = (i + 1) % 3, (i + 2) % 3
        s = math.sqrt(m[i][i] - m[j][j] - m[k][k] + 1.0)
        t[i] = s * 0.5
        s = 0.5 / s
        t[3] = (m[k][j] - m[j][k]) * s
        t[j] = (m[j][i] + m[i][j]) * s
        t[k] = (m[k][i] + m[i][k]) * s
    return t


def pose_round_trip(q):
    """What getBasePositionAndOrientation returns for an orientation stored as ``q``:
    btTransform::setRotation(q) then btTransform::getRotation() -- i.e. normalised, sign-canonical."""
    return quaternion_from_matrix(matrix_from_quaternion(q))


def _matvec(m, v):
    return [m[0][0] * v[0] + m[0][1] * v[1] + m[0][2] * v[2],
            m[1][0] * v[0] + m[1][1] * v[1] + m[1][2] * v[2],
            m[2][0] * v[0] + m[2][1] * v[1] + m[2][2] * v[2]]


def _matTvec(m, v):
    return [m[0][0] * v[0] + m[1][0] * v[1] + m[2][0] * v[2],
            m[0][1] * v[0] + m[1][1] * v[1] + m[2][1] * v[2],
            m[0][2] * v[0] + m[1][2] * v[1] + m[2][2] * v[2]]


def _cross(a, b):
    return [a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]]


def free_body_step(pos, quat, vel, omega, force_w, torque_w, mass, inertia, dt,
                   gravity=(0.0, 0.0, -9.8), lin_damp=BULLET_LINEAR_DAMPING, ang_damp=BULLET_ANGULAR_DAMPING):
    """One btMultiBodyDynamicsWorld step of a floating base with no moving links and no contact.

    ``force_w`` / ``torque_w``: accumulated external wrench about the centre of mass, world frame
    (what addBaseForce/addLinkForce/addLinkTorque sum to). ``omega``: world-frame angular velocity
    (what getBaseVelocity reports). Sequence (btMultiBody::computeAccelerationsArticulatedBody-
    AlgorithmMultiDof for the base, then stepPositionsMultiDof):
      local wrench -> damping (k + k|v|) on both parts -> gyroscopic term -> solve -> rotate back;
      v += dt a ; w += dt alpha ; p += dt v_new ; q = exp(w_new dt) * q ; normalise.
    """
    R = matrix_from_quaternion(quat)           # base -> world
    f = [force_w[0] + mass * gravity[0], force_w[1] + mass * gravity[1], force_w[2] + mass * gravity[2]]
    f_l, t_l = _matTvec(R, f), _matTvec(R, torque_w)
    v_l, w_l = _matTvec(R, vel), _matTvec(R, omega)
    vn = math.sqrt(v_l[0] ** 2 + v_l[1] ** 2 + v_l[2] ** 2)
    wn = math.sqrt(w_l[0] ** 2 + w_l[1] ** 2 + w_l[2] ** 2)
    Jw = [inertia[0] * w_l[0], inertia[1] * w_l[1], inertia[2] * w_l[2]]
    gyro = _cross(w_l, Jw)
    kd_l = lin_damp + lin_damp * vn
    kd_a = ang_damp + ang_damp * wn
    # zero-acceleration force = -(ext) + damping + gyro (+ m w x v, cancelled when mapped back to world)
    a_l = [(f_l[i] - mass * v_l[i] * kd_l) / mass for i in range(3)]
    al_l = [(t_l[i] - Jw[i] * kd_a - gyro[i]) / inertia[i] for i in range(3)]
    a_w, al_w = _matvec(R, a_l), _matvec(R, al_l)
    vel = [vel[i] + dt * a_w[i] for i in range(3)]
    omega = [omega[i] + dt * al_w[i] for i in range(3)]
    pos = [pos[i] + dt * vel[i] for i in range(3)]
    # exponential map (btMultiBody::stepPositionsMultiDof functor, base body branch)
    ang = math.sqrt(omega[0] ** 2 + omega[1] ** 2 + omega[2] ** 2)
    if ang * dt > ANGULAR_MOTION_THRESHOLD:
        ang = 0.5 * (math.pi / 2) / dt
    if ang < 0.001:
        k = 0.5 * dt - (dt * dt * dt) * 0.020833333333 * ang * ang
    else:
        k = math.sin(0.5 * ang * dt) / ang
    ax = [omega[0] * k, omega[1] * k, omega[2] * k]
    cw = math.cos(ang * dt * 0.5)
    x, y, z, w = quat
    # dq * q  (Hamilton product, xyzw)
    nq = [cw * x + ax[0] * w + ax[1] * z - ax[2] * y,
          cw * y + ax[1] * w + ax[2] * x - ax[0] * z,
          cw * z + ax[2] * w + ax[0] * y - ax[1] * x,
          cw * w - ax[0] * x - ax[1] * y - ax[2] * z]
    n = math.sqrt(nq[0] ** 2 + nq[1] ** 2 + nq[2] ** 2 + nq[3] ** 2)
    quat = [c / n for c in nq]
    return pos, quat, vel, omega
