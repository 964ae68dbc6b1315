"""TEST INFRASTRUCTURE -- a stand-in ``pybullet`` module, just big enough to run the reference's
own Python step code (BaseAviary / RL aviaries / MultiRaceAviary) in a container without PyBullet.

It is only ever imported by ``tests/golden/make_golden.py`` (with this directory prepended to
``sys.path``) to produce the committed golden vectors.  Everything Bullet would compute is delegated
to the declared restatements in ``oracle/bullet_model.py`` (one free body, default damping, no
contact impulses) and ``oracle/track_geometry.py`` (rays / contacts / closest points).  Bodies other
than drones are static.  Function names, argument names and return layouts follow the PyBullet API
as used at the call sites listed in ``oracle/bullet_model.py``.
"""
from __future__ import annotations

import math
import os
import sys
import xml.etree.ElementTree as ET

_HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(_HERE)))
from oracle import bullet_model as bm        # noqa: E402
from oracle import track_geometry as tg      # noqa: E402

GUI, DIRECT = 1, 2
LINK_FRAME, WORLD_FRAME = 1, 2
URDF_USE_INERTIA_FROM_FILE = 2
COV_ENABLE_RGB_BUFFER_PREVIEW = COV_ENABLE_DEPTH_BUFFER_PREVIEW = COV_ENABLE_SEGMENTATION_MARK_PREVIEW = 0
STATE_LOGGING_VIDEO_MP4 = 3

#: set by the golden generator: 'simple' (declared ground model) or 'none'
GROUND_MODEL = "simple"
GROUND_FRICTION = 0.5


class _World:
    def __init__(self):
        self.reset()

    def reset(self):
        self.bodies = []
        self.gravity = (0.0, 0.0, 0.0)
        self.dt = 1.0 / 240.0
        self.no_collide = set()


_W = _World()


def connect(mode, options=None):
    return 0


def disconnect(physicsClientId=0):
    _W.reset()


def resetSimulation(physicsClientId=0):
    _W.reset()


def setGravity(x, y, z, physicsClientId=0):
    _W.gravity = (float(x), float(y), float(z))


def setRealTimeSimulation(flag, physicsClientId=0):
    pass


def setTimeStep(dt, physicsClientId=0):
    _W.dt = float(dt)


def setAdditionalSearchPath(path, physicsClientId=0):
    pass


def configureDebugVisualizer(*a, **k):
    pass


getMatrixFromQuaternion = lambda q, physicsClientId=0: tuple(v for row in bm.matrix_from_quaternion([float(c) for c in q]) for v in row)  # noqa: E731
getEulerFromQuaternion = lambda q, physicsClientId=0: tuple(bm.euler_from_quaternion([float(c) for c in q]))  # noqa: E731
getQuaternionFromEuler = lambda e, physicsClientId=0: tuple(bm.quaternion_from_euler([float(c) for c in e]))  # noqa: E731


def _floats(s):
    return [float(v) for v in s.split()]


def _parse_urdf(path):
    root = ET.parse(path).getroot()
    links = []
    for link in root.findall("link"):
        ine = link.find("inertial")
        mass, inertia, origin = 0.0, (0.0, 0.0, 0.0), (0.0, 0.0, 0.0)
        if ine is not None:
            mass = float(ine.find("mass").attrib["value"])
            i = ine.find("inertia").attrib
            inertia = (float(i["ixx"]), float(i["iyy"]), float(i["izz"]))
            o = ine.find("origin")
            if o is not None:
                origin = tuple(_floats(o.attrib.get("xyz", "0 0 0")))
        shapes = []
        for col in link.findall("collision"):
            o = col.find("origin")
            xyz = _floats(o.attrib.get("xyz", "0 0 0")) if o is not None else [0.0, 0.0, 0.0]
            rpy = _floats(o.attrib.get("rpy", "0 0 0")) if o is not None else [0.0, 0.0, 0.0]
            g = col.find("geometry")
            if g.find("box") is not None:
                sx, sy, sz = _floats(g.find("box").attrib["size"])
                if abs(rpy[1] - math.pi / 2) < 1e-2:          # upright edge (E20: taken as exactly pi/2)
                    sx, sz = sz, sx
                shapes.append(("box", xyz, (sx / 2, sy / 2, sz / 2)))
            elif g.find("cylinder") is not None:
                c = g.find("cylinder").attrib
                shapes.append(("cyl", xyz, (float(c["radius"]), float(c["length"]) / 2)))
        links.append(dict(name=link.attrib["name"], mass=mass, inertia=inertia, origin=origin, shapes=shapes))
    return links


def loadURDF(fileName, basePosition=(0, 0, 0), baseOrientation=(0, 0, 0, 1), useFixedBase=0, flags=0,
             physicsClientId=0, globalScaling=1.0):
    name = os.path.basename(fileName)
    body = dict(name=name, pos=[float(v) for v in basePosition], quat=[float(v) for v in baseOrientation],
                vel=[0.0] * 3, omega=[0.0] * 3, force=[0.0] * 3, torque=[0.0] * 3, contacts=set())
    if name == "plane.urdf":
        body["kind"] = "plane"
    else:
        if not os.path.isabs(fileName) and not os.path.exists(fileName):
            fileName = os.path.join(os.environ.get("ADRP_REFERENCE_ROOT", "/root/reference"), fileName)
        links = _parse_urdf(fileName)
        base = links[0]
        body.update(mass=base["mass"], inertia=list(base["inertia"]), links=links[1:])
        if name.startswith(("cf2", "racer")):
            body["kind"] = "drone"
            cyl = base["shapes"][0]
            body["coll_r"], body["coll_hh"] = cyl[2]
        else:
            body["kind"] = "static"
            yaw = bm.euler_from_quaternion(body["quat"])[2]
            x, y, z = body["pos"]
            prisms = []
            for l in links:
                for kind, xyz, dims in l["shapes"]:
                    c, s = math.cos(yaw), math.sin(yaw)
                    cx, cy, cz = x + c * xyz[0] - s * xyz[1], y + s * xyz[0] + c * xyz[1], z + xyz[2]
                    if kind == "box":
                        prisms.append(("rect", cx, cy, yaw, dims[0], dims[1], cz, dims[2]))
                    else:
                        prisms.append(("circ", cx, cy, dims[0], cz, dims[1]))
            # portal.urdf / obstacle.urdf carry each collision element twice; duplicates are harmless
            body["prisms"] = prisms
    _W.bodies.append(body)
    return len(_W.bodies) - 1


def changeDynamics(bodyUniqueId, linkIndex, mass=None, localInertiaDiagonal=None, physicsClientId=0, **kw):
    b = _W.bodies[int(bodyUniqueId)]
    if mass is not None:
        b["mass"] = float(mass)
    if localInertiaDiagonal is not None:
        b["inertia"] = [float(v) for v in localInertiaDiagonal]


def setCollisionFilterPair(a, b, la, lb, enable, physicsClientId=0):
    if not enable:
        _W.no_collide.add((int(a), int(b)))
        _W.no_collide.add((int(b), int(a)))


def getBasePositionAndOrientation(bodyUniqueId, physicsClientId=0):
    b = _W.bodies[int(bodyUniqueId)]
    return tuple(b["pos"]), tuple(bm.pose_round_trip(b["quat"]))


def getBaseVelocity(bodyUniqueId, physicsClientId=0):
    b = _W.bodies[int(bodyUniqueId)]
    return tuple(b["vel"]), tuple(b["omega"])


def resetBasePositionAndOrientation(bodyUniqueId, posObj, ornObj, physicsClientId=0):
    b = _W.bodies[int(bodyUniqueId)]
    b["pos"] = [float(v) for v in posObj]
    b["quat"] = [float(v) for v in ornObj]


def resetBaseVelocity(objectUniqueId, linearVelocity=None, angularVelocity=None, physicsClientId=0):
    b = _W.bodies[int(objectUniqueId)]
    if linearVelocity is not None:
        b["vel"] = [float(v) for v in linearVelocity]
    if angularVelocity is not None:
        b["omega"] = [float(v) for v in angularVelocity]


def _link_origin(b, linkIndex):
    return (0.0, 0.0, 0.0) if linkIndex < 0 else b["links"][linkIndex]["origin"]


def applyExternalForce(objectUniqueId, linkIndex, forceObj, posObj, flags, physicsClientId=0):
    b = _W.bodies[int(objectUniqueId)]
    R = bm.matrix_from_quaternion(b["quat"])
    f = [float(v) for v in forceObj]
    pl = [float(v) for v in posObj]
    lo = _link_origin(b, linkIndex)
    arm_l = bm._matvec(R, list(lo))          # link frame origin relative to the base CoM, world axes
    if flags == LINK_FRAME:
        fw = bm._matvec(R, f)                # fixed links share the base orientation
        rel = bm._matvec(R, pl)
    else:
        fw = f
        link_w = [b["pos"][i] + arm_l[i] for i in range(3)]
        rel = [pl[i] - link_w[i] for i in range(3)]
    r = [arm_l[i] + rel[i] for i in range(3)]
    t = bm._cross(r, fw)
    for i in range(3):
        b["force"][i] += fw[i]
        b["torque"][i] += t[i]


def applyExternalTorque(objectUniqueId, linkIndex, torqueObj, flags, physicsClientId=0):
    b = _W.bodies[int(objectUniqueId)]
    t = [float(v) for v in torqueObj]
    if flags == LINK_FRAME:
        t = bm._matvec(bm.matrix_from_quaternion(b["quat"]), t)
    for i in range(3):
        b["torque"][i] += t[i]


def getLinkStates(bodyUniqueId, linkIndices, computeLinkVelocity=0, computeForwardKinematics=0, physicsClientId=0):
    b = _W.bodies[int(bodyUniqueId)]
    R = bm.matrix_from_quaternion(b["quat"])
    out = []
    for li in linkIndices:
        o = bm._matvec(R, list(_link_origin(b, li)))
        wp = tuple(b["pos"][i] + o[i] for i in range(3))
        out.append((wp, tuple(b["quat"]), (0, 0, 0), (0, 0, 0, 1), wp, tuple(b["quat"]), tuple(b["vel"]), tuple(b["omega"])))
    return out


def _drone_hz(b):
    R22 = bm.matrix_from_quaternion(b["quat"])[2][2]
    return tg.drone_half_height(R22, b["coll_r"], 2 * b["coll_hh"])


def ground_response(pos, vel, hz):
    """Declared plane contact (no Bullet equivalent is reproducible): push the lowest point back to
    z=0, zero the approaching normal velocity, Coulomb friction (mu=0.5) on the tangential velocity."""
    pen = hz - pos[2]
    if pen > 0.0:
        pos[2] += pen
        if vel[2] < 0.0:
            jn = -vel[2]
            vel[2] = 0.0
            vt = math.sqrt(vel[0] ** 2 + vel[1] ** 2)
            if vt > 0.0:
                k = max(0.0, 1.0 - GROUND_FRICTION * jn / vt)
                vel[0] *= k
                vel[1] *= k


def stepSimulation(physicsClientId=0):
    for b in _W.bodies:
        if b["kind"] != "drone":
            continue
        b["pos"], b["quat"], b["vel"], b["omega"] = bm.free_body_step(
            b["pos"], b["quat"], b["vel"], b["omega"], b["force"], b["torque"], b["mass"], b["inertia"],
            _W.dt, _W.gravity)
        b["force"], b["torque"] = [0.0] * 3, [0.0] * 3
        if GROUND_MODEL == "simple":
            ground_response(b["pos"], b["vel"], _drone_hz(b))


def _distance(a, b):
    A, B = _W.bodies[a], _W.bodies[b]
    if B["kind"] != "drone":
        A, B = B, A
    if B["kind"] != "drone":
        return math.inf
    hz, r = _drone_hz(B), B["coll_r"]
    px, py, pz = B["pos"]
    if A["kind"] == "plane":
        return tg.drone_plane_distance(pz, hz)
    if A["kind"] == "static":
        return tg.drone_body_distance(px, py, pz, hz, r, A["prisms"])
    if A is B:
        return math.inf
    return tg.drone_drone_distance(A["pos"], _drone_hz(A), B["pos"], hz, r)


def getContactPoints(bodyA=None, bodyB=None, physicsClientId=0):
    a, b = int(bodyA), int(bodyB)
    if (a, b) in _W.no_collide:
        return ()
    return ((0,),) if _distance(a, b) < bm.CONTACT_BREAKING_THRESHOLD else ()


def getClosestPoints(bodyA, bodyB, distance, physicsClientId=0):
    return ((0,),) if _distance(int(bodyA), int(bodyB)) < distance else ()


def rayTestBatch(rayFromPositions, rayToPositions, physicsClientId=0):
    out = []
    drones = [(i, b) for i, b in enumerate(_W.bodies) if b["kind"] == "drone"]
    for fr, to in zip(rayFromPositions, rayToPositions):
        d = [to[i] - fr[i] for i in range(3)]
        best, best_id = None, -1
        for i, b in drones:
            R = bm.matrix_from_quaternion(b["quat"])
            t = tg.ray_cylinder(fr, d, b["pos"], [R[0][2], R[1][2], R[2][2]], b["coll_r"], b["coll_hh"])
            if t is not None and (best is None or t < best):
                best, best_id = t, i
        out.append((best_id, -1, 1.0 if best is None else best, (0, 0, 0), (0, 0, 0)))
    return out
