"""TEST INFRASTRUCTURE -- stand-in for ``pybullet_data`` (only ``plane.urdf`` is ever asked for)."""


def getDataPath():
    return "/pybullet_data"
