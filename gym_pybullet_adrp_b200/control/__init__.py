"""Controllers with the reference's names (gym_pybullet_adrp/control/__init__.py).  MellingerControl has no host
class: it runs per physics substep inside the race env's step kernel (csrc/adrp_mellinger.cuh)."""
from .BaseControl import BaseControl
from .DSLPIDControl import DSLPIDControl

__all__ = ["BaseControl", "DSLPIDControl"]
