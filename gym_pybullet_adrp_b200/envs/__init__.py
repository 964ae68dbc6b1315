"""Env classes with the reference's names (gym_pybullet_adrp/envs/__init__.py), batched over ``num_envs``."""
from .BaseAviary import BaseAviary
from .BaseRLAviary import BaseRLAviary
from .CtrlAviary import CtrlAviary
from .HoverAviary import HoverAviary
from .MultiHoverAviary import MultiHoverAviary
from .MultiRaceAviary import MultiRaceAviary
from .VelocityAviary import VelocityAviary

__all__ = ["BaseAviary", "BaseRLAviary", "CtrlAviary", "HoverAviary", "MultiHoverAviary", "MultiRaceAviary", "VelocityAviary"]
