"""Controller interface of ``scripts/sim.py`` (user_controller/__init__.py)."""
from .BaseController import BaseController
from .HoverController import HoverController

__all__ = ["BaseController", "HoverController"]
