"""Utilities: enums, constants and config helpers of the reference's API surface."""
from .enums import (ActionType, Command, DroneModel, ImageType, ObservationType, Physics,  # noqa: F401
                    RaceMode, State)
