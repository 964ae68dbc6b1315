"""TEST INFRASTRUCTURE -- ``gymnasium.spaces.Box`` stand-in (shape/dtype/bounds only)."""
import numpy as np


class Box:
    def __init__(self, low, high, shape=None, dtype=np.float32):
        self.low, self.high, self.dtype = np.asarray(low), np.asarray(high), dtype
        self.shape = tuple(shape) if shape is not None else self.low.shape
