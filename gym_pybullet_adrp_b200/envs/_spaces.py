"""``spaces.Box`` for the env classes: gymnasium's when it is installed, else a minimal stand-in with the
attributes callers of the reference read (low, high, shape, dtype, sample, contains)."""
from __future__ import annotations

import numpy as np

try:                                            # pragma: no cover - gymnasium is absent on the build box
    from gymnasium import spaces                # type: ignore
    Box = spaces.Box
except Exception:                               # noqa: BLE001
    class Box:
        def __init__(self, low, high, shape=None, dtype=np.float32):
            dtype = np.dtype(np.float64 if dtype is float else dtype)
            if shape is None:
                shape = np.broadcast(np.asarray(low), np.asarray(high)).shape
            self.low = np.broadcast_to(np.asarray(low, dtype=dtype), shape).copy()
            self.high = np.broadcast_to(np.asarray(high, dtype=dtype), shape).copy()
            self.shape, self.dtype = tuple(shape), dtype
            self._rng = np.random.default_rng()

        def seed(self, seed=None):
            self._rng = np.random.default_rng(seed)

        def sample(self):
            lo = np.where(np.isfinite(self.low), self.low, -1e6)
            hi = np.where(np.isfinite(self.high), self.high, 1e6)
            return self._rng.uniform(lo, hi).astype(self.dtype)

        def contains(self, x):
            x = np.asarray(x)
            return x.shape == self.shape and bool(np.all(x >= self.low) and np.all(x <= self.high))

        def __repr__(self):
            return f"Box({self.low.min()}, {self.high.max()}, {self.shape}, {self.dtype})"
