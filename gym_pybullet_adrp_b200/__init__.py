"""B200-native batched implementation of gym-pybullet-adrp's per-control-step drone physics.

The hot path is hand-written CUDA (csrc/) behind the C ABI of include/adrp.h; this package is the
host-side mirror of the reference's env / controller interface on top of it.
"""
__version__ = "0.1.0"
