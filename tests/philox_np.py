"""Philox4x32-10 (Salmon, Moraes, Dror, Shaw: "Parallel random numbers: as easy as 1, 2, 3", SC'11) in NumPy, vectorised over the
counter words -- the checker of the library's in-kernel draws (csrc/adrp_math.cuh:philox4x32; oracle/adrp_oracle_race.h).
Test infrastructure only."""
import numpy as np

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = 0x9E3779B9, 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)


def philox4x32(c0, c1, c2, c3, k0, k1):
    """Counter words (broadcastable integer arrays) and key words (ints) -> four uint32 arrays."""
    c0, c1, c2, c3 = np.broadcast_arrays(*[np.asarray(c, np.uint64) & MASK for c in (c0, c1, c2, c3)])
    k0, k1 = int(k0) & 0xFFFFFFFF, int(k1) & 0xFFFFFFFF
    for _ in range(10):
        p0, p1 = M0 * c0, M1 * c2
        c0, c1, c2, c3 = (p1 >> np.uint64(32)) ^ c1 ^ np.uint64(k0), p1 & MASK, (p0 >> np.uint64(32)) ^ c3 ^ np.uint64(k1), p0 & MASK
        k0, k1 = (k0 + W0) & 0xFFFFFFFF, (k1 + W1) & 0xFFFFFFFF
    return [c.astype(np.uint32) for c in (c0, c1, c2, c3)]


def random_policy_actions(seed, env_offset, N, M, A, step0, K):
    """The (K, N, M, A) actions of adrp_step_n's ADRP_ACTIONS_PHILOX (include/adrp.h): component k of drone i of env n at
    rollout step s is ((x_k >> 8) - 2^23) * 2^-23 with x = Philox4x32-10(key = seed; counter = (env_offset + n, 192 + i, s, 0))."""
    s = (step0 + np.arange(K, dtype=np.uint64))[:, None, None]
    n = (env_offset + np.arange(N, dtype=np.uint64))[None, :, None]
    i = (192 + np.arange(M, dtype=np.uint64))[None, None, :]
    o = philox4x32(n, i, s, 0, seed & 0xFFFFFFFF, seed >> 32)
    a = np.stack([((x >> np.uint32(8)).astype(np.int64) - (1 << 23)) * 2.0 ** -23 for x in o], axis=-1)
    return a[..., :A]
