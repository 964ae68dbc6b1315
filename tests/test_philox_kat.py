"""CPU: the NumPy Philox4x32-10 of tests/philox_np.py against the known-answer vectors of the Random123 distribution
(kat_vectors, philox4x32 10 rounds), so that the GPU tests which use it as their checker stand on a pinned generator."""
import numpy as np

from tests.philox_np import philox4x32, random_policy_actions

KAT = [((0x00000000, 0x00000000, 0x00000000, 0x00000000), (0x00000000, 0x00000000), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
       ((0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff), (0xffffffff, 0xffffffff), (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
       ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0), (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]


def test_known_answers():
    for ctr, key, want in KAT:
        got = philox4x32(*ctr, *key)
        assert tuple(int(g) for g in got) == want, [hex(int(g)) for g in got]


def test_vectorised_equals_scalar_and_actions_are_in_range():
    c = np.arange(7, dtype=np.uint64)
    vec = philox4x32(c, 192, 5, 0, 1337, 0)
    for j in range(7):
        one = philox4x32(int(c[j]), 192, 5, 0, 1337, 0)
        assert [int(v[j]) for v in vec] == [int(o) for o in one]
    a = random_policy_actions(1337, 40, 64, 3, 4, 0, 50)
    assert a.shape == (50, 64, 3, 4) and a.min() >= -1.0 and a.max() < 1.0
    assert abs(a.mean()) < 0.02 and abs(a.std() - 1 / np.sqrt(3)) < 0.02
    assert np.array_equal(a.astype(np.float32).astype(np.float64), a)          # exact in fp32: the same draw in both builds
    # a shard draws the rows of the whole batch (counter word 0 is the GLOBAL env index)
    assert np.array_equal(random_policy_actions(1337, 40 + 16, 8, 3, 4, 0, 50), a[:, 16:24])
