"""CPU: "the generating script reproduces the vectors".  Runs every mode of tests/golden/make_golden.py and
make_golden_race.py (the reference's own Python over oracle/refstub) into a scratch directory and compares each
committed tests/golden/*.npz with what the generators produce today, array by array, bit for bit.

Needs the reference tree (build container only); skipped where /root/reference is absent (the GPU box)."""
import glob
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
REF = os.environ.get("ADRP_REFERENCE_ROOT", "/root/reference")
MODES = [("make_golden.py", []), ("make_golden.py", ["--new"]), ("make_golden.py", ["--r2"]),
         ("make_golden_race.py", []), ("make_golden_race.py", ["--new"]), ("make_golden_race.py", ["--fullstate"]),
         ("make_golden_race.py", ["--commander"]), ("make_golden_race.py", ["--levels"])]


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "gym_pybullet_adrp")), reason="reference tree not present")
def test_generators_reproduce_every_committed_fixture(tmp_path):
    env = dict(os.environ, ADRP_GOLDEN_OUT=str(tmp_path), ADRP_REFERENCE_ROOT=REF)
    procs = [subprocess.Popen([sys.executable, os.path.join(GOLDEN, script)] + args, env=env, stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for script, args in MODES]
    for (script, args), p in zip(MODES, procs):
        out, _ = p.communicate(timeout=900)
        assert p.returncode == 0, f"{script} {args} failed:\n{out[-2000:]}"
    committed = sorted(os.path.basename(f) for f in glob.glob(os.path.join(GOLDEN, "*.npz")))
    produced = sorted(os.path.basename(f) for f in glob.glob(os.path.join(str(tmp_path), "*.npz")))
    assert committed == produced, (set(committed) ^ set(produced))
    for name in committed:
        a, b = np.load(os.path.join(GOLDEN, name)), np.load(os.path.join(str(tmp_path), name))
        assert sorted(a.files) == sorted(b.files), name
        for k in a.files:
            x, y = a[k], b[k]
            assert x.dtype == y.dtype and x.shape == y.shape, (name, k)
            assert x.tobytes() == y.tobytes(), f"{name}:{k} is not what its generator produces"
