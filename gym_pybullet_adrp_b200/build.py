"""In-tree build of the CUDA library: nvcc, sm_100a only, -lineinfo.

    python -m gym_pybullet_adrp_b200.build [--force]

Produces gym_pybullet_adrp_b200/libadrp_b200.so (git-ignored; travels to the GPU box with the snapshot).
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
OUT = os.path.join(PKG_DIR, "libadrp_b200.so")
# translation unit -> extra flags.  The fp64 parity build contracts no multiply-add (the oracle is compiled
# with -ffp-contract=off); the fp32 product build keeps nvcc's default FMA contraction.
SOURCES = {"adrp_api.cu": [], "adrp_kernels_f32.cu": [], "adrp_race_f32_phy16.cu": [], "adrp_race_f32_phy30.cu": [],
           "adrp_step_f32_plain.cu": [], "adrp_step_f32_forces.cu": [],
           "adrp_race_f32_phy16_multi.cu": [], "adrp_race_f32_phy30_multi.cu": [],
           "adrp_race_f32_phy16_dense.cu": [], "adrp_race_f32_phy30_dense.cu": [],
           "adrp_step_f32_plain_multi.cu": [], "adrp_step_f32_forces_multi.cu": [],
           "adrp_kernels_f64.cu": ["-fmad=false"]}
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "--cudart", "shared"]
OBJ_DIR = os.path.join(PKG_DIR, "build")


def _stale():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(os.path.dirname(PKG_DIR), "include", "adrp.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not _stale():
        return OUT
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found: cannot build libadrp_b200.so")
    os.makedirs(OBJ_DIR, exist_ok=True)
    procs = []
    for src, extra in SOURCES.items():
        obj = os.path.join(OBJ_DIR, src.replace(".cu", ".o"))
        cmd = [nvcc] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", "-o", obj, src]
        procs.append((obj, subprocess.Popen(cmd, cwd=CSRC, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)))
    objs = []
    for obj, pr in procs:
        out, err = pr.communicate()
        if pr.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + out + err)
        if verbose:
            print(err)
        objs.append(obj)
    res = subprocess.run([nvcc, "-shared", "--cudart", "shared", "-o", OUT] + objs, cwd=CSRC, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc link failed:\n" + res.stdout + res.stderr)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
