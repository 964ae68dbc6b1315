#!/bin/bash
# Builds a VARIANT of libadrp_b200.so with extra preprocessor flags (launch bounds, block sizes: ADRP_STEP_BLOCK, ADRP_STEP_MINB,
# ADRP_RACE_BLOCK, ADRP_RACE_MINB, ADRP_RACE_MAXNREG, ADRP_RING_BATCH) into gym_pybullet_adrp_b200/variants/<name>/, for
# measurements with ADRP_LIB_PATH=<that file> (gym_pybullet_adrp_b200/_lib.py).  The product build is build.py's, untouched.
#   tools/variant_build.sh <name> -DADRP_STEP_MINB=7 ...
set -e
NAME=$1; shift
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
SRC=$ROOT/gym_pybullet_adrp_b200/csrc
OUT=$ROOT/gym_pybullet_adrp_b200/variants/$NAME
mkdir -p $OUT
cd $SRC
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --cudart shared"
pids=()
for f in adrp_api adrp_kernels_f32 adrp_race_f32_phy16 adrp_race_f32_phy30 adrp_step_f32_plain adrp_step_f32_forces \
         adrp_race_f32_phy16_multi adrp_race_f32_phy30_multi adrp_race_f32_phy16_dense adrp_race_f32_phy30_dense adrp_step_f32_plain_multi adrp_step_f32_forces_multi; do
  nvcc $FLAGS "$@" -c -o $OUT/$f.o $f.cu & pids+=($!)
done
nvcc $FLAGS -fmad=false "$@" -c -o $OUT/adrp_kernels_f64.o adrp_kernels_f64.cu & pids+=($!)
for p in "${pids[@]}"; do wait $p; done
nvcc -shared --cudart shared -o $OUT/libadrp_b200.so $OUT/*.o
rm -f $OUT/*.o
echo $OUT/libadrp_b200.so
