#!/bin/bash
# A/B of tuning builds (mitsuba_customization_b200/variants/*.so, selected with MERL_B200_LIB): the named kernel's
# time on 2^26 queries for the main build and every variant, then the sampler parity tests on every variant.
# usage: gpu_r2_ab.sh <kernel> [pytest -k expression]
set -u
K=${1:-sample}; EXPR=${2:-sample}
mkdir -p gpurun_out
echo "== base" | tee gpurun_out/ab_$K.log
for r in 1 2; do timeout 200 python tools/one_kernel.py $K 26 20 2>&1 | tail -1 | tee -a gpurun_out/ab_$K.log; done
for so in mitsuba_customization_b200/variants/*.so; do
  [ -e "$so" ] || continue
  echo "== $so" | tee -a gpurun_out/ab_$K.log
  for r in 1 2; do MERL_B200_LIB=$PWD/$so timeout 200 python tools/one_kernel.py $K 26 20 2>&1 | tail -1 | tee -a gpurun_out/ab_$K.log; done
  MERL_B200_LIB=$PWD/$so timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -m gpu -x -q -k "$EXPR" 2>&1 | tail -2 | tee -a gpurun_out/ab_$K.log
done
