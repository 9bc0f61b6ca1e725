#!/bin/bash
# A/B on one box: library before the float|float2 refactor (old) / refactored scalar / refactored packed
set -u
mkdir -p gpurun_out
V=$PWD/mitsuba_customization_b200/variants
for k in eval_pdf eval bins sample; do
  for r in 1 2 3; do
    echo "old    $k: $(MERL_B200_LIB=$V/libmerl_old.so timeout 200 python tools/one_kernel.py $k 26 20 2>&1 | tail -1)"
    echo "scalar $k: $(MERL_B200_LIB=$V/libmerl_scalar.so timeout 200 python tools/one_kernel.py $k 26 20 2>&1 | tail -1)"
    echo "packed $k: $(timeout 200 python tools/one_kernel.py $k 26 20 2>&1 | tail -1)"
  done
done | tee gpurun_out/ab_packed3.log
