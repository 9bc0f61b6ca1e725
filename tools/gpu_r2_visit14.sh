#!/bin/bash
# A/B of the packed fp32x2 bin math (main build) against the scalar build (variants/libmerl_scalar.so):
# parity suite on the main build, every query kernel at 2^26 on both, and the sustained state of eval+pdf on both
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
S=$PWD/mitsuba_customization_b200/variants/libmerl_scalar.so
for k in eval_pdf eval pdf bins sample backward; do
  for r in 1 2; do
    echo "packed $k: $(timeout 200 python tools/one_kernel.py $k 26 20 2>&1 | tail -1)"
    echo "scalar $k: $(MERL_B200_LIB=$S timeout 200 python tools/one_kernel.py $k 26 20 2>&1 | tail -1)"
  done
done | tee gpurun_out/ab_packed.log
timeout 300 python tools/sustained_probe.py eval_pdf 600 50 > gpurun_out/sustained_packed.log 2>&1; tail -5 gpurun_out/sustained_packed.log | cut -c1-160
MERL_B200_LIB=$S timeout 300 python tools/sustained_probe.py eval_pdf 600 50 > gpurun_out/sustained_scalar.log 2>&1; tail -5 gpurun_out/sustained_scalar.log | cut -c1-160
