#!/bin/bash
# A/B visit: parity tests on the main build, then per-kernel timings of the main build and of
# every tuning build under mitsuba_customization_b200/variants/ (selected with MERL_B200_LIB).
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_gpu.log
echo "== base" | tee gpurun_out/ab.log
timeout 300 python tools/quick_perf.py 26 2>&1 | grep -v "^{" | tee -a gpurun_out/ab.log
for so in mitsuba_customization_b200/variants/*.so; do
  [ -e "$so" ] || continue
  echo "== $so" | tee -a gpurun_out/ab.log
  MERL_B200_LIB=$PWD/$so timeout 300 python tools/quick_perf.py 26 2>&1 | grep -v "^{" | grep -v "copy_40B\|torch_gather\|search1" | tee -a gpurun_out/ab.log
done
