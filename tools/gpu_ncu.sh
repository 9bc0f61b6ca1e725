#!/bin/bash
# ncu --set full capture of one launch of each named kernel (after its plain run exited 0).
# usage: gpu_ncu.sh eval_pdf sample ...
set -u
mkdir -p gpurun_out
for k in "$@"; do
  timeout 300 python tools/one_kernel.py $k 26 5 || { echo "$k plain run failed"; continue; }
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:stream_kernel -s 3 -c 1 \
    -o gpurun_out/${k}_full -f python tools/one_kernel.py $k 26 3 > gpurun_out/ncu_${k}.log 2>&1; echo "ncu $k rc=$?"
done
