#!/bin/bash
mkdir -p gpurun_out
for m in 3 4; do echo "== sample/backward MINB=$m"; MERL_B200_LIB=$PWD/build/libmerl_sb$m.so timeout 300 python tools/quick_perf.py 26 2>&1 | grep -E "^(sample_search0|backward)"; done | tee gpurun_out/quick5.log
