#!/bin/bash
# Does the bench context slow one rank's share of config 5 (33.5 M paths, one pass) on ONE GPU too?
set -u
mkdir -p gpurun_out
RD_EMULATE_WORLD=8 timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_emu8.json 2> gpurun_out/bench_emu8.err; echo "bench rc=$?"
python - <<PY
import json
b=json.loads([l for l in open("gpurun_out/bench_emu8.json") if l.startswith('{"metric"')][-1])
for k,v in b["extras"]["render"].items(): print("in bench:", k, v.get("ms"))
PY
for c in cfg2 "cfg5 100"; do echo "standalone $c: $(RD_EMULATE_WORLD=8 timeout 300 python tools/render_bench.py $c 2>/dev/null | tail -1 | cut -c1-160)"; done
