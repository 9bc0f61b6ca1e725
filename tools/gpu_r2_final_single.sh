#!/bin/bash
# Final 1-GPU record of the round on the final build: smoke(), full GPU suite, bench line, reference arm
set -u
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
python - <<PY
import json
b=json.loads([l for l in open("gpurun_out/bench.json") if l.startswith('{"metric"')][-1])
print(b["value"], b["roofline"]["frac"], b["roofline"].get("frac_burst"), b["e2e"]["value"], b["cpu_baseline"].get("value"), b["gpu_launches"])
print({k:(v.get("ms"), v.get("spp_mpix_per_s")) for k,v in b["extras"]["render"].items()})
print(b["extras"].get("sample_states"))
PY
