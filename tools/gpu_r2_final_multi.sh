#!/bin/bash
# Final N-GPU record of the round: exchange table by size and transport, then the bench line (both arms)
set -u
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29515"
timeout 300 $TR tools/exchange_bench.py > gpurun_out/exchange_n$N.log 2> gpurun_out/exchange_n$N.err; echo "exchange rc=$?"
grep -v "^{" gpurun_out/exchange_n$N.log | cut -c1-260
timeout 900 $TR bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"
timeout 300 $TR bench.py --impl reference --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err; echo "ref rc=$?"
grep '^{' gpurun_out/bench_ref_n$N.json | tail -1 | cut -c1-400
python - <<PY
import json
b=json.loads([l for l in open("gpurun_out/bench_n$N.json") if l.startswith('{"metric"')][-1])
print(b["value"], b["roofline"]["frac"], b["e2e"]["value"], b["e2e"].get("host_link_ceiling",{}).get("value"))
g=b["extras"].get("grad_scaling",{})
for k,v in g.get("transports",{}).items(): print(k, {a:round(c,4) for a,c in v.items() if isinstance(c,float)})
print(g.get("config4_2p26_total")); print(g.get("allreduce_check"), g.get("nvls_unavailable"))
r=b["extras"].get("render",{})
for k,v in r.items(): print(k, v.get("ms"), v.get("spp_mpix_per_s"), v.get("ratio_to_sample_sharding"))
PY
