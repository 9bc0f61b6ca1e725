#!/bin/bash
# N-GPU visit: exchange microbenchmark (peer / nvls / nccl), NCCL's own algorithm choice, then the full bench line
set -u
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513"
timeout 300 $TR tools/exchange_bench.py > gpurun_out/exchange_n$N.log 2> gpurun_out/exchange_n$N.err; echo "exchange rc=$?"
grep -v "^{" gpurun_out/exchange_n$N.log | cut -c1-260
MERL_NVLS_CTAS_PER_SM=2 timeout 300 $TR tools/exchange_bench.py > gpurun_out/exchange_n${N}_nvls2.log 2>&1; echo "exchange(2 CTAs/SM) rc=$?"
grep "MB" gpurun_out/exchange_n${N}_nvls2.log | cut -c1-260
NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,COLL,TUNING timeout 300 $TR tools/exchange_bench.py > gpurun_out/exchange_n${N}_ncclinfo.log 2>&1
grep -i "symmetric\|nvls\|algo" gpurun_out/exchange_n${N}_ncclinfo.log | grep -v "^\s*[0-9.]* MB" | sort | uniq -c | sort -rn | head -40 | cut -c1-260 > gpurun_out/nccl_info_n$N.txt
head -12 gpurun_out/nccl_info_n$N.txt
rm -f gpurun_out/exchange_n${N}_ncclinfo.log
timeout 900 $TR bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"
tail -3 gpurun_out/bench_n$N.err | cut -c1-300
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
