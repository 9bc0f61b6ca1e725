"""Exchange-step microbenchmark under torchrun: in-place sum of a float buffer over the ranks, by size and transport.
  peer             merl_peerbuf_allreduce, MERL_XCHG_PEER (one kernel over NVLink peer memory)
  nvls             merl_peerbuf_allreduce, MERL_XCHG_NVLS (own multimem.ld_reduce / multimem.st kernel through the switch)
  nccl_registered  merl_peerbuf_allreduce, MERL_XCHG_NCCL (ncclMemAlloc + symmetric window)
  nccl_plain       merl_allreduce_sum_f32 on a torch tensor (round 1's form)
Prints per size: ms and bus bandwidth 2 (R-1)/R bytes / time, max over ranks."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from mitsuba_customization_b200 import distributed as D

world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", 1), ("RANK", 0), ("LOCAL_RANK", 0)))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
comm = D.Comm(rank, world, local)
sizes = [1 << 18, 1 << 20, 1 << 22, 5832000, 1 << 24, 1 << 26]   # floats: 1 MB ... 256 MB; 5 832 000 = one gradient table
only = os.environ.get("XCHG_ONLY", "")            # e.g. XCHG_ONLY=nvls: that transport alone, three sizes (tuning sweeps)
if only: sizes = [1 << 20, 5832000, 1 << 24]
cap = max(sizes)
bufs = {} if only else {"peer": D.PeerBuffer(comm, cap, "peer"), "nccl_registered": D.PeerBuffer(comm, cap, "nccl")}
try:
    bufs["nvls"] = D.PeerBuffer(comm, cap, "nvls")
except Exception as e:
    if rank == 0: print("nvls unavailable:", e, flush=True)
plain = torch.zeros(cap, device=dev)

def timed(fn, it=20):
    for _ in range(5): fn()
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / it], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)

rows = []
for n in sizes:
    row = {"floats": n, "bytes": 4 * n}
    for name, pb in bufs.items():
        ms = timed(lambda: pb.allreduce_(count=n))
        row[name + "_ms"] = ms
        row[name + "_bus_gbs"] = 2 * (world - 1) / world * 4 * n / (ms * 1e-3) / 1e9
    if only:
        rows.append(row)
        if rank == 0: print(f"{4 * n / 1e6:8.1f} MB  " + "  ".join(f"{k[:-3]} {v * 1e3:7.1f} us" for k, v in row.items() if k.endswith("_ms")), flush=True)
        continue
    ms = timed(lambda: comm.allreduce_sum_(plain[:n]))
    row["nccl_plain_ms"] = ms
    row["nccl_plain_bus_gbs"] = 2 * (world - 1) / world * 4 * n / (ms * 1e-3) / 1e9
    rows.append(row)
    if rank == 0:
        nv = f"   nvls {row['nvls_ms'] * 1e3:8.1f} us ({row['nvls_bus_gbs']:6.0f})" if "nvls_ms" in row else ""
        print(f"{4 * n / 1e6:8.1f} MB  peer {row['peer_ms'] * 1e3:8.1f} us ({row['peer_bus_gbs']:6.0f} GB/s){nv}   nccl registered "
              f"{row['nccl_registered_ms'] * 1e3:8.1f} us ({row['nccl_registered_bus_gbs']:6.0f})   nccl plain {row['nccl_plain_ms'] * 1e3:8.1f} us "
              f"({row['nccl_plain_bus_gbs']:6.0f})", flush=True)
# correctness at the largest size
for name, pb in bufs.items():
    t = pb.tensor((cap,))
    t.fill_(float(rank + 1)); pb.allreduce_()
    torch.cuda.synchronize()
    ok = float(t.min()) == float(t.max()) == world * (world + 1) / 2 and pb.status() == 0
    if rank == 0: print(f"{name}: sum check {'ok' if ok else 'FAILED'} ({pb.registration})")
if rank == 0: print(json.dumps({"world": world, "rows": rows}))
for pb in bufs.values(): pb.close()
comm.close()
dist.destroy_process_group()
