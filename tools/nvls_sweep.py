"""Tuning sweep of the switch exchange kernel under torchrun: total CTAs x loads in flight per thread, three sizes.
The knobs (MERL_NVLS_GRID, MERL_NVLS_UNROLL) are read when a buffer is created, so one process walks all of them."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from mitsuba_customization_b200 import distributed as D

world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", 1), ("RANK", 0), ("LOCAL_RANK", 0)))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
comm = D.Comm(rank, world, local)
sizes = [1 << 20, 5832000, 1 << 24]
cap = max(sizes)

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

ref = D.PeerBuffer(comm, cap, "nccl")
line = "nccl registered".ljust(24) + "".join(f"{timed(lambda: ref.allreduce_(count=n)) * 1e3:9.1f}" for n in sizes)
if rank == 0: print("transport / grid x unroll".ljust(24) + "".join(f"{4 * n / 1e6:7.1f}MB" for n in sizes) + "   (us)\n" + line, flush=True)
ref.close()
for unroll in (2, 4, 8):
    for grid in (8, 16, 32, 64, 148):
        os.environ["MERL_NVLS_GRID"], os.environ["MERL_NVLS_UNROLL"] = str(grid), str(unroll)
        pb = D.PeerBuffer(comm, cap, "nvls")
        line = f"nvls {grid:4d} x {unroll}".ljust(24) + "".join(f"{timed(lambda: pb.allreduce_(count=n)) * 1e3:9.1f}" for n in sizes)
        ok = pb.status() == 0
        pb.close()
        if rank == 0: print(line + ("" if ok else "   STATUS != 0"), flush=True)
comm.close()
dist.destroy_process_group()
