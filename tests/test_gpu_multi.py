"""Two-GPU check of the exchange step: sharded backward + merl_allreduce_sum_f32 (NCCL) equals the
single-GPU gradient.  Skipped on boxes with fewer than two GPUs."""
import os
import socket
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from mitsuba_customization_b200 import bsdf, distributed as D, synth
        plug = bsdf.MERL(data=synth.synthetic_table(0), device=f"cuda:{rank}")
        comm = D.Comm(rank, world, rank)
        n = (1 << 20) + 12
        wi, wo = synth.random_queries(n, seed=5)
        g = np.random.default_rng(9).standard_normal((3, n)).astype(np.float32)
        lo, hi = D.shard_bounds(n, world, rank)
        dev = torch.device("cuda", rank)
        cut = lambda a: torch.from_numpy(np.ascontiguousarray(a[:, lo:hi])).to(dev)
        grad = D.sharded_backward(plug, comm, cut(wi), cut(wo), cut(g))
        torch.cuda.synchronize()
        if rank == 0:
            full = plug.backward(bsdf.SurfaceInteraction3f(wi=torch.from_numpy(wi).to(dev)),
                                 torch.from_numpy(wo).to(dev), torch.from_numpy(g).to(dev))
            # same REDs, different summation order across the two tables: fp32 reassociation only
            torch.testing.assert_close(grad, full, rtol=1e-5, atol=1e-5)
            assert float(grad[:, 3].sum()) == float(full[:, 3].sum())   # query counts are exact integers
        comm.close()
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_gpu_sharded_backward_allreduce(tmp_path):
    import torch.multiprocessing as mp
    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(world))
