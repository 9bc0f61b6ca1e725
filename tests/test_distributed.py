"""Host-side multi-GPU logic on CPU: shard bounds, id exchange and the shard -> sum structure of the
backward, run as two gloo ranks.  (The NCCL path itself is covered by tests/test_gpu_multi.py.)"""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from mitsuba_customization_b200 import distributed as D  # noqa: E402


@pytest.mark.parametrize("n", [0, 1, 3, 4, 5, 1023, 1024, (1 << 20) + 5])
@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_shard_bounds_cover_exactly(n, world):
    cuts = [D.shard_bounds(n, world, r) for r in range(world)]
    assert cuts[0][0] == 0 and cuts[-1][1] == n
    for (lo, hi), (lo2, _) in zip(cuts, cuts[1:]):
        assert hi == lo2 and lo <= hi
    for lo, hi in cuts[:-1]:
        assert (lo % 4 == 0 or lo == n) and (hi % 4 == 0 or hi == n)
    assert sum(hi - lo for lo, hi in cuts) == n


def test_shard_bounds_rejects_bad_rank():
    with pytest.raises(ValueError):
        D.shard_bounds(10, 2, 2)


def test_table_owner_round_robin():
    owners = [D.table_owner(t, 8) for t in range(100)]
    counts = np.bincount(owners, minlength=8)
    assert counts.min() == 12 and counts.max() == 13   # 100 tables over 8 GPUs


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # 1. the communicator id travels from rank 0 to everyone
        marker = bytes(range(128))
        got = D.exchange_id(lambda: marker)
        assert got == marker, "id exchange corrupted the bytes"

        # 2. shard -> private gradient table -> sum == unsharded gradient (oracle as the kernel)
        from mitsuba_customization_b200 import synth
        from oracle import merl_oracle as O
        raw = synth.synthetic_table(0)
        t = O.OracleTable(raw)
        n = 20001
        wi, wo = synth.random_queries(n, seed=99)
        g = np.random.default_rng(5).standard_normal((3, n)).astype(np.float32)
        lo, hi = D.shard_bounds(n, world, rank)
        part = t.backward(np.ascontiguousarray(wi[:, lo:hi]), np.ascontiguousarray(wo[:, lo:hi]),
                          np.ascontiguousarray(g[:, lo:hi]))
        tot = torch.from_numpy(part.copy())
        dist.all_reduce(tot)
        if rank == 0:
            full = t.backward(wi, wo, g)
            np.testing.assert_allclose(tot.numpy(), full, rtol=1e-12, atol=1e-12)
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_sharded_backward(tmp_path):
    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(world))


def _render_worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # every rank renders its share of the sample indices of every pixel (render.spp_share, the split
        # merl_render's spp_begin / spp_end express); one allreduce of the film gives the full image
        from mitsuba_customization_b200 import render as RD, synth
        from oracle import merl_oracle as O, render_oracle as R
        t = O.OracleTable(synth.synthetic_table(0))
        t.build_cdf()
        kw = dict(width=20, height=16, spp=5, max_depth=3, seed=11, cam_pos=(0.0, -4.0, 0.0), cam_target=(0.0, 0.0, 0.0),
                  fov_x_deg=40.0)
        lo, hi = RD.spp_share(kw["spp"], world, rank)
        film = torch.from_numpy(R.render([t], [(0, 0, 0, 1.0, 0)], spp_begin=lo, spp_end=hi, **kw).copy())
        dist.all_reduce(film)
        if rank == 0:
            full = R.render([t], [(0, 0, 0, 1.0, 0)], **kw)
            np.testing.assert_allclose(film.numpy(), full, rtol=1e-12, atol=1e-15)
            np.testing.assert_allclose(film.numpy()[..., 3], 1.0, atol=1e-12)      # every sample rendered exactly once
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_gloo_sample_sharded_render_sums_to_the_full_film(tmp_path, world):
    port = _free_port()
    mp.spawn(_render_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(world))
