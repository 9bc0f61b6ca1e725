"""Two-GPU checks of the exchange step.  Skipped on boxes with fewer than two GPUs (the driver's 1-GPU
test box): tools/gpu_multi.sh runs this file on a 2-GPU box and keeps the full output under profiles/.
  - sharded backward + merl_allreduce_sum_f32 (NCCL) equals the single-GPU gradient
  - the same through the exchange buffers (merl_peerbuf_*): the peer-memory kernel and the registered-
    buffer NCCL transport, with the validity mask fused into the exchange; both ranks end bit-identical
  - merl_render sharded by table (every rank traces, evaluates its own tables, one exchange per bounce)
    equals the replicated render
  - one process holding tables on TWO devices (the occupancy / smem-attribute cache was keyed
    without the device)"""
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


def _peer_worker(rank, world, port, out_dir):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from mitsuba_customization_b200 import bsdf, distributed as D, render as RD, synth
        plug = bsdf.MERL(data=synth.synthetic_table(0), device=f"cuda:{rank}")
        comm = D.Comm(rank, world, rank)
        ncell = plug.info.ncell
        # n >= ncell: the big-batch (unmasked scatter) path, whose mask the exchange applies
        n = 3 * ncell + 8
        wi, wo = synth.random_queries(n, seed=5)
        g = np.random.default_rng(9).standard_normal((3, n)).astype(np.float32)
        lo, hi = D.shard_bounds(n, world, rank)
        cut = lambda a: torch.from_numpy(np.ascontiguousarray(a[:, lo:hi])).to(dev)
        full = plug.backward(bsdf.SurfaceInteraction3f(wi=torch.from_numpy(wi).to(dev)), torch.from_numpy(wo).to(dev),
                             torch.from_numpy(g).to(dev))
        invalid = plug.cells()[:, 3] != 7
        assert int(invalid.sum()) > 0, "the synthetic table has below-horizon cells"
        sums = {}
        for transport in ("peer", "nccl", "nvls"):
            try:
                pb = D.PeerBuffer(comm, 4 * ncell, transport)
            except bsdf.MerlError as e:
                # the switch transport needs NCCL >= 2.28 + NVSwitch multicast; the verdict is the same on every rank
                assert transport == "nvls" and "error -5" in str(e), e
                open(os.path.join(out_dir, f"{transport}{rank}"), "w").write("unsupported: " + str(e)[:120])
                continue
            gt = pb.tensor((ncell, 4))
            for rep in range(3):                       # repeated exchanges: the epochs advance, the buffer is reused
                gt.zero_()
                plug.backward(bsdf.SurfaceInteraction3f(wi=cut(wi)), cut(wo), cut(g), grad_cells=gt, defer_mask=True)
                pb.allreduce_(mask_plug=plug)
            torch.cuda.synchronize()
            assert pb.status() == 0
            torch.testing.assert_close(gt, full, rtol=1e-5, atol=1e-5)
            assert float(gt[:, 3].double().sum()) == float(n)
            ch = gt[invalid][:, :3]
            bits = plug.cells()[invalid][:, 3].to(torch.int32)
            for c in range(3):                          # masked channels of invalid cells are exactly zero
                assert float(ch[:, c][(bits & (1 << c)) == 0].abs().sum()) == 0.0
            t = torch.tensor([float(gt.double().sum())], device=dev, dtype=torch.float64)
            both = [torch.zeros_like(t) for _ in range(world)]
            dist.all_gather(both, t)
            sums[transport] = [float(x) for x in both]
            # a partial exchange (the first 4096 floats of a film-like buffer), no mask
            gt.zero_()
            gt.view(-1)[:4096] = float(rank + 1)
            pb.allreduce_(count=4096)
            torch.cuda.synchronize()
            assert float(gt.view(-1)[:4096].min()) == float(gt.view(-1)[:4096].max()) == world * (world + 1) / 2
            assert float(gt.view(-1)[4096:8192].abs().sum()) == 0.0
            reg = pb.registration
            pb.close()
            open(os.path.join(out_dir, f"{transport}{rank}"), "w").write(reg)
        assert sums["peer"][0] == sums["peer"][1], "peer exchange: ranks must end with identical bits"
        if "nvls" in sums:
            assert sums["nvls"][0] == sums["nvls"][1], "switch exchange: ranks must end with identical bits"

        # by-table sharded render == replicated render (same samples, same seed)
        raws = [synth.synthetic_table(k) for k in range(4)]
        plugs = [bsdf.MERL(data=r, device=f"cuda:{rank}") for r in raws]
        kw = dict(width=96, height=64, spp=4, max_depth=4, seed=23, wavefront=8192)
        ref = RD.render(RD.grid_scene(plugs, nx=3, ny=3), **kw)
        own = [p if k % world == rank else None for k, p in enumerate(plugs)]
        xb = D.PeerBuffer(comm, 12 * 8192, "peer")
        got = RD.render(RD.grid_scene(own, nx=3, ny=3), table_shard=(world, rank), stage_exchange=xb, **kw)
        torch.cuda.synchronize()
        assert xb.status() == 0
        torch.testing.assert_close(got, ref, rtol=1e-5, atol=1e-6)
        xb.close()
        comm.close()
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_gpu_peer_exchange_and_by_table_render(tmp_path):
    import torch.multiprocessing as mp
    world, port = 2, _free_port()
    mp.spawn(_peer_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(world))
    print("registrations:", {f.name: f.read_text() for f in tmp_path.iterdir() if f.name[:4] in ("peer", "nccl", "nvls")})


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_one_process_two_devices():
    """Tables on cuda:0 and cuda:1 in ONE process: the > 48 KB dynamic-smem launches (staged pdf, big-batch
    backward) must work on both devices, whichever ran first."""
    from mitsuba_customization_b200 import bsdf, synth
    from oracle import merl_oracle as O
    raw = synth.synthetic_table(0)
    plugs = [bsdf.MERL(data=raw, device=f"cuda:{d}") for d in (0, 1)]
    orc = O.OracleTable(raw, cdf=plugs[0].cdf().cpu().numpy())
    ncell = plugs[0].info.ncell
    n = ncell + 64                                # >= ncell: the staged big-batch backward (55.6 KB of smem)
    wi, wo = synth.random_queries(n, seed=3)
    g = np.ones((3, n), dtype=np.float32)
    ref_rgb, ref_pdf, _ = orc.eval_pdf(wi, wo)
    ref_grad = orc.backward(wi, wo, g)
    ctx = bsdf.BSDFContext()
    for d in (1, 0, 1):                           # device 1 first: the cache's first entry is not device 0's
        dev = torch.device("cuda", d)
        with torch.cuda.device(dev):
            si = bsdf.SurfaceInteraction3f(wi=torch.from_numpy(wi).to(dev))
            two = torch.from_numpy(wo).to(dev)
            pdf = plugs[d].pdf(ctx, si, two)
            grad = plugs[d].backward(si, two, torch.from_numpy(g).to(dev))
            torch.cuda.synchronize(dev)
        np.testing.assert_allclose(pdf.cpu().numpy(), ref_pdf, rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(grad[:, :3].cpu().numpy(), ref_grad[:, :3], rtol=1e-5, atol=1e-5)
