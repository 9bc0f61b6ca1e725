"""Render throughput of the wavefront harness (BASELINE configs 2 and 5), spp*Mpix/s.

  python tools/render_bench.py cfg2            512x512 sphere, constant envmap, 64 spp, one table
  python tools/render_bench.py cfg5 [ntab]     1024x1024, 10x10 spheres, `ntab` tables (100), 256 spp
Under torchrun every rank renders its share of the sample indices and the films are summed with
one allreduce (libmerl_b200 communicator).  --cpu times the CPU oracle on a bounded sample."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from mitsuba_customization_b200 import bsdf, render as RD, synth


def main():
    cfg = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", 1), ("RANK", 0), ("LOCAL_RANK", 0)))
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    comm = None
    if world > 1:
        import torch.distributed as dist
        from mitsuba_customization_b200 import distributed as D
        dist.init_process_group("nccl", device_id=torch.device(dev))
        comm = D.Comm(rank, world, local)
    if cfg == "cfg2":
        w = h = 512; spp = 64; depth = 8
        sc = RD.sphere_scene(bsdf.MERL(data=synth.synthetic_table(0), device=dev))
    else:
        ntab = int(sys.argv[2]) if len(sys.argv) > 2 and sys.argv[2].isdigit() else 100
        w = h = 1024; spp = 256; depth = 8
        t0 = time.time()
        plugs = [bsdf.MERL(data=synth.synthetic_table(k), device=dev) for k in range(ntab)]
        sc = RD.grid_scene(plugs, 10, 10)
        if rank == 0:
            print(f"# {ntab} tables built in {time.time() - t0:.1f} s, "
                  f"{sum(p.info.table_bytes for p in plugs) / 1e9:.2f} GB of float4 cells", file=sys.stderr)
    rng = RD.spp_share(spp, world, rank)
    emu = int(os.environ.get("RD_EMULATE_WORLD", "0"))   # one GPU doing rank 0's share of an `emu`-GPU render (time split of a rank)
    if emu > 1 and world == 1:
        rng = RD.spp_share(spp, emu, 0)
    wf = int(os.environ.get("MERL_WAVEFRONT", "0"))   # paths per pass; 0 = the library's default
    pb = None
    if comm is not None and os.environ.get("RD_FILM_PEER", "0") == "1":   # as bench.py: the film lives in an exchange buffer
        pb = D.PeerBuffer(comm, w * h * 4, "peer")
        film = pb.tensor((h, w, 4))
    else:
        film = torch.zeros(h, w, 4, device=dev)
    if rank == 0:
        fr, tot = torch.cuda.mem_get_info()
        print(f"# free device memory before the first render: {fr / 2**30:.1f} of {tot / 2**30:.1f} GiB", file=sys.stderr)

    def step():
        film.zero_()
        RD.render(sc, w, h, spp, max_depth=depth, seed=1, spp_range=rng, film=film, wavefront=wf)
        if pb is not None:
            pb.allreduce_()
        elif comm is not None:
            comm.allreduce_sum_(film)

    for _ in range(2):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 3
    e0.record()
    for _ in range(reps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    _, st = RD.render(sc, w, h, spp, max_depth=depth, seed=1, spp_range=rng, want_stats=True, wavefront=wf)
    out = {"config": cfg, "n_gpus": world, "wavefront": wf or "library default (up to 2^26 paths per pass)", "width": w, "height": h, "spp": spp, "max_depth": depth, "ms": ms,
           "spp_mpix_per_s": spp * w * h / 1e6 / (ms * 1e-3), "paths_per_rank": st["paths"],
           "bsdf_queries_per_rank": st["bsdf_queries"], "launches_per_rank": st["launches"],
           "mean_radiance": [float(x) for x in film[..., :3].mean(dim=(0, 1))],
           "sample_share_min_max": [float(film[..., 3].min()), float(film[..., 3].max())]}
    if "--cpu" in sys.argv and rank == 0:
        from tests.render_cpu_baseline import cpu_render_baseline   # the oracle is test infrastructure
        gpu = (lambda r: RD.render(sc, w, h, spp, max_depth=depth, seed=1, spp_range=r).cpu().numpy()) if world == 1 else None
        out.update(cpu_render_baseline(sc, w, h, spp, depth, 1, gpu))
    if rank == 0:
        print(json.dumps(out), flush=True)
    if comm is not None:
        comm.close()
        dist.destroy_process_group()


main()
