"""CPU leg of tools/render_bench.py --cpu: times the render oracle (oracle/render_oracle.c, all host
threads) on a bounded share of the samples and compares that share with the GPU image.  It lives
under tests/ because only tests/, smoke() and bench.py's CPU baseline may execute the oracle."""
import time


def cpu_render_baseline(sc, width, height, spp, depth, seed, gpu_render):
    """gpu_render(spp_range) -> (h, w, 4) numpy film of the GPU path for the same sample range."""
    from mitsuba_customization_b200 import synth
    from oracle import merl_oracle as O, render_oracle as RO
    orcs = [O.OracleTable(synth.synthetic_table(k), cdf=b.cdf().cpu().numpy()) for k, b in enumerate(sc.bsdfs)]
    cs = max(1, spp // 16)
    t0 = time.time()
    ref = RO.render(orcs, sc.spheres, width=width, height=height, spp=spp, max_depth=depth, seed=seed,
                    cam_pos=sc.cam_pos, cam_target=sc.cam_target, cam_up=sc.cam_up, fov_x_deg=sc.fov_x_deg,
                    env=sc.env, spp_begin=0, spp_end=cs)
    dt = time.time() - t0
    out = {"cpu_oracle_spp_mpix_per_s": cs * width * height / 1e6 / dt, "cpu_oracle_threads": O.max_threads(),
           "cpu_oracle_sample": f"{cs} of {spp} spp, {dt:.2f} s"}
    if gpu_render is not None:   # same seed, same samples: the first cs spp must agree as an image
        part = gpu_render((0, cs))
        out["relmse_vs_oracle_first_spp"] = float(((part[..., :3] - ref[..., :3]) ** 2).mean() / (ref[..., :3] ** 2).mean())
    return out
