import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from mitsuba_customization_b200 import bsdf, render as RD, synth
from oracle import merl_oracle as O, render_oracle as RO

def run(ntab, nx, w, h, spp, depth):
    plugs = [bsdf.MERL(data=synth.synthetic_table(k), device="cuda:0") for k in range(ntab)]
    orcs = [O.OracleTable(synth.synthetic_table(k), cdf=p.cdf().cpu().numpy()) for k, p in enumerate(plugs)]
    sc = RD.grid_scene(plugs, nx, nx)
    got = RD.render(sc, w, h, spp, max_depth=depth, seed=1).cpu().numpy()[..., :3]
    ref = RO.render(orcs, sc.spheres, width=w, height=h, spp=spp, max_depth=depth, seed=1, cam_pos=sc.cam_pos,
                    cam_target=sc.cam_target, cam_up=sc.cam_up, fov_x_deg=sc.fov_x_deg, env=sc.env)[..., :3]
    d = np.abs(got - ref).max(axis=2)
    rel = ((got - ref) ** 2).mean() / (ref ** 2).mean()
    print(f"ntab={ntab} grid={nx} {w}x{h} spp={spp} depth={depth}: relMSE={rel:.3e}  pixels>1e-3: {(d > 1e-3).mean():.4f} "
          f">1e-2: {(d > 1e-2).mean():.4f} >1e-1: {(d > 1e-1).mean():.5f}  max={d.max():.3f}", flush=True)

run(1, 10, 256, 256, 16, 8)
run(1, 10, 256, 256, 16, 1)
run(1, 10, 256, 256, 16, 2)
run(4, 10, 256, 256, 16, 8)
run(100, 10, 256, 256, 16, 8)
run(4, 3, 256, 256, 16, 8)
