"""Where a GPU render and the CPU oracle differ: per-pixel outliers of the small multi-material test scene.
usage: python tools/render_diff.py [seeds...]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from mitsuba_customization_b200 import bsdf, render as RD, synth
from oracle import merl_oracle as O, render_oracle as RO

plugs, orcs = [], []
for k in range(3):
    raw = synth.synthetic_table(k)
    p = bsdf.MERL(data=raw, device="cuda:0"); plugs.append(p)
    orcs.append(O.OracleTable(raw, cdf=p.cdf().cpu().numpy()))
dims = (48, 40, 64)
rawc = synth.synthetic_table(9, dims=dims)
pc = bsdf.CustomizedMeasurement(data=rawc, resolution=dims, theta_h_mapping="linear", scale=(1 / 1500.0,) * 3, ni=16, device="cuda:0")
plugs.append(pc)
orcs.append(O.OracleTable(rawc, dims=dims, theta_h_sqrt=False, ni=16, scale=(1 / 1500.0,) * 3, cdf=pc.cdf().cpu().numpy()))
sc = RD.grid_scene(plugs, nx=3, ny=3)
seeds = [int(a) for a in sys.argv[1:]] or [23, 1, 2, 3, 4, 5]
for depth in (1, 2, 5):
    for seed in seeds:
        kw = dict(width=96, height=64, spp=8, max_depth=depth, seed=seed)
        film = RD.render(sc, **kw).cpu().numpy().astype(np.float64)
        ref = RO.render(orcs, sc.spheres, cam_pos=sc.cam_pos, cam_target=sc.cam_target, cam_up=sc.cam_up, fov_x_deg=sc.fov_x_deg, env=sc.env, **kw)
        d = np.abs(film[..., :3] - ref[..., :3]).max(axis=2)
        rel = float(((film[..., :3] - ref[..., :3]) ** 2).mean() / (ref[..., :3] ** 2).mean())
        big = np.argwhere(d > 1e-3)
        rest = d[d <= 1e-3]
        print(f"depth {depth} seed {seed}: relMSE {rel:.3e}  pixels off by > 1e-3: {len(big)}  max diff {d.max():.4f}  max of the rest {rest.max():.2e}  "
              + " ".join(f"({y},{x}: gpu {film[y, x, 0]:.3f} ref {ref[y, x, 0]:.3f})" for y, x in big[:4]))
