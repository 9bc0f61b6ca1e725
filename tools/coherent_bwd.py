"""Backward on COHERENT batches: the gradient queries a differentiable renderer produces, in
wavefront order, instead of uniform random pairs.  Camera rays of BASELINE config 2 (unit sphere, 512x512, pinhole at
(0,-4,0), fov 40 deg, jittered, sample-major order = merl_render's path order), compacted to the rays that hit:
  render       wo = uniformly random direction per path (what merl_render's emitter sampling queries): wi coherent, wo not
  directional  wo = one fixed light direction: wi and wo both coherent (direct lighting from a distant light)
  random       uniform random (wi, wo), the bench's workload, for reference
Times merl_backward plain / warp-aggregated / CTA-aggregated and checks the three gradients against each other.
usage: python tools/coherent_bwd.py [spp=64]"""
import sys, os, json, math
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mitsuba_customization_b200 import bsdf, synth

spp = int(sys.argv[1]) if len(sys.argv) > 1 else 64
W = H = 512
dev = "cuda"
gen = torch.Generator(device=dev); gen.manual_seed(7)
plug = bsdf.MERL(data=synth.synthetic_table(0))

def frame(n):
    sign = torch.where(n[2] >= 0, 1.0, -1.0)
    a = -1.0 / (sign + n[2]); b = n[0] * n[1] * a
    s = torch.stack([1 + sign * n[0] * n[0] * a, sign * b, -sign * n[0]])
    t = torch.stack([b, sign + n[1] * n[1] * a, -n[1]])
    return s, t

def camera_hits():
    npix = W * H
    pix = torch.arange(npix, device=dev).repeat(spp)                       # sample-major: consecutive paths = consecutive pixels
    x, y = (pix % W).float(), (pix // W).float()
    jx, jy = torch.rand(npix * spp, device=dev, generator=gen), torch.rand(npix * spp, device=dev, generator=gen)
    sx = (x + jx) / W * 2 - 1; sy = 1 - (y + jy) / H * 2
    th = math.tan(math.radians(20.0))
    d = torch.stack([sx * th, torch.ones_like(sx), sy * th]); d = d / d.norm(dim=0)   # forward +y, right +x, up +z
    o = torch.tensor([0.0, -4.0, 0.0], device=dev)[:, None]
    b = (o * d).sum(0); c = (o * o).sum() - 1.0
    disc = b * b - c
    hit = disc > 0
    t = -b - torch.sqrt(disc.clamp_min(0))
    p = o + t * d
    n = p / p.norm(dim=0).clamp_min(1e-20)
    s, tt = frame(n)
    wi = torch.stack([-(d * s).sum(0), -(d * tt).sum(0), -(d * n).sum(0)])
    keep = hit & (wi[2] > 0)
    return wi[:, keep].contiguous(), n[:, keep].contiguous(), s[:, keep].contiguous(), tt[:, keep].contiguous()

wi, n, s, tt = camera_hits()
m = wi.shape[1] // 4 * 4
wi, n, s, tt = wi[:, :m].contiguous(), n[:, :m], s[:, :m], tt[:, :m]
def to_local(v):
    return torch.stack([(v * s).sum(0), (v * tt).sum(0), (v * n).sum(0)]).contiguous()
ez = 1 - 2 * torch.rand(m, device=dev, generator=gen); er = torch.sqrt((1 - ez * ez).clamp_min(0)); ph = 2 * math.pi * torch.rand(m, device=dev, generator=gen)
wo_render = to_local(torch.stack([er * torch.cos(ph), er * torch.sin(ph), ez]))
L = torch.tensor([0.3, -0.6, 0.74], device=dev); L = (L / L.norm())[:, None].expand(3, m)
wo_dir = to_local(L)
def hemi(k):
    z = torch.rand(k, device=dev, generator=gen).clamp_min(1e-6); p = torch.rand(k, device=dev, generator=gen) * 2 * math.pi
    r = torch.sqrt(1 - z * z); return torch.stack([r * torch.cos(p), r * torch.sin(p), z]).contiguous()
# extreme reuse: a flat surface seen from far away under a distant light -- every query in one of a handful of cells
flat_wi = torch.tensor([0.3, 0.1, 0.95], device=dev); flat_wi = (flat_wi / flat_wi.norm())[:, None].expand(3, m).contiguous()
flat_wo = torch.tensor([-0.2, 0.4, 0.89], device=dev); flat_wo = (flat_wo / flat_wo.norm())[:, None].expand(3, m)
flat_wo = (flat_wo + 2e-3 * torch.randn(3, m, device=dev, generator=gen)); flat_wo = (flat_wo / flat_wo.norm(dim=0)).contiguous()
sets = {"render (wo random per path)": (wi, wo_render), "directional light": (wi, wo_dir), "random pairs": (hemi(m), hemi(m)),
        "flat surface, distant light": (flat_wi, flat_wo)}
g = torch.rand(3, m, device=dev, generator=gen)
modes = {"plain": {}, "warp": {"warp_aggregate": True}, "cta": {"cta_aggregate": True}}
def timeit(fn, it=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it
out = {"queries": m, "spp": spp}
print(f"{m} queries ({spp} spp of the 512x512 sphere view, hits only)")
for name, (a, b) in sets.items():
    si = bsdf.SurfaceInteraction3f(wi=a)
    idx = plug.bins(si, b)
    valid = idx >= 0
    distinct = int(torch.unique(idx[valid]).numel())
    same_as_prev = float((idx[1:] == idx[:-1])[valid[1:]].float().mean())
    row = {"valid": int(valid.sum()), "distinct_cells": distinct, "same_cell_as_previous_query": same_as_prev}
    grads = {}
    for mode, kw in modes.items():
        grad = torch.zeros(plug.info.ncell, 4, device=dev)
        ms = timeit(lambda: plug.backward(si, b, g, grad_cells=grad, **kw))
        row[mode + "_gquery_s"] = m / ms / 1e6
        grad.zero_(); plug.backward(si, b, g, grad_cells=grad, **kw); grads[mode] = grad
    scale = float(grads["plain"][:, :3].abs().max())
    row["max_diff_warp_vs_plain_over_max"] = float((grads["warp"] - grads["plain"])[:, :3].abs().max()) / scale
    row["max_diff_cta_vs_plain_over_max"] = float((grads["cta"] - grads["plain"])[:, :3].abs().max()) / scale
    row["hit_counts_equal"] = bool(torch.equal(grads["cta"][:, 3], grads["plain"][:, 3]) and torch.equal(grads["warp"][:, 3], grads["plain"][:, 3]))
    out[name] = row
    print(f"{name:30s} distinct cells {distinct:8d}  same-as-previous {same_as_prev:5.3f}  plain {row['plain_gquery_s']:7.2f}  "
          f"warp {row['warp_gquery_s']:7.2f}  cta {row['cta_gquery_s']:7.2f} Gq/s  diff {row['max_diff_cta_vs_plain_over_max']:.1e}  counts equal {row['hit_counts_equal']}")
print(json.dumps(out))
