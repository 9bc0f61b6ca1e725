"""Run ONE query kernel a few times on 2^logn random queries (for ncu captures and A/B timing).
usage: one_kernel.py <eval_pdf|eval|pdf|bins|sample|backward|backward_agg> [logn=26] [iters=5]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from mitsuba_customization_b200 import bsdf, synth

def hemi(n, gen):
    z = torch.rand(n, device="cuda", generator=gen).clamp_min(1e-6)
    phi = torch.rand(n, device="cuda", generator=gen) * (2 * np.pi)
    r = torch.sqrt(1 - z * z)
    return torch.stack([r * torch.cos(phi), r * torch.sin(phi), z]).contiguous()

kind = sys.argv[1]
logn = int(sys.argv[2]) if len(sys.argv) > 2 else 26
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 5
n = 1 << logn
gen = torch.Generator(device="cuda"); gen.manual_seed(1234)
plug = bsdf.MERL(data=synth.synthetic_table(0))
ctx = bsdf.BSDFContext()
wi, wo = hemi(n, gen), hemi(n, gen)
si = bsdf.SurfaceInteraction3f(wi=wi)
out = torch.empty(4, n, device="cuda")
u1 = torch.rand(n, device="cuda"); u2 = torch.rand(2, n, device="cuda")
g = torch.randn(3, n, device="cuda"); grad = torch.zeros(plug.info.ncell, 4, device="cuda")
fn = {
    "eval_pdf": lambda: plug.eval_pdf(ctx, si, wo, out=out),
    "eval": lambda: plug.eval(ctx, si, wo),
    "pdf": lambda: plug.pdf(ctx, si, wo),
    "bins": lambda: plug.bins(si, wo),
    "sample": lambda: plug.sample(ctx, si, u1, u2),
    "backward": lambda: plug.backward(si, wo, g, grad_cells=grad, warp_aggregate=False),
    "backward_agg": lambda: plug.backward(si, wo, g, grad_cells=grad, warp_aggregate=True),
}[kind]
for _ in range(3): fn()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(iters): fn()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / iters
print(f"{kind}: {ms:.3f} ms  {n / ms / 1e6:.2f} Gq/s")
