"""A/B of the persisting-L2 window on the GRADIENT table for the gather-free backward.
The switch is read once per process (MERL_BWD_PIN_GRAD, default 1), so each arm is its own process:

    for v in 0 1 0 1; do MERL_BWD_PIN_GRAD=$v python tools/bwd_window_ab.py; done
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mitsuba_customization_b200 import bsdf, synth  # noqa: E402
from tools.l2_ab import hemi, time_ms  # noqa: E402

gen = torch.Generator(device="cuda")
gen.manual_seed(9)
n = 1 << 26
plug = bsdf.MERL(data=synth.synthetic_table(0), device="cuda:0")
wi, wo = hemi(n, gen), hemi(n, gen)
g = torch.randn(3, n, device="cuda", generator=gen)
si = bsdf.SurfaceInteraction3f(wi=wi)
grad = torch.zeros(plug.info.ncell, 4, device="cuda")
fn = lambda: plug.backward(si, wo, g, grad_cells=grad, defer_mask=True)   # noqa: E731  (unmasked scatter, no gather)
for _ in range(60):
    fn()
ms = [time_ms(fn, 40) for _ in range(3)]
print(f"MERL_BWD_PIN_GRAD={os.environ.get('MERL_BWD_PIN_GRAD', '(default 1)')}: " +
      "  ".join(f"{m:.4f} ms {n / m / 1e6:.1f} Gq/s" for m in ms), "checksum", float(grad[:, 3].sum()))
