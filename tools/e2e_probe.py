"""End-to-end rate of MERL.eval_pdf_host on 2^26 pinned host queries (the bench's e2e leg alone), 8 timed calls.
MERL_HOST_CHUNK_LOG2 selects the pipeline's chunk size (tuning)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mitsuba_customization_b200 import bsdf, synth

n = 1 << 26
plug = bsdf.MERL(data=synth.synthetic_table(0))
g = torch.Generator(); g.manual_seed(3)
def hemi():
    z = torch.rand(n, generator=g).clamp_min(1e-6); ph = torch.rand(n, generator=g) * 6.2831853
    r = torch.sqrt(1 - z * z)
    return torch.stack([r * torch.cos(ph), r * torch.sin(ph), z]).contiguous().pin_memory()
hwi, hwo = hemi(), hemi()
hout = torch.empty(4, n).pin_memory()
for _ in range(2): plug.eval_pdf_host(hwi, hwo, hout)
t0 = time.perf_counter()
for _ in range(8): plug.eval_pdf_host(hwi, hwo, hout)
dt = (time.perf_counter() - t0) / 8
print(f"chunk 2^{os.environ.get('MERL_HOST_CHUNK_LOG2', 'default')}: {dt * 1e3:.2f} ms per call  {n / dt / 1e9:.3f} Gquery/s  checksum {float(hout[:, ::65537].double().sum()):.6f}")
