"""Per-window power / clocks of a kernel held back to back (120 -> 106 Gquery/s over 1000
launches with NVML still reporting 1965 MHz).  usage: sustained_probe.py <eval_pdf|sample|backward|copy> [launches=1000] [window=50]
Prints one line per window: mean launch ms, Gquery/s, power (W), SM / memory clock (MHz), throttle reasons."""
import sys, os, json, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from mitsuba_customization_b200 import bsdf, synth

def hemi(n, gen):
    z = torch.rand(n, device="cuda", generator=gen).clamp_min(1e-6)
    phi = torch.rand(n, device="cuda", generator=gen) * (2 * np.pi)
    r = torch.sqrt(1 - z * z)
    return torch.stack([r * torch.cos(phi), r * torch.sin(phi), z]).contiguous()

kind = sys.argv[1] if len(sys.argv) > 1 else "eval_pdf"
total = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
win = int(sys.argv[3]) if len(sys.argv) > 3 else 50
n = 1 << 26
gen = torch.Generator(device="cuda"); gen.manual_seed(1234)
plug = bsdf.MERL(data=synth.synthetic_table(0))
ctx = bsdf.BSDFContext()
wi, wo = hemi(n, gen), hemi(n, gen)
si = bsdf.SurfaceInteraction3f(wi=wi)
out = torch.empty(4, n, device="cuda")
if kind == "sample":
    u1 = torch.rand(n, device="cuda"); u2 = torch.rand(2, n, device="cuda")
    fn = lambda: plug.sample(ctx, si, u1, u2)
elif kind == "backward":
    g = torch.randn(3, n, device="cuda"); grad = torch.zeros(plug.info.ncell, 4, device="cuda")
    fn = lambda: plug.backward(si, wo, g, grad_cells=grad)
elif kind == "copy":
    a = torch.empty(10 * n // 2, device="cuda"); b = torch.empty_like(a)
    fn = lambda: b.copy_(a)
else:
    fn = lambda: plug.eval_pdf(ctx, si, wo, out=out)
import pynvml as nv
nv.nvmlInit(); h = nv.nvmlDeviceGetHandleByIndex(0)
def snap():
    try: r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
    except Exception: r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
    return (nv.nvmlDeviceGetPowerUsage(h) / 1000.0, nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM),
            nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_MEM), nv.nvmlDeviceGetTemperature(h, nv.NVML_TEMPERATURE_GPU), r)
for _ in range(3): fn()
torch.cuda.synchronize(); time.sleep(1.0)
print(f"# {kind}: {total} launches, windows of {win}; idle snapshot: {snap()}")
rows = []
for w in range(total // win):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(win): fn()
    e1.record()
    mid = snap()                      # taken while the window's launches are still running
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / win
    rows.append({"window": w, "ms": ms, "gquery_s": n / ms / 1e6, "power_w": mid[0], "sm_mhz": mid[1], "mem_mhz": mid[2],
                 "temp_c": mid[3], "reasons": hex(mid[4])})
    print(f"{w:3d}  {ms:7.4f} ms  {n / ms / 1e6:7.2f} Gq/s  {mid[0]:6.1f} W  sm {mid[1]} MHz  mem {mid[2]} MHz  {mid[3]} C  reasons {hex(mid[4])}")
print(json.dumps({"kind": kind, "rows": rows}))
