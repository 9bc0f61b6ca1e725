"""Quick device-side timing of every query kernel (CUDA events, inputs >> L2)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from mitsuba_customization_b200 import bsdf, synth, _cabi

def hemi(n, gen):
    z = torch.rand(n, device="cuda", generator=gen).clamp_min(1e-6)
    phi = torch.rand(n, device="cuda", generator=gen) * (2 * np.pi)
    r = torch.sqrt(1 - z * z)
    return torch.stack([r * torch.cos(phi), r * torch.sin(phi), z]).contiguous()

def timeit(fn, iters=10, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters

def main():
    logn = int(sys.argv[1]) if len(sys.argv) > 1 else 26
    n = 1 << logn
    gen = torch.Generator(device="cuda"); gen.manual_seed(1234)
    plug = bsdf.MERL(data=synth.synthetic_table(0))
    ctx = bsdf.BSDFContext()
    wi, wo = hemi(n, gen), hemi(n, gen)
    si = bsdf.SurfaceInteraction3f(wi=wi)
    out = torch.empty(4, n, device="cuda")
    res = {}
    ms = timeit(lambda: plug.eval_pdf(ctx, si, wo, out=out)); res["eval_pdf"] = (ms, n / ms / 1e6, 40)
    ms = timeit(lambda: plug.eval(ctx, si, wo)); res["eval"] = (ms, n / ms / 1e6, 36)
    ms = timeit(lambda: plug.pdf(ctx, si, wo)); res["pdf"] = (ms, n / ms / 1e6, 28)
    ms = timeit(lambda: plug.bins(si, wo)); res["bins"] = (ms, n / ms / 1e6, 28)
    u1 = torch.rand(n, device="cuda"); u2 = torch.rand(2, n, device="cuda")
    for mode in (0, 1):
        _cabi.lib().merl_set_sample_search(mode)
        ms = timeit(lambda: plug.sample(ctx, si, u1, u2)); res[f"sample_search{mode}"] = (ms, n / ms / 1e6, 56)
    _cabi.lib().merl_set_sample_search(0)
    g = torch.randn(3, n, device="cuda"); grad = torch.zeros(plug.info.ncell, 4, device="cuda")
    for agg in (True, False):
        ms = timeit(lambda: plug.backward(si, wo, g, grad_cells=grad, warp_aggregate=agg)); res[f"backward_agg{int(agg)}"] = (ms, n / ms / 1e6, 36)
    # reference points: plain copy of the same 40 B/query, and a raw random float4 gather
    a = torch.empty(10 * n // 2, device="cuda"); b = torch.empty_like(a)
    ms = timeit(lambda: b.copy_(a)); res["copy_40B_per_q"] = (ms, n / ms / 1e6, 40)
    idx = torch.randint(0, plug.info.ncell, (n,), device="cuda")
    cells = plug.cells()
    ms = timeit(lambda: cells.index_select(0, idx)); res["torch_gather_float4"] = (ms, n / ms / 1e6, 36)
    for k, (ms, gq, bpq) in res.items():
        print(f"{k:22s} {ms:8.3f} ms  {gq:8.2f} Gq/s  {gq * bpq:8.1f} GB/s algorithmic")
    print(json.dumps({k: v[1] for k, v in res.items()}))

main()
