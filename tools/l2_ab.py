"""A/B of the persisting-L2 access window (merl_table_params_t.l2_persist) -- VERDICT r2 item 8.

Same process, same queries: eval+pdf over 2^26 queries against ONE table (23 MB, fits L2 with or without the
window) and the 100-table BSDF stage of config 5 (2.33 GB >> 126 MB L2: the window of whichever table a launch
uses is all that can stay).  Interleaved A/B/A/B, CUDA events, burst state avoided by 60 untimed launches first.

    python tools/l2_ab.py > gpurun_out/l2_ab.txt
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mitsuba_customization_b200 import bsdf, synth  # noqa: E402


def hemi(n, gen):
    z = torch.rand(n, device="cuda", generator=gen).clamp_(1e-4, 1.0)
    phi = torch.rand(n, device="cuda", generator=gen) * 6.2831853
    r = (1 - z * z).sqrt()
    return torch.stack([r * phi.cos(), r * phi.sin(), z]).contiguous()


def time_ms(fn, reps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    gen = torch.Generator(device="cuda")
    gen.manual_seed(5)
    n = 1 << 26
    wi, wo = hemi(n, gen), hemi(n, gen)
    out = torch.empty(4, n, device="cuda")
    ctx = bsdf.BSDFContext()
    si = bsdf.SurfaceInteraction3f(wi=wi)
    raw = synth.synthetic_table(0)
    # a table created with l2_persist=1 raises the device's persisting-L2 carve-out for the whole process, so the
    # l2_persist=0 table is timed FIRST, before any carve-out exists, and again afterwards next to the other one
    plugs = {0: bsdf.MERL(data=raw, device="cuda:0", l2_persist=0)}
    for _ in range(60):
        plugs[0].eval_pdf(ctx, si, wo, out=out)
    print("one table, eval+pdf, 2^26 queries, l2_persist=0 before any carve-out exists:")
    for rnd in range(3):
        ms = time_ms(lambda: plugs[0].eval_pdf(ctx, si, wo, out=out), 40)
        print(f"  round {rnd} l2_persist=0 (no carve-out): {ms:.4f} ms  {n / ms / 1e6:.1f} Gquery/s")
    plugs[1] = bsdf.MERL(data=raw, device="cuda:0", l2_persist=1)
    print("one table, eval+pdf, 2^26 queries (ms per launch, Gquery/s), interleaved, carve-out set:")
    for rnd in range(3):
        for k in (0, 1):
            ms = time_ms(lambda: plugs[k].eval_pdf(ctx, si, wo, out=out), 40)
            print(f"  round {rnd} l2_persist={k}: {ms:.4f} ms  {n / ms / 1e6:.1f} Gquery/s")
    # 100 tables, 2^19 queries per table per launch (the segmented stage of config 5 launches per material)
    m = 1 << 19
    many = {k: [bsdf.MERL(data=synth.synthetic_table(t % 10) * (0.55 + 0.05 * (t // 10)), device="cuda:0", l2_persist=k)
                for t in range(100)] for k in (0, 1)}
    sis = bsdf.SurfaceInteraction3f(wi=wi[:, :m].contiguous())
    wos, outs = wo[:, :m].contiguous(), torch.empty(4, m, device="cuda")

    def sweep(k):
        for p in many[k]:
            p.eval_pdf(ctx, sis, wos, out=outs)
    for k in (0, 1):
        sweep(k)
    print("100 tables (2.33 GB), 2^19 queries per table, one launch per table (ms per 100-launch sweep, Gquery/s):")
    for rnd in range(3):
        for k in (0, 1):
            ms = time_ms(lambda: sweep(k), 5)
            print(f"  round {rnd} l2_persist={k}: {ms:.3f} ms  {100 * m / ms / 1e6:.1f} Gquery/s")


if __name__ == "__main__":
    main()
