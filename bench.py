#!/usr/bin/env python
"""bench.py -- MERL BSDF eval+pdf throughput on B200 (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

A "step" is one pass of the hot path (batched eval+pdf through the plugin's eval_pdf, i.e.
through the C ABI) over one batch of 2^26 synthetic (wi, wo) pairs per GPU against one
synthetic MERL table.  One JSON line is printed by rank 0:

  value      whole-job Gquery/s with the batch resident in HBM (CUDA events, max over ranks)
  e2e        the same metric through merl_eval_pdf_host with PINNED HOST buffers: every step
             copies the 24 B/query of inputs host->device and the 16 B/query of results back
  roofline   achieved algorithmic HBM GB/s (40 B/query) of the eval_pdf kernel
             against the measured copy bandwidth in MEASURED_PEAKS.json
  cpu_baseline  the CPU oracle (kind "port": the reference has no source to run) on a bounded
             sample of the same workload, all host threads

--impl reference times that CPU port alone (the reference's own plugins cannot be built:
/root/reference holds no source).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

LOG2_BATCH = 26
BYTES_PER_QUERY = 40          # 24 B in (wi, wo) + 16 B out (rgb, pdf): SURVEY.md §8d
METRIC = "merl_eval_pdf_throughput"
UNIT = "Gquery/s"
WORKLOAD = ("MERL eval+pdf, 2^26 random upper-hemisphere (wi,wo) pairs per GPU, one synthetic "
            "90x90x180x3 MERL table (float4, L2-resident)")


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--log2-batch", type=int, default=LOG2_BATCH)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true")
    ap.add_argument("--no-render", action="store_true", help="skip the render legs of the extras")
    ap.add_argument("--render-tables", type=int, default=100, help="tables of render config 5 (0 = skip it)")
    return ap.parse_args()


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md: 6.65 TB/s)"


def recorded_traffic():
    """dram bytes per launch of the eval_pdf kernel from the committed ncu capture, or None."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["eval_pdf_dram_bytes_per_launch"]
    except Exception:
        return None


# --------------------------------------------------------------------------------------------
# CPU arm: the oracle port, all host threads, bounded sample
# --------------------------------------------------------------------------------------------
def cpu_port_rate(budget_s=10.0, log2_sample=22):
    import numpy as np
    from mitsuba_customization_b200 import synth
    from oracle import merl_oracle as O

    raw = synth.synthetic_table(0)
    t = O.OracleTable(raw)
    t.build_cdf()
    n = 1 << log2_sample
    wi, wo = synth.random_queries(n, seed=1234)
    t.eval_pdf(wi[:, :4096], wo[:, :4096])  # thread pool warm-up
    reps, t0 = 0, time.perf_counter()
    while True:
        t.eval_pdf(wi, wo)
        reps += 1
        dt = time.perf_counter() - t0
        if dt >= budget_s or reps >= 256:
            break
    return n * reps / dt / 1e9, O.max_threads(), f"{reps} x 2^{log2_sample} eval+pdf queries of the same workload, {dt:.1f} s"


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import numpy as np  # noqa: F401
    from mitsuba_customization_b200 import synth
    from oracle import merl_oracle as O

    raw = synth.synthetic_table(0)
    t = O.OracleTable(raw)
    t.build_cdf()
    n = 1 << 22  # bounded sample of the 2^26-query step
    wi, wo = synth.random_queries(n, seed=1234)
    steps = max(1, min(args.steps, 40))
    warm = max(1, min(args.warmup, 3))
    for _ in range(warm):
        t.eval_pdf(wi, wo)
    t0 = time.perf_counter()
    for _ in range(steps):
        t.eval_pdf(wi, wo)
    dt = time.perf_counter() - t0
    v = n * steps / dt / 1e9
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warm, "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "step_sample": "2^22 of the step's 2^26 queries per step (bounded CPU sample)"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": O.max_threads(), "kind": "port",
                         "sample": f"{steps} steps x 2^22 eval+pdf queries, OpenMP over all host threads; the "
                                   "reference's own plugins are absent from /root/reference, so this is the CPU "
                                   "restatement (oracle/merl_oracle.c), not the reference binary"},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------
# clocks during the run (NVML; the recipe's nvidia-smi line samples the same counters)
# --------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None
        self.power_w = []
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if not self.nv:
            return
        nv = self.nv
        names = {"hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
                 "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                 "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
                 "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4)}
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    self.power_w.append(nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
                except Exception:
                    pass
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.002)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "note": "NVML unavailable"}
        s = sorted(self.samples)
        out = {"sm_mhz": s[len(s) // 2], "sm_min_mhz": s[0], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
               "samples": len(s), "window": "the timed region of the headline metric only"}
        if self.power_w:
            out["power_w_max"] = max(self.power_w)
        return out


# --------------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------------
def hemisphere(n, gen, device):
    import math
    import torch
    z = torch.rand(n, device=device, generator=gen).clamp_min(1e-6)
    phi = torch.rand(n, device=device, generator=gen) * (2 * math.pi)
    r = torch.sqrt(1 - z * z)
    return torch.stack([r * torch.cos(phi), r * torch.sin(phi), z]).contiguous()


def run_b200(args):
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    from mitsuba_customization_b200 import _cabi, bsdf, synth

    plug = bsdf.MERL(data=synth.synthetic_table(0), device=f"cuda:{local}")
    ctx = bsdf.BSDFContext()
    n = 1 << args.log2_batch
    gen = torch.Generator(device=dev)
    gen.manual_seed(1234 + rank)  # each rank owns a different shard of the query stream
    wi, wo = hemisphere(n, gen, dev), hemisphere(n, gen, dev)
    si = bsdf.SurfaceInteraction3f(wi=wi)
    out = torch.empty(4, n, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local)

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        l0 = _cabi.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        launches = _cabi.launch_count() - l0
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, launches

    # ---- headline: device-resident eval+pdf ----------------------------------------------
    W = max(args.warmup, 3)
    for _ in range(W):                      # warm up before the clock sampler's window opens
        plug.eval_pdf(ctx, si, wo, out=out)
    torch.cuda.synchronize()
    sampler.start()
    ms, launches = timed(lambda: plug.eval_pdf(ctx, si, wo, out=out), args.steps, 0)
    sampler.stop_flag = True
    sampler.join(timeout=1.0)
    ms_step = ms / args.steps
    value = world * n / (ms_step * 1e-3) / 1e9
    kernel_ms = ms_step  # one kernel per step (n is a multiple of 4 and buffers are 16-byte aligned)

    # ---- end to end: pinned host buffers through merl_eval_pdf_host ----------------------
    hwi, hwo = wi.cpu().pin_memory(), wo.cpu().pin_memory()
    hout = torch.empty(4, n).pin_memory()
    e2e_steps = max(1, min(args.steps, 8))
    barrier()
    for _ in range(2):
        plug.eval_pdf_host(hwi, hwo, hout)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        plug.eval_pdf_host(hwi, hwo, hout)  # returns when results are in host memory
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e_value = world * n * e2e_steps / e2e_s / 1e9
    checksum = float(hout[3, :: max(1, n // 4096)].double().sum())

    # ---- other kernels of the path (reported, not the headline) --------------------------
    extras = {}
    if not args.no_extras:
        # per-step times of the headline kernel over 200 back-to-back steps (one event pair per
        # step): the first ~50 steps (~30 ms) run at the burst rate, after which the board's power
        # management stretches each step by 7-14 %
        torch.cuda.synchronize()
        time.sleep(0.5)
        ks = 200
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(ks + 1)]
        evs[0].record()
        for i in range(ks):
            plug.eval_pdf(ctx, si, wo, out=out)
            evs[i + 1].record()
        torch.cuda.synchronize()
        per = [evs[i].elapsed_time(evs[i + 1]) for i in range(ks)]
        extras["eval_pdf_step_ms"] = {"first5_mean": sum(per[:5]) / 5, "last5_mean": sum(per[-5:]) / 5,
                                      "min": min(per), "median": sorted(per)[ks // 2], "steps": ks}
        u1, u2 = torch.rand(n, device=dev, generator=gen), torch.rand(2, n, device=dev, generator=gen)
        k = max(3, min(args.steps, 20))
        ms_s, _ = timed(lambda: plug.sample(ctx, si, u1, u2), k, 3)
        extras["sample_gquery_s"] = world * n / (ms_s / k * 1e-3) / 1e9
        extras["sample_hbm_gbs_per_gpu"] = 56 * n / (ms_s / k * 1e-3) / 1e9
        # BASELINE config 3: customized_measurment plugin, 2^24-query eval / sample sweep, L2-resident table
        cdims = (64, 96, 192)
        cplug = bsdf.CustomizedMeasurement(data=synth.synthetic_table(7, dims=cdims, theta_h_sqrt=False), resolution=cdims,
                                           theta_h_mapping="linear", device=f"cuda:{local}")
        m = min(n, 1 << 24)
        csi = bsdf.SurfaceInteraction3f(wi=wi[:, :m].contiguous())
        cwo, cu1, cu2 = wo[:, :m].contiguous(), u1[:m].contiguous(), u2[:, :m].contiguous()
        cout = torch.empty(4, m, device=dev)
        ms_ce, _ = timed(lambda: cplug.eval_pdf(ctx, csi, cwo, out=cout), k, 3)
        ms_cs, _ = timed(lambda: cplug.sample(ctx, csi, cu1, cu2), k, 3)
        extras["customized_2p24"] = {"resolution": list(cdims), "theta_h_mapping": "linear", "queries_per_gpu": m,
                                     "table_bytes": int(cplug.info.table_bytes),
                                     "eval_pdf_gquery_s": world * m / (ms_ce / k * 1e-3) / 1e9,
                                     "sample_gquery_s": world * m / (ms_cs / k * 1e-3) / 1e9}
        del cplug, csi, cwo, cu1, cu2, cout
        del u1, u2
        g = torch.randn(3, n, device=dev, generator=gen)
        grad = torch.zeros(plug.info.ncell, 4, device=dev)
        ms_b, _ = timed(lambda: plug.backward(si, wo, g, grad_cells=grad), k, 3)
        extras["backward_gquery_s_no_allreduce"] = world * n / (ms_b / k * 1e-3) / 1e9
        comm = None
        if world > 1:
            from mitsuba_customization_b200 import distributed as D
            comm = D.Comm(rank, world, local)   # libmerl_b200's NCCL communicator (id shipped over torch.distributed)

            def bwd_ar():
                plug.backward(si, wo, g, grad_cells=grad)
                comm.allreduce_sum_(grad)
            ms_ba, _ = timed(bwd_ar, k, 3)
            extras["backward_gquery_s_with_allreduce"] = world * n / (ms_ba / k * 1e-3) / 1e9
            ms_ar, _ = timed(lambda: comm.allreduce_sum_(grad), k, 3)
            extras["allreduce_ms"] = ms_ar / k
            extras["allreduce_bytes"] = grad.numel() * 4
            # BASELINE config 4: 2^26 queries in TOTAL, sharded over the ranks (strong scaling), one allreduce
            lo, hi = D.shard_bounds(1 << 26, world, rank)
            m = min(hi - lo, n)
            sw = bsdf.SurfaceInteraction3f(wi=wi[:, :m].contiguous())
            swo, sg = wo[:, :m].contiguous(), g[:, :m].contiguous()

            def bwd_strong():
                plug.backward(sw, swo, sg, grad_cells=grad)
                comm.allreduce_sum_(grad)
            ms_bs, _ = timed(bwd_strong, k, 3)
            extras["backward_2p26_total_ms_with_allreduce"] = ms_bs / k
            extras["backward_2p26_total_gquery_s"] = (1 << 26) / (ms_bs / k * 1e-3) / 1e9
            del sw, swo, sg
        del g, grad
        # BASELINE configs 2 and 5 (the metric's spp*Mpix/s @ 1/2/4/8): the wavefront harness around the
        # BSDF stage; every rank renders its share of the sample indices, films summed with one allreduce.
        # A failure here is reported in the line, it never takes the headline down.
        if not args.no_render:
            try:
                extras["render"] = render_legs(plug, world, rank, local, comm, timed, args.render_tables)
            except Exception as e:  # noqa: BLE001
                extras["render"] = {"error": f"{type(e).__name__}: {e}"}
        if comm is not None:
            comm.close()

    if rank == 0:
        peak, peak_src = measured_peak()
        achieved = BYTES_PER_QUERY * n / (kernel_ms * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": W,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "queries_per_gpu_per_step": n, "table": "synthetic MERL id 0",
                       "sharding": f"query batch x{world}, table replicated, no data-path collective",
                       "l2": "no flush needed: 1.6 GB in + 1.07 GB out per step >> 126 MB L2; the 23 MB table is "
                             "meant to stay L2-resident (persisting access window)"},
            "clocks": sampler.summary(),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 24 * n, "d2h_bytes_per_step": 16 * n,
                    "steps": e2e_steps, "call": "MERL.eval_pdf_host -> merl_eval_pdf_host (pinned host buffers)",
                    "result_checksum": checksum},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": recorded_traffic(), "peak_source": peak_src,
                         "algorithmic_bytes_per_query": BYTES_PER_QUERY, "kernel": "stream_kernel<QueryOp<EVAL_PDF,4>,3>",
                         "kernel_ms": kernel_ms},
            "extras": extras,
        }
        if world == 1 and not args.no_cpu_baseline:
            v, cores, sample = cpu_port_rate()
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def render_legs(plug, world, rank, local, comm, timed, ntab):
    """spp*Mpix/s of BASELINE configs 2 (512^2, 64 spp, one sphere, one table) and 5 (1024^2, 256 spp,
    10 x 10 spheres, `ntab` tables, all replicated) through merl_render; whole-job figure over all ranks."""
    import torch
    from mitsuba_customization_b200 import bsdf, render as RD, synth
    dev = f"cuda:{local}"
    out = {}

    def leg(name, sc, w, h, spp, depth, reps):
        rng = RD.spp_share(spp, world, rank)
        film = torch.zeros(h, w, 4, device=dev)

        def step():
            film.zero_()
            RD.render(sc, w, h, spp, max_depth=depth, seed=1, spp_range=rng, film=film)
            if comm is not None:
                comm.allreduce_sum_(film)
        ms, _ = timed(step, reps, 2)
        ms /= reps
        out[name] = {"width": w, "height": h, "spp": spp, "max_depth": depth, "ms": ms,
                     "spp_mpix_per_s": spp * w * h / 1e6 / (ms * 1e-3),
                     "mean_radiance": [float(x) for x in film[..., :3].mean(dim=(0, 1))],
                     "sample_share_min_max": [float(film[..., 3].min()), float(film[..., 3].max())]}
        del film

    leg("cfg2", RD.sphere_scene(plug), 512, 512, 64, 8, 5)
    if ntab > 0:
        t0 = time.time()
        plugs = [bsdf.MERL(data=synth.synthetic_table(k), device=dev) for k in range(ntab)]
        build_s = time.time() - t0
        leg("cfg5", RD.grid_scene(plugs, 10, 10), 1024, 1024, 256, 8, 3)
        out["cfg5"].update({"tables": ntab, "table_bytes_per_gpu": int(sum(p.info.table_bytes for p in plugs)),
                            "table_build_s": build_s})
        del plugs
    return out


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
