#!/usr/bin/env python
"""bench.py -- MERL BSDF eval+pdf throughput on B200 (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

A "step" is one pass of the hot path (batched eval+pdf through the plugin's eval_pdf, i.e.
through the C ABI) over one batch of 2^26 synthetic (wi, wo) pairs per GPU against one
synthetic MERL table.  One JSON line is printed by rank 0:

  value      whole-job Gquery/s with the batch resident in HBM (CUDA events, max over ranks), timed
             over exactly K steps that follow a barrier and then W + `--preheat` untimed steps, i.e. in the SUSTAINED
             state a renderer holds the kernel in (the first ~30 ms after an idle period run 7-14 %
             faster: that burst figure is extras.eval_pdf_burst)
  e2e        the same metric through merl_eval_pdf_host with PINNED HOST buffers: every step
             copies the 24 B/query of inputs host->device and the 16 B/query of results back
  roofline   achieved algorithmic HBM GB/s (40 B/query) of the eval_pdf kernel
             against the measured copy bandwidth in MEASURED_PEAKS.json
  cpu_baseline  the CPU oracle (kind "port": the reference has no source to run) on a bounded
             sample of the same workload, all host threads, plus BASELINE config 1 (2^20 pairs,
             eval+pdf+sample, single thread and all threads); run in a SUBPROCESS so that the GPU
             arm's process never loads anything under oracle/
  extras.grad_scaling   the sharded gradient pass (backward + exchange): peer-memory kernel vs NCCL,
             one exchange per 1 / 4 / 16 batches, BASELINE config 4, and an on-box correctness check
  extras.render         BASELINE configs 2 and 5 (spp*Mpix/s), and config 5 sharded by table

--impl reference times that CPU port alone on the SAME step (all 2^26 queries per step, --steps and
--warmup as given; the reference's own plugins cannot be built: /root/reference holds no source,
see DESIGN.md §7).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
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
    ap.add_argument("--preheat", type=int, default=150,
                    help="extra untimed steps right before the timed ones (sustained state); 0 = burst figure")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--log2-batch", type=int, default=LOG2_BATCH)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true")
    ap.add_argument("--no-render", action="store_true", help="skip the render legs of the extras")
    ap.add_argument("--render-tables", type=int, default=100, help="tables of render config 5 (0 = skip it)")
    ap.add_argument("--cpu-legs", action="store_true", help="(internal) print the CPU legs as JSON and exit")
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
# CPU arm: the oracle port on the box's host cores (bounded samples)
# --------------------------------------------------------------------------------------------
def host_threads():
    """Threads this process may use: the affinity mask, NOT OMP_NUM_THREADS (torchrun exports
    OMP_NUM_THREADS=1 to its workers, which made the round-1 reference arm single-threaded at N > 1)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


def _rate(fn, n, budget_s, max_reps=256):
    fn()
    reps, t0 = 0, time.perf_counter()
    while True:
        fn()
        reps += 1
        dt = time.perf_counter() - t0
        if dt >= budget_s or reps >= max_reps:
            return n * reps / dt / 1e9, reps, dt


def cpu_legs(budget_s=8.0):
    """All CPU figures of the line.  Runs in its own process (bench.py --cpu-legs)."""
    import numpy as np
    from mitsuba_customization_b200 import synth
    from oracle import merl_oracle as O

    cores = host_threads()
    raw = synth.synthetic_table(0)
    t = O.OracleTable(raw)
    t.build_cdf()
    out = {"cores": cores}
    # headline sample: 2^22 of the step's 2^26 queries, all threads
    n = 1 << 22
    wi, wo = synth.random_queries(n, seed=1234)
    O.set_threads(cores)
    v, reps, dt = _rate(lambda: t.eval_pdf(wi, wo), n, budget_s)
    out["eval_pdf_all_threads"] = {"gquery_s": v, "sample": f"{reps} x 2^22 eval+pdf queries of the same workload, {dt:.1f} s"}
    # BASELINE config 1: 2^20 pairs, eval + pdf + sample, single thread and all threads
    m = 1 << 20
    rng = np.random.default_rng(3)
    u1 = rng.random(m, dtype=np.float32)
    u2 = rng.random((2, m), dtype=np.float32)
    a, b = wi[:, :m].copy(), wo[:, :m].copy()
    cfg1 = {"pairs": m}
    for name, nt in (("single_thread", 1), ("all_threads", cores)):
        O.set_threads(nt)
        ve, _, _ = _rate(lambda: t.eval_pdf(a, b), m, 2.0, 64)
        vs, _, _ = _rate(lambda: t.sample(a, u1, u2), m, 2.0, 64)
        cfg1[name] = {"threads": nt, "eval_pdf_gquery_s": ve, "sample_gquery_s": vs,
                      "eval_pdf_sample_ms": m / ve / 1e6 + m / vs / 1e6}
    out["config1_2p20"] = cfg1
    out["render"] = cpu_render_legs(t, cores)
    return out


def cpu_render_legs(t0, cores, budget_s=6.0):
    """The render half of the metric on the host cores: BASELINE configs 2 and 5 through oracle/render_oracle.c
    (the CPU statement of the algorithm merl_render runs), a bounded share of the samples, all threads."""
    import numpy as np
    from mitsuba_customization_b200 import render as RD, synth
    from oracle import merl_oracle as O, render_oracle as RO

    def leg(tables, sc, w, h, spp, depth, nthreads, budget):
        O.set_threads(nthreads)
        kw = dict(width=w, height=h, spp=spp, max_depth=depth, seed=1, cam_pos=sc.cam_pos, cam_target=sc.cam_target,
                  cam_up=sc.cam_up, fov_x_deg=sc.fov_x_deg, env=sc.env)
        a = time.perf_counter()
        RO.render(tables, sc.spheres, spp_begin=0, spp_end=1, **kw)        # 1 spp: sizes the sample (and warms up)
        per_spp = time.perf_counter() - a
        cs = int(max(1, min(spp, budget / max(per_spp, 1e-6))))
        a = time.perf_counter()
        film = RO.render(tables, sc.spheres, spp_begin=0, spp_end=cs, **kw)
        dt = time.perf_counter() - a
        return {"spp_mpix_per_s": cs * w * h / 1e6 / dt, "threads": nthreads, "width": w, "height": h, "max_depth": depth,
                "sample": f"{cs} of {spp} spp, {dt:.2f} s", "mean_radiance_of_sample": float(film[..., :3].mean()) * spp / cs}

    out = {}
    sc2 = RD.sphere_scene(None)
    out["cfg2_all_threads"] = leg([t0], sc2, 512, 512, 64, 8, cores, budget_s)
    out["cfg2_single_thread"] = leg([t0], sc2, 512, 512, 64, 8, 1, 2.0)
    # config 5: 100 tables that are distinct in memory (35 MB of doubles each).  Synthesising 100 takes minutes on
    # the host, so: 10 synthetic ids x 10 brightness factors, each its own array and its own CDF.
    base = [synth.synthetic_table(k) for k in range(10)]
    tabs = []
    for k in range(100):
        tk = O.OracleTable(np.ascontiguousarray(base[k % 10] * (0.55 + 0.05 * (k // 10))))
        tk.build_cdf()
        tabs.append(tk)
    sc5 = RD.grid_scene([None] * 100, 10, 10)
    out["cfg5_all_threads"] = leg(tabs, sc5, 1024, 1024, 256, 8, cores, budget_s)
    out["cfg5_all_threads"]["tables"] = "100 distinct arrays: 10 synthetic ids x 10 brightness factors"
    O.set_threads(cores)
    return out


def cpu_legs_subprocess():
    env = dict(os.environ)
    for k in ("OMP_NUM_THREADS", "RANK", "LOCAL_RANK", "WORLD_SIZE"):
        env.pop(k, None)
    r = subprocess.run([sys.executable, os.path.abspath(__file__), "--cpu-legs"], capture_output=True, text=True, env=env,
                       cwd=ROOT, timeout=600)
    if r.returncode != 0:
        raise RuntimeError("cpu legs failed: " + r.stderr[-400:])
    return json.loads(r.stdout.strip().splitlines()[-1])


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import numpy as np  # noqa: F401
    from mitsuba_customization_b200 import synth
    from oracle import merl_oracle as O

    cores = host_threads()
    O.set_threads(cores)
    raw = synth.synthetic_table(0)
    t = O.OracleTable(raw)
    t.build_cdf()
    # one step = the GPU arm's whole step (2^log2_batch queries), walked in 2^22-query calls so that the
    # oracle's fp64 result arrays stay at 150 MB per call instead of 2.4 GB of fresh pages per step
    n = 1 << args.log2_batch
    chunk = min(n, 1 << 22)
    wi, wo = synth.random_queries(n, seed=1234)

    def step():
        for o in range(0, n, chunk):
            t.eval_pdf(wi[:, o:o + chunk], wo[:, o:o + chunk])

    steps = max(1, args.steps)
    warm = max(0, args.warmup)
    for _ in range(warm):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = time.perf_counter() - t0
    v = n * steps / dt / 1e9
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warm, "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "queries_per_gpu_per_step": n, "table": "synthetic MERL id 0",
                   "step_sample": f"the whole step: 2^{args.log2_batch} queries per step, in {n // chunk} calls of 2^22"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{steps} steps x 2^{args.log2_batch} eval+pdf queries (the full step), OpenMP over all "
                                   f"{cores} host threads of the affinity mask (OMP_NUM_THREADS ignored); the reference's own "
                                   "plugins are absent from /root/reference, so this is the CPU restatement "
                                   "(oracle/merl_oracle.c), not the reference binary"},
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
               "samples": len(s), "window": "the preheat + timed region of the headline metric"}
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
        # barrier FIRST, then the untimed steps, then the start event with no host round trip in between:
        # a barrier (dist.barrier + device sync) idles the board long enough to leave its power cap, so
        # placed after the preheat it made N >= 2 report the burst rate and N = 1 the sustained one
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(warmup):
            fn()
        l0 = _cabi.launch_count()
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

    # ---- headline: device-resident eval+pdf, sustained ------------------------------------------
    W = max(args.warmup, 3)
    for _ in range(W):                      # warm up before the clock sampler's window opens
        plug.eval_pdf(ctx, si, wo, out=out)
    torch.cuda.synchronize()
    sampler.start()
    ms, launches = timed(lambda: plug.eval_pdf(ctx, si, wo, out=out), args.steps, max(0, args.preheat))
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
    # what the host link alone allows: the same bytes (24 B/query up, 16 B/query down) as plain pinned copies on two
    # streams, no kernel -- if e2e sits at this figure the path is bound by PCIe / host memory, not by the library
    s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()
    d_in, d_out = torch.empty(6, n, device=dev), torch.empty(4, n, device=dev)
    h_in = torch.empty(6, n).pin_memory()

    def copies():
        with torch.cuda.stream(s_up):
            d_in.copy_(h_in, non_blocking=True)
        with torch.cuda.stream(s_dn):
            hout.copy_(d_out, non_blocking=True)
    copies(); torch.cuda.synchronize(); barrier()
    t0 = time.perf_counter()
    for _ in range(4):
        copies()
    torch.cuda.synchronize()
    cp_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([cp_s], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        cp_s = float(t.item())
    e2e_copy_ceiling = world * n * 4 / cp_s / 1e9
    del hwi, hwo, hout, h_in, d_in, d_out

    # ---- other kernels of the path (reported, not the headline) --------------------------
    extras = {}
    peak, peak_src = measured_peak()
    if not args.no_extras:
        # burst: what a caller sees for the first ~20 launches after an idle period
        torch.cuda.synchronize()
        time.sleep(0.5)
        ms_b, _ = timed(lambda: plug.eval_pdf(ctx, si, wo, out=out), 20, 0)
        extras["eval_pdf_burst"] = {"gquery_s_per_gpu": n / (ms_b / 20 * 1e-3) / 1e9, "steps": 20, "after_idle_s": 0.5,
                                    "frac_of_hbm": BYTES_PER_QUERY * n / (ms_b / 20 * 1e-3) / 1e9 / peak}
        # per-step times over 200 back-to-back steps after the same idle period (one event pair per step)
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

        def sm_clock():
            try:
                return int(sampler.nv.nvmlDeviceGetClockInfo(sampler.h, sampler.nv.NVML_CLOCK_SM))
            except Exception:
                return None
        # the sampler is issue-bound, so it follows the SM clock: measured (a) right after the 200-step eval+pdf
        # hold above, i.e. under the power cap that hold provokes (this is the state round 1's 57 Gquery/s on the
        # scaling node came from), and (b) after an idle period, as tools/one_kernel.py measures it
        ms_h, _ = timed(lambda: plug.sample(ctx, si, u1, u2), k, 3)
        clk_h = sm_clock()
        torch.cuda.synchronize()
        time.sleep(1.0)
        ms_s, _ = timed(lambda: plug.sample(ctx, si, u1, u2), k, 3)
        clk_s = sm_clock()
        extras["sample_gquery_s"] = world * n / (ms_s / k * 1e-3) / 1e9
        extras["sample_hbm_gbs_per_gpu"] = 56 * n / (ms_s / k * 1e-3) / 1e9
        extras["sample_frac_of_hbm"] = extras["sample_hbm_gbs_per_gpu"] / peak
        extras["sample_states"] = {"after_idle": {"gquery_s_per_gpu": n / (ms_s / k * 1e-3) / 1e9, "sm_mhz_at_end": clk_s},
                                   "after_200_step_eval_pdf_hold": {"gquery_s_per_gpu": n / (ms_h / k * 1e-3) / 1e9,
                                                                    "sm_mhz_at_end": clk_h}}
        # BASELINE config 1 on the GPU: 2^20 pairs, eval+pdf then sample (launch-latency dominated: 42 MB is
        # ~6 us at the roofline; reported, not used for any roofline fraction)
        m20 = 1 << 20
        s20 = bsdf.SurfaceInteraction3f(wi=wi[:, :m20].contiguous())
        wo20, u120, u220, out20 = wo[:, :m20].contiguous(), u1[:m20].contiguous(), u2[:, :m20].contiguous(), torch.empty(4, m20, device=dev)
        ms_e20, _ = timed(lambda: plug.eval_pdf(ctx, s20, wo20, out=out20), 50, 5)
        ms_s20, _ = timed(lambda: plug.sample(ctx, s20, u120, u220), 50, 5)
        extras["config1_2p20_gpu"] = {"pairs": m20, "eval_pdf_us": ms_e20 / 50 * 1e3, "sample_us": ms_s20 / 50 * 1e3,
                                      "eval_pdf_gquery_s": m20 / (ms_e20 / 50 * 1e-3) / 1e9,
                                      "sample_gquery_s": m20 / (ms_s20 / 50 * 1e-3) / 1e9}
        del s20, wo20, u120, u220, out20
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
                                     "eval_pdf_frac_of_hbm": BYTES_PER_QUERY * m / (ms_ce / k * 1e-3) / 1e9 / peak,
                                     "sample_gquery_s": world * m / (ms_cs / k * 1e-3) / 1e9}
        del cplug, csi, cwo, cu1, cu2, cout
        del u1, u2
        comm = None
        if world > 1:
            from mitsuba_customization_b200 import distributed as D
            comm = D.Comm(rank, world, local)   # libmerl_b200's NCCL communicator (id shipped over torch.distributed)
        try:
            extras["grad_scaling"] = grad_legs(plug, si, wo, n, world, rank, local, comm, timed, gen, k)
        except Exception as e:  # noqa: BLE001
            extras["grad_scaling"] = {"error": f"{type(e).__name__}: {e}"}
        # BASELINE configs 2 and 5 (the metric's spp*Mpix/s @ 1/2/4/8): the wavefront harness around the
        # BSDF stage; every rank renders its share of the sample indices, films summed with one exchange.
        # A failure here is reported in the line, it never takes the headline down.
        if not args.no_render:
            try:
                extras["render"] = render_legs(plug, world, rank, local, comm, timed, args.render_tables)
            except Exception as e:  # noqa: BLE001
                extras["render"] = {"error": f"{type(e).__name__}: {e}"}
        if comm is not None:
            comm.close()

    if rank == 0:
        achieved = BYTES_PER_QUERY * n / (kernel_ms * 1e-3) / 1e9
        roof = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": recorded_traffic(), "peak_source": peak_src,
                "algorithmic_bytes_per_query": BYTES_PER_QUERY, "kernel": "stream_kernel<QueryOp<EVAL_PDF,4>,3>",
                "kernel_ms": kernel_ms,
                "state": f"sustained: timed after {W} + {max(0, args.preheat)} untimed steps" if args.preheat > 0 else "burst"}
        if args.preheat > 0:
            roof["frac_sustained"] = achieved / peak
        if "eval_pdf_burst" in extras:
            roof["frac_burst"] = extras["eval_pdf_burst"]["frac_of_hbm"]
        if "customized_2p24" in extras:
            roof["frac_config3_customized_2p24"] = extras["customized_2p24"]["eval_pdf_frac_of_hbm"]
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": W, "preheat": max(0, args.preheat),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "queries_per_gpu_per_step": n, "table": "synthetic MERL id 0",
                       "sharding": f"query batch x{world}, table replicated, no data-path collective",
                       "l2": "no flush needed: 1.6 GB in + 1.07 GB out per step >> 126 MB L2; the 23 MB table is "
                             "meant to stay L2-resident (persisting access window)"},
            "clocks": sampler.summary(),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 24 * n, "d2h_bytes_per_step": 16 * n,
                    "steps": e2e_steps, "call": "MERL.eval_pdf_host -> merl_eval_pdf_host (pinned host buffers)",
                    "result_checksum": checksum,
                    "host_link_ceiling": {"value": e2e_copy_ceiling, "unit": UNIT,
                                          "what": "the same 24 B/query up + 16 B/query down as plain pinned cudaMemcpyAsync on "
                                                  "two streams per rank, all ranks at once, no kernel"}},
            "gpu_launches": int(launches),
            "roofline": roof,
            "extras": extras,
        }
        if world == 1 and not args.no_cpu_baseline:
            try:
                c = cpu_legs_subprocess()
                line["cpu_baseline"] = {"value": c["eval_pdf_all_threads"]["gquery_s"], "unit": UNIT, "cores": c["cores"],
                                        "kind": "port", "sample": c["eval_pdf_all_threads"]["sample"],
                                        "single_thread": c["config1_2p20"]["single_thread"]["eval_pdf_gquery_s"],
                                        "sample_gquery_s": c["config1_2p20"]["all_threads"]["sample_gquery_s"],
                                        "config1_2p20": c["config1_2p20"],
                                        "render_cfg2_spp_mpix_s": c["render"]["cfg2_all_threads"]["spp_mpix_per_s"],
                                        "render_cfg5_spp_mpix_s": c["render"]["cfg5_all_threads"]["spp_mpix_per_s"],
                                        "render": c["render"],
                                        "process": "subprocess (bench.py --cpu-legs): the GPU arm never loads oracle/"}
            except Exception as e:  # noqa: BLE001
                line["cpu_baseline"] = {"error": f"{type(e).__name__}: {e}"}
        # last key of the line, so that a tail of the output still holds the multi-GPU record
        line["scaling_summary"] = scaling_summary(extras, world)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def scaling_summary(extras, world):
    """A compact (< 1200 characters) digest of extras.grad_scaling / extras.render: what a reader of the last
    1500 characters of the line needs to judge the gradient pass and the render legs."""
    def r3(x):
        return None if x is None else round(float(x), 4)
    s = {"n": world}
    g = extras.get("grad_scaling") or {}
    if "error" in g:
        s["grad_error"] = str(g["error"])[:120]
    elif g:
        s["bwd_gq_s_per_gpu"] = r3(g.get("backward_gquery_s_per_gpu_no_exchange"))
        s["allreduce_check"] = g.get("allreduce_check", "n/a (1 GPU)")
        tr = g.get("transports") or {}
        best = max((t for t in tr if "weak_efficiency_vs_no_exchange_every_1" in tr[t]),
                   key=lambda t: tr[t]["weak_efficiency_vs_no_exchange_every_1"], default=None)
        if best:
            b = tr[best]
            s["grad"] = {"transport": best, "xchg_us": r3(b.get("exchange_ms", 0) * 1e3),
                         "eff_every": [r3(b.get(f"weak_efficiency_vs_no_exchange_every_{m}")) for m in (1, 4, 16)]}
        c4 = g.get("config4_2p26_total") or {}
        ms = [v for k2, v in c4.items() if k2.endswith("_ms") and k2 != "scatter_only_ms"]
        if ms:
            s["cfg4_ms"] = r3(min(ms))
    r = extras.get("render") or {}
    if "error" in r:
        s["render_error"] = str(r["error"])[:120]
    for name in ("cfg2", "cfg5"):
        if name in r:
            s[name] = {"spp_mpix_s": round(r[name]["spp_mpix_per_s"], 1),
                       "film_check": r[name].get("film_check", {}).get("verdict"),
                       "rel_mse": r[name].get("film_check", {}).get("rel_mse_exchanged_vs_one_gpu_render")}
    if "cfg5_by_table" in r:
        s["cfg5_by_table_rel_mse"] = r["cfg5_by_table"].get("rel_mse_vs_replicated_same_samples")
    return s


def grad_legs(plug, si, wo, n, world, rank, local, comm, timed, gen, k):
    """The sharded gradient pass (BASELINE config 4; north_star's scaling target): scatter-add of this rank's
    query shard into its private gradient table + ONE exchange that sums the tables over the ranks."""
    import torch
    import torch.distributed as dist
    from mitsuba_customization_b200 import bsdf
    from mitsuba_customization_b200 import distributed as D
    dev = torch.device("cuda", local)
    ncell = plug.info.ncell
    g = torch.randn(3, n, device=dev, generator=gen)
    out = {"queries_per_gpu_per_batch": n, "exchange_bytes": ncell * 16}
    grad_plain = torch.zeros(ncell, 4, device=dev)
    ms_b, _ = timed(lambda: plug.backward(si, wo, g, grad_cells=grad_plain), k, 3)
    out["backward_gquery_s_no_exchange"] = world * n / (ms_b / k * 1e-3) / 1e9
    out["backward_gquery_s_per_gpu_no_exchange"] = n / (ms_b / k * 1e-3) / 1e9
    out["backward_frac_of_hbm"] = 36 * n / (ms_b / k * 1e-3) / 1e9 / measured_peak()[0]
    if comm is None:
        # one GPU: the deferred-mask form the multi-GPU path uses (mask applied once per exchange)
        ms_d, _ = timed(lambda: (plug.backward(si, wo, g, grad_cells=grad_plain, defer_mask=True),
                                 plug.apply_grad_mask(grad_plain)), k, 3)
        out["backward_gquery_s_deferred_mask"] = n / (ms_d / k * 1e-3) / 1e9
        return out

    transports = {}
    bufs = {"peer": D.PeerBuffer(comm, 4 * ncell, "peer"), "nccl_registered": D.PeerBuffer(comm, 4 * ncell, "nccl")}
    try:   # the library's own multimem kernel through the switch; every rank gets the same verdict
        bufs["nvls"] = D.PeerBuffer(comm, 4 * ncell, "nvls")
    except bsdf.MerlError as e:
        out["nvls_unavailable"] = str(e)[:160]
    for name, pb in bufs.items():
        gt = pb.tensor((ncell, 4))
        res = {"registration": pb.registration}
        ms_x, _ = timed(lambda: pb.allreduce_(mask_plug=plug), k, 3)
        res["exchange_ms"] = ms_x / k
        res["exchange_bus_gbs"] = 2 * (world - 1) / world * ncell * 16 / (ms_x / k * 1e-3) / 1e9
        for m in (1, 4, 16):
            def step():
                for _ in range(m):
                    plug.backward(si, wo, g, grad_cells=gt, defer_mask=True)
                pb.allreduce_(mask_plug=plug)
            reps = max(2, k // m)
            ms_w, _ = timed(step, reps, 2)
            v = world * n * m / (ms_w / reps * 1e-3) / 1e9
            res[f"weak_gquery_s_exchange_every_{m}"] = v
            res[f"weak_efficiency_vs_no_exchange_every_{m}"] = v / out["backward_gquery_s_no_exchange"]
        transports[name] = res
    # round 1's form, for comparison: ncclAllReduce on an unregistered torch tensor + separate mask pass
    def plain():
        plug.backward(si, wo, g, grad_cells=grad_plain)
        comm.allreduce_sum_(grad_plain)
    ms_p, _ = timed(plain, k, 3)
    ms_px, _ = timed(lambda: comm.allreduce_sum_(grad_plain), k, 3)
    transports["nccl_unregistered"] = {"exchange_ms": ms_px / k,
                                       "exchange_bus_gbs": 2 * (world - 1) / world * ncell * 16 / (ms_px / k * 1e-3) / 1e9,
                                       "weak_gquery_s_exchange_every_1": world * n / (ms_p / k * 1e-3) / 1e9,
                                       "weak_efficiency_vs_no_exchange_every_1":
                                           world * n / (ms_p / k * 1e-3) / 1e9 / out["backward_gquery_s_no_exchange"]}
    out["transports"] = transports

    # BASELINE config 4: 2^26 queries in TOTAL, sharded over the ranks (strong scaling), one exchange
    lo, hi = D.shard_bounds(1 << 26, world, rank)
    m4 = min(hi - lo, n)
    sw = bsdf.SurfaceInteraction3f(wi=si.wi[:, :m4].contiguous())
    swo, sg = wo[:, :m4].contiguous(), g[:, :m4].contiguous()
    cfg4 = {"total_queries": 1 << 26, "queries_per_gpu": m4}
    for name, pb in bufs.items():
        gt = pb.tensor((ncell, 4))

        def strong():
            plug.backward(sw, swo, sg, grad_cells=gt, defer_mask=True)
            pb.allreduce_(mask_plug=plug)
        ms_s, _ = timed(strong, k, 3)
        cfg4[f"{name}_ms"] = ms_s / k
        cfg4[f"{name}_gquery_s"] = (1 << 26) / (ms_s / k * 1e-3) / 1e9
    ms_c, _ = timed(lambda: plug.backward(sw, swo, sg, grad_cells=grad_plain, defer_mask=True), k, 3)
    cfg4["scatter_only_ms"] = ms_c / k
    out["config4_2p26_total"] = cfg4

    # correctness of the exchange on this box: every rank scatters ITS shard of 2^22 x world seeded queries
    # and exchanges; rank-local recomputation of ALL shards must agree, hit counts must sum exactly
    mq = 1 << 22
    checks = {}
    full = torch.zeros(ncell, 4, device=dev)
    shards = []
    for r in range(world):
        gg = torch.Generator(device=dev)
        gg.manual_seed(777 + r)
        a, b = hemisphere(mq, gg, dev), hemisphere(mq, gg, dev)
        c = torch.randn(3, mq, device=dev, generator=gg)
        plug.backward(bsdf.SurfaceInteraction3f(wi=a), b, c, grad_cells=full)
        if r == rank:
            shards = [a, b, c]
    for name, pb in bufs.items():
        gt = pb.tensor((ncell, 4))
        gt.zero_()
        plug.backward(bsdf.SurfaceInteraction3f(wi=shards[0]), shards[1], shards[2], grad_cells=gt, defer_mask=True)
        pb.allreduce_(mask_plug=plug)
        torch.cuda.synchronize()
        hits = float(gt[:, 3].double().sum())
        scale = float(full[:, :3].abs().max())
        err = float((gt[:, :3] - full[:, :3]).abs().max()) / max(scale, 1e-30)
        sums = torch.tensor([float(gt.double().sum())], device=dev, dtype=torch.float64)
        allsums = [torch.zeros_like(sums) for _ in range(world)]
        dist.all_gather(allsums, sums)
        same = all(float(x) == float(allsums[0]) for x in allsums)
        ok = hits == float(world * mq) and err <= 2e-5 and pb.status() == 0 and (same or name != "peer")
        checks[name] = {"ok": bool(ok), "hit_count_sum": hits, "expected_hits": world * mq, "max_err_over_max": err,
                        "bitwise_same_on_all_ranks": bool(same)}
    flag = torch.tensor([1.0 if all(c["ok"] for c in checks.values()) else 0.0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    out["allreduce_check"] = "ok" if float(flag) == 1.0 else "FAILED"
    out["allreduce_check_detail"] = checks
    for pb in bufs.values():
        pb.close()
    if out["allreduce_check"] != "ok":
        raise RuntimeError(f"gradient exchange check failed: {checks}")
    return out


def render_legs(plug, world, rank, local, comm, timed, ntab):
    """spp*Mpix/s of BASELINE configs 2 (512^2, 64 spp, one sphere, one table) and 5 (1024^2, 256 spp,
    10 x 10 spheres, `ntab` tables, all replicated) through merl_render; whole-job figure over all ranks.
    Multi-GPU: films live in an exchange buffer and are summed by the peer-memory kernel."""
    import torch
    from mitsuba_customization_b200 import bsdf, render as RD, synth
    from mitsuba_customization_b200 import distributed as D
    dev = f"cuda:{local}"
    out = {}

    def leg(name, sc, w, h, spp, depth, reps):
        rng = RD.spp_share(spp, world, rank)
        emu = int(os.environ.get("RD_EMULATE_WORLD", "0"))   # diagnostics: one GPU doing rank 0's share of an emu-GPU render
        if emu > 1 and world == 1:
            rng = RD.spp_share(spp, emu, 0)
        pb = D.PeerBuffer(comm, w * h * 4, "peer") if comm is not None else None
        film = pb.tensor((h, w, 4)) if pb is not None else torch.zeros(h, w, 4, device=dev)

        def step():
            film.zero_()
            RD.render(sc, w, h, spp, max_depth=depth, seed=1, spp_range=rng, film=film)
            if pb is not None:
                pb.allreduce_()
        ms, _ = timed(step, reps, 2)
        ms /= reps
        # the film as the last timed frame left it, i.e. AFTER the exchange: taken before the render-only loops
        # below overwrite it with this rank's partial film
        torch.cuda.synchronize()
        mean_radiance = [float(x) for x in film[..., :3].mean(dim=(0, 1))]
        share = [float(film[..., 3].min()), float(film[..., 3].max())]
        film_check = {"verdict": "n/a (1 GPU: nothing exchanged)"}
        if comm is not None:
            # this rank alone renders ALL sample indices of the same seed; the exchanged film must equal it up to
            # the order of the float additions
            solo = torch.zeros(h, w, 4, device=dev)
            RD.render(sc, w, h, spp, max_depth=depth, seed=1, film=solo)
            torch.cuda.synchronize()
            num = float(((film[..., :3] - solo[..., :3]) ** 2).mean())
            den = float((solo[..., :3] ** 2).mean())
            rel = num / max(den, 1e-30)
            film_check = {"rel_mse_exchanged_vs_one_gpu_render": rel, "bar": 1e-6, "verdict": "ok" if rel <= 1e-6 else "FAILED"}
            del solo
        per_rank = None
        if comm is not None:
            # each rank's own render time, no exchange and no barrier in the loop: is the frame time the slowest
            # rank's render, or the waiting at the exchange?
            import torch.distributed as dist
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                film.zero_()
                RD.render(sc, w, h, spp, max_depth=depth, seed=1, spp_range=rng, film=film)
            e1.record()
            torch.cuda.synchronize()
            t_default = e0.elapsed_time(e1) / reps
            e0.record()
            for _ in range(reps):       # the same in 8 Mi-path passes (CUDA-graph replay path of merl_render)
                film.zero_()
                RD.render(sc, w, h, spp, max_depth=depth, seed=1, spp_range=rng, film=film, wavefront=1 << 23)
            e1.record()
            torch.cuda.synchronize()
            mine = torch.tensor([t_default, e0.elapsed_time(e1) / reps], device=dev, dtype=torch.float64)
            allr = [torch.zeros_like(mine) for _ in range(world)]
            dist.all_gather(allr, mine)
            per_rank = {"default_pass_size": [round(float(x[0]), 4) for x in allr],
                        "passes_of_8Mi": [round(float(x[1]), 4) for x in allr],
                        "free_gib": round(torch.cuda.mem_get_info()[0] / 2**30, 1)}
        out[name] = {"width": w, "height": h, "spp": spp, "max_depth": depth, "ms": ms,
                     "render_only_ms_per_rank": per_rank,
                     "spp_mpix_per_s": spp * w * h / 1e6 / (ms * 1e-3),
                     "mean_radiance": mean_radiance, "sample_share_min_max": share, "film_check": film_check,
                     "film_exchange": "peer-memory kernel" if pb is not None else "none (1 GPU)"}
        del film
        if pb is not None:
            pb.close()

    leg("cfg2", RD.sphere_scene(plug), 512, 512, 64, 8, 5)
    if ntab > 0:
        t0 = time.time()
        plugs = [bsdf.MERL(data=synth.synthetic_table(k), device=dev) for k in range(ntab)]
        build_s = time.time() - t0
        sc5 = RD.grid_scene(plugs, 10, 10)
        leg("cfg5", sc5, 1024, 1024, 256, 8, 3)
        out["cfg5"].update({"tables": ntab, "table_bytes_per_gpu": int(sum(p.info.table_bytes for p in plugs)),
                            "table_build_s": build_s, "sharding": "samples (tables replicated)"})
        if comm is not None:
            # the by-table variant (BASELINE config 5 "sharded by table"): every rank traces everything, evaluates
            # only its own tables, one exchange of the stage outputs per bounce.  Fewer samples: it is ~world x slower.
            wave = 1 << 21
            xb = D.PeerBuffer(comm, 12 * wave, "peer")
            own = [p if kk % world == rank else None for kk, p in enumerate(plugs)]
            scb = RD.grid_scene(own, 10, 10)
            film_b = torch.zeros(1024, 1024, 4, device=dev)
            film_r = torch.zeros(1024, 1024, 4, device=dev)
            spp_b = 16

            def step_b():
                film_b.zero_()
                RD.render(scb, 1024, 1024, spp_b, max_depth=8, seed=1, film=film_b, wavefront=wave,
                          table_shard=(world, rank), stage_exchange=xb)
            ms_b, _ = timed(step_b, 2, 1)
            ms_b /= 2
            RD.render(sc5, 1024, 1024, spp_b, max_depth=8, seed=1, film=film_r, wavefront=wave)   # replicated, this rank alone
            torch.cuda.synchronize()
            num = float(((film_b[..., :3] - film_r[..., :3]) ** 2).mean())
            den = float((film_r[..., :3] ** 2).mean())
            out["cfg5_by_table"] = {"spp": spp_b, "ms": ms_b, "spp_mpix_per_s": spp_b * 1024 * 1024 / 1e6 / (ms_b * 1e-3),
                                    "tables_per_gpu": sum(1 for p in own if p is not None),
                                    "table_bytes_per_gpu": int(sum(p.info.table_bytes for p in own if p is not None)),
                                    "rel_mse_vs_replicated_same_samples": num / max(den, 1e-30),
                                    "stage_exchange_bytes_per_bounce": 48 * wave,
                                    "ratio_to_sample_sharding": (spp_b * 1024 * 1024 / 1e6 / (ms_b * 1e-3)) / out["cfg5"]["spp_mpix_per_s"]}
            xb.close()
        del plugs
    return out


def main():
    args = parse()
    if args.cpu_legs:
        print(json.dumps(cpu_legs()), flush=True)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
