"""What the tabulated half-vector lobe buys over the cosine lobe alone (north_star asked for
per-(theta_d, phi_d) CDFs; this build conditions a theta_h CDF on the incident elevation instead -- DESIGN.md section 4).
CPU only: the estimator's variance is a property of the sampling scheme, so the oracle measures it.
For each table and incident elevation, X = luminance(f cos / pdf) with wo ~ sampler; reports mean (the albedo, the
same for every scheme), the variance per sample, and the ratio to cosine-only sampling (spec_weight = 0).
usage: python tools/sampler_variance.py [log2_samples=20]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from mitsuba_customization_b200 import synth
from oracle import merl_oracle as O

logn = int(sys.argv[1]) if len(sys.argv) > 1 else 20
n = 1 << logn
rng = np.random.default_rng(2026)
u1 = rng.random(n, dtype=np.float32)
u2 = rng.random((2, n), dtype=np.float32)
rows = []
# table 0: alpha 0.3-ish glossy paint; a second table picked for a sharp lobe (small GGX alpha)
alphas = {}
for tid in range(12):
    g = np.random.Generator(np.random.PCG64(synth.TABLE_SEED + tid))
    g.uniform(0.05, 0.6, size=3)
    alphas[tid] = float(g.uniform(0.05, 0.5))
sharp = min(alphas, key=alphas.get)
for tid in (0, sharp):
    raw = synth.synthetic_table(tid)
    for w in (0.0, 0.5, 0.8):
        t = O.OracleTable(raw, spec_weight=w)
        t.build_cdf()
        for cz in (0.95, 0.7, 0.4, 0.15):
            wi = np.zeros((3, n), dtype=np.float32)
            wi[0], wi[2] = np.float32(np.sqrt(1 - cz * cz)), np.float32(cz)
            wo, pdf, weight, lobe = t.sample(wi, u1, u2)
            x = 0.2126 * weight[0] + 0.7152 * weight[1] + 0.0722 * weight[2]
            rows.append({"table": tid, "ggx_alpha": round(alphas[tid], 3), "spec_weight": w, "cos_theta_i": cz,
                         "mean": float(x.mean()), "variance": float(x.var()), "max_weight": float(x.max())})
base = {(r["table"], r["cos_theta_i"]): r["variance"] for r in rows if r["spec_weight"] == 0.0}
print(f"# {n} samples per cell; variance of luminance(f cos / pdf) per sample, ratio = variance / cosine-only variance")
print("| table (GGX alpha) | cos theta_i | albedo estimate | var, cosine only | var, w = 0.5 (default) | ratio | var, w = 0.8 | ratio |")
print("|---|---|---|---|---|---|---|---|")
for tid in (0, sharp):
    for cz in (0.95, 0.7, 0.4, 0.15):
        r = {x["spec_weight"]: x for x in rows if x["table"] == tid and x["cos_theta_i"] == cz}
        print(f"| {tid} ({alphas[tid]:.3f}) | {cz} | {r[0.5]['mean']:.4f} | {r[0.0]['variance']:.4g} | {r[0.5]['variance']:.4g} | "
              f"{r[0.5]['variance'] / base[(tid, cz)]:.3f} | {r[0.8]['variance']:.4g} | {r[0.8]['variance'] / base[(tid, cz)]:.3f} |")
print(json.dumps(rows))
