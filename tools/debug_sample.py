import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from mitsuba_customization_b200 import bsdf, synth
from oracle import merl_oracle as O
raw = synth.synthetic_table(0)
plug = bsdf.MERL(data=raw)
orc = O.OracleTable(raw, cdf=plug.cdf().cpu().numpy())
n = 1 << 18
wi, _ = synth.random_queries(n, seed=n + 17)
rng = np.random.default_rng(n)
u1 = rng.random(n, dtype=np.float32); u2 = rng.random((2, n), dtype=np.float32)
bs, w = plug.sample(bsdf.BSDFContext(), bsdf.SurfaceInteraction3f(wi=torch.from_numpy(wi).cuda()), torch.from_numpy(u1).cuda(), torch.from_numpy(u2).cuda())
ref_wo, ref_pdf, ref_w, ref_lobe = orc.sample(wi, u1, u2)
wo = bs.wo.cpu().numpy().astype(np.float64); pdf = bs.pdf.cpu().numpy().astype(np.float64); lobe = bs.lobe.cpu().numpy()
print("lobe mismatches", (lobe != ref_lobe).sum())
err = np.abs(wo - ref_wo); bad = (err > 1e-6 + 1e-5 * np.abs(ref_wo)).any(0)
print("wo bad", bad.sum(), "by lobe", [(bad & (ref_lobe == l)).sum() for l in (0, 1)], "max err", err.max())
for i in np.where(bad)[0][:8]:
    print(" q", i, "lobe", lobe[i], "u", u1[i], u2[:, i], "wi", wi[:, i], "gpu", wo[:, i], "ref", ref_wo[:, i])
rgb_at, pdf_at, _ = orc.eval_pdf(wi, wo.astype(np.float32))
e = np.abs(pdf - pdf_at); badp = e > 1e-6 + 1e-5 * np.abs(pdf_at)
print("pdf bad", badp.sum(), "max", e.max())
for i in np.where(badp)[0][:8]:
    print(" q", i, "lobe", lobe[i], "gpu pdf", pdf[i], "orc pdf@gpu wo", pdf_at[i], "ref pdf", ref_pdf[i], "wo", wo[:, i], "wi", wi[:, i])
ok = pdf_at > 0
ww = w.cpu().numpy().astype(np.float64)
tgt = rgb_at[:, ok] / pdf_at[ok]
ew = np.abs(ww[:, ok] - tgt); badw = (ew > 1e-6 + 1e-5 * np.abs(tgt)).any(0)
print("weight bad", badw.sum(), "max", ew.max())
