import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from mitsuba_customization_b200 import bsdf, synth
from oracle import merl_oracle as O
raw = synth.synthetic_table(0)
plug = bsdf.MERL(data=raw)
orc = O.OracleTable(raw)
for name, (wi, wo) in {"edge": synth.edge_stress_queries(1 << 20), "rand": synth.random_queries(1 << 24, seed=5)}.items():
    ref = orc.bins(wi, wo)
    got = plug.bins(bsdf.SurfaceInteraction3f(wi=torch.from_numpy(wi).cuda()), torch.from_numpy(wo).cuda()).cpu().numpy()
    gh, gd, gp = got // (90 * 180), (got // 180) % 90, got % 180
    st = ref["stable"]
    print(name, "n", wi.shape[1], "unstable", (~st).sum(), "mismatch(stable)", ((got != ref["idx"]) & st).sum(),
          "mismatch(unstable)", ((got != ref["idx"]) & ~st).sum())
    for c, g, r in (("ih", gh, ref["ih"]), ("id", gd, ref["id"]), ("ip", gp, ref["ip"])):
        bad = np.where((g != r) & st)[0]
        print("  ", c, len(bad))
        for i in bad[:6]:
            hd = O.half_diff(wi[:, i].astype(np.float64), wo[:, i].astype(np.float64))
            print("     q", i, "gpu", g[i], "ref", r[i], "wi", wi[:, i], "wo", wo[:, i], "hd(deg)", np.degrees(hd))
