"""Generates tests/golden/merl_oracle_v1.npz from the CPU oracle (oracle/merl_oracle.c).

The reference repository ships no golden vectors and no source to generate them from
(/root/reference/.SUBMODULES.json:8-23), so these fixtures pin the ORACLE, not the reference:
they freeze its answers on a seeded batch so later rounds cannot drift silently, and they give
the GPU tests a committed known-answer set that does not depend on rebuilding the oracle.
Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from mitsuba_customization_b200 import synth  # noqa: E402
from oracle import merl_oracle as O  # noqa: E402

n = 4096
raw = synth.synthetic_table(0)
t = O.OracleTable(raw)
cdf, _ = t.build_cdf()
wi, wo = synth.random_queries(n, seed=2024)
ewi, ewo = synth.edge_stress_queries(2048, seed=7)
st = t.bins(ewi, ewo)["stable"]
wi = np.ascontiguousarray(np.concatenate([wi, ewi[:, st][:, :1024]], axis=1))
wo = np.ascontiguousarray(np.concatenate([wo, ewo[:, st][:, :1024]], axis=1))
n = wi.shape[1]
rgb, pdf, idx = t.eval_pdf(wi, wo)
rng = np.random.default_rng(11)
u1, u2 = rng.random(n, dtype=np.float32), rng.random((2, n), dtype=np.float32)
s_wo, s_pdf, s_w, lobe = t.sample(wi, u1, u2)
np.savez_compressed(os.path.join(os.path.dirname(__file__), "merl_oracle_v1.npz"), wi=wi, wo=wo, rgb=rgb, pdf=pdf,
                    idx=idx, u1=u1, u2=u2, s_wo=s_wo, s_pdf=s_pdf, s_w=s_w, lobe=lobe, cdf=cdf,
                    table_checksum=np.float64(raw.sum()))
print("wrote", n, "queries")
