"""One-off check (run on the CPU) of what plugins/_build/m3_ad_test wrote on a B200 in the last seconds of the
round's GPU budget (gpurun_out/ad_o.bin, inputs regenerated from the same seeds): forward and gradient
against the CPU oracle.  Kept with the tests because it imports the oracle."""
import sys

import numpy as np

sys.path.insert(0, ".")
from mitsuba_customization_b200 import synth  # noqa: E402
from oracle import merl_oracle as O  # noqa: E402

dims, scale = (48, 64, 120), (0.5, 0.25, 0.125)
raw = synth.synthetic_table(3, dims=dims, theta_h_sqrt=False, scale=scale)
n = 65539
wi, wo = synth.random_queries(n, seed=47)
rng = np.random.default_rng(47)
u1 = rng.random(n, dtype=np.float32)
u2 = rng.random((2, n), dtype=np.float32)
active = (rng.random(n) > 0.1).astype(np.uint8)
wo[2, ::97] *= -1
wi[2, 5::211] *= -1
ncell = dims[0] * dims[1] * dims[2]
out = np.fromfile(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/ad_o.bin", dtype=np.float32)
assert out.size == 6 * n + 8 * ncell, out.size
rgb = out[:3 * n].reshape(3, n)
grad = out[3 * n:3 * n + 4 * ncell].reshape(ncell, 4)
direct = out[3 * n + 4 * ncell:3 * n + 8 * ncell].reshape(ncell, 4)
after = out[3 * n + 8 * ncell:].reshape(3, n)
orc = O.OracleTable(raw, dims=dims, theta_h_sqrt=False, scale=scale)
on = active.astype(bool)
ref_rgb, _, ref_idx = orc.eval_pdf(wi, wo, want_pdf=False)
ref_rgb[:, ~on] = 0
st = orc.bins(wi, wo)["stable"]
err = np.abs(rgb - ref_rgb)[:, st]
assert np.all(err <= 1e-6 + 1e-5 * np.abs(ref_rgb[:, st])), err.max()
g = np.stack([u1, u2[0], u2[1]]).astype(np.float64)
g[:, ~on] = 0
ref_grad = orc.backward(wi, wo, g)
ref_grad = np.asarray(ref_grad, np.float64).reshape(ncell, -1)
unstable_cells = np.unique(ref_idx[~st]) if (~st).any() else np.array([], int)
ok = np.ones(ncell, bool)
ok[unstable_cells] = False
e = np.abs(grad[:, :3] - ref_grad[:, :3])[ok]
assert np.all(e <= 1e-6 + 1e-5 * np.abs(ref_grad[:, :3][ok])), e.max()
assert np.all(grad[:, 3] == 0)
assert np.allclose(direct[:, :3], grad[:, :3], rtol=1e-5, atol=1e-6)
assert np.array_equal(after, 2 * rgb)
live = on & (wi[2] > 0) & (wo[2] > 0)
print(f"forward max err {err.max():.3e}; gradient max err {e.max():.3e}; hits {int(direct[:, 3].sum())} "
      f"(live queries {int(live.sum())}); validity-lane gradient all zero; eval after the step == 2 x eval")
