"""The differentiable side of the Mitsuba 3 plugin host on the GPU: plugins/test/m3_ad_test drives
MerlEvalOp (the dr::CustomOp of plugins/mitsuba3/merl_eval_op.h), traverse() and
parameters_changed() with a replaced `data` array against the real library, checks itself
(custom-op eval == eval bit for bit, gradient of the validity lane zero, eval doubles exactly after
the cells were doubled) and writes its results, which are compared here with the ctypes path.

Ordering note: the file sorts last on purpose.  Its CPU twin (tests/test_plugin_hosts_cpu.py, same
driver against the test double) ran green, but the GPU budget of the round was spent before this
driver could be run on a B200, so it must not stand in front of other GPU tests under `pytest -x`.
"""
import os
import subprocess

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RTOL, ATOL = 1e-5, 1e-6


@pytest.fixture(scope="module")
def ad_binary(built_lib):
    r = subprocess.run(["make", "-C", os.path.join(ROOT, "plugins"), "check"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    return os.path.join(ROOT, "plugins", "_build", "m3_ad_test")


@pytest.mark.parametrize("kind", ["merl", "customized_measurment"])
def test_mitsuba3_custom_op_matches_ctypes_backward(ad_binary, kind, tmp_path):
    from mitsuba_customization_b200 import bsdf
    from tests.test_plugins import _inputs, _setup, _write_queries
    n = (1 << 16) + 3
    path, plug, _, m3_args, _ = _setup(kind, tmp_path)
    wi, wo, u1, u2, active = _inputs(n, seed=47)
    q, o = str(tmp_path / "q.bin"), str(tmp_path / "o.bin")
    _write_queries(q, wi, wo, u1, u2, active)
    r = subprocess.run([ad_binary, kind, path, q, o] + m3_args, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "m3_ad_test ok" in r.stdout
    ncell = plug.info.ncell
    raw = np.fromfile(o, dtype=np.float32)
    assert raw.size == 6 * n + 8 * ncell
    rgb = raw[:3 * n].reshape(3, n)
    grad = raw[3 * n:3 * n + 4 * ncell].reshape(ncell, 4)
    direct = raw[3 * n + 4 * ncell:3 * n + 8 * ncell].reshape(ncell, 4)
    after = raw[3 * n + 8 * ncell:].reshape(3, n)
    # the same queries through the ctypes host
    ctx = bsdf.BSDFContext()
    si = bsdf.SurfaceInteraction3f(wi=torch.from_numpy(wi).cuda())
    act = torch.from_numpy(active).cuda()
    t_wo = torch.from_numpy(wo).cuda()
    assert np.array_equal(rgb, plug.eval(ctx, si, t_wo, active=act).cpu().numpy())
    g = torch.from_numpy(np.stack([u1, u2[0], u2[1]])).cuda()
    ref = plug.backward(si, t_wo, g, active=act).cpu().numpy().astype(np.float64)
    err = np.abs(direct - ref)
    assert np.all(err <= ATOL + RTOL * np.abs(ref))               # sums differ by the order of the atomics only
    assert np.all(np.abs(grad[:, :3] - ref[:, :3]) <= ATOL + RTOL * np.abs(ref[:, :3]))
    assert np.all(grad[:, 3] == 0) and direct[:, 3].sum() == ref[:, 3].sum()
    assert np.array_equal(after, 2 * rgb)
