"""The C++ plugin hosts under plugins/ (Mitsuba 3 and Mitsuba 0.6 BSDF subclasses on the C ABI).

CPU part: they compile and link against the mock host headers and the real library, and fail
the way the host renderer expects (exception text from merl_last_error) when the table file is bad.
GPU part: the test drivers run the plugins as an integrator would and their output is compared
with the CPU oracle (BASELINE.json tolerances) and, bit for bit, with the ctypes path.
"""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "plugins", "_build")
RTOL, ATOL = 1e-5, 1e-6


@pytest.fixture(scope="module")
def binaries(built_lib):
    r = subprocess.run(["make", "-C", os.path.join(ROOT, "plugins"), "check"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    return {k: os.path.join(BUILD, k) for k in ("m3_host_test", "m06_host_test")}


def _write_queries(path, wi, wo, u1, u2, active):
    with open(path, "wb") as f:
        f.write(np.int64(wi.shape[1]).tobytes())
        for a in (wi, wo, u1, u2):
            f.write(np.ascontiguousarray(a, dtype=np.float32).tobytes())
        f.write(np.ascontiguousarray(active, dtype=np.uint8).tobytes())


def _close(a, b, what):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    bad = np.abs(a - b) > ATOL + RTOL * np.abs(b)
    assert not bad.any(), f"{what}: {bad.sum()} of {bad.size} outside tolerance, worst {np.abs(a - b).max():.3e}"


def test_plugin_hosts_build(binaries):
    for p in binaries.values():
        assert os.access(p, os.X_OK)


@pytest.mark.parametrize("exe", ["m3_host_test", "m06_host_test"])
def test_plugin_reports_missing_table_like_the_host(binaries, exe, tmp_path):
    """Construction fails with the library's message, where the reference plugin would Throw / Log(EError)."""
    q = tmp_path / "q.bin"
    z = np.zeros((3, 1), np.float32)
    _write_queries(q, z, z, np.zeros(1), np.zeros((2, 1)), np.ones(1))
    r = subprocess.run([binaries[exe], "merl", str(tmp_path / "missing.binary"), str(q), str(tmp_path / "o.bin")],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 1
    assert "merl" in r.stderr and "missing.binary" in r.stderr


@pytest.mark.parametrize("exe", ["m3_host_test", "m06_host_test"])
def test_plugin_rejects_bad_parameter(binaries, exe, tmp_path):
    q = tmp_path / "q.bin"
    z = np.zeros((3, 1), np.float32)
    _write_queries(q, z, z, np.zeros(1), np.zeros((2, 1)), np.ones(1))
    key = "theta_h_mapping" if exe == "m3_host_test" else "thetaHMapping"
    r = subprocess.run([binaries[exe], "customized_measurment", str(tmp_path / "x.binary"), str(q),
                        str(tmp_path / "o.bin"), f"{key}=cubic"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 1 and "sqrt" in r.stderr


# ---- GPU ----------------------------------------------------------------------------------------
def _inputs(n, seed):
    from mitsuba_customization_b200 import synth
    wi, wo = synth.random_queries(n, seed=seed)
    rng = np.random.default_rng(seed)
    u1 = rng.random(n, dtype=np.float32)
    u2 = rng.random((2, n), dtype=np.float32)
    active = (rng.random(n) > 0.1).astype(np.uint8)
    # a few lanes below the horizon, as an integrator produces them
    wo[2, ::97] *= -1
    wi[2, 5::211] *= -1
    return wi, wo, u1, u2, active


def _setup(kind, tmp_path):
    """(table file, python plugin on the same file, oracle on the plugin's CDF, extra argv for m3, for m06)."""
    import torch  # noqa: F401
    from mitsuba_customization_b200 import bsdf, synth
    from oracle import merl_oracle as O
    path = str(tmp_path / f"{kind}.binary")
    if kind == "merl":
        raw = synth.synthetic_table(0)
        synth.write_merl_binary(path, raw)
        plug = bsdf.load_dict({"type": "merl", "filename": path})
        orc = O.OracleTable(raw, cdf=plug.cdf().cpu().numpy())
        return path, plug, orc, [], []
    dims, scale = (48, 64, 120), (0.5, 0.25, 0.125)
    raw = synth.synthetic_table(3, dims=dims, theta_h_sqrt=False, scale=scale)
    synth.write_merl_binary(path, raw, dims)
    plug = bsdf.load_dict({"type": kind, "filename": path, "theta_h_mapping": "linear", "scale": scale})
    orc = O.OracleTable(raw, dims=dims, theta_h_sqrt=False, scale=scale, cdf=plug.cdf().cpu().numpy())
    m3 = ["theta_h_mapping=linear", "scale_r=0.5", "scale_g=0.25", "scale_b=0.125"]
    m06 = ["thetaHMapping=linear", "scaleR=0.5", "scaleG=0.25", "scaleB=0.125"]
    return path, plug, orc, m3, m06


def _python_path(plug, wi, wo, u1, u2, active):
    import torch
    from mitsuba_customization_b200 import bsdf
    ctx = bsdf.BSDFContext()
    si = bsdf.SurfaceInteraction3f(wi=torch.from_numpy(wi).cuda())
    act = torch.from_numpy(active).cuda()
    rgb, pdf = plug.eval_pdf(ctx, si, torch.from_numpy(wo).cuda(), active=act)
    bs, w = plug.sample(ctx, si, torch.from_numpy(u1).cuda(), torch.from_numpy(u2).cuda(), active=act)
    return (rgb.cpu().numpy(), pdf.cpu().numpy(), bs.wo.cpu().numpy(), bs.pdf.cpu().numpy(), w.cpu().numpy(),
            bs.lobe.cpu().numpy())


def _check_against_oracle(orc, wi, wo, u1, u2, active, rgb, pdf, s_wo, s_pdf, s_w, lobe):
    on = active.astype(bool)
    ref_rgb, ref_pdf, _ = orc.eval_pdf(wi, wo)
    ref_rgb[:, ~on] = 0
    ref_pdf[~on] = 0
    _close(rgb, ref_rgb, "eval")
    _close(pdf, ref_pdf, "pdf")
    ref_wo, _, _, ref_lobe = orc.sample(wi, u1, u2)
    live = on & (wi[2] > 0)
    assert np.array_equal(lobe[live], ref_lobe[live]), "sampled lobe must be bit-exact"
    _close(s_wo[:, live], ref_wo[:, live], "sample.wo")
    assert np.all(s_wo[:, ~live] == 0) and np.all(s_pdf[~live] == 0) and np.all(s_w[:, ~live] == 0)
    rgb_at, pdf_at, _ = orc.eval_pdf(wi, s_wo.astype(np.float32))
    _close(s_pdf[live], pdf_at[live], "sample.pdf")
    ok = live & (pdf_at > 0)
    _close(s_w[:, ok], rgb_at[:, ok] / pdf_at[ok], "sample.weight")


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["merl", "customized_measurment"])
def test_mitsuba3_plugin_matches_oracle(binaries, kind, tmp_path):
    n = (1 << 16) + 3
    path, plug, orc, m3_args, _ = _setup(kind, tmp_path)
    wi, wo, u1, u2, active = _inputs(n, seed=41)
    q, o = str(tmp_path / "q.bin"), str(tmp_path / "o.bin")
    _write_queries(q, wi, wo, u1, u2, active)
    r = subprocess.run([binaries["m3_host_test"], kind, path, q, o] + m3_args, capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "kernel launches" in r.stdout
    raw = np.fromfile(o, dtype=np.float32)
    blk = raw[:16 * n].reshape(16, n)
    rgb, pdf, rgb_only, pdf_only = blk[0:3], blk[3], blk[4:7], blk[7]
    s_wo, s_pdf, s_w, lobe = blk[8:11], blk[11], blk[12:15], blk[15].view(np.uint32).astype(np.int32)
    assert np.array_equal(rgb, rgb_only) and np.array_equal(pdf, pdf_only)      # eval_pdf == (eval, pdf)
    _check_against_oracle(orc, wi, wo, u1, u2, active, rgb, pdf, s_wo, s_pdf, s_w, lobe)
    # the C++ host and the ctypes host drive the same kernels: identical bits
    for got, want in zip((rgb, pdf, s_wo, s_pdf, s_w, lobe), _python_path(plug, wi, wo, u1, u2, active)):
        assert np.array_equal(got, want)
    # llvm variant: whole packets through the *_host entry points -> the same bits as the cuda variant
    lb = np.fromfile(o + ".llvm", dtype=np.float32).reshape(16, n)
    assert np.array_equal(lb, blk)
    assert "recording guard ok" in r.stdout and "scalar warning lines: 1" in r.stdout
    # scalar variant: first 64 queries one by one through the *_host entry points
    m = 64
    sc = raw[16 * n:16 * n + 15 * m].reshape(15, m)
    sl = raw[16 * n + 15 * m:].view(np.uint32).astype(np.int32)
    assert sl.size == m
    assert np.array_equal(sc[0:3], rgb[:, :m]) and np.array_equal(sc[3], pdf[:m])
    assert np.array_equal(sc[4:7], rgb[:, :m]) and np.array_equal(sc[7], pdf[:m])
    assert np.array_equal(sc[8:11], s_wo[:, :m]) and np.array_equal(sc[11], s_pdf[:m])
    assert np.array_equal(sc[12:15], s_w[:, :m])
    on = active[:m].astype(bool)
    assert np.array_equal(sl[on], lobe[:m][on])


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["merl", "customized_measurment"])
def test_mitsuba06_plugin_matches_oracle(binaries, kind, tmp_path):
    n = (1 << 16) + 1
    path, plug, orc, _, m06_args = _setup(kind, tmp_path)
    wi, wo, _, u2, active = _inputs(n, seed=43)
    # 0.6 passes two numbers: the plugin derives the lobe choice from sample.x (splitSample)
    w = np.float32(plug.info.spec_weight)
    u1 = u2[0].copy()
    u2x = np.where(u1 < w, u1 / w, (u1 - w) / (np.float32(1) - w)).astype(np.float32)
    u2s = np.stack([u2x, u2[1]])
    q, o = str(tmp_path / "q.bin"), str(tmp_path / "o.bin")
    _write_queries(q, wi, wo, u1, u2, active)
    r = subprocess.run([binaries["m06_host_test"], kind, path, q, o] + m06_args, capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    raw = np.fromfile(o, dtype=np.float32)
    ep = raw[:4 * n].reshape(n, 4).T
    swo = raw[4 * n:8 * n].reshape(n, 4).T
    sw = raw[8 * n:12 * n].reshape(n, 4).T
    _check_against_oracle(orc, wi, wo, u1, u2s, active, ep[0:3], ep[3], swo[0:3], swo[3], sw[0:3],
                          sw[3].astype(np.int32))
    m = 64
    rows = raw[12 * n:].reshape(m, 15)
    on = active[:m].astype(bool)
    assert np.array_equal(rows[:, 0:3].T, ep[0:3, :m]) and np.array_equal(rows[:, 3], ep[3, :m])
    good = on & (swo[3, :m] > 0)
    assert np.array_equal(rows[good, 4:7].T, swo[0:3, :m][:, good]) and np.array_equal(rows[:, 7], swo[3, :m])
    assert np.array_equal(rows[:, 8:11].T, sw[0:3, :m])
    assert np.all(rows[good, 11] == 8) and np.all(rows[good, 12] == 0)         # EGlossyReflection, component 0
    assert np.all(rows[:, 13] == 0)                                              # EDiscrete measure evaluates to 0
    assert np.array_equal(rows[:, 14], rows[:, 0])                               # serialize() -> clone gives the same
