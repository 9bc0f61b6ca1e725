"""Host plumbing of the C++ plugin hosts WITHOUT a GPU (-m "not gpu").

The drivers under plugins/test are linked against plugins/test/fake_device, a test double for the
CUDA runtime and libmerl_b200 that computes no BSDF: every query entry point writes a position-coded
function of its arguments (see its header).  What is checked here is everything the plugin sources do
around the C ABI: which array goes into which argument, broadcast of width-1 operands, mask bytes,
result wrapping (BSDFSample3f fields, sampled_type/component), scalar variants, the CustomOp
(forward, backward, validity lane), traverse / parameters_changed with a replaced array, and for
Mitsuba 0.6 the wavefront stage's buffers, splitSample and the serialize() round trip.
"""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CPU = os.path.join(ROOT, "plugins", "_build", "cpu")
DIMS = (6, 5, 8)
NCELL = DIMS[0] * DIMS[1] * DIMS[2]


@pytest.fixture(scope="module")
def cpu_binaries():
    r = subprocess.run(["make", "-C", os.path.join(ROOT, "plugins"), "check-cpu"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    return {k: os.path.join(CPU, k) for k in ("m3_host_test", "m06_host_test", "m3_ad_test")}


def _table(path):
    with open(path, "wb") as f:
        f.write(np.asarray(DIMS, np.int32).tobytes())
        f.write(np.zeros(3 * NCELL, np.float64).tobytes())
    return str(path)


def _queries(path, n, seed):
    rng = np.random.default_rng(seed)
    wi = rng.random((3, n), dtype=np.float32)
    wo = rng.random((3, n), dtype=np.float32)
    u1 = rng.random(n, dtype=np.float32)
    u2 = rng.random((2, n), dtype=np.float32)
    active = (rng.random(n) > 0.2).astype(np.uint8)
    with open(path, "wb") as f:
        f.write(np.int64(n).tobytes())
        for a in (wi, wo, u1, u2):
            f.write(np.ascontiguousarray(a, dtype=np.float32).tobytes())
        f.write(active.tobytes())
    return wi, wo, u1, u2, active


def _cells(n):
    """The double's table: cells[k].c = 1 + (k % 1024) + 0.25 c; query i reads cell i % ncell."""
    k = (np.arange(n) % NCELL) % 1024
    return np.stack([1 + k + 0.25 * c for c in range(3)]).astype(np.float32)


def _code_pdf(wi, wo):
    f = np.float32
    return ((((wi[0] + f(2) * wi[1]) + f(4) * wi[2]) + f(8) * wo[0]) + f(16) * wo[1]) + f(32) * wo[2]


ARGS = ["nh=6", "nd=5", "np=8"]


def test_mitsuba3_host_marshals_every_argument(cpu_binaries, tmp_path):
    n = 777
    tab = _table(tmp_path / "t.binary")
    wi, wo, u1, u2, active = _queries(tmp_path / "q.bin", n, 5)
    o = str(tmp_path / "o.bin")
    r = subprocess.run([cpu_binaries["m3_host_test"], "customized_measurment", tab, str(tmp_path / "q.bin"), o,
                        "resolution_theta_h=6", "resolution_theta_d=5", "resolution_phi_d=8"],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    raw = np.fromfile(o, dtype=np.float32)
    blk = raw[:16 * n].reshape(16, n)
    on = active.astype(bool)
    want_rgb = np.where(on, _cells(n) * wo[2], 0).astype(np.float32)
    want_pdf = np.where(on, _code_pdf(wi, wo), 0).astype(np.float32)
    assert np.array_equal(blk[0:3], want_rgb) and np.array_equal(blk[4:7], want_rgb)
    np.testing.assert_allclose(blk[3], want_pdf, rtol=1e-6)
    assert np.array_equal(blk[3], blk[7])
    # sample: wo = (u1, u2x, u2y), pdf = wi code, weight = 2 (u1, u2x, u2y), lobe = i & 1
    assert np.array_equal(blk[8], np.where(on, u1, 0)) and np.array_equal(blk[9], np.where(on, u2[0], 0))
    assert np.array_equal(blk[10], np.where(on, u2[1], 0))
    np.testing.assert_allclose(blk[11], np.where(on, wi[0] + 2 * wi[1] + 4 * wi[2], 0), rtol=1e-6)
    assert np.array_equal(blk[12], np.where(on, 2 * u1, 0)) and np.array_equal(blk[14], np.where(on, 2 * u2[1], 0))
    assert np.array_equal(blk[15].view(np.uint32), np.where(on, np.arange(n) & 1, 0).astype(np.uint32))
    # llvm variant: the WHOLE packet per virtual through the *_host entry points (one library call each,
    # not one per lane), masked lanes zeroed afterwards -> the same position codes as the cuda block
    lb = np.fromfile(o + ".llvm", dtype=np.float32).reshape(16, n)
    assert np.array_equal(lb, blk)
    assert f"llvm variant: {n} lanes per call, 4 library launches for 4 virtuals" in r.stdout
    # JIT variants refuse a symbolic recording with a message naming wavefront mode; scalar variants warn ONCE
    assert "recording guard ok" in r.stdout and "scalar warning lines: 1" in r.stdout
    assert r.stderr.count("[WARN]") == 1 and "round trip" in r.stderr
    # scalar variant: one query per call -> always cell 0, lobe 0; inactive lanes stay zero
    m = 64
    sc = raw[16 * n:16 * n + 15 * m].reshape(15, m)
    cell0 = np.array([1.0, 1.25, 1.5], np.float32)[:, None]
    assert np.array_equal(sc[0:3], np.where(on[:m], cell0 * wo[2, :m], 0).astype(np.float32))
    np.testing.assert_allclose(sc[3], want_pdf[:m], rtol=1e-6)
    assert np.array_equal(sc[8], np.where(on[:m], u1[:m], 0))


def test_mitsuba3_custom_op_and_optimiser_step(cpu_binaries, tmp_path):
    n = 2000          # > ncell: several queries per cell, so the gradient sums
    tab = _table(tmp_path / "t.binary")
    wi, wo, u1, u2, active = _queries(tmp_path / "q.bin", n, 6)
    o = str(tmp_path / "o.bin")
    r = subprocess.run([cpu_binaries["m3_ad_test"], "customized_measurment", tab, str(tmp_path / "q.bin"), o,
                        "resolution_theta_h=6", "resolution_theta_d=5", "resolution_phi_d=8"],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "m3_ad_test ok" in r.stdout
    raw = np.fromfile(o, dtype=np.float32)
    assert raw.size == 3 * n + 8 * NCELL + 3 * n
    rgb = raw[:3 * n].reshape(3, n)
    grad = raw[3 * n:3 * n + 4 * NCELL].reshape(NCELL, 4)
    direct = raw[3 * n + 4 * NCELL:3 * n + 8 * NCELL].reshape(NCELL, 4)
    after = raw[3 * n + 8 * NCELL:].reshape(3, n)
    on = active.astype(bool)
    assert np.array_equal(rgb, np.where(on, _cells(n) * wo[2], 0).astype(np.float32))
    # backward of the double: grad[i % ncell] += (u1, u2x, u2y) * wo.z, one hit per active query
    want = np.zeros((NCELL, 4), np.float64)
    g = np.stack([u1, u2[0], u2[1], np.ones(n, np.float32)]) * np.where(on, wo[2], 0)
    g[3] = on
    np.add.at(want, np.arange(n) % NCELL, g.T)
    np.testing.assert_allclose(direct, want, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(grad[:, :3], want[:, :3], rtol=1e-5, atol=1e-6)
    assert np.all(grad[:, 3] == 0)                       # validity lane: no gradient
    assert np.array_equal(after, 2 * rgb)                 # the replaced parameter reached the table


def test_mitsuba06_host_marshals_every_argument(cpu_binaries, tmp_path):
    n = 513
    tab = _table(tmp_path / "t.binary")
    wi, wo, u1, u2, active = _queries(tmp_path / "q.bin", n, 7)
    o = str(tmp_path / "o.bin")
    r = subprocess.run([cpu_binaries["m06_host_test"], "customized_measurment", tab, str(tmp_path / "q.bin"), o,
                        "thetaHRes=6", "thetaDRes=5", "phiDRes=8"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    # the scalar virtuals say ONCE that they are one PCIe round trip per call and name the batched path
    assert r.stderr.count("round trip") == 1 and "WavefrontStage" in r.stderr
    raw = np.fromfile(o, dtype=np.float32)
    ep = raw[:4 * n].reshape(n, 4).T
    on = active.astype(bool)
    # the wavefront stage has no mask: a record whose component is switched off is queued with
    # wi.z = -1 (below the horizon -> the kernels return zero); the double just shows what arrived
    wi_q = wi.copy()
    wi_q[2] = np.where(on, wi[2], -1)
    assert np.array_equal(ep[0:3], (_cells(n) * wo[2]).astype(np.float32))
    np.testing.assert_allclose(ep[3], _code_pdf(wi_q, wo), rtol=1e-6)
    # sample stage: 0.6 passes two numbers; the plugin derives the lobe choice from sample.x
    w = np.float32([l for l in r.stdout.splitlines() if "specWeight" in l][0].split("=")[1].strip())
    sx = u2[0]
    u2x = np.where(sx < w, sx / w, (sx - w) / (np.float32(1) - w)).astype(np.float32)
    swo = raw[4 * n:8 * n].reshape(n, 4).T
    sw = raw[8 * n:12 * n].reshape(n, 4).T
    assert np.array_equal(swo[0], sx) and np.array_equal(swo[2], u2[1])
    np.testing.assert_allclose(swo[1], u2x, rtol=1e-6)
    np.testing.assert_allclose(swo[3], wi_q[0] + 2 * wi_q[1] + 4 * wi_q[2], rtol=1e-6)
    assert np.array_equal(sw[0], 2 * sx) and np.array_equal(sw[3], (np.arange(n) & 1).astype(np.float32))
    # scalar virtuals (one query per call -> cell 0), typeMask honoured on the host side, and the
    # clone rebuilt from serialize() answers like the original
    m = 64
    rows = raw[12 * n:].reshape(m, 15)
    cell0 = np.array([1.0, 1.25, 1.5], np.float32)
    assert np.array_equal(rows[:, 0:3], np.where(on[:m, None], cell0[None, :] * wo[2, :m, None], 0).astype(np.float32))
    np.testing.assert_allclose(rows[:, 3], np.where(on[:m], _code_pdf(wi, wo)[:m], 0), rtol=1e-6)
    good = on[:m] & (rows[:, 7] > 0)
    assert np.array_equal(rows[good, 4], sx[:m][good]) and np.array_equal(rows[good, 6], u2[1, :m][good])
    assert np.all(rows[good, 11] == 8) and np.all(rows[good, 12] == 0)           # EGlossyReflection, component 0
    assert np.all(rows[~on[:m], 7] == 0)
    assert np.all(rows[:, 13] == 0)                                                # EDiscrete measure
    assert np.array_equal(rows[:, 14], rows[:, 0])                                 # serialize() -> clone
