"""Table files either side of the loader: the host-only file report (CPU) and the packed sidecar (GPU)."""
import numpy as np
import pytest


def test_file_report_counts_markers(tmp_path, raw_table, built_lib):
    from mitsuba_customization_b200 import synth, tabletool
    path = str(tmp_path / "t.binary")
    raw = raw_table.copy()
    raw[1, 5] = np.nan
    raw[2, 7] = np.inf
    synth.write_merl_binary(path, raw)
    r = tabletool.report(path)
    assert r["dims"] == [90, 90, 180] and r["is_merl_layout"] and r["ncell"] == 90 * 90 * 180
    assert r["file_bytes"] == 12 + 3 * 8 * 90 * 90 * 180
    assert r["nonfinite"] == [0, 1, 1]
    for c in range(3):
        v = raw[c][np.isfinite(raw[c])]
        assert r["negative"][c] == int((v < 0).sum())
        assert r["min_valid"][c] == v[v >= 0].min() and r["max_valid"][c] == v[v >= 0].max()


def test_file_report_rejects_bad_files(tmp_path, raw_table, built_lib):
    from mitsuba_customization_b200 import _cabi, synth, tabletool
    with pytest.raises(_cabi.MerlError):
        tabletool.report(str(tmp_path / "missing.binary"))
    trunc = str(tmp_path / "trunc.binary")
    synth.write_merl_binary(trunc, raw_table)
    with open(trunc, "r+b") as f:
        f.truncate(12 + 1000)
    with pytest.raises(_cabi.MerlError, match="needs"):
        tabletool.report(trunc)
    bad = tmp_path / "bad.binary"
    bad.write_bytes(np.array([-3, 90, 180], np.int32).tobytes() + b"\0" * 64)
    with pytest.raises(_cabi.MerlError, match="implausible"):
        tabletool.report(str(bad))
    assert tabletool.main(["report", str(bad)]) == 1


def test_packed_loader_rejects_garbage(tmp_path, built_lib):
    import ctypes as C
    from mitsuba_customization_b200 import _cabi
    junk = tmp_path / "junk.merlb200"
    junk.write_bytes(b"not a table" * 20)
    h = C.c_void_p()
    assert _cabi.lib().merl_table_create_from_packed(str(junk).encode(), 0, C.byref(h)) == -2   # MERL_ERR_IO
    assert b"not a packed" in _cabi.lib().merl_last_error()


@pytest.mark.gpu
def test_packed_sidecar_round_trip(tmp_path):
    """pack -> load gives a bit-identical table, CDF and results; a flipped byte is caught."""
    import torch
    from mitsuba_customization_b200 import bsdf, synth, tabletool
    dims, scale = (48, 64, 120), (0.5, 0.25, 0.125)
    src = str(tmp_path / "c.binary")
    synth.write_merl_binary(src, synth.synthetic_table(3, dims=dims, theta_h_sqrt=False, scale=scale), dims)
    out = str(tmp_path / "c.merlb200")
    assert tabletool.main(["pack", src, out, "--linear", "--scale", "0.5", "0.25", "0.125"]) == 0
    a = bsdf.load_dict({"type": "customized_measurment", "filename": src, "theta_h_mapping": "linear", "scale": scale})
    b = bsdf.MERL(packed=out)
    assert (b.info.nh, b.info.nd, b.info.np, b.info.theta_h_sqrt) == (48, 64, 120, 0)
    assert torch.equal(a.cells(), b.cells()) and torch.equal(a.cdf(), b.cdf()) and torch.equal(a.dens(), b.dens())
    wi, wo = synth.random_queries(1 << 16, seed=2)
    ctx, si, two = bsdf.BSDFContext(), bsdf.SurfaceInteraction3f(wi=torch.from_numpy(wi).cuda()), torch.from_numpy(wo).cuda()
    ra, pa = a.eval_pdf(ctx, si, two)
    rb, pb = b.eval_pdf(ctx, si, two)
    assert torch.equal(ra, rb) and torch.equal(pa, pb)
    u1, u2 = torch.rand(1 << 16, device="cuda"), torch.rand(2, 1 << 16, device="cuda")
    sa, wa = a.sample(ctx, si, u1, u2)
    sb, wb = b.sample(ctx, si, u1, u2)
    assert torch.equal(sa.wo, sb.wo) and torch.equal(wa, wb)
    blob = bytearray(open(out, "rb").read())
    blob[len(blob) // 2] ^= 0x40
    bad = str(tmp_path / "flipped.merlb200")
    open(bad, "wb").write(bytes(blob))
    with pytest.raises(bsdf.MerlError, match="checksum"):
        bsdf.MERL(packed=bad)
