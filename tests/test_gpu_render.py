"""GPU parity of the wavefront path tracer harness (merl_render) against oracle/render_oracle.c.

BASELINE.json: "rendered images must have relative MSE <= 1e-4 at equal spp and seed".  relMSE
here = mean((gpu - oracle)^2) / mean(oracle^2) over all pixels and channels."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

REL_MSE = 1e-4


def _relmse(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(((a - b) ** 2).mean() / (b ** 2).mean())


def _relmse_trimmed(a, b, frac=1e-3):
    """relMSE with the worst `frac` of the pixels left out.  The harness traces in fp32 and the oracle in double:
    a ray that grazes a silhouette, or a wi within an ulp of a bin edge, takes another branch in a handful of
    pixels per seed (tools/render_diff.py: 1-20 of 6144 pixels differ by more than 1e-3 at every seed, the rest
    agree to < 1e-3), and on a 96x64x8-spp image ONE such path with a large f/pdf weight moves the plain relMSE by
    1e-4.  At BASELINE's sizes the same flips are diluted to 1e-8..1e-7 (the full-size tests below use the plain
    metric); the small scenes bound the plain metric loosely and the trimmed one tightly."""
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    e = ((a - b) ** 2).sum(axis=-1).ravel()
    k = int(np.ceil(frac * e.size))
    e = np.sort(e)[:e.size - k]
    return float(e.sum() / a.shape[-1] / e.size / (b ** 2).mean())


@pytest.fixture(scope="module")
def pair(raw_table):
    from mitsuba_customization_b200 import bsdf
    from oracle import merl_oracle as O
    plug = bsdf.MERL(data=raw_table, device="cuda:0")
    return plug, O.OracleTable(raw_table, cdf=plug.cdf().cpu().numpy())


def test_sphere_image_matches_oracle(pair):
    """BASELINE config 2 at test size: unit sphere, constant envmap, path integrator."""
    from mitsuba_customization_b200 import render as RD
    from oracle import render_oracle as RO
    plug, orc = pair
    sc = RD.sphere_scene(plug)
    sc.env = (1.0, 0.8, 0.6)
    kw = dict(width=96, height=96, spp=16, max_depth=4, seed=11)
    film, st = RD.render(sc, want_stats=True, **kw)
    ref = RO.render([orc], sc.spheres, cam_pos=sc.cam_pos, cam_target=sc.cam_target, cam_up=sc.cam_up,
                    fov_x_deg=sc.fov_x_deg, env=sc.env, **kw)
    got = film.cpu().numpy()
    np.testing.assert_allclose(got[..., 3], 1.0, atol=1e-5)          # every path reached the film exactly once
    np.testing.assert_allclose(got[0, 0, :3], sc.env, rtol=1e-5)     # background
    assert st["paths"] == 96 * 96 * 16 and st["bsdf_queries"] > 0
    assert _relmse(got[..., :3], ref[..., :3]) <= REL_MSE


def test_small_wavefront_and_sample_split(pair):
    """Many passes (wavefront << paths) and a 3-way sample split add up to the one-pass image."""
    from mitsuba_customization_b200 import render as RD
    plug, _ = pair
    sc = RD.sphere_scene(plug)
    kw = dict(width=64, height=48, spp=6, max_depth=3, seed=5)
    one = RD.render(sc, **kw)
    many = RD.render(sc, wavefront=4096, **kw)
    torch.testing.assert_close(many, one, rtol=1e-5, atol=1e-6)
    split = torch.zeros_like(one)
    for r in range(3):
        RD.render(sc, spp_range=RD.spp_share(6, 3, r), film=split, **kw)
    torch.testing.assert_close(split, one, rtol=1e-5, atol=1e-6)


def test_multi_material_grid_matches_oracle():
    """BASELINE config 5 at test size: 3x3 spheres, 4 tables (one of them customized, linear bins),
    inter-reflection, queries sorted by material and served by the segmented BSDF stage."""
    from mitsuba_customization_b200 import bsdf, render as RD, synth
    from oracle import merl_oracle as O, render_oracle as RO
    plugs, orcs = [], []
    for k in range(3):
        raw = synth.synthetic_table(k)
        p = bsdf.MERL(data=raw, device="cuda:0")
        plugs.append(p)
        orcs.append(O.OracleTable(raw, cdf=p.cdf().cpu().numpy()))
    dims = (48, 40, 64)
    rawc = synth.synthetic_table(9, dims=dims)
    pc = bsdf.CustomizedMeasurement(data=rawc, resolution=dims, theta_h_mapping="linear", scale=(1 / 1500.0,) * 3,
                                    ni=16, device="cuda:0")
    plugs.append(pc)
    orcs.append(O.OracleTable(rawc, dims=dims, theta_h_sqrt=False, ni=16, scale=(1 / 1500.0,) * 3,
                              cdf=pc.cdf().cpu().numpy()))
    sc = RD.grid_scene(plugs, nx=3, ny=3)
    kw = dict(width=96, height=64, spp=8, max_depth=5, seed=23)
    film, st = RD.render(sc, want_stats=True, **kw)
    ref = RO.render(orcs, sc.spheres, cam_pos=sc.cam_pos, cam_target=sc.cam_target, cam_up=sc.cam_up,
                    fov_x_deg=sc.fov_x_deg, env=sc.env, **kw)
    got = film.cpu().numpy()
    np.testing.assert_allclose(got[..., 3], 1.0, atol=1e-5)
    assert st["bsdf_queries"] > st["paths"] * 0.2          # inter-reflections happen
    assert _relmse(got[..., :3], ref[..., :3]) <= 10 * REL_MSE
    assert _relmse_trimmed(got[..., :3], ref[..., :3]) <= REL_MSE / 10
    # the trimmed metric must not hide a bug that hits a small share of the pixels: the pixels it leaves out are
    # COUNTED (fp32-vs-double branch flips were measured at 1-20 of 6144 per seed, tools/render_diff.py) ...
    flipped = int((np.abs(got[..., :3] - ref[..., :3]).max(axis=-1) > 1e-3).sum())
    print(f"pixels off by more than 1e-3: {flipped} of {got.shape[0] * got.shape[1]}")
    assert flipped <= 24
    # ... and at 64 spp, where one flipped path weighs 1/64 of a pixel, the PLAIN metric holds BASELINE's bar
    kw64 = dict(kw, spp=64)
    film64 = RD.render(sc, **kw64).cpu().numpy()
    ref64 = RO.render(orcs, sc.spheres, cam_pos=sc.cam_pos, cam_target=sc.cam_target, cam_up=sc.cam_up,
                      fov_x_deg=sc.fov_x_deg, env=sc.env, **kw64)
    assert _relmse(film64[..., :3], ref64[..., :3]) <= REL_MSE


def test_render_argument_errors(pair):
    from mitsuba_customization_b200 import render as RD
    from mitsuba_customization_b200._cabi import MerlError
    plug, _ = pair
    sc = RD.sphere_scene(plug)
    sc.spheres = [(0, 0, 0, 1.0, 3)]          # material out of range
    with pytest.raises(MerlError):
        RD.render(sc, 8, 8, 1)
    sc = RD.sphere_scene(plug)
    with pytest.raises(MerlError):
        RD.render(sc, 8, 8, 4, spp_range=(3, 9))


def test_graph_replay_equals_plain_launches(pair):
    """The captured-pass CUDA graph (default) and kernel-by-kernel launches give the same film, with a
    ragged last pass (paths % wavefront != 0), on the default stream and on a side stream."""
    from mitsuba_customization_b200 import _cabi, render as RD
    plug, _ = pair
    sc = RD.sphere_scene(plug)
    kw = dict(width=50, height=30, spp=7, max_depth=3, seed=9, wavefront=2048)   # 10500 paths, 6 passes
    try:
        _cabi.check(_cabi.lib().merl_set_render_graphs(0))
        plain, st0 = RD.render(sc, want_stats=True, **kw)
    finally:
        _cabi.check(_cabi.lib().merl_set_render_graphs(1))
    graph, st1 = RD.render(sc, want_stats=True, **kw)
    assert st0 == st1 and st0["paths"] == 10500
    torch.testing.assert_close(graph, plain, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(graph.cpu().numpy()[..., 3], 1.0, atol=1e-5)
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        for _ in range(3):   # back-to-back calls reuse the side stream and retire their graphs
            again = RD.render(sc, **kw)
    side.synchronize()
    torch.testing.assert_close(again, plain, rtol=1e-5, atol=1e-6)


# ---- BASELINE.json's own sizes: "rendered images must have relative MSE <= 1e-4 at
# equal spp and seed".  The CPU oracle renders config 2 in well under a second on 16 threads and config 5
# at 16 spp in a few seconds, so both run at full image size.
def _record(name, payload):
    import json
    import os
    out_dir = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    if os.path.isdir(out_dir):
        json.dump(payload, open(os.path.join(out_dir, name), "w"), indent=1)


def test_config2_full_size_matches_oracle(pair):
    """BASELINE config 2 exactly: 512x512 unit sphere under a constant envmap, 64 spp, depth 8."""
    from mitsuba_customization_b200 import render as RD
    from oracle import render_oracle as RO
    plug, orc = pair
    sc = RD.sphere_scene(plug)
    kw = dict(width=512, height=512, spp=64, max_depth=8, seed=1)
    film, st = RD.render(sc, want_stats=True, **kw)
    ref = RO.render([orc], sc.spheres, cam_pos=sc.cam_pos, cam_target=sc.cam_target, cam_up=sc.cam_up,
                    fov_x_deg=sc.fov_x_deg, env=sc.env, **kw)
    got = film.cpu().numpy()
    np.testing.assert_allclose(got[..., 3], 1.0, atol=2e-5)
    r = _relmse(got[..., :3], ref[..., :3])
    _record("render_cfg2_fullsize.json", {"config": kw, "rel_mse": r, "paths": st["paths"], "bsdf_queries": st["bsdf_queries"],
                                          "max_abs_pixel_diff": float(np.abs(got[..., :3] - ref[..., :3]).max())})
    assert st["paths"] == 512 * 512 * 64
    assert r <= REL_MSE, r


def test_config5_full_image_100_tables_matches_oracle():
    """BASELINE config 5 at full image size and table count: 1024x1024, 10x10 spheres, 100 synthetic MERL
    tables (2.33 GB of float4 on the device), depth 8; 16 of the 256 spp (the CPU oracle's budget)."""
    from mitsuba_customization_b200 import bsdf, render as RD, synth
    from oracle import merl_oracle as O, render_oracle as RO
    plugs, orcs = [], []
    for k in range(100):
        raw = synth.synthetic_table(k)
        p = bsdf.MERL(data=raw, device="cuda:0")
        plugs.append(p)
        orcs.append(O.OracleTable(raw, cdf=p.cdf().cpu().numpy()))
    sc = RD.grid_scene(plugs, 10, 10)
    kw = dict(width=1024, height=1024, spp=16, max_depth=8, seed=1)
    film, st = RD.render(sc, want_stats=True, **kw)
    ref = RO.render(orcs, sc.spheres, cam_pos=sc.cam_pos, cam_target=sc.cam_target, cam_up=sc.cam_up,
                    fov_x_deg=sc.fov_x_deg, env=sc.env, **kw)
    got = film.cpu().numpy()
    np.testing.assert_allclose(got[..., 3], 1.0, atol=2e-5)
    r = _relmse(got[..., :3], ref[..., :3])
    _record("render_cfg5_fullsize.json", {"config": kw, "tables": 100, "rel_mse": r, "paths": st["paths"],
                                          "bsdf_queries": st["bsdf_queries"],
                                          "max_abs_pixel_diff": float(np.abs(got[..., :3] - ref[..., :3]).max())})
    assert st["bsdf_queries"] > st["paths"] * 0.2
    assert r <= REL_MSE, r
