"""CPU checks of the render oracle: Philox known answers (Random123 kat_vectors), exact background,
sample-range additivity.  No GPU."""
import numpy as np
import pytest

from oracle import render_oracle as R


@pytest.mark.parametrize("ctr,key,want", [
    ((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
    ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
    ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
     (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)),
])
def test_philox4x32_10_known_answers(ctr, key, want):
    assert R.philox(ctr, key) == want


CAM = dict(cam_pos=(0.0, -4.0, 0.0), cam_target=(0.0, 0.0, 0.0), fov_x_deg=40.0)


def test_sphere_render_background_and_range(oracle_table):
    img = R.render([oracle_table], [(0, 0, 0, 1.0, 0)], width=32, height=32, spp=4, max_depth=3, seed=7,
                   env=(1.0, 0.5, 0.25), **CAM)
    assert np.isfinite(img).all()
    np.testing.assert_allclose(img[..., 3], 1.0, rtol=0, atol=1e-12)          # every sample accounted for
    np.testing.assert_allclose(img[0, 0, :3], (1.0, 0.5, 0.25), rtol=1e-12)    # corner pixel sees only the env
    centre = img[16, 16, :3]
    assert (centre > 0).all() and (centre < np.array([1.0, 0.5, 0.25]) * 1.5).all()


def test_sample_ranges_add_up(oracle_table):
    kw = dict(width=24, height=24, spp=6, max_depth=3, seed=3, **CAM)
    full = R.render([oracle_table], [(0, 0, 0, 1.0, 0)], **kw)
    a = R.render([oracle_table], [(0, 0, 0, 1.0, 0)], spp_begin=0, spp_end=2, **kw)
    b = R.render([oracle_table], [(0, 0, 0, 1.0, 0)], spp_begin=2, spp_end=6, **kw)
    np.testing.assert_allclose(a + b, full, rtol=1e-12, atol=1e-15)


def test_render_refuses_a_table_without_sampler(raw_table):
    from oracle import merl_oracle as O
    with pytest.raises(ValueError):
        R.render([O.OracleTable(raw_table)], [(0, 0, 0, 1.0, 0)], width=4, height=4, spp=1, max_depth=1, seed=1, **CAM)
