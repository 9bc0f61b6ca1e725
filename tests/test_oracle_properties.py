"""Property tests (hypothesis) of the oracle's index functions: the statements of the public MERL reader
 that hold for EVERY angle, not only the hand-picked ones of test_oracle.py --
ranges, monotonicity, the literal formulas evaluated in Python doubles, the phi_d fold, and that the flat
index a direction pair resolves to is the one the three angle indices compose."""
import math

import numpy as np
import pytest

hypothesis = pytest.importorskip("hypothesis")
from hypothesis import given, settings, strategies as st  # noqa: E402

from oracle import merl_oracle as O  # noqa: E402

RES = st.sampled_from([(90, 90, 180), (48, 64, 120), (64, 96, 192), (7, 5, 3)])
ANGLE = st.floats(min_value=-0.5, max_value=math.pi, allow_nan=False)


@settings(max_examples=400, deadline=None)
@given(RES, ANGLE, ANGLE)
def test_theta_indices_range_formula_and_monotonicity(res, a, b):
    L = O.lib()
    nh, nd, _ = res
    lo, hi = min(a, b), max(a, b)
    for sqrt_map in (0, 1):
        ia, ib = L.orc_theta_half_index(lo, nh, sqrt_map), L.orc_theta_half_index(hi, nh, sqrt_map)
        assert 0 <= ia <= ib <= nh - 1
        if lo <= 0:
            want = 0
        elif sqrt_map:                                   # BRDFRead.cpp: deg = th / (pi/2) * n; (int) sqrt(deg * n)
            want = min(max(int(math.sqrt(lo / (math.pi / 2) * nh * nh)), 0), nh - 1)
        else:
            want = min(max(int(lo / (math.pi / 2) * nh), 0), nh - 1)
        assert ia == want
    ja, jb = L.orc_theta_diff_index(lo, nd), L.orc_theta_diff_index(hi, nd)
    assert 0 <= ja <= jb <= nd - 1
    assert ja == min(max(int(lo / (math.pi * 0.5) * nd), 0), nd - 1)


@settings(max_examples=400, deadline=None)
@given(RES, st.floats(min_value=-math.pi, max_value=math.pi, allow_nan=False))
def test_phi_index_folds_by_pi(res, phi):
    L = O.lib()
    n_p = res[2]
    i = L.orc_phi_diff_index(phi, n_p)
    assert 0 <= i <= n_p - 1
    folded = phi + math.pi if phi < 0 else phi          # BRDFRead.cpp: if (phi < 0) phi += M_PI
    assert i == min(max(int(folded / math.pi * n_p), 0), n_p - 1)


@settings(max_examples=60, deadline=None)
@given(st.integers(min_value=0, max_value=2**31 - 1))
def test_flat_index_composes_the_three_indices(seed):
    from mitsuba_customization_b200 import synth
    dims = (12, 10, 16)
    t = O.OracleTable(synth.synthetic_table(1, dims=dims), dims=dims)
    wi, wo = synth.random_queries(64, seed=seed)
    b = t.bins(wi, wo)
    assert np.array_equal(b["idx"], b["ip"] + dims[2] * b["id"] + dims[2] * dims[1] * b["ih"])
    assert b["idx"].min() >= 0 and b["idx"].max() < dims[0] * dims[1] * dims[2]
    rgb, _, idx = t.eval_pdf(wi, wo, want_pdf=False)
    assert np.array_equal(idx, b["idx"])
    # eval = raw cell * channel scale * cos(theta_o) in doubles, negatives clamped (assumptions B2, B3)
    raw = np.asarray(t.raw, np.float64).reshape(3, -1)
    want = np.maximum(raw[:, idx] * np.asarray(O.MERL_SCALE)[:, None], 0.0) * wo[2].astype(np.float64)
    np.testing.assert_allclose(rgb, want, rtol=1e-14, atol=0)


def test_f32_twin_of_the_binning_mostly_agrees_and_only_moves_to_neighbours(raw_table):
    """bins_f32 (every step in float) is a measuring stick, not a parity target: it must
    agree with the double path on the bulk of random queries and, where it differs, sit in a neighbouring cell."""
    from mitsuba_customization_b200 import synth
    from oracle import merl_oracle as O
    from tests.parity_risk_report import measure
    orc = O.OracleTable(raw_table)
    wi, wo = synth.random_queries(1 << 20, seed=77)
    r = measure(orc, wi, wo)
    assert 0 < r["cells_differ"] < 0.01 * r["n"]
    assert r["neighbour_phi"] + r["neighbour_theta_d"] + r["neighbour_theta_h"] >= 0.9 * r["cells_differ"]
    assert r["ld_unstable"] < r["cells_differ"]          # double is far closer to long double than float is to double
