"""merl_eval_pdf_host_static: eval+pdf on host buffers for hosts that must not allocate device memory.

Who needs it: a renderer that has handed the whole device to its own memory pool (Dr.Jit's or
torch's caching allocator, or a scene of a hundred tables) so that the staging buffers
merl_eval_pdf_host would cudaMalloc on first use may not be there to get; a frame or optimiser loop
that cannot take the device-wide synchronisation cudaMalloc / cudaFree bring; a scalar-variant
plugin queried from many worker threads.  The entry point stages through a fixed 2.5 MiB array that
is part of libmerl_b200.so itself, so it is limited to MERL_STATIC_STAGE_MAX_QUERIES per call.

The array is shared by all calls on a device and keeps the previous call's data, so the test makes
two calls in a row -- the second one shorter, with different queries, so that stale queries and
results from the first sit right behind its own -- and wants both to equal merl_eval_pdf_host bit
for bit (the two entry points launch the same kernel; the suite's 1e-5 tolerance is not needed).

This file is named to be collected last: under `pytest -x` a failure here hides nothing.  It
builds its own table and depends on no state of the other tests, so it also runs alone.
"""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _ptr(a):
    return C.c_void_p(a.ctypes.data)


def _call(fn, handle, wi, wo, n, *tail):
    out = np.full((4, max(n, 1)), np.nan, dtype=np.float32)
    from mitsuba_customization_b200 import _cabi
    _cabi.check(fn(handle, _ptr(wi[0]), _ptr(wi[1]), _ptr(wi[2]), _ptr(wo[0]), _ptr(wo[1]), _ptr(wo[2]), n,
                   _ptr(out[0]), _ptr(out[1]), _ptr(out[2]), _ptr(out[3]), *tail))
    return out[:, :n]


def test_static_staging_matches_allocating_entry_twice_in_a_row(raw_table):
    from mitsuba_customization_b200 import _cabi, bsdf, synth
    L = _cabi.lib()
    plug = bsdf.MERL(data=raw_table, device="cuda:0")
    nmax = _cabi.STATIC_STAGE_MAX_QUERIES
    # first call: the full capacity; second: shorter, odd (vector body + scalar tail), other queries
    for n, seed in ((nmax, 71), (nmax // 3 + 1, 72)):
        wi, wo = synth.random_queries(n, seed=seed)
        wi, wo = np.ascontiguousarray(wi), np.ascontiguousarray(wo)
        # some below-horizon lanes: the kernel must store zeros there, not leave the last call's values
        wi[2, ::17] *= -1.0
        want = _call(L.merl_eval_pdf_host, plug._h, wi, wo, n)
        got = _call(L.merl_eval_pdf_host_static, plug._h, wi, wo, n, None)
        assert not np.isnan(got).any()
        assert np.array_equal(got, want), f"n={n}: {np.sum(got != want)} values differ"
        assert (got[:, ::17] == 0).all()
        assert got[:3].max() > 0 and got[3].max() > 0


def test_static_staging_rejects_oversize_and_accepts_empty(raw_table):
    from mitsuba_customization_b200 import _cabi, bsdf
    L = _cabi.lib()
    plug = bsdf.MERL(data=raw_table, device="cuda:0")
    n = _cabi.STATIC_STAGE_MAX_QUERIES + 1
    wi = np.zeros((3, n), dtype=np.float32)
    out = np.zeros((4, n), dtype=np.float32)
    rc = L.merl_eval_pdf_host_static(plug._h, _ptr(wi[0]), _ptr(wi[1]), _ptr(wi[2]), _ptr(wi[0]), _ptr(wi[1]),
                                     _ptr(wi[2]), n, _ptr(out[0]), _ptr(out[1]), _ptr(out[2]), _ptr(out[3]), None)
    assert rc == -1 and b"MERL_STATIC_STAGE_MAX_QUERIES" in L.merl_last_error()
    assert L.merl_eval_pdf_host_static(plug._h, None, None, None, None, None, None, 0, None, None, None, None, None) == 0
