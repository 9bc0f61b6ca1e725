"""ctypes front end of the CPU oracle (oracle/merl_oracle.c).  TEST INFRASTRUCTURE ONLY.

Importable from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
``--impl reference`` legs -- never from the product package.  PARITY UNPINNED: the
reference repository holds no plugin source (see the header of merl_oracle.c).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libmerl_oracle.so")

MERL_DIMS = (90, 90, 180)
MERL_SCALE = (1.0 / 1500.0, 1.15 / 1500.0, 1.66 / 1500.0)


def build(force: bool = False) -> str:
    """Compile the oracle with gcc via oracle/Makefile (a few seconds)."""
    src = os.path.join(_HERE, "merl_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)
    return _SO


class _Table(C.Structure):
    _fields_ = [
        ("nh", C.c_int32), ("nd", C.c_int32), ("np", C.c_int32),
        ("theta_h_sqrt", C.c_int32), ("ni", C.c_int32), ("clamp_negative", C.c_int32),
        ("scale", C.c_double * 3), ("spec_weight", C.c_double),
        ("raw", C.c_void_p), ("cdf", C.c_void_p), ("dens", C.c_void_p),
    ]


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.orc_max_threads.restype = C.c_int
        _lib.orc_theta_half_index.argtypes = [C.c_double, C.c_int, C.c_int]
        _lib.orc_theta_diff_index.argtypes = [C.c_double, C.c_int]
        _lib.orc_phi_diff_index.argtypes = [C.c_double, C.c_int]
        _lib.orc_build_cdf.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_void_p, C.c_void_p]
        _lib.orc_lookup_brdf_val.argtypes = [C.c_void_p] + [C.c_double] * 4 + [C.c_void_p]
    return _lib


def _p(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


def _dirs(w):
    w = np.ascontiguousarray(w, dtype=np.float32)
    assert w.ndim == 2 and w.shape[0] == 3, "directions are SoA (3, n) float32"
    return w


class OracleTable:
    """A MERL-layout table plus the sampler state, evaluated by the C oracle."""

    def __init__(self, raw, dims=MERL_DIMS, theta_h_sqrt=True, ni=32, scale=MERL_SCALE,
                 spec_weight=0.5, clamp_negative=True, cdf=None):
        self.dims = tuple(int(d) for d in dims)
        self.ncell = self.dims[0] * self.dims[1] * self.dims[2]
        self.raw = np.ascontiguousarray(raw, dtype=np.float64).reshape(3, self.ncell)
        self.t = _Table()
        self.t.nh, self.t.nd, self.t.np = self.dims
        self.t.theta_h_sqrt = int(theta_h_sqrt)
        self.t.ni = int(ni)
        self.t.clamp_negative = int(clamp_negative)
        self.t.scale = (C.c_double * 3)(*scale)
        self.t.spec_weight = float(spec_weight)
        self.t.raw = self.raw.ctypes.data
        self.cdf = None
        self.dens = None
        if cdf is not None:
            self.set_cdf(cdf)

    # -- sampler tables ---------------------------------------------------------------
    def set_cdf(self, cdf):
        nh = self.dims[0]
        self.cdf = np.ascontiguousarray(cdf, dtype=np.float32).reshape(self.t.ni, nh + 1)
        self.t.cdf = self.cdf.ctypes.data
        self.t.dens = None
        dens = np.empty((self.t.ni, nh), dtype=np.float64)
        lib().orc_build_dens(C.byref(self.t), _p(dens))
        self.dens = dens
        self.t.dens = dens.ctypes.data

    def build_cdf(self, nm=2, nq=64, floor_frac=0.01, install=True):
        nh = self.dims[0]
        cdf = np.empty((self.t.ni, nh + 1), dtype=np.float32)
        mass = np.empty((self.t.ni, nh), dtype=np.float64)
        lib().orc_build_cdf(C.addressof(self.t), nm, nq, floor_frac, cdf.ctypes.data, mass.ctypes.data)
        if install:
            self.set_cdf(cdf)
        return cdf, mass

    def m_edges(self):
        m = np.empty(self.dims[0] + 1, dtype=np.float64)
        lib().orc_m_edges(C.byref(self.t), _p(m))
        return m

    # -- hot path ---------------------------------------------------------------------
    def bins(self, wi, wo):
        wi, wo = _dirs(wi), _dirs(wo)
        n = wi.shape[1]
        out = {k: np.empty(n, dtype=np.int32) for k in ("ih", "id", "ip", "idx")}
        stable = np.empty(n, dtype=np.uint8)
        lib().orc_bins(C.byref(self.t), _p(wi[0]), _p(wi[1]), _p(wi[2]), _p(wo[0]), _p(wo[1]), _p(wo[2]),
                       C.c_int64(n), _p(out["ih"]), _p(out["id"]), _p(out["ip"]), _p(out["idx"]), _p(stable))
        out["stable"] = stable.astype(bool)
        return out

    def bins_f32(self, wi, wo):
        """Flat cell index from the all-float twin of the binning (what a float variant of a straight port of the
        public reader would compute): for measuring the double <-> float disagreement, not a parity target."""
        wi, wo = _dirs(wi), _dirs(wo)
        n = wi.shape[1]
        idx = np.empty(n, dtype=np.int32)
        lib().orc_bins_f32(C.byref(self.t), _p(wi[0]), _p(wi[1]), _p(wi[2]), _p(wo[0]), _p(wo[1]), _p(wo[2]),
                           C.c_int64(n), _p(idx))
        return idx

    def eval_pdf(self, wi, wo, want_pdf=True):
        if want_pdf and self.cdf is None:
            # the C side would return the cosine lobe's share only: a silently incomplete density
            raise ValueError("pdf needs the sampler CDF (set_cdf / build_cdf); pass want_pdf=False for eval alone")
        wi, wo = _dirs(wi), _dirs(wo)
        n = wi.shape[1]
        rgb = np.empty((3, n), dtype=np.float64)
        pdf = np.empty(n, dtype=np.float64) if want_pdf else None
        idx = np.empty(n, dtype=np.int32)
        lib().orc_eval_pdf(C.byref(self.t), _p(wi[0]), _p(wi[1]), _p(wi[2]), _p(wo[0]), _p(wo[1]), _p(wo[2]),
                           C.c_int64(n), _p(rgb), _p(pdf), _p(idx))
        return rgb, pdf, idx

    def sample(self, wi, u1, u2):
        assert self.cdf is not None, "install a CDF first (set_cdf / build_cdf)"
        wi = _dirs(wi)
        n = wi.shape[1]
        u1 = np.ascontiguousarray(u1, dtype=np.float32)
        u2 = np.ascontiguousarray(u2, dtype=np.float32).reshape(2, n)
        wo = np.empty((3, n), dtype=np.float64)
        pdf = np.empty(n, dtype=np.float64)
        weight = np.empty((3, n), dtype=np.float64)
        lobe = np.empty(n, dtype=np.int32)
        lib().orc_sample(C.byref(self.t), _p(wi[0]), _p(wi[1]), _p(wi[2]), _p(u1), _p(u2[0]), _p(u2[1]),
                         C.c_int64(n), _p(wo), _p(pdf), _p(weight), _p(lobe))
        return wo, pdf, weight, lobe

    def backward(self, wi, wo, grad):
        wi, wo = _dirs(wi), _dirs(wo)
        g = _dirs(grad)
        n = wi.shape[1]
        out = np.zeros((self.ncell, 4), dtype=np.float64)
        lib().orc_backward(C.byref(self.t), _p(wi[0]), _p(wi[1]), _p(wi[2]), _p(wo[0]), _p(wo[1]), _p(wo[2]),
                           _p(g[0]), _p(g[1]), _p(g[2]), C.c_int64(n), _p(out))
        return out

    def lookup_brdf_val(self, theta_in, fi_in, theta_out, fi_out):
        rgb = np.empty(3, dtype=np.float64)
        lib().orc_lookup_brdf_val(C.addressof(self.t), theta_in, fi_in, theta_out, fi_out, rgb.ctypes.data)
        return rgb

    def prescaled_float4(self):
        """The device table this oracle corresponds to: [ncell][4] fp32, clamped."""
        out = np.zeros((self.ncell, 4), dtype=np.float32)
        for c in range(3):
            v = self.raw[c] * self.t.scale[c]
            if self.t.clamp_negative:
                v = np.maximum(v, 0.0)
            out[:, c] = v.astype(np.float32)
        return out


def half_diff(wi, wo):
    hd = np.empty(4, dtype=np.float64)
    a = np.ascontiguousarray(wi, dtype=np.float64)
    b = np.ascontiguousarray(wo, dtype=np.float64)
    lib().orc_half_diff(_p(a), _p(b), _p(hd))
    return hd


def read_merl(path):
    dims = (C.c_int32 * 3)()
    ptr = C.c_void_p()
    rc = lib().orc_read_merl(path.encode(), dims, C.byref(ptr))
    if rc != 0:
        raise IOError(f"orc_read_merl({path}) failed: {rc}")
    n = dims[0] * dims[1] * dims[2]
    arr = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_double)), shape=(3 * n,)).copy()
    lib().orc_free(ptr)
    return tuple(dims), arr.reshape(3, n)


def write_merl(path, dims, raw):
    d = (C.c_int32 * 3)(*dims)
    raw = np.ascontiguousarray(raw, dtype=np.float64)
    rc = lib().orc_write_merl(path.encode(), d, _p(raw))
    if rc != 0:
        raise IOError(f"orc_write_merl({path}) failed: {rc}")


def max_threads():
    return lib().orc_max_threads()


def set_threads(n):
    lib().orc_set_threads(int(n))
