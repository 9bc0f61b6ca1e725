"""ctypes front end of oracle/render_oracle.c (the CPU restatement of merl_render).
TEST INFRASTRUCTURE ONLY -- see the header of render_oracle.c."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from . import merl_oracle as O

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libmerl_render_oracle.so")


class _Params(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("spp_total", C.c_int32), ("spp_begin", C.c_int32),
                ("spp_end", C.c_int32), ("max_depth", C.c_int32), ("seed", C.c_uint64),
                ("cam_pos", C.c_float * 3), ("cam_target", C.c_float * 3), ("cam_up", C.c_float * 3),
                ("fov_x_deg", C.c_float), ("env", C.c_float * 3)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        srcs = [os.path.join(_HERE, f) for f in ("render_oracle.c", "merl_oracle.c")]
        if not os.path.exists(_SO) or os.path.getmtime(_SO) < max(os.path.getmtime(s) for s in srcs):
            subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)
        _lib = C.CDLL(_SO)
    return _lib


def philox(ctr, key):
    c = (C.c_uint32 * 4)(*ctr)
    k = (C.c_uint32 * 2)(*key)
    out = (C.c_uint32 * 4)()
    lib().orc_philox(c, k, out)
    return tuple(int(x) for x in out)


def render(tables, spheres, *, width, height, spp, max_depth, seed, cam_pos, cam_target, cam_up=(0, 0, 1),
           fov_x_deg=40.0, env=(1.0, 1.0, 1.0), spp_begin=0, spp_end=None):
    """tables: list of OracleTable (CDF installed); spheres: [(cx, cy, cz, radius, material)].
    Returns the film, (height, width, 4) float64: rgb and the sample-count share."""
    for k, t in enumerate(tables):
        if not t.t.cdf or not t.t.dens:
            raise ValueError(f"table {k} has no sampler CDF installed (build_cdf() / set_cdf() first)")
    arr = (O._Table * len(tables))(*[t.t for t in tables])
    sph = np.ascontiguousarray([[s[0], s[1], s[2], s[3]] for s in spheres], dtype=np.float64)
    mat = np.ascontiguousarray([int(s[4]) for s in spheres], dtype=np.int32)
    p = _Params(width, height, spp, spp_begin, spp if spp_end is None else spp_end, max_depth, seed,
                (C.c_float * 3)(*cam_pos), (C.c_float * 3)(*cam_target), (C.c_float * 3)(*cam_up), fov_x_deg,
                (C.c_float * 3)(*env))
    film = np.zeros((height * width, 4), dtype=np.float64)
    lib().orc_render(arr, len(tables), C.c_void_p(sph.ctypes.data), C.c_void_p(mat.ctypes.data), len(spheres),
                     C.byref(p), C.c_void_p(film.ctypes.data))
    return film.reshape(height, width, 4)
