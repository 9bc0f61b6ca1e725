"""Host side of the wavefront path tracer harness (libmerl_b200's merl_render).

Scenes are spheres under a constant environment map seen by a pinhole camera -- the callers either
side of the BSDF stage for BASELINE configs 2 (one sphere, one `merl` BSDF) and 5 (many spheres,
many tables).  Nothing here computes: the film is a torch CUDA tensor handed to the library by
pointer.  Multi-GPU: every rank renders a disjoint range of the sample indices of every pixel and
the films are summed with one allreduce (distributed.Comm), BASELINE.json's "film accumulation".
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple

import torch

from . import _cabi
from ._cabi import MerlError, check


class _Sphere(C.Structure):
    _fields_ = [("cx", C.c_float), ("cy", C.c_float), ("cz", C.c_float), ("radius", C.c_float),
                ("material", C.c_int32), ("pad", C.c_int32 * 3)]


class _Params(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("spp_total", C.c_int32), ("spp_begin", C.c_int32),
                ("spp_end", C.c_int32), ("max_depth", C.c_int32), ("wavefront", C.c_int32), ("reserved", C.c_int32),
                ("seed", C.c_uint64), ("cam_pos", C.c_float * 3), ("cam_target", C.c_float * 3),
                ("cam_up", C.c_float * 3), ("fov_x_deg", C.c_float), ("env", C.c_float * 3),
                ("table_shard_world", C.c_int32), ("table_shard_rank", C.c_int32), ("stage_exchange", C.c_void_p)]


class _Stats(C.Structure):
    _fields_ = [("paths", C.c_int64), ("bsdf_queries", C.c_int64), ("launches", C.c_int64)]


@dataclass
class Scene:
    """spheres: (cx, cy, cz, radius, material index into `bsdfs`)."""
    bsdfs: Sequence            # MERL / CustomizedMeasurement plugin objects, all on one device
    spheres: List[Tuple[float, float, float, float, int]]
    cam_pos: Tuple[float, float, float] = (0.0, -4.0, 0.0)
    cam_target: Tuple[float, float, float] = (0.0, 0.0, 0.0)
    cam_up: Tuple[float, float, float] = (0.0, 0.0, 1.0)
    fov_x_deg: float = 40.0
    env: Tuple[float, float, float] = (1.0, 1.0, 1.0)


def sphere_scene(bsdf) -> Scene:
    """BASELINE config 2: a unit sphere under a constant environment map."""
    return Scene(bsdfs=[bsdf], spheres=[(0.0, 0.0, 0.0, 1.0, 0)])


def grid_scene(bsdfs, nx=10, ny=10, spacing=2.2) -> Scene:
    """BASELINE config 5: an nx x ny grid of unit spheres, one material each (table k on sphere k),
    close enough to inter-reflect, seen from above at an angle."""
    sph = []
    for k in range(nx * ny):
        i, j = k % nx, k // nx
        sph.append(((i - (nx - 1) / 2) * spacing, (j - (ny - 1) / 2) * spacing, 0.0, 1.0, k % len(bsdfs)))
    ext = max(nx, ny) * spacing
    return Scene(bsdfs=list(bsdfs), spheres=sph, cam_pos=(0.0, -1.6 * ext, 1.3 * ext), cam_target=(0.0, 0.0, 0.0),
                 fov_x_deg=40.0)


def render(scene: Scene, width: int, height: int, spp: int, max_depth: int = 8, seed: int = 1,
           spp_range: Optional[Tuple[int, int]] = None, film: Optional[torch.Tensor] = None, wavefront: int = 0,
           want_stats: bool = False, table_shard: Optional[Tuple[int, int]] = None, stage_exchange=None):
    """Adds this call's samples to `film` ((height, width, 4) float32 CUDA, zeroed when None) and
    returns it (and the stats when asked; asking synchronises the stream).

    table_shard = (world, rank) + stage_exchange (a distributed.PeerBuffer of >= 12 * paths-per-pass
    floats): the by-table variant -- every rank traces the same samples, evaluates only the tables
    k % world == rank (scene.bsdfs[k] may be None for the others) and the stage outputs are summed
    over the ranks once per bounce; every rank ends with the full film."""
    dev = next(b for b in scene.bsdfs if b is not None).device
    if film is None:
        film = torch.zeros(height, width, 4, dtype=torch.float32, device=dev)
    if film.shape != (height, width, 4) or film.dtype != torch.float32 or not film.is_contiguous():
        raise MerlError("film must be a contiguous (height, width, 4) float32 CUDA tensor")
    handles = (C.c_void_p * len(scene.bsdfs))(*[(b._h if b is not None else None) for b in scene.bsdfs])
    sph = (_Sphere * len(scene.spheres))(*[_Sphere(s[0], s[1], s[2], s[3], int(s[4])) for s in scene.spheres])
    b, e = spp_range if spp_range is not None else (0, spp)
    p = _Params(width, height, spp, b, e, max_depth, wavefront, 0, seed, (C.c_float * 3)(*scene.cam_pos),
                (C.c_float * 3)(*scene.cam_target), (C.c_float * 3)(*scene.cam_up), scene.fov_x_deg,
                (C.c_float * 3)(*scene.env), table_shard[0] if table_shard else 0, table_shard[1] if table_shard else 0,
                stage_exchange._h if stage_exchange is not None else None)
    st = _Stats()
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    check(_cabi.lib().merl_render(handles, len(scene.bsdfs), sph, len(scene.spheres), C.byref(p),
                                  C.c_void_p(film.data_ptr()), C.byref(st) if want_stats else None, stream))
    if want_stats:
        return film, {"paths": st.paths, "bsdf_queries": st.bsdf_queries, "launches": st.launches}
    return film


def spp_share(spp: int, world: int, rank: int) -> Tuple[int, int]:
    """Sample-index range of `rank` when `world` GPUs split the samples of every pixel."""
    per = -(-spp // world)
    return min(rank * per, spp), min((rank + 1) * per, spp)
