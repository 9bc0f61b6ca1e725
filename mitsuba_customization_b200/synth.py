"""Synthetic MERL-format tables and query batches.

No measured data ships with the reference (its /data and /scene are git-ignored,
/root/reference/.gitignore:31,34), so every table used by tests and benchmarks is
generated here: an analytic Lambert + GGX BRDF evaluated at the bin centres of the MERL
half/difference grid, divided by the channel scale so raw magnitudes look like a real
file, perturbed by log-normal noise, with below-horizon cells set to -1 as real files do.
"""
from __future__ import annotations

import functools

import numpy as np

MERL_DIMS = (90, 90, 180)
MERL_SCALE = (1.0 / 1500.0, 1.15 / 1500.0, 1.66 / 1500.0)
TABLE_SEED = 0x4D45524C  # "MERL"


def bin_centres(dims=MERL_DIMS, theta_h_sqrt=True):
    """(theta_h, theta_d, phi_d) at the centre of every cell, flat MERL order."""
    nh, nd, np_ = dims
    uh = (np.arange(nh) + 0.5) / nh
    th = (np.pi / 2) * (uh * uh if theta_h_sqrt else uh)
    td = (np.arange(nd) + 0.5) / nd * (np.pi / 2)
    pd = (np.arange(np_) + 0.5) / np_ * np.pi
    TH, TD, PD = np.meshgrid(th, td, pd, indexing="ij")
    return TH.ravel(), TD.ravel(), PD.ravel()


def half_diff_to_dirs(th, td, pd):
    """Inverse of the half/difference map with phi_half = 0: returns (wi, wo), each (3, n)."""
    d = np.stack([np.sin(td) * np.cos(pd), np.sin(td) * np.sin(pd), np.cos(td)])
    c, s = np.cos(th), np.sin(th)
    wi = np.stack([d[0] * c + d[2] * s, d[1], d[2] * c - d[0] * s])  # rotate about y by +theta_h
    h = np.stack([s, np.zeros_like(s), c])
    wo = 2.0 * (wi * h).sum(0) * h - wi
    return wi, wo


def analytic_brdf(wi, wo, albedo, alpha, f0):
    """Lambert + GGX microfacet lobe, RGB, without the cosine factor.  (3, n)."""
    h = wi + wo
    h = h / np.maximum(np.linalg.norm(h, axis=0), 1e-30)
    ci, co = np.maximum(wi[2], 1e-4), np.maximum(wo[2], 1e-4)
    ch = np.clip(h[2], 0.0, 1.0)
    cd = np.clip((wi * h).sum(0), 0.0, 1.0)
    a2 = alpha * alpha
    D = a2 / (np.pi * (ch * ch * (a2 - 1.0) + 1.0) ** 2)
    G = (2 * ci / (ci + np.sqrt(a2 + (1 - a2) * ci * ci))) * (2 * co / (co + np.sqrt(a2 + (1 - a2) * co * co)))
    F = f0 + (1.0 - f0) * (1.0 - cd) ** 5
    spec = D * G * F / (4.0 * ci * co)
    return np.asarray(albedo)[:, None] / np.pi + spec[None, :]


@functools.lru_cache(maxsize=4)
def _cell_directions(dims, theta_h_sqrt):
    """(wi, wo) at every cell centre: the same for every table of a layout, so built once."""
    th, td, pd = bin_centres(dims, theta_h_sqrt)
    wi, wo = half_diff_to_dirs(th, td, pd)
    wi.setflags(write=False)
    wo.setflags(write=False)
    return wi, wo


def synthetic_table(table_id=0, dims=MERL_DIMS, theta_h_sqrt=True, scale=MERL_SCALE, noise=0.1):
    """Raw planar table (3, ncell) float64 in .binary units (value / channel scale)."""
    rng = np.random.Generator(np.random.PCG64(TABLE_SEED + int(table_id)))
    albedo = rng.uniform(0.05, 0.6, size=3)
    alpha = float(rng.uniform(0.05, 0.5))
    f0 = float(rng.uniform(0.02, 0.2))
    wi, wo = _cell_directions(tuple(int(d) for d in dims), bool(theta_h_sqrt))
    f = analytic_brdf(wi, wo, albedo, alpha, f0)
    f = f * np.exp(noise * rng.standard_normal(f.shape))
    raw = f / np.asarray(scale)[:, None]
    below = (wi[2] <= 0.0) | (wo[2] <= 0.0)
    raw[:, below] = -1.0
    return np.ascontiguousarray(raw)


def write_merl_binary(path, raw, dims=MERL_DIMS):
    with open(path, "wb") as f:
        f.write(np.asarray(dims, dtype="<i4").tobytes())
        f.write(np.ascontiguousarray(raw, dtype="<f8").tobytes())


def read_merl_binary(path):
    with open(path, "rb") as f:
        dims = tuple(int(x) for x in np.frombuffer(f.read(12), dtype="<i4"))
        raw = np.frombuffer(f.read(), dtype="<f8")
    n = dims[0] * dims[1] * dims[2]
    if raw.size != 3 * n:
        raise ValueError(f"{path}: expected {3 * n} doubles, found {raw.size}")
    return dims, raw.reshape(3, n).copy()


def uniform_hemisphere(n, rng):
    """Directions uniform on the upper hemisphere, fp32 SoA (3, n)."""
    z = rng.random(n, dtype=np.float64)
    z = np.maximum(z, 1e-6)
    phi = rng.random(n, dtype=np.float64) * (2 * np.pi)
    r = np.sqrt(np.maximum(0.0, 1.0 - z * z))
    return np.ascontiguousarray(np.stack([r * np.cos(phi), r * np.sin(phi), z]).astype(np.float32))


def random_queries(n, seed=1234):
    rng = np.random.Generator(np.random.Philox(seed))
    return uniform_hemisphere(n, rng), uniform_hemisphere(n, rng)


def edge_stress_queries(n, dims=MERL_DIMS, theta_h_sqrt=True, seed=99, spread=(1e-3, 1e-5, 1e-7, 1e-9, 0.0)):
    """Queries whose (theta_h, theta_d, phi_d) sit within `spread` radians of a bin edge in
    one of the three coordinates.  fp32 SoA."""
    rng = np.random.Generator(np.random.Philox(seed))
    nh, nd, np_ = dims
    k = rng.integers(0, 3, size=n)
    uh = rng.random(n)
    th = (np.pi / 2) * (uh * uh if theta_h_sqrt else uh) * 0.98
    td = rng.random(n) * (np.pi / 2) * 0.98
    pd = rng.random(n) * np.pi
    eh = rng.integers(1, nh, size=n) / nh
    eh = (np.pi / 2) * (eh * eh if theta_h_sqrt else eh)
    ed = rng.integers(1, nd, size=n) / nd * (np.pi / 2)
    ep = rng.integers(0, np_ + 1, size=n) / np_ * np.pi
    off = rng.choice(np.asarray(spread), size=n) * rng.choice([-1.0, 1.0], size=n)
    th = np.where(k == 0, eh + off, th)
    td = np.where(k == 1, ed + off, td)
    pd = np.where(k == 2, ep + off, pd)
    wi, wo = half_diff_to_dirs(th, td, pd)
    # spin about the normal so phi_half is not always 0
    a = rng.random(n) * 2 * np.pi
    ca, sa = np.cos(a), np.sin(a)

    def spin(w):
        return np.stack([w[0] * ca - w[1] * sa, w[0] * sa + w[1] * ca, w[2]])

    wi, wo = spin(wi), spin(wo)
    keep = (wi[2] > 1e-4) & (wo[2] > 1e-4)
    return (np.ascontiguousarray(wi[:, keep].astype(np.float32)),
            np.ascontiguousarray(wo[:, keep].astype(np.float32)))
