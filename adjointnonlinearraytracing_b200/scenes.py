"""Seeded synthetic IOR grids and ray batches.

Shapes follow the reference's experiment drivers: a cubic grid over a unit span
(``h = 1/(N_g - 1)``, ``ds = h / step_res``; reference ``core/luneburg_opt.py:85-89``),
a square sensor of rays on the ``z = 0`` face (reference ``src/test.cpp:25-32``).
Everything is numpy float32 on the host; callers move it where they need it.
Tensor axes are ``rif[iz, iy, ix]`` so that position component x indexes the
fastest axis, matching ``idx = (z*H + y)*W + x`` (reference ``src/volume.cpp:134``).
"""
from __future__ import annotations

import numpy as np


def grid_spacing(n_g: int, span: float = 1.0) -> np.float32:
    return np.float32(span / (n_g - 1))


def _coords(n_g: int):
    h = 1.0 / (n_g - 1)
    c = np.arange(n_g, dtype=np.float64) * h
    return np.meshgrid(c, c, c, indexing="ij")  # z, y, x


def field_bump(n_g: int, amp: float = 0.2, sigma: float = 0.2) -> np.ndarray:
    """Field A: n = 1 + amp * exp(-|p - c|^2 / (2 sigma^2)), c = (.5, .5, .5)."""
    z, y, x = _coords(n_g)
    r2 = (x - 0.5) ** 2 + (y - 0.5) ** 2 + (z - 0.5) ** 2
    return (1.0 + amp * np.exp(-r2 / (2 * sigma * sigma))).astype(np.float32)


def field_luneburg(n_g: int, radius: float = 0.45) -> np.ndarray:
    """Field B: n = sqrt(2 - (r/R)^2) inside the lens, 1 outside (reference core/luneburg_opt.py)."""
    z, y, x = _coords(n_g)
    r2 = ((x - 0.5) ** 2 + (y - 0.5) ** 2 + (z - 0.5) ** 2) / (radius * radius)
    return np.where(r2 < 1.0, np.sqrt(np.maximum(2.0 - r2, 1.0)), 1.0).astype(np.float32)


def field_noise(n_g: int, seed: int = 0, eps: float = 0.01) -> np.ndarray:
    """Field C: bump + eps * U(-1, 1) per voxel."""
    rng = np.random.default_rng(seed)
    return (field_bump(n_g) + eps * rng.uniform(-1, 1, size=(n_g,) * 3)).astype(np.float32)


def field_uniform(n_g: int, value: float = 1.0) -> np.ndarray:
    return np.full((n_g,) * 3, value, dtype=np.float32)


def sdf_sphere(n_g: int, radius: float = 0.4) -> np.ndarray:
    """Signed distance to a centred sphere (negative inside), for trace_sdf."""
    z, y, x = _coords(n_g)
    return (np.sqrt((x - 0.5) ** 2 + (y - 0.5) ** 2 + (z - 0.5) ** 2) - radius).astype(np.float32)


def rays_collimated(s: int, inset: float = 0.005, z0: float = 0.0):
    """S x S rays on the plane z = z0, direction +z (reference src/test.cpp:25-32)."""
    u = np.linspace(inset, 1.0 - inset, s, dtype=np.float32)
    yy, xx = np.meshgrid(u, u, indexing="ij")
    pos = np.stack([xx.ravel(), yy.ravel(), np.full(s * s, z0, dtype=np.float32)], axis=1)
    vel = np.tile(np.array([0, 0, 1], dtype=np.float32), (s * s, 1))
    return np.ascontiguousarray(pos, dtype=np.float32), vel


def rays_cone(s: int, spread: float = 0.15, seed: int = 0, inset: float = 0.005, z0: float = 0.0,
              backoff: float = 0.0):
    """Same origins, direction normalize((0,0,1) + spread * N(0, I)); `backoff` moves each origin
    back along its direction (rays then start outside the box and fly straight to it)."""
    pos, vel = rays_collimated(s, inset, z0)
    rng = np.random.default_rng(seed)
    vel = vel + spread * rng.standard_normal(vel.shape).astype(np.float32)
    vel = (vel / np.linalg.norm(vel, axis=1, keepdims=True)).astype(np.float32)
    if backoff:
        pos = (pos - np.float32(backoff) * vel).astype(np.float32)
    return pos, vel


def seeds_ones(n: int):
    """dx = dv = 1 (reference src/test.cpp:142-144)."""
    one = np.ones((n, 3), dtype=np.float32)
    return one, one.copy()


def seeds_normal(n: int, seed: int = 1):
    rng = np.random.default_rng(seed)
    return (rng.standard_normal((n, 3)).astype(np.float32),
            rng.standard_normal((n, 3)).astype(np.float32))


def cable_profile(res: int = 129, n0: float = 1.5, delta: float = 0.05) -> np.ndarray:
    """Parabolic graded-index fibre profile n(r) = n0 * (1 - delta * (r/R)^2)."""
    r = np.linspace(0.0, 1.0, res)
    return (n0 * (1.0 - delta * r * r)).astype(np.float32)


def cable_rays(s: int, radius: float = 1.0, seed: int = 0, spread: float = 0.05):
    """S x S rays entering the fibre face y = 0 inside 0.8 R, mostly along +y."""
    rng = np.random.default_rng(seed)
    n = s * s
    rr = 0.8 * radius * np.sqrt(rng.uniform(0, 1, n))
    th = rng.uniform(0, 2 * np.pi, n)
    pos = np.stack([radius + rr * np.cos(th), np.full(n, 1e-3), radius + rr * np.sin(th)], axis=1)
    vel = np.array([0, 1, 0]) + spread * rng.standard_normal((n, 3))
    vel /= np.linalg.norm(vel, axis=1, keepdims=True)
    return pos.astype(np.float32), vel.astype(np.float32)


def rays_multiview(n_views: int, s: int, spp: int, seed: int = 0):
    """n_views collimated s x s sensors around the volume centre (rotation about y), spp jittered samples
    per pixel, launched outside the box: the batch shape of the reference's tomography driver
    (path_matrix/run_fuel_injection_2008.py:80-92, core/source.py:345-357)."""
    rng = np.random.default_rng(seed)
    u = (np.arange(s) + 0.5) / s
    yy, xx = np.meshgrid(u, u, indexing="ij")
    P, V = [], []
    for k in range(n_views):
        th = np.pi * k / n_views
        d = np.array([np.sin(th), 0.0, np.cos(th)])
        e1 = np.array([np.cos(th), 0.0, -np.sin(th)]); e2 = np.array([0.0, 1.0, 0.0])
        for _ in range(spp):
            jx = (rng.uniform(-0.5, 0.5, xx.shape)) / s; jy = (rng.uniform(-0.5, 0.5, xx.shape)) / s
            a = (xx + jx - 0.5) * 0.7; b = (yy + jy - 0.5) * 0.7
            p = 0.5 + a[..., None] * e1 + b[..., None] * e2 - 0.9 * d
            P.append(p.reshape(-1, 3)); V.append(np.tile(d, (s * s, 1)))
    return np.concatenate(P).astype(np.float32), np.concatenate(V).astype(np.float32)


# The shapes of BASELINE.json's five configurations (SURVEY.md section 8 size table) plus the reference's own
# measurement protocol (src/test.cpp:148-206), as synthetic seeded workloads.  Used by bench.py (`configs`
# entries of the JSON line), tools/gpu_configs.py and the reference-parity tests.
CONFIG_NAMES = ("luneburg32", "fiber64", "cable129", "fuel65", "caustic256", "caustic256_cone", "focal512")


def config_case(name: str) -> dict:
    """-> dict(kind="grid"|"cable", label, rif, h / radius / length, ds, pos, vel[, target])."""
    def grid(label, ng, field, pos, vel, step_res=2):
        h = grid_spacing(ng)
        return dict(kind="grid", label=label, ng=ng, rif=field(ng), h=h, ds=np.float32(h / step_res), pos=pos, vel=vel)
    if name == "luneburg32":
        return grid("configs[0] Luneburg lens: 32^3 lens field, 256^2 collimated rays", 32, field_luneburg, *rays_collimated(256))
    if name == "fiber64":
        return grid("configs[1] fibre, grid form: 64^3, 512^2 collimated rays", 64, field_bump, *rays_collimated(512))
    if name == "cable129":
        res, radius, length, s = 129, 1.0, 5.0, 512
        pos, vel = cable_rays(s, radius)
        tgt = np.tile(np.array([radius, length, radius], np.float32), (s * s, 1))
        return dict(kind="cable", label="configs[1] fibre, cable form: 129-sample profile, R = 1, L = 5, 512^2 rays, ds = R/res/2",
                    rif=cable_profile(res), radius=radius, length=length, ds=np.float32(radius / res / 2), pos=pos, vel=vel,
                    target=tgt)
    if name == "fuel65":
        return grid("configs[2] fuel-injection tomography: 65^3, 32 views x 64^2 x 16 spp", 65, field_bump, *rays_multiview(32, 64, 16))
    if name == "caustic256":
        return grid("configs[3] caustic design: 256^3, 2048^2 collimated rays", 256, field_bump, *rays_collimated(2048))
    if name == "caustic256_cone":
        return grid("configs[3'] caustic design, oblique batch: 256^3, 2048^2 rays, cone spread 0.15", 256, field_bump,
                    *rays_cone(2048, spread=0.15))
    if name == "focal512":
        return grid("configs[4] focal stack: 512^3, 1448^2 = 2.1 M rays (one GPU's share of 16.8 M over 8)", 512, field_bump,
                    *rays_collimated(1448))
    raise KeyError(name)


def subsample(n_total: int, n: int) -> np.ndarray:
    """Indices of a strided subsample of n rays spread over the whole batch."""
    if n >= n_total:
        return np.arange(n_total)
    return (np.arange(n, dtype=np.int64) * n_total) // n
