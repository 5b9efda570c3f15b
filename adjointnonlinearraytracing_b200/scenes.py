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
