"""On-device ray generation with the names and call signatures of the reference's
``core/source.py``.

The reference builds every batch with CPU torch ops (``torch.rand`` + meshgrid + matmul)
and copies it to the GPU each optimiser iteration (reference core/luneburg_opt.py:53-59);
here one CUDA kernel (``drrt_plane_source_rand``) writes x, v and the plane triples
straight into HBM.  Extra keyword arguments that the reference does not have:
``device`` (default ``"cuda"``), ``seed`` (in-kernel Philox jitter; default: drawn from
torch's global generator so ``torch.manual_seed`` makes runs repeatable) and ``offset``
(explicit uniforms laid out like the reference's ``torch.rand(2*spp, P0, P1)`` -- used by
the parity tests to reproduce the reference's batch exactly).
"""
from __future__ import annotations

import ctypes
from ctypes import c_void_p

import numpy as np
import torch

from . import _capi

__all__ = ["rotation_matrix", "plane_source3_rand", "point_source3", "point_source3_rand", "cone_source3_rand",
           "area_source3_cone", "area_source3_rand_bias", "rand_cone_in_sphere", "rand_area_in_sphere",
           "rand_rays_in_sphere", "rand_ptrays_in_sphere", "rand_ptcone_in_sphere", "rand_rays_cube",
           "random_rotmat", "random_rotate_ic", "rotate_ray3", "sum_norm"]


def rotation_matrix(angle, vert=False) -> np.ndarray:
    """The matrix of ``rotate_ray3`` (reference core/source.py:303-312): `angle` in degrees about z,
    or about x when `vert`."""
    theta = np.radians(float(angle))
    c, s = np.cos(theta), np.sin(theta)
    if vert:
        return np.array(((1, 0, 0), (0, c, -s), (0, s, c)), dtype=np.float64)
    return np.array(((c, -s, 0), (s, c, 0), (0, 0, 1)), dtype=np.float64)


def rotate_ray3(x, angle, vert=False):
    """reference core/source.py:303-312"""
    R = torch.from_numpy(rotation_matrix(angle, vert)).to(device=x.device, dtype=x.dtype)
    return torch.matmul(x, R.T)


def _fresh_seed() -> int:
    return int(torch.randint(0, 2 ** 62, (1,), dtype=torch.int64).item())


PLANE, POINT, CONE, AREA_CONE, AREA_BIAS, POINT_LATTICE = 0, 1, 2, 3, 4, 5   # DRRT_SOURCE_* of include/drrt_b200.h


def _generate(view_mats, pixels, spp, width, circle, sensor_dist, independent, offset, seed, device, post=None,
              post_span=0.0, kind=PLANE, cone_angle=90.0):
    nviews = len(view_mats)
    P0, P1 = int(pixels[0]), int(pixels[1])
    per_view = int(spp) * P0 * P1
    n = nviews * per_view
    dev = torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError("ray generation runs on the GPU only (there is no CPU path)")
    x = torch.empty((n, 3), dtype=torch.float32, device=dev)
    v = torch.empty((n, 3), dtype=torch.float32, device=dev)
    planes = torch.empty((n, 3, 3), dtype=torch.float32, device=dev)
    keep = torch.empty(n, dtype=torch.uint8, device=dev) if circle else None
    area = kind in (AREA_CONE, AREA_BIAS)
    tpv = torch.empty(n, dtype=torch.float32, device=dev) if area else None
    target = torch.empty((n, 3), dtype=torch.float32, device=dev) if kind == AREA_BIAS else None
    planes_per_ray = 4 if area else 2
    if offset is not None:
        offset = offset.to(device=dev, dtype=torch.float32).contiguous()
        if offset.numel() != planes_per_ray * n:
            raise RuntimeError(f"offset: expected {planes_per_ray * n} uniforms ([views][{planes_per_ray}][spp*P0*P1]), "
                               f"got {offset.numel()}")
    elif seed is None:
        seed = _fresh_seed()
    mats = np.ascontiguousarray(np.stack(view_mats).astype(np.float32)) if nviews else np.zeros((0, 3, 3), np.float32)
    post_arr = None if post is None else np.ascontiguousarray(np.asarray(post, dtype=np.float32).reshape(9))
    with torch.cuda.device(dev):
        _capi.check(_capi.load().drrt_source_rand(
            int(kind), c_void_p(offset.data_ptr() if offset is not None else 0), ctypes.c_uint64(int(seed or 0)),
            mats.ctypes.data_as(c_void_p), nviews, P0, P1, int(spp), float(width), float(sensor_dist),
            float(cone_angle), 1 if independent else 0,
            post_arr.ctypes.data_as(c_void_p) if post_arr is not None else c_void_p(0), float(post_span),
            c_void_p(x.data_ptr()), c_void_p(v.data_ptr()), c_void_p(planes.data_ptr()),
            c_void_p(keep.data_ptr() if keep is not None else 0),
            c_void_p(tpv.data_ptr() if tpv is not None else 0),
            c_void_p(target.data_ptr() if target is not None else 0),
            c_void_p(torch.cuda.current_stream(dev).cuda_stream)))
    if circle:
        # one host sync (the per-view survivor counts = the reference's `nrays`), then a sync-free
        # compaction: nonzero_static knows its output size, index_select gathers the three arrays
        counts = [int(c) for c in keep.view(nviews, per_view).sum(dim=1, dtype=torch.int64).tolist()]
        idx = torch.nonzero_static(keep, size=sum(counts)).squeeze(1)
        out = (x.index_select(0, idx), v.index_select(0, idx), planes.index_select(0, idx), counts)
        if area:
            out += (tpv.index_select(0, idx), target.index_select(0, idx) if target is not None else None)
        return out
    if area:
        return x, v, planes, [per_view] * nviews, tpv, target
    return x, v, planes, [per_view] * nviews


def plane_source3_rand(angle, pixels, spp, width, circle=False, xaxis=False, sensor_dist=1.0, independent=False,
                       *, offset=None, seed=None, device="cuda"):
    """reference core/source.py:54-69 -> (x, v, planes)."""
    x, v, planes, _ = _generate([rotation_matrix(angle, xaxis)], pixels, spp, width, circle, sensor_dist,
                                independent, offset, seed, device)
    return x, v, planes


def point_source3_rand(angle, pixels, spp, width, circle=False, xaxis=False, sensor_dist=1.0, *, offset=None,
                       seed=None, device="cuda"):
    """reference core/source.py:72-104 -> (x, v, planes): rays from one point through a jittered pixel grid."""
    x, v, planes, _ = _generate([rotation_matrix(angle, xaxis)], pixels, spp, width, circle, sensor_dist, False,
                                offset, seed, device, kind=POINT)
    return x, v, planes


def _lattice(pixels, spp):
    """point_source3's direction lattice (reference core/source.py:30-33): per axis pixels * floor(sqrt(spp))."""
    k = max(int(np.floor(np.sqrt(spp))), 1)
    return (int(pixels[0]) * k, int(pixels[1]) * k)


def point_source3(angle, pixels, spp, width, cone_angle=90, xaxis=False, sensor_dist=0.0, circle=False, *,
                  device="cuda"):
    """reference core/source.py:29-51 -> (x, v, planes): a point source on the volume face with a regular lattice of
    (pixels[0] k) x (pixels[1] k) directions, k = floor(sqrt(spp)); `circle` is accepted and ignored, as in the
    reference."""
    x, v, planes, _ = _generate([rotation_matrix(angle, xaxis)], _lattice(pixels, spp), 1, width, False, sensor_dist,
                                False, None, 0, device, kind=POINT_LATTICE, cone_angle=cone_angle)
    return x, v, planes


def cone_source3_rand(angle, pixels, spp, width, circle=False, xaxis=False, sensor_dist=1.0, cone_angle=100.0, *,
                      offset=None, seed=None, device="cuda"):
    """reference core/source.py:186-202 (+ hatbox_sample :531-545) -> (x, v, planes): pixels[0]*pixels[1]*spp
    directions uniform in a cone; `circle` is accepted and ignored, as in the reference."""
    x, v, planes, _ = _generate([rotation_matrix(angle, xaxis)], pixels, spp, width, False, sensor_dist, False,
                                offset, seed, device, kind=CONE, cone_angle=cone_angle)
    return x, v, planes


def rand_ptrays_in_sphere(nviews, im_res, spp, width, angle_span=360, circle=False, xaxis=False, sensor_dist=0.0, *,
                          offset=None, seed=None, device="cuda"):
    """reference core/source.py:360-365 -> ((x, v, planes), rays_per_view)."""
    angles = torch.linspace(0, angle_span, nviews + 1)
    mats = [rotation_matrix(angles[i], xaxis) for i in range(nviews)]
    x, v, planes, nrays = _generate(mats, im_res, spp, width, circle, sensor_dist, False, offset, seed, device,
                                    kind=POINT)
    return (x, v, planes), nrays


def rand_ptcone_in_sphere(nviews, im_res, spp, width, angle_span=360, circle=False, xaxis=False, sensor_dist=1.0,
                          cone_angle=90.0, *, offset=None, seed=None, device="cuda"):
    """reference core/source.py:388-396 -> ((x, v, planes), dists, rays_per_view); dists = zeros(nviews)."""
    angles = torch.linspace(0, angle_span, nviews + 1)
    mats = [rotation_matrix(angles[i], xaxis) for i in range(nviews)]
    x, v, planes, nrays = _generate(mats, im_res, spp, width, False, sensor_dist, False, offset, seed, device,
                                    kind=CONE, cone_angle=cone_angle)
    return (x, v, planes), torch.zeros(nviews), nrays


def area_source3_cone(angle, pixels, spp, width, circle=False, xaxis=False, sensor_dist=1.0, cone_angle=90, *,
                      offset=None, seed=None, device="cuda"):
    """reference core/source.py:153-183 -> ((x, v, planes), tpv): a cone of directions from every jittered
    pixel of the launch plane."""
    x, v, planes, _, tpv, _t = _generate([rotation_matrix(angle, xaxis)], pixels, spp, width, circle, sensor_dist,
                                         False, offset, seed, device, kind=AREA_CONE, cone_angle=cone_angle)
    return (x, v, planes), tpv


def area_source3_rand_bias(angle, pixels, spp, width, circle=False, xaxis=False, sensor_dist=1.0, *, offset=None,
                           seed=None, device="cuda"):
    """reference core/source.py:107-150 -> ((x, v, planes), xt, tpv): each jittered pixel aims at a random
    point of the sensor."""
    x, v, planes, _, tpv, xt = _generate([rotation_matrix(angle, xaxis)], pixels, spp, width, circle, sensor_dist,
                                         False, offset, seed, device, kind=AREA_BIAS)
    return (x, v, planes), xt, tpv


def rand_cone_in_sphere(nviews, im_res, spp, width, angle_span=360, circle=False, xaxis=False, sensor_dist=1.0,
                        cone_angle=90.0, *, offset=None, seed=None, device="cuda"):
    """reference core/source.py:379-386 -> ((x, v, planes), tpv, rays_per_view)."""
    angles = torch.linspace(0, angle_span, nviews + 1)
    mats = [rotation_matrix(angles[i], xaxis) for i in range(nviews)]
    x, v, planes, nrays, tpv, _t = _generate(mats, im_res, spp, width, circle, sensor_dist, False, offset, seed,
                                             device, kind=AREA_CONE, cone_angle=cone_angle)
    return (x, v, planes), tpv, nrays


def rand_area_in_sphere(nviews, im_res, spp, width, angle_span=360, circle=False, xaxis=False, sensor_dist=1.0, *,
                        offset=None, seed=None, device="cuda"):
    """reference core/source.py:368-376 -> ((x, v, planes), targets, tpv, rays_per_view)."""
    angles = torch.linspace(0, angle_span, nviews + 1)
    mats = [rotation_matrix(angles[i], xaxis) for i in range(nviews)]
    x, v, planes, nrays, tpv, xt = _generate(mats, im_res, spp, width, circle, sensor_dist, False, offset, seed,
                                             device, kind=AREA_BIAS)
    return (x, v, planes), xt, tpv, nrays


def rand_rays_in_sphere(nviews, im_res, spp, width, angle_span=360, circle=False, xaxis=False, sensor_dist=1.0,
                        indep=False, *, offset=None, seed=None, device="cuda"):
    """reference core/source.py:352-357 -> ((x, v, planes), rays_per_view)."""
    angles = torch.linspace(0, angle_span, nviews + 1)
    mats = [rotation_matrix(angles[i], xaxis) for i in range(nviews)]
    x, v, planes, nrays = _generate(mats, im_res, spp, width, circle, sensor_dist, indep, offset, seed, device)
    return (x, v, planes), nrays


def _cube_views():
    angles = torch.linspace(0, 360, 5)
    mats = [rotation_matrix(angles[i], False) for i in range(4)]
    mats += [rotation_matrix(va, True) for va in (90, -90)]
    return mats


def rand_rays_cube(im_res, spp, width, circle=False, src_type="plane", cone_ang=90, *, offset=None, seed=None,
                   device="cuda", rotmat=None):
    """reference core/source.py:398-412: 'plane' -> plane_source3_rand, 'point' -> point_source3(cone_angle=cone_ang)
    (the regular direction lattice, :29-51), anything else -> cone_source3_rand(cone_angle=cone_ang), as in the
    reference.  `rotmat` fuses the ``random_rotate_ic`` that always follows in the reference's drivers
    (core/luneburg_opt.py:54-56) into the same kernel."""
    kind = PLANE if src_type == "plane" else POINT_LATTICE if src_type == "point" else CONE
    if kind == POINT_LATTICE:
        im_res, spp, offset, seed = _lattice(im_res, spp), 1, None, 0
    x, v, planes, nrays = _generate(_cube_views(), im_res, spp, width, circle and kind == PLANE, 0.0, False, offset,
                                    seed, device, post=rotmat, post_span=width, kind=kind, cone_angle=cone_ang)
    return (x, v, planes), nrays


def random_rotmat():
    """reference core/source.py:548-552"""
    from scipy.spatial.transform import Rotation as R
    return torch.from_numpy(R.random().as_matrix())


def random_rotate_ic(x, v, planes, span, rotmat=None):
    """reference core/source.py:555-563 (torch ops; pass `rotmat` to `rand_rays_cube` to fuse it)."""
    rotmat = (random_rotmat() if rotmat is None else torch.as_tensor(rotmat)).to(device=x.device, dtype=x.dtype)
    xn = torch.matmul(x - (span / 2), rotmat.T) + (span / 2)
    vn = torch.matmul(v, rotmat.T)
    sp = torch.matmul(planes[:, 0, :] - (span / 2), rotmat.T) + (span / 2)
    sn = torch.matmul(planes[:, 1, :], rotmat.T)
    st = torch.matmul(planes[:, 2, :], rotmat.T)
    return xn, vn, torch.stack([sp, sn, st], dim=1)


def sum_norm(im, scale=False):
    """reference core/source.py:415-420"""
    scalar = torch.numel(im) / im.sum()
    if scale:
        return scalar * im, scalar
    return scalar * im
