#!/usr/bin/env python
"""Generates tests/golden/source.npz by running the REFERENCE's own core/source.py ray
generators on the CPU in float64, capturing the uniforms they drew (by re-seeding torch's
generator and drawing the same tensors again).

    python tests/golden/make_golden_source.py        # needs /root/reference (build container only)
"""
import os
import sys

import numpy as np
import torch

REF = os.environ.get("DRRT_REFERENCE", "/root/reference")
sys.path.insert(0, os.path.join(REF, "core"))
import source as ref_source  # noqa: E402  (the reference's module)

HERE = os.path.dirname(os.path.abspath(__file__))


def uniforms(seed, nviews, spp, pixels):
    torch.manual_seed(seed)
    return np.stack([torch.rand(2 * spp, *pixels).numpy() for _ in range(nviews)])


def main():
    torch.set_default_dtype(torch.float64)
    out = {}
    pixels, spp, width = (6, 5), 3, 2.0
    # single view, both jitter modes
    for name, indep, circle, xaxis, ang in (("single", False, False, False, 35.0),
                                            ("indep", True, False, True, -50.0),
                                            ("circle", False, True, False, 120.0)):
        torch.manual_seed(7)
        x, v, pl = ref_source.plane_source3_rand(torch.tensor(ang), pixels, spp, width, circle=circle, xaxis=xaxis,
                                                 sensor_dist=0.25, independent=indep)
        out[name + "_x"] = x.numpy(); out[name + "_v"] = v.numpy(); out[name + "_planes"] = pl.numpy()
        out[name + "_offset"] = uniforms(7, 1, spp, pixels)
        out[name + "_angle"] = ang
    # rand_rays_in_sphere, 5 views over 180 degrees
    torch.manual_seed(11)
    (x, v, pl), nr = ref_source.rand_rays_in_sphere(5, pixels, spp, width, angle_span=180, circle=False, xaxis=False,
                                                    sensor_dist=0.0)
    out["sphere_x"] = x.numpy(); out["sphere_v"] = v.numpy(); out["sphere_planes"] = pl.numpy()
    out["sphere_nrays"] = np.array(nr); out["sphere_offset"] = uniforms(11, 5, spp, pixels)
    # rand_rays_cube with the circle filter + random_rotate_ic (the Luneburg driver's batch)
    torch.manual_seed(13)
    (x, v, pl), nr = ref_source.rand_rays_cube(pixels, spp, width, circle=True)
    np.random.seed(3)
    R = ref_source.random_rotmat().numpy()
    np.random.seed(3)
    xr, vr, plr = ref_source.random_rotate_ic(x, v, pl, width)
    out["cube_x"] = xr.numpy(); out["cube_v"] = vr.numpy(); out["cube_planes"] = plr.numpy()
    out["cube_nrays"] = np.array(nr); out["cube_offset"] = uniforms(13, 6, spp, pixels); out["cube_rotmat"] = R
    # point source through a jittered pixel grid, 3 views, circle filter (rand_ptrays_in_sphere)
    torch.manual_seed(17)
    (x, v, pl), nr = ref_source.rand_ptrays_in_sphere(3, pixels, spp, width, angle_span=270, circle=True,
                                                      sensor_dist=0.5)
    out["pt_x"] = x.numpy(); out["pt_v"] = v.numpy(); out["pt_planes"] = pl.numpy()
    out["pt_nrays"] = np.array(nr); out["pt_offset"] = uniforms(17, 3, spp, pixels)
    # cone source, 2 views (rand_ptcone_in_sphere): per view torch.rand(n) for z, then torch.rand(n) for theta
    torch.manual_seed(19)
    (x, v, pl), dists, nr = ref_source.rand_ptcone_in_sphere(2, pixels, spp, width, angle_span=180, sensor_dist=0.3,
                                                             cone_angle=70.0)
    out["cone_x"] = x.numpy(); out["cone_v"] = v.numpy(); out["cone_planes"] = pl.numpy()
    n = pixels[0] * pixels[1] * spp
    torch.manual_seed(19)
    out["cone_offset"] = np.stack([np.stack([torch.rand(n).numpy(), torch.rand(n).numpy()]) for _ in range(2)])
    # area sources (circle=False): per view rand(2*spp, P0, P1), then two more planes of n uniforms
    def uniforms4(seed, nviews, second):
        torch.manual_seed(seed)
        res = []
        for _ in range(nviews):
            a = torch.rand(2 * spp, *pixels).numpy().reshape(2, -1)
            b = second()
            res.append(np.concatenate([a, b]))
        return np.stack(res)

    torch.manual_seed(23)
    (x, v, pl), tpv, nr = ref_source.rand_cone_in_sphere(2, pixels, spp, width, angle_span=180, sensor_dist=0.4,
                                                         cone_angle=50.0)
    out["acone_x"] = x.numpy(); out["acone_v"] = v.numpy(); out["acone_planes"] = pl.numpy(); out["acone_tpv"] = tpv.numpy()
    out["acone_offset"] = uniforms4(23, 2, lambda: np.stack([torch.rand(n).numpy(), torch.rand(n).numpy()]))
    torch.manual_seed(29)
    (x, v, pl), xt, tpv, nr = ref_source.rand_area_in_sphere(2, pixels, spp, width, angle_span=180, sensor_dist=0.7)
    out["abias_x"] = x.numpy(); out["abias_v"] = v.numpy(); out["abias_planes"] = pl.numpy()
    out["abias_xt"] = xt.numpy(); out["abias_tpv"] = tpv.numpy()
    out["abias_offset"] = uniforms4(29, 2, lambda: torch.rand(2, n).numpy())
    out.update(pixels=np.array(pixels), spp=spp, width=width)
    np.savez_compressed(os.path.join(HERE, "source.npz"), **out)
    print("wrote source.npz:", {k: np.shape(a) for k, a in out.items()})


if __name__ == "__main__":
    main()
