#!/usr/bin/env python
"""Generates tests/golden/source_point.npz by running the REFERENCE's own core/source.py:
point_source3 (the regular direction lattice, :29-51) for one view and rand_rays_cube(src_type='point')
(:398-412), float64 on the CPU.

    python tests/golden/make_golden_source_point.py        # needs /root/reference (build container only)
"""
import os
import sys

import numpy as np
import torch

REF = os.environ.get("DRRT_REFERENCE", "/root/reference")
sys.path.insert(0, os.path.join(REF, "core"))
import source as ref_source  # noqa: E402  (the reference's module)

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    torch.set_default_dtype(torch.float64)
    out = {}
    pixels, spp, width = (6, 5), 5, 2.0   # floor(sqrt(5)) = 2 -> a 12 x 10 lattice
    x, v, pl = ref_source.point_source3(torch.tensor(35.0), pixels, spp, width, cone_angle=70.0, xaxis=True, sensor_dist=0.25)
    out.update(one_x=x.numpy(), one_v=v.numpy(), one_planes=pl.numpy(), one_angle=35.0, one_cone=70.0, one_sd=0.25)
    (x, v, pl), nr = ref_source.rand_rays_cube(pixels, spp, width, circle=True, src_type="point", cone_ang=50)
    out.update(cube_x=x.numpy(), cube_v=v.numpy(), cube_planes=pl.numpy(), cube_nrays=np.array(nr), cube_cone=50.0)
    out.update(pixels=np.array(pixels), spp=spp, width=width)
    np.savez_compressed(os.path.join(HERE, "source_point.npz"), **out)
    print("wrote source_point.npz:", {k: np.shape(a) for k, a in out.items()})


if __name__ == "__main__":
    main()
