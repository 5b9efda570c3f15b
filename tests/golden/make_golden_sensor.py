#!/usr/bin/env python
"""Generates tests/golden/sensor.npz by running the REFERENCE's own sensor epilogue
(core/sensor.py generate_sensor / generate_inf_sensor over core/grid.py Grid.Splat) on
seeded ray states, in float64 on the CPU, with torch autograd for the backward.

    python tests/golden/make_golden_sensor.py        # needs /root/reference (build container only)
"""
import os
import sys

import numpy as np
import torch

REF = os.environ.get("DRRT_REFERENCE", "/root/reference")
sys.path.insert(0, os.path.join(REF, "core"))
import sensor as ref_sensor  # noqa: E402  (the reference's module)

HERE = os.path.dirname(os.path.abspath(__file__))


def rays(n, seed, span):
    rng = np.random.default_rng(seed)
    # exit states of a +z-ish beam leaving a unit volume, some of them outside the sensor
    x = np.stack([rng.uniform(-0.1, 1.1, n) * span, rng.uniform(-0.1, 1.1, n) * span,
                  np.full(n, span)], axis=1)
    v = np.array([0.0, 0.0, 1.0]) + 0.25 * rng.standard_normal((n, 3))
    v *= rng.uniform(0.8, 1.5, (n, 1))   # |v| = n(x) != 1 in general
    e = rng.uniform(0.5, 2.0, n)
    return x, v, e


def main():
    torch.set_default_dtype(torch.float64)
    out = {}
    span, res, n = 1.0, 16, 400
    x, v, e = rays(n, 0, span)
    p = np.array([[0.5, 0.5, 1.3]]); nn = np.array([[0.0, 0.0, 1.0]]); tg = np.array([[0.0, 1.0, 0.0]])
    rng = np.random.default_rng(5)
    dimg = rng.standard_normal((res, res))
    for name, use_e, tangent in (("near", True, tg), ("near_notan", False, None)):
        X = torch.tensor(x, requires_grad=True); V = torch.tensor(v, requires_grad=True)
        E = torch.tensor(e) if use_e else 1
        T = None if tangent is None else torch.tensor(tangent)
        img = ref_sensor.generate_sensor((X, V), E, (torch.tensor(p), torch.tensor(nn)), res, span, T)
        (img * torch.tensor(dimg)).sum().backward()
        out[name + "_image"] = img.detach().numpy()
        out[name + "_dx"] = X.grad.numpy(); out[name + "_dv"] = V.grad.numpy()
    # tilted plane
    nt = np.array([[0.2, -0.1, 1.0]]); nt /= np.linalg.norm(nt)
    tt = np.cross(nt, np.array([[1.0, 0.0, 0.0]])); tt /= np.linalg.norm(tt)
    X = torch.tensor(x, requires_grad=True); V = torch.tensor(v, requires_grad=True)
    img = ref_sensor.generate_sensor((X, V), torch.tensor(e), (torch.tensor(p), torch.tensor(nt)), res, span,
                                     torch.tensor(tt))
    (img * torch.tensor(dimg)).sum().backward()
    out["tilt_image"] = img.detach().numpy(); out["tilt_dx"] = X.grad.numpy(); out["tilt_dv"] = V.grad.numpy()
    out["tilt_n"] = nt; out["tilt_t"] = tt
    # far field
    X = torch.tensor(x); V = torch.tensor(v, requires_grad=True)
    img = ref_sensor.generate_inf_sensor((X, V), 1, (torch.tensor(p), torch.tensor(nn)), res, angle_span=60,
                                         tangent=torch.tensor(tg))
    (img * torch.tensor(dimg)).sum().backward()
    out["far_image"] = img.detach().numpy(); out["far_dv"] = V.grad.numpy()
    # texture look-ups (get_sdf_vals_near / _far) and their autograd backward
    tex = rng.standard_normal((res, res))
    df = rng.standard_normal(n)
    X = torch.tensor(x, requires_grad=True); V = torch.tensor(v, requires_grad=True)
    f = ref_sensor.get_sdf_vals_near((X, V), torch.tensor(tex), (torch.tensor(p), torch.tensor(nn)), span,
                                     torch.tensor(tg))
    (f * torch.tensor(df)).sum().backward()
    out["gnear_f"] = f.detach().numpy(); out["gnear_dx"] = X.grad.numpy(); out["gnear_dv"] = V.grad.numpy()
    vs = v * 0.4   # keep the raw direction coordinates on the far-field texture
    X = torch.tensor(x); V = torch.tensor(vs, requires_grad=True)
    f = ref_sensor.get_sdf_vals_far((X, V), torch.tensor(tex), (torch.tensor(p), torch.tensor(nn)), 60,
                                    torch.tensor(tg))
    (f * torch.tensor(df)).sum().backward()
    out["gfar_f"] = f.detach().numpy(); out["gfar_dv"] = V.grad.numpy(); out["gfar_v"] = vs
    out.update(tex=tex, df=df)
    out.update(x=x, v=v, e=e, p=p, n=nn, tangent=tg, dimage=dimg, res=res, span=span, far_angle_span=60)
    np.savez_compressed(os.path.join(HERE, "sensor.npz"), **out)
    print("wrote sensor.npz:", {k: np.shape(a) for k, a in out.items()})


if __name__ == "__main__":
    main()
