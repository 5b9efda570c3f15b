"""Generates tests/golden/*_ad.npz by running the REFERENCE's own CPU tape-autodiff tracer
(Tracer<true,false> on enoki's DiffArray, compiled from /root/reference by
oracle/build_ref.sh -> oracle/_ref/libdrrt_refad_v3.so; reference src/tracer.cpp:569,
ext/enoki/src/autodiff/autodiff.cpp:1229-1230) exactly as the reference's ADTracerC does
(core/tracer.py:17-66): trace, set_gradient(xt), set_gradient(vt), backward(), then
gradient(rif), gradient(pos), gradient(vel).  Run in the build container:

    python tests/golden/make_golden_ad.py

Cases (VERDICT r1 "pin f4"):
  bump9_ad / lune17_ad      9^3 bump / 17^3 Luneburg, h = 1/(N-1), rays launched OUTSIDE the box
  bump9_face_ad             rays launched just inside the first cell layer (no straight prefix)
  bump9_h1_ad / lune17_h1_ad  the same fields with h = 1: the reference's splat omits the 1/h of
                            eval_grad (src/volume.cpp:227-243 vs :178), so its adjoint d_rif equals
                            the tape's only for h = 1 -- this is the paper's adjoint-vs-tape check
                            on reference code, and it pins d_rif of the adjoint too
  bump9_sdf_ad              ADSDFTracerC (trace_sdf), rays starting inside the sphere
Each file also stores the adjoint gradient of the reference's NON-AD tracer (Tracer<false,false>
::backtrace on the tape run's endpoints) so a reader can see the tape/adjoint agreement directly.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
from adjointnonlinearraytracing_b200 import scenes as sc  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def rel(a, b):
    return float(np.linalg.norm(np.asarray(a, np.float64) - b) / max(np.linalg.norm(b), 1e-30))


def case(name, ng, field, s, backoff_steps, unit_h=False, face=False, sdf=False, seed=3):
    ad, ref = oracle.refad, oracle.ref
    rif = field(ng)
    span = float(ng - 1) if unit_h else 1.0
    h = np.float32(1.0) if unit_h else sc.grid_spacing(ng)
    ds = np.float32(h / 2)
    if sdf:
        rng = np.random.default_rng(5)
        n = s * s
        pos = (0.5 + 0.15 * rng.uniform(-1, 1, (n, 3))).astype(np.float32)
        vel = rng.standard_normal((n, 3)).astype(np.float32)
        vel /= np.linalg.norm(vel, axis=1, keepdims=True)
        sd = sc.sdf_sphere(ng, 0.4)
    else:
        pos, vel = sc.rays_cone(s, spread=0.12, seed=seed)
        pos = (pos * np.float32(span)).astype(np.float32)
        if face:
            pos[:, 2] = np.float32(0.3) * ds
        else:
            pos = (pos - np.float32(backoff_steps) * ds * vel).astype(np.float32)
        sd = None
    rng = np.random.default_rng(seed + 10)
    wx = rng.standard_normal(pos.shape).astype(np.float32)
    wv = rng.standard_normal(pos.shape).astype(np.float32)
    xt, vt, d_rif, d_pos, d_vel = ad.trace_backward(rif, pos, vel, h, ds, wx, wv, sdf=sd)
    # the NON-AD reference on the same inputs: forward must agree, adjoint is stored beside the tape
    if sdf:
        xr, vr = ref.trace_sdf(rif, sd, pos, vel, h, ds)
        g_adj = ref.backtrace_sdf(rif, sd, xt, vt, wx, wv, h, ds)
    else:
        xr, vr = ref.trace(rif, pos, vel, h, ds)
        g_adj = ref.backtrace(rif, xt, vt, wx, wv, h, ds)
    print(f"{name:16s} N={pos.shape[0]:4d} fwd AD vs non-AD max {np.abs(xt - xr).max():.1e}/{np.abs(vt - vr).max():.1e}  "
          f"d_rif adjoint vs tape rel L2 {rel(g_adj, d_rif):.2e}  |d_pos| {np.abs(d_pos).max():.2e} |d_vel| {np.abs(d_vel).max():.2e}")
    d = dict(rif=rif, h=h, ds=ds, pos=pos, vel=vel, wx=wx, wv=wv, xt=xt, vt=vt, d_rif=d_rif, d_pos=d_pos,
             d_vel=d_vel, grad_adjoint=g_adj, unit_h=np.uint8(unit_h))
    if sdf:
        d["sdf"] = sd
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **d)


def main():
    assert oracle.refad is not None and oracle.ref is not None, "run oracle/build_ref.sh first (needs /root/reference)"
    case("bump9_ad", 9, sc.field_bump, 12, 3.37)
    case("lune17_ad", 17, sc.field_luneburg, 16, 2.6)
    case("bump9_face_ad", 9, sc.field_bump, 12, 0, face=True)
    case("bump9_h1_ad", 9, sc.field_bump, 12, 3.37, unit_h=True)
    case("lune17_h1_ad", 17, sc.field_luneburg, 16, 2.6, unit_h=True)
    case("bump9_sdf_ad", 9, sc.field_bump, 12, 0, sdf=True)


if __name__ == "__main__":
    main()
