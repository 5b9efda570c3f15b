"""Dev script (GPU box): parity of every kernel against the oracle + first timings."""
import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import oracle
from adjointnonlinearraytracing_b200 import ops, scenes as sc

dev = torch.device("cuda:0")
T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)

def stats_err(name, a, b):
    e = np.abs(a - b)
    if e.ndim > 1: e = e.max(axis=1)
    print(f"  {name}: median {np.median(e):.2e} p99 {np.percentile(e,99):.2e} p99.9 {np.percentile(e,99.9):.2e} max {e.max():.2e} n>1e-5 {(e>1e-5).sum()}/{e.size}")

for Ng, S in [(9, 16), (32, 64), (64, 128)]:
    rif = sc.field_bump(Ng); h = sc.grid_spacing(Ng); ds = np.float32(h / 2)
    pos, vel = sc.rays_cone(S, backoff=0.37 * float(ds))
    xr, vr = oracle.port.trace(rif, pos, vel, h, ds)
    xR, vR = oracle.ref.trace(rif, pos, vel, h, ds) if oracle.ref else (xr, vr)
    print(f"grid {Ng} rays {S*S}: port-vs-ref max {np.abs(xr-xR).max():.2e}")
    for nm, fl in [("ref-arith", ops.make_flags(reference_arith=True)), ("fast", 0)]:
        st = ops.new_stats(dev)
        xg, vg = ops.trace(T(rif), T(pos), T(vel), h, ds, flags=fl, stats=st)
        print(f" fwd {nm}: steps gpu {st[0].item()} port {oracle.port.last_steps if False else ''} not_exited {st[1].item()}")
        stats_err("xt vs port", xg.cpu().numpy(), xr); stats_err("vt vs port", vg.cpu().numpy(), vr)
    dx, dv = sc.seeds_normal(S * S)
    gp = oracle.port.backtrace(rif, xr, vr, dx, dv, h, ds); steps_p = oracle.port.last_steps
    g64 = oracle.port64.backtrace(rif, xr, vr, dx, dv, h, ds)
    scale = np.abs(gp).max(); nrm = np.linalg.norm(gp)
    print(f" adj: port-vs-64 relmax {np.abs(gp-g64).max()/scale:.2e} relL2 {np.linalg.norm(gp-g64)/nrm:.2e}")
    for nm, fl in [("ref-arith", ops.make_flags(reference_arith=True)), ("ref-arith+det", ops.make_flags(True, True)),
                   ("fast", 0), ("fast+det", ops.make_flags(False, True))]:
        st = ops.new_stats(dev)
        gg = ops.backtrace(T(rif), T(xr), T(vr), T(dx), T(dv), h, ds, flags=fl, stats=st).cpu().numpy()
        print(f"  adj {nm}: steps {st[0].item()} (port {steps_p}) relmax {np.abs(gg-gp).max()/scale:.2e} relL2 {np.linalg.norm(gg-gp)/nrm:.2e}  vs64 relL2 {np.linalg.norm(gg-g64)/nrm:.2e}")
    if nm:
        g1 = ops.backtrace(T(rif), T(xr), T(vr), T(dx), T(dv), h, ds, flags=ops.make_flags(False, True))
        g2 = ops.backtrace(T(rif), T(xr), T(vr), T(dx), T(dv), h, ds, flags=ops.make_flags(False, True))
        print("  deterministic bit-stable:", bool(torch.equal(g1, g2)))

# timing
def timeit(fn, iters=3):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return min(ts)

for Ng, S in [(256, 1024), (256, 2048)]:
    rif = T(sc.field_bump(Ng)); h = sc.grid_spacing(Ng); ds = np.float32(h / 2)
    for rays_name, (pos, vel) in [("collimated", sc.rays_collimated(S)), ("cone", sc.rays_cone(S))]:
        pos, vel = T(pos), T(vel)
        dx = torch.ones_like(pos); dv = torch.ones_like(pos)
        for nm, fl in [("fast", 0), ("ref-arith", 1)]:
            st = ops.new_stats(dev)
            xt, vt = ops.trace(rif, pos, vel, h, ds, flags=fl, stats=st); sf = st[0].item()
            tf = timeit(lambda: ops.trace(rif, pos, vel, h, ds, flags=fl))
            st = ops.new_stats(dev)
            g = ops.backtrace(rif, xt, vt, dx, dv, h, ds, flags=fl, stats=st); sb = st[0].item()
            tb = timeit(lambda: ops.backtrace(rif, xt, vt, dx, dv, h, ds, flags=fl))
            print(f"{Ng}^3 {S*S} rays {rays_name} {nm}: fwd {sf/tf/1e6:.1f} Grs/s ({tf:.2f} ms, {sf} steps) adj {sb/tb/1e6:.1f} Grs/s ({tb:.2f} ms, {sb} steps) combined {(sf+sb)/(tf+tb)/1e6:.1f} Grs/s  alg GB/s {(32*sf+64*sb)/(tf+tb)/1e6:.0f}")
        if S == 1024 and rays_name == "cone":
            tb = timeit(lambda: ops.backtrace(rif, xt, vt, dx, dv, h, ds, flags=2))
            print(f"   fast+det adj {sb/tb/1e6:.1f} Grs/s ({tb:.2f} ms)")
