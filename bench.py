#!/usr/bin/env python
"""bench.py -- fwd+adjoint ray-steps/s of the eikonal tracer on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--scaling strong|weak] [--workload caustic256|focal512] [--no-cpu] [--no-configs]

One "step" = one forward trace + one adjoint backtrace (+ gradient all-reduce when N > 1) over one
ray batch.  Workload (config.workload): BASELINE.json configs[3] "caustic design" AS WRITTEN: 256^3 IOR
grid (Gaussian-bump field A of SURVEY.md section 8d), 2048^2 = 4 194 304 collimated rays IN TOTAL, ds = h/2,
adjoint seeds dx = dv = 1 (reference src/test.cpp:142-144).  N > 1: the batch's sensor rows are dealt out cyclically
(dist.shard_cyclic), i.e. STRONG scaling; the weak figure (4 M rays per rank) is measured too and reported
under `secondary.weak`.  Every N > 1 run checks the all-reduced gradient against a single-GPU gradient of
the whole batch (`grad_check`) and reports the collective's own time.

At N = 1 the line also carries
  parity        the GPU against the reference's own CPU tracer (oracle/_ref) on the SAME strided subsample of
                the benchmarked batch that the cpu_baseline leg times -- endpoints and gradient -- plus the
                reference's own AVX2-vs-AVX-512 disagreement on it (the noise floor of this lattice-exact
                batch) and a backed-off batch (z0 = -0.37 ds) where no lattice excuse applies; the run FAILS
                (exit 3, after printing) when endpoints leave 1e-5 or the gradient leaves 1e-4,
  configs       fwd / adjoint times of all five BASELINE.json configurations + the reference's own
                step-size sweep (src/test.cpp:148-206),
  roofline      against the roof that actually binds the dominant kernel (instruction issue; the grids are
                L2-resident), with the section-8d algorithmic-bytes-vs-HBM figure beside it; counts taken
                from committed ncu captures carry the hash of the kernel sources they were captured on and
                `stale: true` is emitted when the built sources differ.
"""
from __future__ import annotations

import argparse
import glob
import hashlib
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

NG = 256
SENSOR = 2048
METRIC = "fwd+adjoint ray-steps/sec at 256^3 grid"
UNIT = "ray-steps/s"
BYTES_FWD, BYTES_ADJ = 32, 64  # algorithmic bytes per ray-step, SURVEY.md section 8d
END_TOL, GRAD_TOL = 1e-5, 1e-4  # north-star tolerances (BASELINE.json)


def workload_name():
    return (f"{'caustic-256 (BASELINE.json configs[3])' if NG == 256 else 'focal-512 (BASELINE.json configs[4])'}: "
            f"{NG}^3 bump IOR grid, {SENSOR}^2 = {SENSOR * SENSOR} collimated rays in total, ds = h/2, seeds dx=dv=1")


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def csrc_sha():
    """Hash of the kernel sources the library is built from; committed ncu counts carry the hash they were
    captured on (tools/ncu_counts.py)."""
    h = hashlib.sha256()
    d = os.path.join(ROOT, "adjointnonlinearraytracing_b200", "csrc")
    for p in sorted(glob.glob(os.path.join(d, "*.cu")) + glob.glob(os.path.join(d, "*.cuh")) + glob.glob(os.path.join(d, "*.h"))):
        h.update(os.path.basename(p).encode())
        with open(p, "rb") as f:
            h.update(f.read())
    return h.hexdigest()[:16]


def workload_host(n_g=None, sensor=None):
    from adjointnonlinearraytracing_b200 import scenes as sc
    n_g = NG if n_g is None else n_g
    sensor = SENSOR if sensor is None else sensor
    rif = sc.field_bump(n_g)
    h = float(sc.grid_spacing(n_g))
    ds = h / 2
    pos, vel = sc.rays_collimated(sensor)
    return rif, h, ds, pos, vel


# --------------------------------------------------------------------------- clocks
class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU during the timed region (NVML)."""

    def __init__(self, index):
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4),
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.01)

    def start(self):
        if self.nv:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()

    def stop(self):
        if self._thr:
            self._stop.set()
            self._thr.join()
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


# --------------------------------------------------------------------------- CPU reference arm
def _cpu_lib(h):
    """The reference's own CPU tracer (oracle/_ref): the build whose rcp(h) is the correctly rounded 1/h when there
    is one (oracle.ref_for_spacing: what parity is stated against), else any build; else the C port."""
    import oracle
    lib = oracle.ref_for_spacing(h) or oracle.ref
    if lib is not None:
        return lib, "reference", oracle
    return oracle.port, "port", oracle


def rel_l2(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


class CpuArm:
    """Times the reference CPU path on `threads` host threads: the sample (a strided subsample of the benchmarked
    batch) is split into one strided sub-batch per thread, each thread owns a private gradient grid (the same
    scheme as the multi-GPU path; the reference itself is single-threaded).  `step()` keeps
    what it computes -- endpoints and the summed gradient -- for the parity leg."""

    def __init__(self, rays_per_thread=None, threads=None, target_s=2.5, z0=0.0, host=None):
        self.rif, self.h, self.ds, pos, vel = host if host is not None else workload_host()
        self.host = (self.rif, self.h, self.ds, pos, vel)
        if z0:
            pos = pos.copy(); pos[:, 2] = np.float32(z0)
        self.lib, self.kind, self.oracle = _cpu_lib(np.float32(self.h))
        self.threads = threads or max(1, len(os.sched_getaffinity(0)))
        if rays_per_thread is None:  # pilot: size the sample so one step takes ~target_s per thread
            p = np.ascontiguousarray(pos[:: max(1, len(pos) // 1024)][:1024])
            v = np.ascontiguousarray(vel[:: max(1, len(vel) // 1024)][:1024])
            t0 = time.perf_counter()
            xt, vt = self.lib.trace(self.rif, p, v, np.float32(self.h), np.float32(self.ds))
            self.lib.backtrace(self.rif, xt, vt, np.ones_like(p), np.ones_like(p), np.float32(self.h), np.float32(self.ds))
            dt = max(time.perf_counter() - t0, 1e-3)
            rays_per_thread = int(min(65536, max(4096, 1024 * target_s / dt)))  # >= 4096 amortises the 64 MB grid copies
        self.rpt = rays_per_thread
        n = self.rpt * self.threads
        stride = max(1, len(pos) // n)
        self.idx = [np.arange(i, len(pos), stride * self.threads)[: self.rpt] for i in range(self.threads)]
        self.pos = [np.ascontiguousarray(pos[ix]) for ix in self.idx]
        self.vel = [np.ascontiguousarray(vel[ix]) for ix in self.idx]
        self.steps = None
        self.last = None

    def count_steps(self):
        port = self.oracle.port
        h, ds = np.float32(self.h), np.float32(self.ds)

        def one(i):  # each thread needs its own wrapper state (last_steps)
            lib = self.oracle._Lib(port.path, "orc_", np.float32, returns_steps=True)
            xt, vt = lib.trace(self.rif, self.pos[i], self.vel[i], h, ds); sf = lib.last_steps
            lib.backtrace(self.rif, xt, vt, np.ones_like(xt), np.ones_like(xt), h, ds); sb = lib.last_steps
            return sf, sb
        r = self.oracle.run_threads(one, self.threads)
        self.steps = (sum(a for a, _ in r), sum(b for _, b in r))
        return self.steps

    def step(self, lib=None, keep=True):
        lib = lib or self.lib
        h, ds = np.float32(self.h), np.float32(self.ds)

        def one(i):
            xt, vt = lib.trace(self.rif, self.pos[i], self.vel[i], h, ds)
            return xt, vt, lib.backtrace(self.rif, xt, vt, np.ones_like(xt), np.ones_like(xt), h, ds)
        t0 = time.perf_counter()
        out = self.oracle.run_threads(one, self.threads)
        dt = time.perf_counter() - t0
        if keep:
            g = np.zeros(out[0][2].shape, np.float64)
            for o in out:
                g += o[2]
            self.last = (np.concatenate([o[0] for o in out]), np.concatenate([o[1] for o in out]), g)
        return dt

    def describe(self):
        return (f"{self.threads} threads x {self.rpt} rays (strided subsample of the {SENSOR}^2 batch), "
                f"{NG}^3 grid, one trace + one backtrace per thread, private gradient grids; library "
                f"{os.path.basename(getattr(self.lib, 'path', '?'))}")


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    arm = CpuArm()
    sf, sb = arm.count_steps()
    for _ in range(args.warmup):
        arm.step(keep=False)
    times = [arm.step(keep=False) for _ in range(args.steps)]
    total = sum(times)
    value = (sf + sb) * args.steps / total
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(), "rays_total": SENSOR * SENSOR, "grid": [NG, NG, NG],
                   "note": "CPU reference arm: each step is a bounded strided sample of the batch (cpu_baseline.sample)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": arm.threads, "kind": arm.kind, "sample": arm.describe()},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _emit(line)


# --------------------------------------------------------------------------- parity leg (N = 1)
def parity_leg(arm, ops, torch, dev, rif_dev):
    """GPU vs the reference library on the sample the cpu_baseline leg just ran (`arm.last`)."""
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    h, ds = arm.h, arm.ds

    def gpu(pos, vel):
        one = torch.ones(len(pos), 3, device=dev)
        xt, vt = ops.trace(rif_dev, T(pos), T(vel), h, ds)
        g = ops.backtrace(rif_dev, xt, vt, one, one, h, ds)
        return xt.cpu().numpy(), vt.cpu().numpy(), g.cpu().numpy()

    def compare(a, b):
        e = np.abs(a[0] - b[0]).max(axis=1)
        return {"rays": int(len(e)), "endpoint_max_abs": float(e.max()), "direction_max_abs": float(np.abs(a[1] - b[1]).max()),
                "frac_rays_within_1e-5": float(np.mean(e <= END_TOL)),
                "bit_identical": bool(np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])),
                "grad_rel_l2": rel_l2(a[2], b[2])}

    pos = np.concatenate(arm.pos); vel = np.concatenate(arm.vel)
    out = compare(gpu(pos, vel), arm.last)
    lib = arm.lib
    out["vs"] = (f"oracle/_ref = the reference's Tracer<false,false> built from /root/reference "
                 f"({os.path.basename(getattr(lib, 'path', 'port'))}, {lib.isa()}; its rcp(h) "
                 f"{'==' if lib.rcp(np.float32(h)) == np.float32(1) / np.float32(h) else '!='} the correctly rounded 1/h the kernels use)")
    out["sample"] = "the cpu_baseline sample: strided subsample of the benchmarked batch (rays ON the z = 0 lattice plane, ds = h/2)"
    out["tol"] = {"endpoint_abs_unit_span": END_TOL, "grad_rel_l2": GRAD_TOL}
    ok = out["endpoint_max_abs"] <= END_TOL and out["grad_rel_l2"] <= GRAD_TOL
    # noise floor: the reference's OTHER ISA build on a quarter of the sample
    try:
        builds = arm.oracle.ref_builds()
        other = [l for l in builds.values() if l.path != getattr(lib, "path", None)]
        if other and arm.kind == "reference":
            sub = CpuArm(rays_per_thread=max(1, arm.rpt // 4), threads=arm.threads, host=arm.host)
            sub.step(lib=lib); a = sub.last
            sub.step(lib=other[0]); b = sub.last
            nf = compare(b, a)
            nf["what"] = (f"the reference against ITSELF: {os.path.basename(other[0].path)} ({other[0].isa()}) vs "
                          f"{os.path.basename(lib.path)} ({lib.isa()}) on a quarter of the sample -- enoki's rcp(h) differs by one ulp "
                          "between the builds and every other sample of this batch lies exactly on a cell face")
            out["noise_floor"] = nf
    except Exception as e:  # the noise floor is context, never a reason to lose the line
        out["noise_floor"] = {"error": repr(e)}
    # backed-off batch: z0 = -0.37 ds, no sample on a lattice plane
    try:
        sub = CpuArm(rays_per_thread=max(1, arm.rpt // 4), threads=arm.threads, z0=-0.37 * ds, host=arm.host)
        sub.step()
        bo = compare(gpu(np.concatenate(sub.pos), np.concatenate(sub.vel)), sub.last)
        bo["z0"] = "-0.37 ds"
        out["backed_off"] = bo
        ok = ok and bo["endpoint_max_abs"] <= END_TOL and bo["grad_rel_l2"] <= GRAD_TOL
    except Exception as e:
        out["backed_off"] = {"error": repr(e)}
        ok = False
    out["ok"] = bool(ok)
    return out


# --------------------------------------------------------------------------- the other configurations (N = 1)
def config_lines(ops, torch, dev):
    """fwd / adjoint CUDA-event times of BASELINE.json's five configurations (synthetic fields of the named shapes,
    adjacent to the headline run so they land in the driver-timed line) + the reference's own measurement
    protocol: 33^3 grid, 512^2 rays, 15 step sizes 0.3 .. 1.2 h (src/test.cpp:148-206)."""
    from adjointnonlinearraytracing_b200 import scenes as sc
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    ev = lambda: torch.cuda.Event(enable_timing=True)

    def timed(fn, iters):
        fn(); fn()
        torch.cuda.synchronize(dev)
        ts = []
        for _ in range(iters):
            a, b = ev(), ev()
            a.record(); fn(); b.record(); torch.cuda.synchronize(dev)
            ts.append(a.elapsed_time(b))
        return float(np.median(ts))

    def bump_dev(ng):  # field A on the device (512^3 takes seconds in numpy)
        c = torch.arange(ng, device=dev, dtype=torch.float64) / (ng - 1)
        r2 = ((c - 0.5) ** 2)
        r2 = r2[:, None, None] + r2[None, :, None] + r2[None, None, :]
        return (1.0 + 0.2 * torch.exp(-r2 / (2 * 0.2 * 0.2))).to(torch.float32).contiguous()

    out = []
    for name in sc.CONFIG_NAMES:
        st = ops.new_stats(dev)
        if name == "focal512":
            ng = 512
            rif = bump_dev(ng); h = float(sc.grid_spacing(ng)); ds = h / 2
            pos, vel = (T(a) for a in sc.rays_collimated(1448))
            label = "configs[4] focal stack: 512^3, 1448^2 = 2.1 M rays (one GPU's share of 16.8 M over 8)"
            kind = "grid"
        else:
            case = sc.config_case(name)
            kind, label = case["kind"], case["label"]
            rif = T(case["rif"]); pos, vel = T(case["pos"]), T(case["vel"]); ds = float(case["ds"])
            if kind == "grid":
                h = float(case["h"])
        one = torch.ones_like(pos)
        iters = 5 if pos.shape[0] > 1_000_000 else 20
        if kind == "grid":
            xt, vt = ops.trace(rif, pos, vel, h, ds, stats=st); sf = int(st[0]); st.zero_()
            g = torch.empty_like(rif)
            ops.backtrace(rif, xt, vt, one, one, h, ds, stats=st, out=g); sb = int(st[0])
            tf = timed(lambda: ops.trace(rif, pos, vel, h, ds), iters)
            tb = timed(lambda: ops.backtrace(rif, xt, vt, one, one, h, ds, out=g), iters)
            grid = list(rif.shape)
        else:
            R, L, tgt = case["radius"], case["length"], T(case["target"])
            xt, vt, _ = ops.trace_cable(rif, R, L, pos, vel, tgt, ds, stats=st); sf = int(st[0]); st.zero_()
            ops.backtrace_cable(rif, R, L, xt, vt, one, one, ds, stats=st); sb = int(st[0])
            tf = timed(lambda: ops.trace_cable(rif, R, L, pos, vel, tgt, ds), iters)
            tb = timed(lambda: ops.backtrace_cable(rif, R, L, xt, vt, one, one, ds), iters)
            grid = [int(rif.numel())]
        entry = {"name": name, "label": label, "grid": grid, "rays": int(pos.shape[0]), "fwd_ms": tf, "adj_ms": tb,
                 "fwd_ray_steps_per_s": sf / (tf * 1e-3), "adj_ray_steps_per_s": sb / (tb * 1e-3),
                 "ray_steps_per_s": (sf + sb) / ((tf + tb) * 1e-3)}
        if kind == "grid" and pos.shape[0] <= 300_000:
            # the small configurations as ONE CUDA graph replay (pipeline.GraphedStep: trace -> seeds -> backtrace), which
            # takes the host's launch path out of a step that lasts 0.1 - 0.6 ms
            try:
                from adjointnonlinearraytracing_b200.pipeline import GraphedStep
                gs = GraphedStep(rif, h, ds, int(pos.shape[0]))
                gs.step(pos, vel)
                tg = timed(lambda: gs.step(), iters)
                entry["graphed_step_ms"] = tg
                entry["graphed_ray_steps_per_s"] = (sf + sb) / (tg * 1e-3)
                del gs
            except Exception as e:
                entry["graphed_step_error"] = repr(e)[:200]
        out.append(entry)
        del rif, pos, vel, xt, vt, one
    # configs[0] the way the reference's driver batches it: SIX cube-face views per iteration, generated on the device with
    # the random rigid rotation fused in (core/luneburg_opt.py:53-59 -> source.rand_rays_cube), through the 32^3 lens field
    try:
        from adjointnonlinearraytracing_b200 import source
        from adjointnonlinearraytracing_b200.pipeline import GraphedStep
        ng = 32
        rif = T(sc.field_luneburg(ng)); h = float(sc.grid_spacing(ng)); ds = h / 2
        rot = np.array([[0.36, 0.48, -0.8], [-0.8, 0.6, 0.0], [0.48, 0.64, 0.6]])  # a fixed rotation matrix
        (pos, vel, _planes), _nr = source.rand_rays_cube((256, 256), 1, 1.0, circle=True, rotmat=rot, seed=0)
        one = torch.ones_like(pos); g = torch.empty_like(rif); st = ops.new_stats(dev)
        xt, vt = ops.trace(rif, pos, vel, h, ds, stats=st); sf = int(st[0]); st.zero_()
        ops.backtrace(rif, xt, vt, one, one, h, ds, stats=st, out=g); sb = int(st[0])
        tf = timed(lambda: ops.trace(rif, pos, vel, h, ds), 20)
        tb = timed(lambda: ops.backtrace(rif, xt, vt, one, one, h, ds, out=g), 20)
        gs = GraphedStep(rif, h, ds, int(pos.shape[0])); gs.step(pos, vel)
        tg = timed(lambda: gs.step(), 20)
        out.insert(1, {"name": "luneburg32_cube6", "label": "configs[0] as the reference's driver batches it: six cube-face views per iteration "
                       "(256^2 pixels each, circle filter, random rotation; generated on the device) through the 32^3 lens field",
                       "grid": [ng, ng, ng], "rays": int(pos.shape[0]), "fwd_ms": tf, "adj_ms": tb,
                       "fwd_ray_steps_per_s": sf / (tf * 1e-3), "adj_ray_steps_per_s": sb / (tb * 1e-3),
                       "ray_steps_per_s": (sf + sb) / ((tf + tb) * 1e-3), "graphed_step_ms": tg,
                       "graphed_ray_steps_per_s": (sf + sb) / (tg * 1e-3)})
        del gs, rif, pos, vel, xt, vt, one, g
    except Exception as e:
        out.insert(1, {"name": "luneburg32_cube6", "error": repr(e)[:200]})
    # the reference's protocol: 33^3, 512^2 collimated rays from z = 0, step sizes 0.3 .. 1.2 h, all-ones seeds
    ng = 33
    rif = T(sc.field_bump(ng)); h = float(sc.grid_spacing(ng))
    pos, vel = (T(a) for a in sc.rays_collimated(512))
    one = torch.ones_like(pos); g = torch.empty_like(rif)
    sweep = []
    for k in range(15):
        ds = (0.3 + 0.9 * k / 14) * h
        st = ops.new_stats(dev)
        xt, vt = ops.trace(rif, pos, vel, h, ds, stats=st); sf = int(st[0]); st.zero_()
        ops.backtrace(rif, xt, vt, one, one, h, ds, stats=st, out=g); sb = int(st[0])
        tf = timed(lambda: ops.trace(rif, pos, vel, h, ds), 10)
        tb = timed(lambda: ops.backtrace(rif, xt, vt, one, one, h, ds, out=g), 10)
        sweep.append({"ds_over_h": round(ds / h, 4), "fwd_ms": tf, "adj_ms": tb, "ray_steps_per_s": (sf + sb) / ((tf + tb) * 1e-3)})
    out.append({"name": "sweep33", "label": "the reference's own protocol (src/test.cpp:148-206): 33^3 grid, 512^2 rays, 15 step sizes "
                "0.3 .. 1.2 h, trace + backtrace with all-ones seeds",
                "grid": [ng, ng, ng], "rays": int(pos.shape[0]), "steps": sweep,
                "ray_steps_per_s_min": min(s["ray_steps_per_s"] for s in sweep),
                "ray_steps_per_s_max": max(s["ray_steps_per_s"] for s in sweep)})
    return out


# --------------------------------------------------------------------------- GPU arm
def run_ours(args):
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge
    ge.build()
    from adjointnonlinearraytracing_b200 import _capi, ops
    from adjointnonlinearraytracing_b200.dist import GradAllReduce, shard_cyclic

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    allreduce = GradAllReduce() if world > 1 else None
    collective = {"used": "ncclAllReduce (drrt_allreduce_grad)"} if world > 1 else None

    if NG >= 512:
        # 512^3: the field is evaluated on the device (the float64 numpy meshgrid takes ~6 GB of host memory per
        # process, x N ranks); same formula as scenes.field_bump, float64 then rounded to float32
        from adjointnonlinearraytracing_b200 import scenes as sc
        c = torch.arange(NG, device=dev, dtype=torch.float64) / (NG - 1)
        r2 = (c - 0.5) ** 2
        rif = (1.0 + 0.2 * torch.exp(-(r2[:, None, None] + r2[None, :, None] + r2[None, None, :]) / (2 * 0.2 * 0.2))
               ).to(torch.float32).contiguous()
        del c, r2
        h = float(sc.grid_spacing(NG)); ds = h / 2
        pos_h, vel_h = sc.rays_collimated(SENSOR)
        host = None  # the cpu_baseline leg (N = 1 only) builds its own host copy
    else:
        host = workload_host()
        rif_h, h, ds, pos_h, vel_h = host
        rif = torch.from_numpy(rif_h).to(dev)
    n_total = pos_h.shape[0]
    strong = args.scaling == "strong"
    flags = 0
    stats = ops.new_stats(dev)
    grad = torch.empty_like(rif)
    ev = lambda: torch.cuda.Event(enable_timing=True)

    # ---- N > 1: which collective.  The in-switch sum (our multimem kernel, dist.SwitchAllReduce) is timed against
    # NCCL's all-reduce on the gradient grid itself and used when it is available, correct and faster. ----------
    if world > 1 and args.collective != "nccl":
        try:
            from adjointnonlinearraytracing_b200.dist import SwitchAllReduce
            sw = SwitchAllReduce()
            gsw = sw.alloc(rif.shape, dev)

            def t_coll(fn, buf, iters=10):
                for _ in range(3):
                    fn(buf)
                dist.barrier(); torch.cuda.synchronize()
                a_, b_ = ev(), ev()
                a_.record()
                for _ in range(iters):
                    fn(buf)
                b_.record(); torch.cuda.synchronize()
                t = torch.tensor([a_.elapsed_time(b_) / iters], device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                return float(t)
            gen = torch.Generator(device=dev).manual_seed(7 + rank)
            src = torch.randn(rif.shape, device=dev, generator=gen)
            ref = src.clone(); allreduce(ref)
            gsw.copy_(src); sw(gsw); torch.cuda.synchronize()
            err = float(((gsw.double() - ref.double()).norm() / ref.double().norm()).item())
            t_nccl, t_sw = t_coll(allreduce, ref), t_coll(sw, gsw)
            okt = torch.tensor([1.0 if err <= 1e-6 else 0.0], device=dev); dist.all_reduce(okt, op=dist.ReduceOp.MIN)
            use = bool(okt.item() > 0) and (args.collective == "switch" or t_sw < t_nccl)
            collective.update({"nccl_ms": t_nccl, "switch_multimem_ms": t_sw, "switch_vs_nccl_rel_l2": err,
                               "note": "64-bit-identical inputs on both: ncclAllReduce vs drrt_allreduce_multimem (multimem.ld_reduce + multimem.st "
                                       "on the multicast mapping of the gradient grid, two stream barriers included), collective alone, max over ranks"})
            del src, ref
            if use:
                allreduce_nccl, allreduce, grad = allreduce, sw, gsw
                collective["used"] = "in-switch sum: drrt_allreduce_multimem (dist.SwitchAllReduce)"
            else:
                del gsw
        except Exception as e:  # no multicast fabric / torch without symmetric memory: NCCL it is
            collective["switch_unavailable"] = repr(e)[:300]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def measure(pos_pin, vel_pin, steps, warmup, want_clocks):
        """Device-resident timing of `steps` steps on this rank's batch -> dict (times are this rank's)."""
        pos, vel = pos_pin.to(dev), vel_pin.to(dev)
        dx = torch.ones_like(pos); dv = torch.ones_like(pos)

        def step(t=None):
            if t is not None:
                t[0].record()
            xt, vt = ops.trace(rif, pos, vel, h, ds, flags=flags, stats=stats)
            if t is not None:
                t[1].record()
            ops.backtrace(rif, xt, vt, dx, dv, h, ds, flags=flags, stats=stats, out=grad)
            if t is not None:
                t[2].record()
            if allreduce is not None:
                allreduce(grad)
            if t is not None:
                t[3].record()
            return xt, vt

        for _ in range(max(warmup, 3)):
            step()
        barrier()
        stats.zero_()
        timers = [[ev(), ev(), ev(), ev()] for _ in range(steps)]
        launches0 = _capi.launch_count()
        sampler = ClockSampler(local) if want_clocks else None
        t0, t1 = ev(), ev()
        barrier()
        if sampler:
            sampler.start()
        t0.record()
        for k in range(steps):
            step(timers[k])
        t1.record()
        barrier()
        clocks = sampler.stop() if sampler else None
        r = {"ms": t0.elapsed_time(t1), "launches": _capi.launch_count() - launches0, "clocks": clocks,
             "steps_total": int(stats[0].item()),
             "fwd_ms": sum(t[0].elapsed_time(t[1]) for t in timers) / steps,
             "adj_ms": sum(t[1].elapsed_time(t[2]) for t in timers) / steps,
             "ar_ms": sum(t[2].elapsed_time(t[3]) for t in timers) / steps}
        # ray-step split fwd / adjoint (one extra untimed pass each)
        stats.zero_(); xt, vt = ops.trace(rif, pos, vel, h, ds, flags=flags, stats=stats); r["s_f"] = int(stats[0].item())
        stats.zero_(); ops.backtrace(rif, xt, vt, dx, dv, h, ds, flags=flags, stats=stats, out=grad); r["s_b"] = int(stats[0].item())
        r["not_exited"] = int(stats[1].item())
        r["dev"] = (pos, vel, dx, dv)
        return r

    def reduce_times(vals):
        """-> (max over ranks, sum over ranks) of every entry."""
        t = torch.tensor(vals, dtype=torch.float64, device=dev)
        if world > 1:
            mx = t.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
            sm = t.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
            return mx.tolist(), sm.tolist()
        return t.tolist(), t.tolist()

    # ---- this rank's batch ---------------------------------------------------------
    if strong and world > 1:  # sensor rows dealt out cyclically: every rank gets the same mix of cheap and expensive rays
        # blocks of one cell's worth of sensor rows (8 rows at 2048^2 / 256^3): neighbouring rows of a rank still share
        # cells (L1 hits in the forward gather); profiles/r12_shard_block_ab.log
        sel = np.concatenate([np.arange(a, b) for a, b in shard_cyclic(n_total, rank, world, SENSOR * max(1, SENSOR // NG))])
    else:
        sel = slice(0, n_total)
    pos_pin = torch.from_numpy(np.ascontiguousarray(pos_h[sel])).pin_memory()
    vel_pin = torch.from_numpy(np.ascontiguousarray(vel_h[sel])).pin_memory()
    n_rays = pos_pin.shape[0]

    m = measure(pos_pin, vel_pin, args.steps, args.warmup, True)
    pos, vel, dx, dv = m.pop("dev")
    s_f, s_b = m["s_f"], m["s_b"]

    # ---- N > 1: the summed gradient against a single-GPU gradient of the whole batch ---------
    grad_check = None
    if world > 1:
        xt, vt = ops.trace(rif, pos, vel, h, ds, flags=flags)
        ops.backtrace(rif, xt, vt, dx, dv, h, ds, flags=flags, out=grad)
        allreduce(grad)
        det = _capi.DRRT_DETERMINISTIC
        gsum = grad.clone()
        ops.backtrace(rif, xt, vt, dx, dv, h, ds, flags=det, out=grad)  # `grad` is the buffer the collective knows
        allreduce(grad)
        gdet = grad.clone()
        torch.cuda.synchronize()
        if rank == 0:
            if strong:
                pf, vf = torch.from_numpy(pos_h).to(dev), torch.from_numpy(vel_h).to(dev)
                scale = 1.0
            else:  # weak: every rank traced the same batch, the sum is world x one gradient
                pf, vf, scale = pos, vel, float(world)
            of = torch.ones_like(pf)
            xf, vf2 = ops.trace(rif, pf, vf, h, ds, flags=flags)
            g1 = ops.backtrace(rif, xf, vf2, of, of, h, ds, flags=flags)
            g1d = ops.backtrace(rif, xf, vf2, of, of, h, ds, flags=det)
            e = float(((gsum.double() - scale * g1.double()).norm() / (scale * g1.double()).norm()).item())
            ed = float(((gdet.double() - scale * g1d.double()).norm() / (scale * g1d.double()).norm()).item())
            grad_check = {"rel_l2": e, "deterministic_mode_rel_l2": ed, "tol": 1e-5, "ok": bool(e <= 1e-5 and ed <= 1e-5),
                          "what": f"all-reduced gradient of {world} ranks vs rank 0's single-GPU gradient of the whole batch "
                                  f"({pf.shape[0]} rays{', x world' if not strong else ''}); float atomics resp. DRRT_DETERMINISTIC on both sides"}
            del pf, vf, of, xf, vf2, g1, g1d
        del xt, vt, gdet, gsum

    # ---- secondary, labelled: opt-in polynomial arithmetic (NOT a parity mode)
    fast = _capi.DRRT_FAST_ARITH
    fe = [ev(), ev()]
    for k in range(4):
        if k == 1:
            fe[0].record()
        xf_, vf_ = ops.trace(rif, pos, vel, h, ds, flags=fast)
        ops.backtrace(rif, xf_, vf_, dx, dv, h, ds, flags=fast, out=grad)
    fe[1].record(); torch.cuda.synchronize()
    fast_ms = fe[0].elapsed_time(fe[1]) / 3
    del xf_, vf_

    # ---- end to end: host ray batch (pinned) -> H2D -> trace -> backtrace -> allreduce -> D2H ----
    grad_pin = torch.empty(rif.shape, dtype=torch.float32).pin_memory()
    stat_pin = torch.empty(2, dtype=torch.int64).pin_memory()

    def e2e_step():
        p = pos_pin.to(dev, non_blocking=True)
        v = vel_pin.to(dev, non_blocking=True)
        stats.zero_()
        xt_, vt_ = ops.trace(rif, p, v, h, ds, flags=flags, stats=stats)
        ops.backtrace(rif, xt_, vt_, dx, dv, h, ds, flags=flags, stats=stats, out=grad)
        if allreduce is not None:
            allreduce(grad)
        grad_pin.copy_(grad, non_blocking=True)
        stat_pin.copy_(stats, non_blocking=True)

    for _ in range(3):
        e2e_step()
    barrier()
    e0, e1 = ev(), ev()
    e0.record()
    for _ in range(args.steps):
        e2e_step()
    e1.record()
    barrier()
    e2e_serial_ms = e0.elapsed_time(e1)

    # the same step through the package's streaming API (pipeline.HostStepPipeline): copies of
    # batch k+1 / gradient k-1 overlap the kernels of batch k; every step's gradient is read on the host
    from adjointnonlinearraytracing_b200.pipeline import HostStepPipeline
    # N > 1: every rank holds the same summed gradient after the collective; each rank lands ITS slab of planes in host
    # memory, so the job as a whole reads the full gradient grid on the host every step (d2h bytes summed over ranks =
    # one grid + the counters) instead of N identical copies of it
    from adjointnonlinearraytracing_b200.dist import shard_bounds
    pa, pb = shard_bounds(NG, rank, world)
    pipe = HostStepPipeline(rif, h, ds, n_rays, flags=flags, allreduce=allreduce,
                            host_planes=(pa, pb - pa) if world > 1 else None)
    for _ in range(3):
        pipe.submit(pos_pin, vel_pin)
    pipe.drain()
    barrier()
    e0, e1 = ev(), ev()
    e0.record()
    checksum = 0
    for k in range(args.steps):
        idx = pipe.submit(pos_pin, vel_pin)
        if k > 0:
            g_host, st_host = pipe.result(idx - 1)
            checksum += int(st_host[0])
    g_host, st_host = pipe.result(idx)
    checksum += int(st_host[0])
    e1.record(pipe.s_out)
    barrier()
    e2e_ms = e0.elapsed_time(e1)
    assert checksum == args.steps * (s_f + s_b), (checksum, s_f, s_b)
    h2d = pos_pin.numel() * 4 + vel_pin.numel() * 4
    d2h = pipe.d2h_bytes
    del pipe

    # ---- reduce over ranks: max time, sum of work ----------------------------------
    mx, sm = reduce_times([m["ms"], e2e_ms, float(m["steps_total"]), m["fwd_ms"], m["adj_ms"], m["ar_ms"], e2e_serial_ms,
                           float(h2d), float(d2h), float(m["launches"])])
    ms, e2e_ms, fwd_ms, adj_ms, ar_ms, e2e_serial_ms = mx[0], mx[1], mx[3], mx[4], mx[5], mx[6]
    steps_all = sm[2]
    value = steps_all / (ms * 1e-3)
    e2e_value = steps_all / (e2e_ms * 1e-3)

    # ---- N > 1, strong: the weak figure beside it (4 M rays on every rank) ----------------
    weak = None
    if world > 1 and strong:
        del pos, vel, dx, dv
        pw = torch.from_numpy(pos_h).pin_memory(); vw = torch.from_numpy(vel_h).pin_memory()
        wsteps = max(3, args.steps // 4)
        mw = measure(pw, vw, wsteps, 3, False)
        mw.pop("dev")
        wmx, wsm = reduce_times([mw["ms"], float(mw["steps_total"]), mw["ar_ms"]])
        weak = {"value": wsm[1] / (wmx[0] * 1e-3), "unit": UNIT, "ms_per_step": wmx[0] / wsteps,
                "rays_per_gpu": n_total, "allreduce_ms_per_step": wmx[2],
                "note": "weak scaling: every rank traces its own 4 194 304-ray batch (the r01 bench line)"}

    line = None
    if rank == 0:
        peak, peak_src = peaks()
        adj_gbs = BYTES_ADJ * s_b / (adj_ms * 1e-3) / 1e9
        fwd_gbs = BYTES_FWD * s_f / (fwd_ms * 1e-3) / 1e9
        sha = csrc_sha()

        def load(name):
            try:
                with open(os.path.join(ROOT, "profiles", name)) as f:
                    return json.load(f)
            except Exception:
                return None

        counts = load("kernel_counts.json")  # tools/ncu_counts.py: per-launch ncu counters of THIS workload + csrc_sha
        l2p = load("l2_peaks.json")
        roof = None
        scaled = world > 1 or not strong  # this rank's launches cover a share of the batch the counts were taken on
        if counts is not None and NG == 256 and (world == 1 or "warp_inst_per_warp_step" in counts["backtrace_kernel"]):
            k = dict(counts["backtrace_kernel"]); kf = dict(counts["trace_kernel"])
            if "warp_inst_per_warp_step" in k:  # same kernel, same batch geometry: instructions scale with the ray-steps
                k["warp_inst"] = k["warp_inst_per_warp_step"] * s_b / 32.0
                kf["warp_inst"] = kf["warp_inst_per_warp_step"] * s_f / 32.0
                if world > 1:
                    for d in (k, kf):
                        for key in ("dram_bytes", "l1_red_sectors", "lts_sectors"):
                            d[key] = None  # per-launch memory counters of the full batch do not carry over
            sm_mhz = (m["clocks"] or {}).get("sm_mhz") or counts.get("sm_mhz") or 1965
            pk_inst = 148 * 4 * sm_mhz * 1e6 / 1e9  # 4 schedulers x 1 warp instruction per clock per SM
            ai = k["warp_inst"] / (adj_ms * 1e-3) / 1e9
            fi = kf["warp_inst"] / (fwd_ms * 1e-3) / 1e9
            roof = {
                "bound": "issue",
                "kernel": "backtrace_kernel (adjoint; dominant: %.0f %% of the step)" % (100 * adj_ms / (fwd_ms + adj_ms)),
                "achieved": ai, "peak": pk_inst, "unit": "G warp-instructions/s", "frac": ai / pk_inst,
                "peak_source": f"148 SMs x 4 schedulers x 1 warp instruction per clock x {sm_mhz} MHz (SM clock sampled during the timed region)",
                "why": "both grids are L2-resident at 256^3 (DRAM traffic is ~0.3 % of the algorithmic bytes, see `traffic`), so the "
                       "kernel is bound by instruction issue / the FP32 pipes; frac = warp instructions per launch (ncu, committed) / "
                       "live kernel time / issue peak",
                "traffic": k.get("dram_bytes"),
                "warp_inst_per_launch": k["warp_inst"], "warp_inst_per_warp_step": k["warp_inst"] / (s_b / 32.0),
                "threads_per_inst": k.get("threads_per_inst"), "pipes_pct": k.get("pipes_pct"),
                "fwd_kernel": {"achieved": fi, "frac": fi / pk_inst, "warp_inst_per_launch": kf["warp_inst"],
                               "warp_inst_per_warp_step": kf["warp_inst"] / (s_f / 32.0),
                               "threads_per_inst": kf.get("threads_per_inst"), "pipes_pct": kf.get("pipes_pct"),
                               "traffic": kf.get("dram_bytes")},
                "counts_source": counts.get("source"), "counts_csrc_sha": counts.get("csrc_sha"), "built_csrc_sha": sha,
                "stale": bool(counts.get("csrc_sha") != sha),
                "counts_scaled_by_ray_steps": bool(scaled),  # N > 1: instructions per warp-step of the captured launch x this rank's ray-steps
            }
            if l2p is not None and k.get("l1_red_sectors"):
                red = k["l1_red_sectors"] / (adj_ms * 1e-3) / 1e9
                roof["l2"] = {"red_gsectors_per_s": red, "red_peak_gsectors_per_s": l2p["red_scattered_gsectors_per_s"],
                              "red_frac": red / l2p["red_scattered_gsectors_per_s"],
                              "lts_sectors_per_ray_step": (k.get("lts_sectors") or 0) / s_b,
                              "note": "RED sectors per launch (ncu) / live kernel time against the measured L2 atomic peak (tools/l2_probe.cu)"}
        hbm = {"bound": "hbm", "achieved": adj_gbs, "peak": peak, "unit": "GB/s", "frac": adj_gbs / peak, "peak_source": peak_src,
               "algorithmic_bytes_per_ray_step": BYTES_ADJ,
               "fwd_kernel": {"achieved": fwd_gbs, "frac": fwd_gbs / peak, "algorithmic_bytes_per_ray_step": BYTES_FWD},
               "step": {"achieved": (BYTES_FWD * s_f + BYTES_ADJ * s_b) / ((fwd_ms + adj_ms) * 1e-3) / 1e9,
                        "frac": (BYTES_FWD * s_f + BYTES_ADJ * s_b) / ((fwd_ms + adj_ms) * 1e-3) / 1e9 / peak},
               "note": "SURVEY section 8d's figure: algorithmic bytes (8 voxel reads [+ 8 atomic adds] x 4 B per ray-step) / kernel "
                       "time against the measured HBM copy bandwidth.  It can exceed 1: the grids never leave L2, HBM is not the roof."}
        if roof is None:  # no committed counts for this workload / multi-GPU line: the section-8d figure only
            roof = dict(hbm)
            roof["kernel"] = "backtrace_kernel"
            roof["traffic"] = None
        else:
            roof["algorithmic_vs_hbm"] = hbm
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {
                "workload": workload_name(),
                "rays_total": n_total * (1 if strong else world), "rays_per_gpu": n_rays, "grid": [NG, NG, NG],
                "parallelism": (f"rays-dp{world} (blocks of 8 sensor rows of the one batch dealt out cyclically, dist.shard_cyclic) + private gradient grids + one gradient sum: {collective['used']}"
                                if world > 1 else "single GPU"),
                "l2": (f"no explicit flush: per-step working set (rays in/out {n_rays * 72 / 1e6:.0f} MB + grid {NG ** 3 * 4 / 1e6:.0f} MB + "
                       f"gradient {NG ** 3 * 4 / 1e6:.0f} MB) exceeds the 126 MB L2"
                       + ("; the 67 MB IOR grid staying L2-resident across the step is the design" if NG == 256 else "")),
                "arithmetic": "exact trajectory (default flags=0): ray path bit-identical to the float32 reference formulas",
                "ray_steps_fwd_rank0": s_f, "ray_steps_adj_rank0": s_b, "rays_not_exited_rank0": m["not_exited"],
            },
            "fwd_ms": fwd_ms, "adj_ms": adj_ms,
            "fwd_ray_steps_per_s": s_f / (fwd_ms * 1e-3), "adj_ray_steps_per_s": s_b / (adj_ms * 1e-3),
            "roofline": roof,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(sm[7]), "d2h_bytes_per_step": int(sm[8]),
                    "ms_per_step": e2e_ms / args.steps,
                    "path": "pipeline.HostStepPipeline: pinned host rays -> H2D -> ops.trace -> ops.backtrace -> gradient sum -> D2H gradient grid + counters into pinned memory, read on the host every step; depth-2 double buffering on copy-in / compute / copy-out streams; bytes are the sum over ranks (N > 1: each rank lands its slab of the summed gradient, the job one full grid per step; serial_ms_per_step copies the whole grid on every rank)",
                    "serial_ms_per_step": e2e_serial_ms / args.steps},
            "gpu_launches": int(sm[9]), "clocks": m["clocks"],
            "secondary": {"fast_arith_ms_per_step": fast_ms, "fast_arith_ray_steps_per_s": (s_f + s_b) / (fast_ms * 1e-3),
                          "note": "DRRT_FAST_ARITH (factored polynomial cell) on rank 0's batch: ray paths differ from the reference at cell faces; not used for value/e2e"},
        }
        if world > 1:
            line["allreduce"] = {"ms_per_step": ar_ms, "bytes": NG ** 3 * 4, "collective": collective,
                                 "note": "the gradient sum inside the timed step, CUDA events around the call on the compute stream, max over ranks "
                                         "(includes waiting for the slowest rank's adjoint)"}
            line["grad_check"] = grad_check
            if weak is not None:
                line["secondary"]["weak"] = weak
        if world == 1 and not args.no_cpu:
            arm = CpuArm(target_s=8.0, host=host)
            sf, sb = arm.count_steps()
            t = arm.step()
            line["cpu_baseline"] = {"value": (sf + sb) / t, "unit": UNIT, "cores": arm.threads, "kind": arm.kind,
                                    "sample": arm.describe(), "seconds": t}
            try:
                line["parity"] = parity_leg(arm, ops, torch, dev, rif)
            except Exception as e:
                line["parity"] = {"ok": False, "error": repr(e)}
        if world == 1 and not args.no_configs:
            try:
                del grad_pin
                line["configs"] = config_lines(ops, torch, dev)
            except Exception as e:
                line["configs"] = {"error": repr(e)}
        _emit(line)
    if allreduce is not None:
        allreduce.close()
        if "allreduce_nccl" in locals():
            allreduce_nccl.close()
    if world > 1:
        dist.destroy_process_group()
    if rank == 0:
        if "parity" in line and not line["parity"].get("ok", False):
            sys.stderr.write("bench.py: PARITY FAILED: " + json.dumps(line["parity"]) + "\n")
            sys.exit(3)
        if grad_check is not None and not grad_check["ok"]:
            sys.stderr.write("bench.py: GRADIENT CHECK FAILED: " + json.dumps(grad_check) + "\n")
            sys.exit(4)


_REAL_STDOUT = None


def _guard_stdout():
    """Everything libraries print to fd 1 (e.g. NCCL's version banner) goes to stderr; the one JSON
    line is written to the original stdout by _emit()."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def _emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"],
                    help="N > 1: strong = the configuration as written (one batch of 2048^2 rays sharded over the ranks, default); "
                         "weak = 2048^2 rays per rank")
    ap.add_argument("--collective", default="auto", choices=["auto", "nccl", "switch"],
                    help="N > 1: the gradient sum -- NCCL all-reduce, this library's in-switch multimem kernel, or (auto) whichever is faster here")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline + parity legs")
    ap.add_argument("--no-configs", action="store_true", help="skip the per-configuration entries")
    ap.add_argument("--workload", default="caustic256", choices=["caustic256", "focal512"],
                    help="caustic256 = BASELINE.json's metric configuration (default, the bench line); focal512 = "
                         "configs[4]: 512^3 grid, 4096^2 = 16.8 M rays in total (strong) / 1448^2 = 2.1 M per GPU (weak), a labelled extra run")
    args = ap.parse_args()
    if args.workload == "focal512":
        global NG, SENSOR, METRIC
        NG = 512
        SENSOR = 4096 if args.scaling == "strong" else 1448
        METRIC = "fwd+adjoint ray-steps/sec at 512^3 grid"
    _guard_stdout()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
