"""GPU: the streaming host pipeline returns, for every submitted batch, exactly what the direct
ops calls return (deterministic mode: bit-identical), also when buffers are recycled."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DET = 2


def test_host_pipeline_matches_direct_calls(cuda_lib):
    from adjointnonlinearraytracing_b200 import ops, scenes as sc
    from adjointnonlinearraytracing_b200.pipeline import HostStepPipeline
    ng, s = 33, 64
    rif = torch.from_numpy(sc.field_noise(ng)).cuda()
    h = float(sc.grid_spacing(ng)); ds = h / 2
    batches = []
    for k in range(5):
        pos, vel = sc.rays_cone(s, seed=k, backoff=0.37 * ds)
        batches.append((torch.from_numpy(pos).pin_memory(), torch.from_numpy(vel).pin_memory()))
    seed_fn = lambda xt, vt: (xt - 0.5, vt)  # a loss gradient that depends on the forward outputs
    pipe = HostStepPipeline(rif, h, ds, s * s, flags=DET, seed_fn=seed_fn)
    got = {}
    for k, (p, v) in enumerate(batches):
        idx = pipe.submit(p, v)
        assert idx == k
        if k > 0:
            g, st = pipe.result(k - 1)
            got[k - 1] = (g.clone(), st.clone())
    g, st = pipe.result(len(batches) - 1)
    got[len(batches) - 1] = (g.clone(), st.clone())
    with pytest.raises(RuntimeError):
        pipe.result(0)  # recycled
    for k, (p, v) in enumerate(batches):
        stats = ops.new_stats("cuda")
        xt, vt = ops.trace(rif, p.cuda(), v.cuda(), h, ds, flags=DET, stats=stats)
        ref = ops.backtrace(rif, xt, vt, xt - 0.5, vt, h, ds, flags=DET, stats=stats)
        assert torch.equal(got[k][0], ref.cpu()), k
        assert torch.equal(got[k][1], stats.cpu()), k


def test_host_pipeline_slab_copy(cuda_lib):
    """host_planes: only the given planes of the gradient grid travel to the host (one slab per rank after the
    all-reduce); they equal the same planes of the full result, and d2h_bytes counts the slab."""
    from adjointnonlinearraytracing_b200 import ops, scenes as sc
    from adjointnonlinearraytracing_b200.pipeline import HostStepPipeline
    ng, s = 33, 48
    rif = torch.from_numpy(sc.field_noise(ng)).cuda()
    h = float(sc.grid_spacing(ng)); ds = h / 2
    pos, vel = sc.rays_cone(s, seed=3, backoff=0.37 * ds)
    p, v = torch.from_numpy(pos).pin_memory(), torch.from_numpy(vel).pin_memory()
    pipe = HostStepPipeline(rif, h, ds, s * s, flags=DET, host_planes=(9, 7))
    for _ in range(3):  # buffers recycled
        k = pipe.submit(p, v)
    g, st = pipe.result(k)
    xt, vt = ops.trace(rif, p.cuda(), v.cuda(), h, ds, flags=DET)
    one = torch.ones_like(xt)
    ref = ops.backtrace(rif, xt, vt, one, one, h, ds, flags=DET)
    assert g.shape == (7, ng, ng) and torch.equal(g, ref[9:16].cpu())
    assert pipe.d2h_bytes == 7 * ng * ng * 4 + 16
    with pytest.raises(RuntimeError):
        HostStepPipeline(rif, h, ds, s * s, host_planes=(30, 7))


@pytest.mark.parametrize("flags", [DET, 0])
def test_graphed_step_replays_what_direct_calls_compute(flags, cuda_lib):
    """One captured CUDA graph, replayed over new ray batches and an in-place-updated grid."""
    from adjointnonlinearraytracing_b200 import ops, scenes as sc
    from adjointnonlinearraytracing_b200.pipeline import GraphedStep
    ng, s = 33, 64
    rif = torch.from_numpy(sc.field_noise(ng)).cuda()
    h = float(sc.grid_spacing(ng)); ds = h / 2
    seed_fn = lambda xt, vt: (xt - 0.5, vt)
    step = GraphedStep(rif, h, ds, s * s, flags=flags, seed_fn=seed_fn)
    for k in range(4):
        pos, vel = sc.rays_cone(s, seed=k, backoff=0.37 * ds)
        p, v = torch.from_numpy(pos).pin_memory(), torch.from_numpy(vel).pin_memory()
        xt, vt, grad, st = step.step(p, v)
        stats = ops.new_stats("cuda")
        xr, vr = ops.trace(rif, p.cuda(), v.cuda(), h, ds, flags=flags, stats=stats)
        ref = ops.backtrace(rif, xr, vr, xr - 0.5, vr, h, ds, flags=flags, stats=stats)
        assert torch.equal(xt, xr) and torch.equal(vt, vr) and torch.equal(st, stats), k
        if flags & DET:
            assert torch.equal(grad, ref), k
        else:
            assert float((grad - ref).norm() / ref.norm()) <= 2e-6, k
        rif.mul_(1.0 + 0.01 * (k + 1)).clamp_(min=1.0)  # the optimiser's in-place update
    # same batch again (no copy): the updated grid is what the replay reads
    xt2, vt2, grad2, _ = step.step()
    xr, vr = ops.trace(rif, p.cuda(), v.cuda(), h, ds, flags=flags)
    assert torch.equal(xt2, xr) and torch.equal(vt2, vr)
    assert step.launches == 5
