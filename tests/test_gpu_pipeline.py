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
