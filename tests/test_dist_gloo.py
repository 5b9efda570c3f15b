"""CPU, world_size 2, gloo: the host logic of the multi-GPU path --
contiguous ray shards, one private gradient grid per rank, one sum all-reduce.  The per-rank
adjoint is computed by the CPU oracle here (test infrastructure); on the GPU box the same
`GradAllReduce` object drives `drrt_allreduce_grad` (NCCL) instead of gloo."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import oracle
        from adjointnonlinearraytracing_b200 import scenes as sc
        from adjointnonlinearraytracing_b200.dist import GradAllReduce, shard_bounds, shard_rays
        ng, s = 17, 24
        rif = sc.field_bump(ng); h = sc.grid_spacing(ng); ds = np.float32(h / 2)
        pos, vel = sc.rays_cone(s, backoff=0.37 * float(ds))
        dx, dv = sc.seeds_normal(s * s)
        a, b = shard_bounds(s * s, rank, world)
        tp, tv, tdx, tdv = shard_rays(rank, world, *(torch.from_numpy(t) for t in (pos, vel, dx, dv)))
        assert tp.shape[0] == b - a
        xt, vt = oracle.port.trace(rif, tp.numpy(), tv.numpy(), h, ds)
        g_local = oracle.port.backtrace(rif, xt, vt, tdx.numpy(), tdv.numpy(), h, ds)
        ar = GradAllReduce()  # gloo group -> backend "torch"
        assert ar.backend == "torch" and ar.world == world and ar.rank == rank
        g = ar(torch.from_numpy(np.ascontiguousarray(g_local)))
        np.save(os.path.join(out_dir, f"grad{rank}.npy"), g.numpy())
        np.save(os.path.join(out_dir, f"xt{rank}.npy"), xt)
        ar.close()
    finally:
        dist.destroy_process_group()


def test_shard_bounds_cover_and_balance():
    sys.path.insert(0, ROOT)
    from adjointnonlinearraytracing_b200.dist import shard_bounds
    for n in (0, 1, 7, 8, 4194304, 1001):
        for world in (1, 2, 3, 8):
            edges = [shard_bounds(n, r, world) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(edges[i][1] == edges[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in edges]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_bounds(10, 2, 2)


def test_two_rank_gradient_allreduce_matches_single_batch(tmp_path, oracle_mod):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    from adjointnonlinearraytracing_b200 import scenes as sc
    ng, s = 17, 24
    rif = sc.field_bump(ng); h = sc.grid_spacing(ng); ds = np.float32(h / 2)
    pos, vel = sc.rays_cone(s, backoff=0.37 * float(ds))
    dx, dv = sc.seeds_normal(s * s)
    xt, vt = oracle_mod.port.trace(rif, pos, vel, h, ds)
    whole = oracle_mod.port.backtrace(rif, xt, vt, dx, dv, h, ds)
    g0 = np.load(tmp_path / "grad0.npy"); g1 = np.load(tmp_path / "grad1.npy")
    assert np.array_equal(g0, g1)  # every rank holds the same summed grid
    assert np.linalg.norm(g0 - whole) / np.linalg.norm(whole) <= 1e-6
    # forward outputs stay sharded: concatenating the shards gives the single-batch result
    xs = np.concatenate([np.load(tmp_path / "xt0.npy"), np.load(tmp_path / "xt1.npy")])
    assert np.array_equal(xs, xt)


def _adam_worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import optim_oracle as oo
        from adjointnonlinearraytracing_b200.dist import ShardedAdam
        rng = np.random.default_rng(3)
        shape = (8, 5, 6)
        n0 = (1.0 + 0.2 * rng.random(shape)).astype(np.float32)
        grads = [[rng.standard_normal(shape).astype(np.float32) for _ in range(world)] for _ in range(3)]

        def step_fn(n_slab, g_slab, m_slab, v_slab, full_shape, i0, step):
            # the CPU restatement on the WHOLE grid (it owns the boundary mask), cut back to this slab
            k = n_slab.shape[0]
            pad = lambda t: np.concatenate([np.zeros((i0,) + t.shape[1:], np.float32), t.numpy(),
                                            np.zeros((full_shape[0] - i0 - k,) + t.shape[1:], np.float32)])
            out, m, v = oo.adam_step(pad(n_slab), pad(g_slab), pad(m_slab), pad(v_slab), step, lr=1e-2)
            n_slab.copy_(torch.from_numpy(out[i0:i0 + k])); m_slab.copy_(torch.from_numpy(m[i0:i0 + k]))
            v_slab.copy_(torch.from_numpy(v[i0:i0 + k]))

        n = torch.from_numpy(n0.copy())
        opt = ShardedAdam(n, 1e-2, step_fn=step_fn)
        assert opt.world == world and opt.planes == shape[0] // world and opt.i0 == rank * opt.planes
        assert opt.exp_avg.shape == (shape[0] // world,) + shape[1:]
        for it in range(3):
            opt.step(torch.from_numpy(grads[it][rank].copy()))
        np.save(os.path.join(out_dir, f"n{rank}.npy"), n.numpy())
        bad = torch.zeros(7, 5, 6)
        try:
            ShardedAdam(bad, 1e-2, step_fn=step_fn)
            raise AssertionError("7 planes over 2 ranks must be refused")
        except RuntimeError:
            pass
    finally:
        dist.destroy_process_group()


def test_two_rank_sharded_adam_matches_single_process(tmp_path, oracle_mod):
    world = 2
    mp.spawn(_adam_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    from oracle import optim_oracle as oo
    rng = np.random.default_rng(3)
    shape = (8, 5, 6)
    n = (1.0 + 0.2 * rng.random(shape)).astype(np.float32)
    grads = [[rng.standard_normal(shape).astype(np.float32) for _ in range(world)] for _ in range(3)]
    m = np.zeros(shape, np.float32); v = np.zeros(shape, np.float32)
    for it in range(3):
        n, m, v = oo.adam_step(n, grads[it][0] + grads[it][1], m, v, it + 1, lr=1e-2)
    n0 = np.load(tmp_path / "n0.npy"); n1 = np.load(tmp_path / "n1.npy")
    assert np.array_equal(n0, n1)  # every rank holds the same updated grid
    assert np.array_equal(n0, n)   # and it is the single-process result


def test_shard_cyclic_covers_every_ray_once():
    from adjointnonlinearraytracing_b200.dist import shard_cyclic
    for n, world, block in ((4194304, 8, 2048), (1000, 3, 64), (5, 4, 2), (7, 1, 3)):
        seen = np.zeros(n, np.int32)
        sizes = []
        for r in range(world):
            tot = 0
            for a, b in shard_cyclic(n, r, world, block):
                assert 0 <= a < b <= n and b - a <= block and a % block == 0
                seen[a:b] += 1; tot += b - a
            sizes.append(tot)
        assert (seen == 1).all()
        assert max(sizes) - min(sizes) <= block
    with pytest.raises(ValueError):
        shard_cyclic(10, 2, 2, 4)
