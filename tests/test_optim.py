"""SURVEY.md section 8 f1 -- optimiser step + multires up-res.

CPU: the numpy restatement (oracle/optim_oracle.py) replays the reference's own
core/optimizer.py::multires_opt run stored in tests/golden/optim5to9.npz.
GPU: the fused CUDA step / up-res (through the C ABI) replay the same run, and the drop-in
`optimizer.multires_opt` reproduces the reference's final grid, Adam state and loss history.
Tolerance: 2e-6 absolute on n ~ 1 (float32 rounding of a different but equivalent operation order)."""
import numpy as np
import pytest
import torch

from conftest import load_golden


def _replay(step_fn, upres_fn, g):
    """Replays the recorded run with the given step / up-res implementations; returns per-iteration n."""
    n = g["eta"].copy()
    m = np.zeros_like(n); v = np.zeros_like(n)
    it, step, out = 0, 0, []
    res_list = [int(r) for r in g["res_list"]]
    for lvl, r in enumerate(res_list):
        T, Wt = g[f"target{r}"], g[f"weight{r}"]
        for _ in range(int(g["iterations"]) * (lvl + 1)):
            out.append(n.copy())
            grad = (2 * Wt * (n - T)).astype(np.float32)
            step += 1
            n, m, v = step_fn(n, grad, m, v, step, float(g["lr"]))
            it += 1
        if lvl < len(res_list) - 1:
            nr = [res_list[lvl + 1]] * 3
            n, m, v = (upres_fn(a, nr).astype(np.float32) for a in (n, m, v))
    return out, n, m, v


def _check(out, n, m, v, g, tol=2e-6):
    for i, snap in enumerate(out):
        assert np.abs(snap - g[f"n_before_iter{i}"]).max() <= tol, i
    assert np.abs(n - g["final"]).max() <= tol
    assert np.abs(m - g["final_exp_avg"]).max() <= 5e-6
    assert np.abs(v - g["final_exp_avg_sq"]).max() <= 1e-6


def test_oracle_replays_reference_run(oracle_mod):
    from oracle import optim_oracle as oo
    g = load_golden("optim5to9")
    assert np.abs(oo.upres_volume(g["eta"], [9, 9, 9]).astype(np.float32) - g["upres_eta_9"]).max() <= 1e-7
    out, n, m, v = _replay(lambda n, gr, m, v, s, lr: oo.adam_step(n, gr, m, v, s, lr), oo.upres_volume, g)
    assert len(out) == 12 and float(g["final_step"]) == 12
    _check(out, n, m, v, g)
    assert (n >= 1).all() and (n == 1).any()  # the clamp was exercised


@pytest.mark.gpu
def test_cuda_step_and_upres_replay_reference_run(cuda_lib):
    from adjointnonlinearraytracing_b200 import ops
    g = load_golden("optim5to9")
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).cuda()

    def step(n, grad, m, v, s, lr):
        tn, tm, tv = T(n), T(m), T(v)
        ops.adam_step_(tn, T(grad), tm, tv, s, lr)
        return tn.cpu().numpy(), tm.cpu().numpy(), tv.cpu().numpy()

    up = lambda a, r: ops.upres_volume(T(a), r).cpu().numpy()
    assert np.abs(up(g["eta"], [9, 9, 9]) - g["upres_eta_9"]).max() <= 1e-7
    out, n, m, v = _replay(step, up, g)
    _check(out, n, m, v, g)


@pytest.mark.gpu
def test_multires_opt_drop_in(cuda_lib, tmp_path):
    from adjointnonlinearraytracing_b200 import ops, optimizer
    g = load_golden("optim5to9")
    tg = {int(r): torch.from_numpy(g[f"target{int(r)}"]).cuda() for r in g["res_list"]}
    wt = {int(r): torch.from_numpy(g[f"weight{int(r)}"]).cuda() for r in g["res_list"]}
    func = lambda n: (wt[n.shape[0]] * (n - tg[n.shape[0]]) ** 2).sum()
    snaps = []
    n, hist = optimizer.multires_opt(func, torch.from_numpy(g["eta"]).cuda(), int(g["iterations"]),
                                     [int(r) for r in g["res_list"]], log_func=lambda i, x: snaps.append(x.detach().cpu().numpy()),
                                     lr=float(g["lr"]), statename=str(tmp_path / "state.pt"))
    assert len(snaps) == 12
    for i, s in enumerate(snaps):
        assert np.abs(s - g[f"n_before_iter{i}"]).max() <= 2e-6, i
    assert np.abs(n.detach().cpu().numpy() - g["final"]).max() <= 2e-6
    assert np.allclose(hist, g["loss_hist"], rtol=1e-5)
    st = torch.load(tmp_path / "state.pt", weights_only=False)
    assert set(st) == {"rif", "opto_state_dict", "loss_hist"}
    sd = st["opto_state_dict"]
    assert int(sd["state"][0]["step"]) == 12 and sd["param_groups"][0]["params"] == [0]
    # the checkpoint is torch.optim.Adam's own layout (reference core/optimizer.py:73-78): torch loads it ...
    p = st["rif"].detach().clone().requires_grad_(True)
    topt = torch.optim.Adam([p], lr=123.0)
    topt.load_state_dict(sd)
    assert topt.param_groups[0]["lr"] == float(g["lr"])
    # ... and the next step of torch.optim.Adam from it equals the fused kernel's next step
    mine = optimizer.FusedAdamState(p.detach(), lr=0.0)
    mine.load_state_dict(topt.state_dict())
    assert mine.step == 12 and mine.lr == float(g["lr"])
    gr = torch.randn_like(p)
    p.grad = gr.clone(); topt.step()
    q = st["rif"].detach().clone()
    mine.step += 1
    ops.adam_step_(q, gr, mine.exp_avg, mine.exp_avg_sq, mine.step, mine.lr, mine.betas, mine.eps,
                   clamp_min=1.0, zero_boundary=False, clamp=False)
    assert (q - p.detach()).abs().max().item() <= 2e-6


@pytest.mark.gpu
def test_adam_step_full_size_properties(cuda_lib):
    """256^3: boundary voxels never move, interior matches torch.optim.Adam + clamp, non-vector path agrees."""
    from adjointnonlinearraytracing_b200 import ops
    gen = torch.Generator(device="cuda").manual_seed(0)
    S = 256
    n0 = 1 + 0.1 * torch.rand(S, S, S, device="cuda", generator=gen)
    grad = torch.randn(S, S, S, device="cuda", generator=gen)
    p = n0.clone().requires_grad_(True)
    opt = torch.optim.Adam([p], lr=1e-2)
    n, m, v = n0.clone(), torch.zeros_like(n0), torch.zeros_like(n0)
    mask = torch.ones_like(n0, dtype=torch.bool); mask[1:-1, 1:-1, 1:-1] = False
    for step in (1, 2, 3):
        p.grad = grad.clone(); p.grad[mask] = 0
        opt.step()
        with torch.no_grad():
            p.clamp_(min=1)
        ops.adam_step_(n, grad, m, v, step, 1e-2)
        grad = grad * 0.5 + 0.1
    assert torch.equal(n[mask], n0[mask])
    assert float((n - p.detach()).abs().max()) <= 2e-6
    # odd innermost size -> scalar kernel
    a = 1 + 0.1 * torch.rand(7, 6, 5, device="cuda", generator=gen); ga = torch.randn(7, 6, 5, device="cuda", generator=gen)
    pa = a.clone().requires_grad_(True); oa = torch.optim.Adam([pa], lr=1e-2)
    ma, va = torch.zeros_like(a), torch.zeros_like(a)
    mk = torch.ones_like(a, dtype=torch.bool); mk[1:-1, 1:-1, 1:-1] = False
    pa.grad = ga.clone(); pa.grad[mk] = 0; oa.step()
    ops.adam_step_(a, ga, ma, va, 1, 1e-2, clamp=False)
    assert float((a - pa.detach()).abs().max()) <= 2e-6


@pytest.mark.gpu
def test_adam_step_slab_equals_whole_grid(cuda_lib):
    """Any split of the grid into slabs of the slowest axis, stepped slab by slab, gives the bits of the one-pass
    step (the boundary that is zeroed is the whole grid's), for vectorised and odd shapes."""
    from adjointnonlinearraytracing_b200 import ops
    gen = torch.Generator(device="cuda").manual_seed(11)
    for shape, cuts in (((16, 12, 20), (0, 4, 8, 12, 16)), ((9, 7, 5), (0, 1, 2, 8, 9)), ((6, 4, 8), (0, 6))):
        n = 1.0 + 0.3 * torch.rand(shape, device="cuda", generator=gen)
        m = torch.rand(shape, device="cuda", generator=gen) * 0.1
        v = torch.rand(shape, device="cuda", generator=gen) * 0.01
        g = torch.randn(shape, device="cuda", generator=gen)
        for kw in ({}, {"zero_boundary": False}, {"clamp": False}):
            a, am, av = n.clone(), m.clone(), v.clone()
            ops.adam_step_(a, g, am, av, 3, 1e-2, **kw)
            b, bm, bv = n.clone(), m.clone(), v.clone()
            for lo, hi in zip(cuts[:-1], cuts[1:]):
                ops.adam_step_slab_(b[lo:hi], g[lo:hi], bm[lo:hi], bv[lo:hi], shape, lo, 3, 1e-2, **kw)
            assert torch.equal(a, b) and torch.equal(am, bm) and torch.equal(av, bv)
    lib_err = None
    try:
        ops.adam_step_slab_(n[0:2], g[0:2], m[0:2], v[0:2], (6, 4, 8), 5, 1, 1e-2)  # planes 5..6 of 6
    except RuntimeError as e:
        lib_err = str(e)
    assert lib_err and "slab" in lib_err


def _nccl_adam_worker(rank, world, port, out_dir):
    import os, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from adjointnonlinearraytracing_b200 import ops
        from adjointnonlinearraytracing_b200.dist import GradAllReduce, ShardedAdam
        shape = (32, 20, 24)
        gen = torch.Generator(device="cuda").manual_seed(7)  # same stream of numbers on every rank
        n0 = 1.0 + 0.2 * torch.rand(shape, device="cuda", generator=gen)
        grads = [[torch.randn(shape, device="cuda", generator=gen) for _ in range(world)] for _ in range(3)]
        ar = GradAllReduce()
        assert ar.backend == "nccl-native"
        n = n0.clone()
        opt = ShardedAdam(n, 1e-2, allreduce=ar)
        ref = n0.clone(); m = torch.zeros_like(ref); v = torch.zeros_like(ref)
        for it in range(3):
            opt.step(grads[it][rank].clone())
            summed = ar(grads[it][rank].clone())
            ops.adam_step_(ref, summed, m, v, it + 1, 1e-2)
        torch.cuda.synchronize()
        err = float((n - ref).abs().max())
        lo, hi = opt.i0, opt.i0 + opt.planes
        merr = float((opt.exp_avg - m[lo:hi]).abs().max())
        np.save(os.path.join(out_dir, f"n{rank}.npy"), n.cpu().numpy())
        np.save(os.path.join(out_dir, f"err{rank}.npy"), np.array([err, merr]))
        ar.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.timeout(300)  # a rendezvous that never completes fails the test instead of hanging the run
def test_sharded_adam_two_gpus_nccl(cuda_lib, tmp_path):
    """reduce-scatter -> slab Adam -> all-gather over NCCL against all-reduce -> whole-grid Adam (needs 2 GPUs)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import socket
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
    mp.spawn(_nccl_adam_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    n0 = np.load(tmp_path / "n0.npy"); n1 = np.load(tmp_path / "n1.npy")
    assert np.array_equal(n0, n1)
    for r in range(2):
        err, merr = np.load(tmp_path / f"err{r}.npy")
        assert err <= 1e-6 and merr <= 1e-6, (r, err, merr)


def _peer_adam_worker(rank, world, port, out_dir):
    import os, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from adjointnonlinearraytracing_b200 import ops
        from adjointnonlinearraytracing_b200.dist import GradAllReduce, PeerShardedAdam
        ar = GradAllReduce()
        errs = []
        for shape in ((32, 20, 24), (6, 7, 5)):  # float4 rows / odd rows (scalar kernel)
            gen = torch.Generator(device="cuda").manual_seed(7)  # the same stream of numbers on every rank
            n0 = 1.0 + 0.2 * torch.rand(shape, device="cuda", generator=gen)
            grads = [[torch.randn(shape, device="cuda", generator=gen) for _ in range(world)] for _ in range(3)]
            n = n0.clone(); gbuf = torch.zeros_like(n)
            opt = PeerShardedAdam(n, gbuf, 1e-2, allreduce=ar)
            ref = n0.clone(); m = torch.zeros_like(ref); v = torch.zeros_like(ref)
            for it in range(3):
                gbuf.copy_(grads[it][rank])           # stands for the adjoint writing this rank's gradient
                opt.step()
                total = grads[it][0].clone()
                for r in range(1, world):
                    total += grads[it][r]             # rank order, like the kernel
                ops.adam_step_(ref, total, m, v, it + 1, 1e-2)
            torch.cuda.synchronize()
            lo, hi = opt.i0, opt.i0 + opt.planes
            errs += [float((n - ref).abs().max()), float((opt.exp_avg - m[lo:hi]).abs().max()),
                     float(torch.equal(n, ref))]
            np.save(os.path.join(out_dir, f"pn{rank}_{shape[0]}.npy"), n.cpu().numpy())
        np.save(os.path.join(out_dir, f"perr{rank}.npy"), np.array(errs))
        opt.close()
        ar.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.timeout(300)  # a rendezvous that never completes fails the test instead of hanging the run
def test_peer_sharded_adam_two_gpus(cuda_lib, tmp_path):
    """The one-kernel sharded step over peer memory (CUDA IPC + NVLink loads / stores) against the whole-grid
    step on the rank-ordered sum: bit-identical, and every replica ends up the same (needs 2 GPUs)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import socket
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
    mp.spawn(_peer_adam_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    for k in (32, 6):
        assert np.array_equal(np.load(tmp_path / f"pn0_{k}.npy"), np.load(tmp_path / f"pn1_{k}.npy"))
    for r in range(2):
        e = np.load(tmp_path / f"perr{r}.npy")
        assert e[0] == 0 and e[1] == 0 and e[2] == 1 and e[3] == 0 and e[4] == 0 and e[5] == 1, (r, e)


@pytest.mark.gpu
def test_peer_sharded_adam_single_rank_equals_whole_grid(cuda_lib):
    """world = 1: the peer-memory kernel with one replica is the whole-grid step, bit for bit (float4 and scalar
    rows), through the same pointer-table path the multi-GPU case uses."""
    from adjointnonlinearraytracing_b200 import ops
    from adjointnonlinearraytracing_b200.dist import PeerShardedAdam
    gen = torch.Generator(device="cuda").manual_seed(21)
    for shape in ((12, 8, 16), (5, 6, 7)):
        n0 = 1.0 + 0.2 * torch.rand(shape, device="cuda", generator=gen)
        n = n0.clone(); gbuf = torch.zeros_like(n)
        opt = PeerShardedAdam(n, gbuf, 1e-2)
        assert opt.world == 1 and opt.planes == shape[0]
        ref = n0.clone(); m = torch.zeros_like(ref); v = torch.zeros_like(ref)
        for it in range(3):
            g = torch.randn(shape, device="cuda", generator=gen)
            gbuf.copy_(g)
            opt.step()
            ops.adam_step_(ref, g, m, v, it + 1, 1e-2)
        assert torch.equal(n, ref) and torch.equal(opt.exp_avg, m) and torch.equal(opt.exp_avg_sq, v)
    with pytest.raises(RuntimeError):
        PeerShardedAdam(n, torch.zeros(5, 6, 8, device="cuda"), 1e-2)


def test_checkpoint_layout_is_torch_adam(tmp_path):
    """CPU: FusedAdamState.state_dict() has torch.optim.Adam's layout and round-trips through it
    (reference core/optimizer.py:73-78 saves opto.state_dict())."""
    from adjointnonlinearraytracing_b200.optimizer import FusedAdamState
    n = torch.ones(5, 5, 5)
    a = FusedAdamState(n, lr=3e-3)
    assert a.state_dict()["state"] == {}  # like torch before the first step
    a.step = 7
    a.exp_avg.normal_(); a.exp_avg_sq.uniform_()
    sd = a.state_dict()
    ref = torch.optim.Adam([torch.ones(5, 5, 5, requires_grad=True)], lr=1.0)
    assert set(sd) == set(ref.state_dict()) and set(sd["param_groups"][0]) == set(ref.state_dict()["param_groups"][0])
    ref.load_state_dict(sd)
    rs = ref.state_dict()
    assert rs["param_groups"][0]["lr"] == 3e-3 and float(rs["state"][0]["step"]) == 7.0
    b = FusedAdamState(n, lr=0.0)
    b.load_state_dict(rs)
    assert b.step == 7 and b.lr == 3e-3 and torch.equal(b.exp_avg, a.exp_avg) and torch.equal(b.exp_avg_sq, a.exp_avg_sq)
    torch.save({"opto_state_dict": sd}, tmp_path / "c.pt")
    back = torch.load(tmp_path / "c.pt", weights_only=False)["opto_state_dict"]
    assert torch.equal(back["state"][0]["exp_avg"], a.exp_avg)


def _switch_allreduce_worker(rank, world, port, out_dir):
    import os, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from adjointnonlinearraytracing_b200.dist import GradAllReduce, SwitchAllReduce
        nccl = GradAllReduce()
        res = [0.0, 0.0, 0.0]
        try:
            sw = SwitchAllReduce()
            buf = sw.alloc((24, 20, 32), torch.device("cuda", rank))
        except Exception:  # noqa: BLE001 -- no multicast fabric on this box
            np.save(os.path.join(out_dir, f"sw{rank}.npy"), np.array([-1.0, 0, 0]))
            return
        gen = torch.Generator(device="cuda").manual_seed(5 + rank)
        for it in range(3):
            src = torch.randn(buf.shape, device="cuda", generator=gen)
            ref = src.clone(); nccl(ref)
            buf.copy_(src); sw(buf)
            torch.cuda.synchronize()
            res[0] = max(res[0], float((buf - ref).abs().max()))
            other = buf.clone(); dist.broadcast(other, 0)
            res[1] = max(res[1], float((other - buf).abs().max()))
        res[2] = 1.0
        try:
            sw(torch.zeros(24, 20, 32, device="cuda"))  # not a buffer of alloc(): refused
            res[2] = 0.0
        except RuntimeError:
            pass
        np.save(os.path.join(out_dir, f"sw{rank}.npy"), np.array(res))
        sw.close(); nccl.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.timeout(300)
def test_switch_allreduce_two_gpus(cuda_lib, tmp_path):
    """The in-switch gradient sum (multimem.ld_reduce / multimem.st kernel over the multicast mapping) against NCCL's
    all-reduce: same values (2 ranks: bit-identical), identical on every rank; needs 2 GPUs behind a multicast-capable
    NVSwitch (skipped otherwise)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import socket
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
    mp.spawn(_switch_allreduce_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    for r in range(2):
        e = np.load(tmp_path / f"sw{r}.npy")
        if e[0] < 0:
            pytest.skip("no multicast mapping on this fabric")
        assert e[0] <= 1e-6 and e[1] == 0 and e[2] == 1, (r, e)


def test_switch_allreduce_argument_checks(cuda_lib):
    """C ABI error paths of drrt_allreduce_multimem (no GPU needed)."""
    import ctypes
    lib = cuda_lib
    buf = ctypes.c_void_p(256)
    assert lib.drrt_allreduce_multimem(None, 16, 0, 2, None) == -1
    assert lib.drrt_allreduce_multimem(buf, 16, 2, 2, None) == -1
    assert lib.drrt_allreduce_multimem(buf, 18, 0, 2, None) == -1
    assert lib.drrt_allreduce_multimem(ctypes.c_void_p(260), 16, 0, 2, None) == -1


def _switch_adam_worker(rank, world, port, out_dir):
    import os, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from adjointnonlinearraytracing_b200 import ops
        from adjointnonlinearraytracing_b200.dist import SwitchAllReduce, SwitchShardedAdam
        try:
            sw = SwitchAllReduce()
            shape = (8, 6, 16)
            n = sw.alloc(shape, torch.device("cuda", rank)); g = sw.alloc(shape, torch.device("cuda", rank))
        except Exception:  # noqa: BLE001
            np.save(os.path.join(out_dir, f"sa{rank}.npy"), np.array([-1.0, 0]))
            return
        gen = torch.Generator(device="cuda").manual_seed(3)  # the same numbers on every rank
        n0 = 1.0 + 0.2 * torch.rand(shape, device="cuda", generator=gen)
        grads = [[torch.randn(shape, device="cuda", generator=gen) for _ in range(world)] for _ in range(3)]
        n.copy_(n0)
        opt = SwitchShardedAdam(sw, n, g, 1e-2)
        ref = n0.clone(); m = torch.zeros_like(ref); v = torch.zeros_like(ref)
        for it in range(3):
            g.copy_(grads[it][rank])
            opt.step()
            total = grads[it][0].clone()
            for r in range(1, world):
                total += grads[it][r]
            ops.adam_step_(ref, total, m, v, it + 1, 1e-2)
        torch.cuda.synchronize()
        lo, hi = opt.i0, opt.i0 + opt.planes
        np.save(os.path.join(out_dir, f"sa{rank}.npy"),
                np.array([float((n - ref).abs().max()), float((opt.exp_avg - m[lo:hi]).abs().max())]))
        np.save(os.path.join(out_dir, f"san{rank}.npy"), n.cpu().numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.timeout(300)
def test_switch_sharded_adam_two_gpus(cuda_lib, tmp_path):
    """The sharded optimiser step with the switch summing the gradients and broadcasting the slab
    (drrt_adam_step_multimem) against the whole-grid step on the summed gradient; every replica identical."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import socket
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
    mp.spawn(_switch_adam_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    for r in range(2):
        e = np.load(tmp_path / f"sa{r}.npy")
        if e[0] < 0:
            pytest.skip("no multicast mapping on this fabric")
        assert e[0] <= 1e-6 and e[1] <= 1e-6, (r, e)
    assert np.array_equal(np.load(tmp_path / "san0.npy"), np.load(tmp_path / "san1.npy"))
