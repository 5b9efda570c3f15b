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
    from adjointnonlinearraytracing_b200 import optimizer
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
    assert set(st) == {"rif", "opto_state_dict", "loss_hist"} and st["opto_state_dict"]["step"] == 12


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
