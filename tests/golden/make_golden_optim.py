"""Generates tests/golden/optim5to9.npz by running the REFERENCE's own coarse-to-fine optimiser
(/root/reference/core/optimizer.py::multires_opt, which uses torch.optim.Adam and
core/grid.py::upres_volume) on a small quadratic objective.  Run in the build container:

    python tests/golden/make_golden_optim.py
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, "/root/reference/core")
import optimizer  # noqa: E402  (the reference's)
import grid  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def main():
    torch.manual_seed(0)
    res_list = [5, 9]
    eta = (1.0 + 0.2 * torch.rand(5, 5, 5)).float()
    targets = {r: (0.9 + 0.4 * torch.rand(r, r, r)).float() for r in res_list}  # some targets < 1: the clamp acts
    weights = {r: (0.5 + torch.rand(r, r, r)).float() for r in res_list}
    func = lambda n: (weights[n.shape[0]] * (n - targets[n.shape[0]]) ** 2).sum()
    snaps = []
    out, hist = optimizer.multires_opt(func, eta, 4, res_list, log_func=lambda i, n: snaps.append(n.detach().clone().numpy()),
                                       lr=2e-2, statename="/tmp/_golden_optim_state.pt")
    st = torch.load("/tmp/_golden_optim_state.pt", weights_only=False)
    s = st["opto_state_dict"]["state"][0]
    up = grid.upres_volume(eta.double(), [9, 9, 9]).float().numpy()
    d = dict(eta=eta.numpy(), final=out.detach().numpy(), loss_hist=np.array(hist, np.float64), lr=2e-2, iterations=4,
             res_list=np.array(res_list), upres_eta_9=up, final_exp_avg=s["exp_avg"].numpy(),
             final_exp_avg_sq=s["exp_avg_sq"].numpy(), final_step=float(s["step"]))
    for r in res_list:
        d[f"target{r}"] = targets[r].numpy(); d[f"weight{r}"] = weights[r].numpy()
    for i, sn in enumerate(snaps):
        d[f"n_before_iter{i}"] = sn
    np.savez_compressed(os.path.join(OUT, "optim5to9.npz"), **d)
    print("iterations:", len(snaps), "final loss", hist[-1], "-> tests/golden/optim5to9.npz")


if __name__ == "__main__":
    main()
