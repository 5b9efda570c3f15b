"""Coarse-to-fine Adam loop with the signature of the reference's ``core/optimizer.py``
(``multires_opt``, :44-83), driving the fused CUDA step (``ops.adam_step_``) and up-res
(``ops.upres_volume``) instead of torch.optim + ~10 elementwise kernels per iteration.

Behaviour kept from the reference, including the parts that look accidental:
  * level ``r`` runs ``iterations * (r + 1)`` steps (:55);
  * boundary voxels never move: their gradient is zeroed before the step (:52-53, :61-62);
  * ``n`` is clamped to >= 1 after every step (:67);
  * between levels ``n`` and both Adam moments are up-sampled trilinearly in float64 (:7-10, :26-29)
    and the step counter carries over (:25);
  * the learning rate does NOT halve between levels: ``reload_opto`` is handed ``0.5**r * lr`` (:81)
    but then copies the old group's ``lr`` over it (:34-35);
  * a checkpoint ``{'rif', 'opto_state_dict', 'loss_hist'}`` is written after every level (:73-78).
"""
from __future__ import annotations

import torch

from . import ops


class FusedAdamState:
    """The state torch.optim.Adam keeps for the single parameter ``n``."""

    def __init__(self, n: torch.Tensor, lr: float, betas=(0.9, 0.999), eps: float = 1e-8):
        self.lr, self.betas, self.eps = float(lr), tuple(betas), float(eps)
        self.step = 0
        self.exp_avg = torch.zeros_like(n)
        self.exp_avg_sq = torch.zeros_like(n)

    def state_dict(self):
        """The layout ``torch.optim.Adam([n]).state_dict()`` has -- what the reference writes into its checkpoint
        (core/optimizer.py:73-78) -- so either side can load the other's file."""
        ref = torch.optim.Adam([torch.zeros(1)], lr=self.lr, betas=self.betas, eps=self.eps)
        group = {k: v for k, v in ref.state_dict()["param_groups"][0].items()}
        group.update(lr=self.lr, betas=self.betas, eps=self.eps, params=[0])
        state = {}
        if self.step > 0:  # torch creates the per-parameter state on the first step
            state[0] = {"step": torch.tensor(float(self.step)), "exp_avg": self.exp_avg, "exp_avg_sq": self.exp_avg_sq}
        return {"state": state, "param_groups": [group]}

    def load_state_dict(self, sd):
        """Accepts a ``torch.optim.Adam`` state dict for one parameter (the reference's checkpoint)."""
        group = sd["param_groups"][0]
        if len(group["params"]) != 1:
            raise ValueError("expected the state of a single parameter (the IOR grid)")
        if group.get("amsgrad", False) or group.get("weight_decay", 0) != 0 or group.get("maximize", False):
            raise ValueError("FusedAdamState implements plain Adam (no amsgrad / weight decay / maximize)")
        self.lr, self.betas, self.eps = float(group["lr"]), tuple(group["betas"]), float(group["eps"])
        st = sd["state"].get(group["params"][0])
        if st is None:
            self.step = 0
            self.exp_avg.zero_(); self.exp_avg_sq.zero_()
            return
        if tuple(st["exp_avg"].shape) != tuple(self.exp_avg.shape):
            raise ValueError("Resolution doesn't match data")
        self.step = int(st["step"])
        self.exp_avg.copy_(st["exp_avg"]); self.exp_avg_sq.copy_(st["exp_avg_sq"])

    def upres(self, res: int) -> "FusedAdamState":
        new = FusedAdamState.__new__(FusedAdamState)
        new.lr, new.betas, new.eps, new.step = self.lr, self.betas, self.eps, self.step
        if self.step == 0:  # "The optimizer hasn't even started yet" (:19-21)
            shape = (res,) * self.exp_avg.dim()
            new.exp_avg = torch.zeros(shape, device=self.exp_avg.device)
            new.exp_avg_sq = torch.zeros(shape, device=self.exp_avg.device)
        else:
            new.exp_avg = ops.upres_volume(self.exp_avg, [res] * 3)
            new.exp_avg_sq = ops.upres_volume(self.exp_avg_sq, [res] * 3)
        return new


def upres_scene(n: torch.Tensor, res: int) -> torch.Tensor:
    """reference core/optimizer.py:7-10"""
    return ops.upres_volume(n.detach(), [res for _ in n.shape])


def multires_opt(func, eta, iterations, res_list, log_func=None, lr=1e-3, statename="result"):
    """reference core/optimizer.py:44-83; ``func(n)`` returns a scalar loss differentiable in ``n``."""
    n = eta.detach().clone().contiguous()
    n.requires_grad = True
    opto = FusedAdamState(n, lr)
    iteration_count = 0
    loss_hist = []
    for res_iter in range(len(res_list)):
        for _ in range(iterations * (res_iter + 1)):
            n.grad = None
            loss = func(n)
            loss.backward()
            with torch.no_grad():
                if log_func is not None:
                    log_func(iteration_count, n)
                opto.step += 1
                ops.adam_step_(n, n.grad.contiguous(), opto.exp_avg, opto.exp_avg_sq, opto.step, opto.lr,
                               opto.betas, opto.eps, clamp_min=1.0, zero_boundary=True, clamp=True)
                loss_hist.append(loss.item())
            iteration_count += 1
        with torch.no_grad():
            if statename is not None:
                torch.save({"rif": n, "opto_state_dict": opto.state_dict(), "loss_hist": torch.tensor(loss_hist)},
                           statename)
            if res_iter < len(res_list) - 1:
                n = upres_scene(n, res_list[res_iter + 1])
                n.requires_grad = True
                opto = opto.upres(n.shape[0])
    return n, loss_hist
