"""``FusedAdam``: torch.optim.Adam semantics (defaults, state layout ``step / exp_avg / exp_avg_sq``) with the
update done by the K4 kernel (``kge_adam_dense``).  ``KGEModel.train_step`` recognises both this class and
``torch.optim.Adam`` and then skips ``.grad`` altogether; ``step()`` here serves every other use."""
from __future__ import annotations

import torch

from . import ops


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        if lr < 0.0 or eps < 0.0 or not (0.0 <= betas[0] < 1.0) or not (0.0 <= betas[1] < 1.0):
            raise ValueError("invalid Adam hyper-parameter")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=0, amsgrad=False, maximize=False))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group in self.param_groups:
            b1, b2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = torch.tensor(0.0, dtype=torch.float32)
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st["step"] += 1
                ops.adam_dense(p.data, p.grad.contiguous(), st["exp_avg"], st["exp_avg_sq"], lr=group["lr"], beta1=b1,
                               beta2=b2, eps=group["eps"], step=int(st["step"]))
        return loss
