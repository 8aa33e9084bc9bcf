"""Checkpoints and learning-rate schedules in the reference's layouts.

* ``save_model`` / ``load_checkpoint`` / ``override_config``: the artefacts of the PyTorch codes' ``run.py``
  (absent submodule; the reference's .gitignore:1-2 and Checker/copy_weights.py:5-30 name the same files and
  keys): ``config.json`` (the argparse dict), ``checkpoint`` (torch.save of ``{**save_variable_list,
  'model_state_dict', 'optimizer_state_dict'}``), ``entity_embedding.npy``, ``relation_embedding.npy``.
  Because ``KGEModel`` keeps the upstream parameter names and updates the optimizer's own ``step / exp_avg /
  exp_avg_sq`` tensors, checkpoints interchange with reference runs in both directions.
* ``WarmupDecay``: upstream's schedule -- at ``step >= warm_up_steps`` the learning rate is divided by 10, Adam is
  re-created (its moments are dropped) and ``warm_up_steps`` is tripled.
* ``lrfn``: the TF trainer's per-epoch schedule (/root/reference/tensorflow_codes/run.py:69-84).
"""
from __future__ import annotations

import json
import os
from typing import Dict, Optional

import numpy as np
import torch


def save_model(model, optimizer, save_variable_list: Dict, args) -> None:
    """Save the parameters of the model and the optimizer, as well as some other variables such as step and
    learning_rate (same files, same keys as upstream)."""
    os.makedirs(args.save_path, exist_ok=True)
    argparse_dict = dict(vars(args))
    with open(os.path.join(args.save_path, "config.json"), "w") as fjson:
        json.dump(argparse_dict, fjson)
    torch.save({**save_variable_list,
                "model_state_dict": model.state_dict(),
                "optimizer_state_dict": optimizer.state_dict()},
               os.path.join(args.save_path, "checkpoint"))
    entity_embedding = model.entity_embedding.detach().cpu().numpy()
    np.save(os.path.join(args.save_path, "entity_embedding"), entity_embedding)
    relation_embedding = model.relation_embedding.detach().cpu().numpy()
    np.save(os.path.join(args.save_path, "relation_embedding"), relation_embedding)


def override_config(args) -> None:
    """Override model and data configuration from ``init_checkpoint/config.json`` (upstream ``override_config``)."""
    with open(os.path.join(args.init_checkpoint, "config.json"), "r") as fjson:
        argparse_dict = json.load(fjson)
    if getattr(args, "data_path", None) is None:
        args.data_path = argparse_dict.get("data_path")
    for key in ("model", "double_entity_embedding", "double_relation_embedding", "hidden_dim", "test_batch_size"):
        if key in argparse_dict:
            setattr(args, key, argparse_dict[key])
    for key in ("countries", "triple_relation_embedding"):  # the second one is the reference's own -tr flag
        if key in argparse_dict:
            setattr(args, key, argparse_dict[key])


def load_checkpoint(model, optimizer, init_checkpoint: str, do_train: bool = True, map_location=None) -> Dict:
    """Restore ``model`` (and, when training, ``optimizer``) from ``init_checkpoint/checkpoint``; returns the saved
    variables (``step``, ``current_learning_rate``, ``warm_up_steps``, ...) -- the body of upstream's
    ``if args.init_checkpoint:`` block."""
    checkpoint = torch.load(os.path.join(init_checkpoint, "checkpoint"), map_location=map_location, weights_only=False)
    model.load_state_dict(checkpoint["model_state_dict"])
    if do_train and optimizer is not None:
        optimizer.load_state_dict(checkpoint["optimizer_state_dict"])
        # torch keeps a non-capturable Adam's `step` wherever torch.load put it: with map_location="cuda" that is the
        # GPU, and every later train_step would read it back with a device sync (and miss the staged fast path, which
        # wants the host-side counters torch.optim.Adam itself creates).  Put them where Adam.step() would.
        for group in optimizer.param_groups:
            if not group.get("capturable", False):
                for p in group["params"]:
                    st = optimizer.state.get(p)
                    if st and torch.is_tensor(st.get("step")) and st["step"].is_cuda:
                        st["step"] = st["step"].detach().to("cpu", torch.float32)
    return {k: v for k, v in checkpoint.items() if k not in ("model_state_dict", "optimizer_state_dict")}


class WarmupDecay:
    """``current_learning_rate`` / ``warm_up_steps`` bookkeeping of upstream's training loop."""

    def __init__(self, model, learning_rate: float, max_steps: int, warm_up_steps: Optional[int] = None,
                 optimizer_cls=torch.optim.Adam):
        self.model, self.optimizer_cls = model, optimizer_cls
        self.current_learning_rate = float(learning_rate)
        self.warm_up_steps = int(warm_up_steps) if warm_up_steps else max_steps // 2
        self.optimizer = self._make()

    def _make(self):
        return self.optimizer_cls(filter(lambda p: p.requires_grad, self.model.parameters()), lr=self.current_learning_rate)

    def after_step(self, step: int):
        """Call with the index of the step just taken; returns the (possibly new) optimizer."""
        if step >= self.warm_up_steps:
            self.current_learning_rate = self.current_learning_rate / 10
            self.optimizer = self._make()
            self.warm_up_steps = self.warm_up_steps * 3
        return self.optimizer


def lrfn(epoch: float, num_replicas_in_sync: int = 1) -> float:
    """/root/reference/tensorflow_codes/run.py:69-84."""
    LR_START = 0.00001
    LR_MAX = 0.00005 * num_replicas_in_sync
    LR_MIN = 0.00001
    LR_RAMPUP_EPOCHS = 5.0
    LR_SUSTAIN_EPOCHS = 0.0
    LR_EXP_DECAY = .8
    if float(epoch) < LR_RAMPUP_EPOCHS:
        lr = (LR_MAX - LR_START) / LR_RAMPUP_EPOCHS * float(epoch) + LR_START
    elif float(epoch) < LR_RAMPUP_EPOCHS + LR_SUSTAIN_EPOCHS:
        lr = LR_MAX
    else:
        lr = (LR_MAX - LR_MIN) * LR_EXP_DECAY ** (float(epoch) - LR_RAMPUP_EPOCHS - LR_SUSTAIN_EPOCHS) + LR_MIN
    return lr
